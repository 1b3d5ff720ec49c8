#!/bin/bash
TAG=${1:-r2f}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_fom.py tests/test_gpu_segments.py tests/test_gpu_surface.py -m gpu -q -x > $OUT/pytest_fom.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_fom.log
tail -5 $OUT/pytest_fom.log
python - <<PY > $OUT/segments.txt 2>&1
import sys; sys.path.insert(0, '.')
import burgers_b200
src = open("tools/fom_one.py").read()
for seg in (0, 10, 25, 50):
    for B in (4096, 2048):
        burgers_b200.set_tuning("fom_segment_steps", seg)
        sys.argv = ["fom_one", str(B), "500", "0", "1", "5"]
        print(f"segment steps {seg}:", end=" ")
        exec(src)
PY
cat $OUT/segments.txt
timeout 600 python tools/fom_warps.py 500 148 296 512 1024 1184 2048 4096 > $OUT/fom_warps.txt 2>&1
cat $OUT/fom_warps.txt
