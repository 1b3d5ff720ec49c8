#!/bin/bash
TAG=${1:-r2e}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/probe_rcp tools/probe_rcp.cu && /tmp/probe_rcp > $OUT/probe_rcp.txt 2>&1
cat $OUT/probe_rcp.txt
timeout 900 python -m pytest tests/test_gpu_fom.py tests/test_gpu_segments.py tests/test_gpu_surface.py -m gpu -q -x > $OUT/pytest_fom.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_fom.log
tail -5 $OUT/pytest_fom.log
python - <<PY > $OUT/segments.txt 2>&1
import sys; sys.path.insert(0, '.')
import burgers_b200
src = open("tools/fom_one.py").read()
for rep in range(2):
    for seg in (0, 25, 50):
        for B in (4096, 2048):
            burgers_b200.set_tuning("fom_segment_steps", seg)
            sys.argv = ["fom_one", str(B), "500", "0", "1", "5"]
            print(f"segment steps {seg}:", end=" ")
            exec(src)
PY
cat $OUT/segments.txt
