#!/bin/bash
TAG=${1:-r2r}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 2400 python -m pytest tests/test_gpu_prom.py tests/test_gpu_segments.py tests/test_gpu_determinism.py tests/test_gpu_surface.py -m gpu -q --durations=5 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
grep -E "^E   .*(Error|assert)|^FAILED|passed|failed|rc=" $OUT/pytest_gpu.log | cut -c1-300 | tail -20
for n in 40 32 36 44 47; do for proj in LSPG Galerkin; do python tools/prom_time.py 4096 $n $proj 100 | tail -1; done; done > $OUT/prom_wide_time.txt 2>&1
python - >> $OUT/prom_wide_time.txt 2>&1 <<PY
import sys; sys.path.insert(0, '.')
import burgers_b200
burgers_b200.set_tuning("prom_warps_per_trajectory", 8)
for proj in ("LSPG", "Galerkin"):
    sys.argv = ["prom_time", "4096", "40", proj, "100"]
    print("CTA kernel:", end=" ")
    exec(open("tools/prom_time.py").read())
PY
cat $OUT/prom_wide_time.txt
