#!/bin/bash
TAG=${1:-r2t}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 2400 python -m pytest tests/test_gpu_prom.py tests/test_gpu_segments.py tests/test_gpu_determinism.py -m gpu -q -k "prom or modes" --durations=5 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
grep -E "^E   .*(Error|assert)|^FAILED|passed|failed|rc=" $OUT/pytest_gpu.log | cut -c1-300 | tail -20
for n in 40 32 36 44 47 48; do for proj in LSPG Galerkin; do python tools/prom_time.py 4096 $n $proj 100 | tail -1; done; done > $OUT/prom_wide_time.txt 2>&1
cat $OUT/prom_wide_time.txt
