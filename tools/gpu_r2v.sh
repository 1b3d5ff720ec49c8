#!/bin/bash
TAG=${1:-r2v}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 2400 python -m pytest tests/test_gpu_prom.py tests/test_gpu_segments.py tests/test_gpu_determinism.py -m gpu -q -k "prom or modes" > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
grep -E "^E   .*(Error|assert)|^FAILED|passed|failed|rc=" $OUT/pytest_gpu.log | cut -c1-300 | tail -20
for n in 40 32 34 36 44 48; do for proj in LSPG Galerkin; do python tools/prom_time.py 4096 $n $proj 100 | tail -1; done; done > $OUT/prom_wide_time.txt 2>&1
for proj in LSPG Galerkin; do python tools/prom_time.py 512 40 $proj 500 | tail -1; python tools/prom_time.py 4096 40 $proj 500 | tail -1; done >> $OUT/prom_wide_time.txt 2>&1
cat $OUT/prom_wide_time.txt
