#!/bin/bash
TAG=${1:-r2m}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 2400 python -m pytest tests/test_gpu_segments.py tests/test_gpu_parity_full.py tests/test_gpu_prom.py tests/test_gpu_qm.py tests/test_gpu_determinism.py tests/test_gpu_surface.py -m gpu -q --durations=5 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
grep -E "^E   .*(Error|assert)|^FAILED|passed|failed|rc=" $OUT/pytest_gpu.log | cut -c1-300 | tail -20
for B in 2048 4096; do for proj in LSPG Galerkin; do python tools/prom_time.py $B 28 $proj 500 | tail -1; done; done > $OUT/prom28_by_batch.txt 2>&1
cat $OUT/prom28_by_batch.txt
for B in 2048 4096; do for proj in Galerkin; do python tools/qm_time.py $B $proj 500 28 | tail -1; done; done > $OUT/qm28_by_batch.txt 2>&1
cat $OUT/qm28_by_batch.txt
