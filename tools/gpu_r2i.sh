#!/bin/bash
TAG=${1:-r2i}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 2400 python -m pytest tests/test_gpu_parity_full.py tests/test_gpu_prom.py tests/test_gpu_qm.py tests/test_gpu_pod.py tests/test_gpu_determinism.py -m gpu -q -x --durations=5 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
tail -12 $OUT/pytest_gpu.log
for B in 512 1024 2048 4096; do for proj in LSPG Galerkin; do python tools/prom_time.py $B 28 $proj 500 | tail -1; done; done > $OUT/prom28_by_batch.txt 2>&1
cat $OUT/prom28_by_batch.txt
