#!/bin/bash
# FOM latency kernel: parity of all instances, then the warps-per-trajectory / kernel table
TAG=${1:-r2b}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_fom.py -m gpu -q -x > $OUT/pytest_fom.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_fom.log
tail -15 $OUT/pytest_fom.log
timeout 600 python tools/fom_warps.py 500 148 296 512 592 1024 2048 4096 > $OUT/fom_warps.txt 2>&1
cat $OUT/fom_warps.txt
