#!/bin/bash
# ncu --set full of the quadratic-manifold kernel (n = 28, Galerkin, v2 schedule and shared-tile schedule)
TAG=${1:-r2o}
OUT=gpurun_out/$TAG
mkdir -p $OUT
python tools/qm_time.py 1184 Galerkin 5 28 > $OUT/time.txt 2>&1 || exit 1
tail -1 $OUT/time.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:qm_kernel -c 1 \
  -o $OUT/qm28_galerkin -f python tools/qm_time.py 1184 Galerkin 5 28 > $OUT/ncu.log 2>&1
tail -2 $OUT/ncu.log
