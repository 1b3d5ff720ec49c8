#!/bin/bash
# ncu --set full of the FOM latency kernel (W = 2 and W = 1) at 512 trajectories x 40 steps
TAG=${1:-r2c}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for W in 2 1; do
python tools/fom_one.py 512 40 1 $W > $OUT/time_w$W.txt 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fom_strong_kernel -c 1 \
  -o $OUT/fom_strong_w$W -f python tools/fom_one.py 512 40 1 $W 1 > $OUT/ncu_w$W.log 2>&1
cat $OUT/time_w$W.txt; tail -2 $OUT/ncu_w$W.log
done
