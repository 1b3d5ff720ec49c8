#!/bin/bash
# ncu --set full of the FOM latency kernel at the strong-scaling shard sizes (512 and 1024 trajectories x 500 steps)
TAG=${1:-r2x}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for B in 512 1024; do
python tools/fom_one.py $B 500 -1 -1 3 > $OUT/time_$B.txt 2>&1 || exit 1
cat $OUT/time_$B.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fom_strong_kernel -c 1 \
  -o $OUT/fom_latency_${B}x500 -f python tools/fom_one.py $B 500 -1 -1 1 > $OUT/ncu_$B.log 2>&1
tail -1 $OUT/ncu_$B.log
done
