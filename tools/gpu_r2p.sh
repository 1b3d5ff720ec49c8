#!/bin/bash
TAG=${1:-r2p}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for proj in LSPG Galerkin; do
python tools/qm_time.py 1184 $proj 5 28 1 > $OUT/time_$proj.txt 2>&1 || exit 1
tail -1 $OUT/time_$proj.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:qm_kernel -c 1 \
  -o $OUT/qm28_ts_$proj -f python tools/qm_time.py 1184 $proj 5 28 1 > $OUT/ncu_$proj.log 2>&1
tail -1 $OUT/ncu_$proj.log
done
