#!/bin/bash
# ncu --set full of the POD-PROM register kernel at n = 40 (LSPG, Galerkin)
TAG=${1:-r2s}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for proj in LSPG Galerkin; do
python tools/prom_time.py 1184 40 $proj 10 > $OUT/time_$proj.txt 2>&1 || exit 1
tail -1 $OUT/time_$proj.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:prom_wide_kernel -c 1 \
  -o $OUT/prom_wide40_$proj -f python tools/prom_time.py 1184 40 $proj 10 > $OUT/ncu_$proj.log 2>&1
tail -1 $OUT/ncu_$proj.log
done
