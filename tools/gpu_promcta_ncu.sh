#!/bin/bash
# one ncu --set full capture of the PROM CTA kernel at n = 40 LSPG (the reference's shipped driver size)
mkdir -p gpurun_out/r03j
python tools/prom_time.py 4096 40 LSPG 10 > gpurun_out/r03j/time.txt 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:prom_cta_kernel -c 1 \
  -o gpurun_out/r03j/promcta40 -f python tools/prom_time.py 4096 40 LSPG 10 > gpurun_out/r03j/ncu.log 2>&1
tail -3 gpurun_out/r03j/time.txt; tail -3 gpurun_out/r03j/ncu.log
