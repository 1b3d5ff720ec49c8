#!/bin/bash
# ncu --set full captures of the final fine-mesh kernel: 2^20 nodes (16 rows per thread) and 131072 nodes (4 rows per thread)
mkdir -p gpurun_out/r4w
timeout 400 ncu --set full --clock-control none --import-source on -k regex:fine_fom_kernel -c 1 -o gpurun_out/r4w/fine20 -f python tools/fine_time.py 20 20 > gpurun_out/r4w/ncu20.log 2>&1; echo ncu20 rc=$?
timeout 400 ncu --set full --clock-control none --import-source on -k regex:fine_fom_kernel -c 1 -o gpurun_out/r4w/fine17 -f python tools/fine_time.py 17 20 > gpurun_out/r4w/ncu17.log 2>&1; echo ncu17 rc=$?
