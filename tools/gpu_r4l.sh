#!/bin/bash
# fine-mesh kernel: tests, rows table, one ncu --set full capture of the 2^20-node launch (source-level)
mkdir -p gpurun_out/r4l
timeout 500 python -m pytest tests/test_gpu_finemesh.py tests/test_zy_finemesh_ranks.py -m gpu -x -q > gpurun_out/r4l/tests.log 2>&1; echo tests rc=$?; tail -3 gpurun_out/r4l/tests.log
timeout 300 python tools/fine_rows.py 100 > gpurun_out/r4l/fine_rows.txt 2>&1; cut -c1-200 gpurun_out/r4l/fine_rows.txt
timeout 200 python tools/fine_time.py 20 500 > gpurun_out/r4l/fine_time.txt 2>&1; cat gpurun_out/r4l/fine_time.txt
timeout 400 ncu --set full --clock-control none --import-source on -k regex:fine_fom_kernel -c 1 -o gpurun_out/r4l/fine_full -f python tools/fine_time.py 20 20 > gpurun_out/r4l/ncu.log 2>&1; echo ncu rc=$?; tail -2 gpurun_out/r4l/ncu.log
