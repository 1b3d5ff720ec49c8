#!/bin/bash
mkdir -p gpurun_out/r4m
timeout 500 python -m pytest tests/test_gpu_finemesh.py tests/test_zy_finemesh_ranks.py -m gpu -x -q > gpurun_out/r4m/tests.log 2>&1; echo tests rc=$?; tail -3 gpurun_out/r4m/tests.log
timeout 300 python tools/fine_rows.py 100 > gpurun_out/r4m/fine_rows.txt 2>&1; cut -c1-200 gpurun_out/r4m/fine_rows.txt
timeout 200 python tools/fine_time.py 20 500 > gpurun_out/r4m/fine_time.txt 2>&1; cat gpurun_out/r4m/fine_time.txt
