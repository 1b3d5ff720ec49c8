#!/bin/bash
# full GPU suite + smoke + default bench line on the final code
mkdir -p gpurun_out/r4p
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r4p/tests.log 2>&1; echo tests rc=$?; tail -4 gpurun_out/r4p/tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r4p/smoke.log 2>&1; echo smoke rc=$?; tail -2 gpurun_out/r4p/smoke.log
timeout 900 python bench.py > gpurun_out/r4p/bench_n1.json 2> gpurun_out/r4p/bench_n1.err; echo bench rc=$?; cut -c1-1500 gpurun_out/r4p/bench_n1.json
