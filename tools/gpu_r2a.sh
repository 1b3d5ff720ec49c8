#!/bin/bash
# Round-2 first GPU session: parity tests (new config-4 / full-length cases, multi-warp FOM, error surface),
# warps-per-trajectory table, latency probe, bench line.
TAG=${1:-r2a}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $OUT/gpu.txt 2>&1
nproc > $OUT/nproc.txt; lscpu | grep -E "Model name|^CPU\(s\)" >> $OUT/nproc.txt
timeout 300 tools/bin/probe_latency > $OUT/probe_latency.txt 2>&1
timeout 600 python tools/fom_warps.py 500 > $OUT/fom_warps.txt 2>&1
timeout 1800 python -m pytest tests -m gpu -q -x --durations=15 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
timeout 900 python bench.py > $OUT/bench_n1.json 2> $OUT/bench_n1.err; echo "bench rc=$?" >> $OUT/bench_n1.err
cat $OUT/probe_latency.txt $OUT/fom_warps.txt; tail -25 $OUT/pytest_gpu.log; tail -5 $OUT/bench_n1.err; cut -c1-1500 $OUT/bench_n1.json
