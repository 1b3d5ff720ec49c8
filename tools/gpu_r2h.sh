#!/bin/bash
# full GPU suite, default bench line, ncu --set full of the headline kernel (config 2, throughput instance)
TAG=${1:-r2h}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 2400 python -m pytest tests -m gpu -q -x --durations=10 > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
tail -18 $OUT/pytest_gpu.log
timeout 1200 python bench.py > $OUT/bench_n1.json 2> $OUT/bench_n1.err; echo "bench rc=$?" >> $OUT/bench_n1.err
tail -5 $OUT/bench_n1.err; cut -c1-600 $OUT/bench_n1.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fom_picard_kernel -c 1 \
  -o $OUT/fom_throughput_4096x500 -f python tools/fom_one.py 4096 500 0 1 1 > $OUT/ncu_fom.log 2>&1
tail -2 $OUT/ncu_fom.log
