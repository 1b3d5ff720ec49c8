#!/bin/bash
# One GPU-box session: parity tests, bench (both arms), ncu launch list + full captures.
# usage (from repo root, on the box): bash tools/gpu_round.sh [tag]
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $OUT/gpu.txt 2>&1
nproc > $OUT/nproc.txt; lscpu | grep -E "Model name|^CPU\(s\)" >> $OUT/nproc.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.log
timeout 900 python bench.py > $OUT/bench_n1.json 2> $OUT/bench_n1.err; echo "bench rc=$?" >> $OUT/bench_n1.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err
timeout 600 python tools/prom_time.py 4096 28 LSPG 500 > $OUT/prom_lspg.txt 2>&1
timeout 600 python tools/prom_time.py 4096 28 Galerkin 500 > $OUT/prom_galerkin.txt 2>&1
# ncu passes (numbers printed under ncu are never bench values)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/ncu_launches.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:fom_picard -s 3 -c 1 -f -o $OUT/fom_full \
  python bench.py --steps 1 --warmup 3 --nsteps 50 --no-cpu-baseline --no-e2e > $OUT/ncu_fom_full.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:prom_kernel -s 1 -c 1 -f -o $OUT/prom_full \
  python tools/prom_time.py 4096 28 LSPG 20 > $OUT/ncu_prom_full.log 2>&1
ls -la $OUT
tail -3 $OUT/pytest_gpu.log; cat $OUT/bench_n1.json; cat $OUT/prom_lspg.txt
