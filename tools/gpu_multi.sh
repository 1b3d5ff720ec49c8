#!/bin/bash
# Multi-GPU session (run with gpurun --gpus N): scaling bench and config 3 over NCCL.
N=${1:-2}
OUT=gpurun_out/multi$N
mkdir -p $OUT
nvidia-smi -L > $OUT/gpus.txt
for n in 1 $N; do
  if [ $n -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu-baseline > $OUT/bench_n1.json 2> $OUT/bench_n1.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $n --steps 3 --warmup 3 --no-cpu-baseline > $OUT/bench_n$n.json 2> $OUT/bench_n$n.err
  fi
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
  tools/config3.py 4096 500 28 > $OUT/config3_n$N.txt 2> $OUT/config3_n$N.err
cat $OUT/bench_n1.json $OUT/bench_n$N.json $OUT/config3_n$N.txt; for f in $OUT/*.err; do tail -n 3 $f; done; true
