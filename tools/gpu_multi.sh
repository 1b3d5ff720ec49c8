#!/bin/bash
# Multi-GPU session (run with gpurun --gpus N): scaling bench at 1, 2, 4, .. N and configs 3 / 4 over NCCL.
N=${1:-2}
OUT=gpurun_out/multi$N
mkdir -p $OUT
nvidia-smi -L > $OUT/gpus.txt
nvidia-smi topo -m > $OUT/topo.txt 2>&1
for n in 1 2 4 8; do
  [ $n -gt $N ] && break
  if [ $n -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu-baseline > $OUT/bench_n1.json 2> $OUT/bench_n1.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n \
      bench.py --gpus $n --steps 3 --warmup 3 --no-cpu-baseline > $OUT/bench_n$n.json 2> $OUT/bench_n$n.err
  fi
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 \
  tools/config3.py 4096 500 28 > $OUT/config3_n$N.txt 2> $OUT/config3_n$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 \
  tools/config4.py $((4096 / N)) 500 28 Galerkin > $OUT/config4_n$N.txt 2> $OUT/config4_n$N.err
for f in $OUT/bench_n*.json; do python -c "import json,sys; d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', d['n_gpus'], d['value'], d['ms_per_step'], d.get('e2e',{}).get('value'))"; done
cat $OUT/config3_n$N.txt $OUT/config4_n$N.txt; for f in $OUT/*.err; do tail -n 2 $f; done; true
