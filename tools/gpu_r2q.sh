#!/bin/bash
# 8-GPU session: topology, the strong-scaling bench line exactly as the driver launches it
TAG=${1:-r2q}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi topo -m > $OUT/topo.txt 2>&1
(numactl -H; lscpu | grep -E "Model name|^CPU\(s\)|NUMA|Socket") > $OUT/numa.txt 2>&1
N=${2:-8}
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err; echo "bench rc=$?" >> $OUT/bench_n$N.err
tail -5 $OUT/bench_n$N.err; grep "^{" $OUT/bench_n$N.json | cut -c1-300
cat $OUT/numa.txt | head -20
