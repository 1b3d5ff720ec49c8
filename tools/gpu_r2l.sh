#!/bin/bash
# 2-GPU session: the strong-scaling bench line under torchrun (as the driver launches it), and the reference arm
TAG=${1:-r2l}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi topo -m > $OUT/topo.txt 2>&1
timeout 600 python -m pytest "tests/test_gpu_parity_full.py::test_qm_config4_path_vs_oracle" -m gpu -q 2>&1 | grep -E "^E   .*(Error|assert)|^FAILED|passed|failed" | cut -c1-300 | tail -8
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > $OUT/bench_n2.json 2> $OUT/bench_n2.err; echo "bench rc=$?" >> $OUT/bench_n2.err
tail -5 $OUT/bench_n2.err; cut -c1-400 $OUT/bench_n2.json
