"""BASELINE config 3 end to end on N GPUs (one process per GPU, torchrun or plain python for N = 1):
  FOM sweep of this rank's block of the 64x64 grid -> snapshot Gram (DMMA) -> NCCL all-reduce of the 512x512
  Gram -> eigh -> Phi (28 modes) -> POD-LSPG PROM sweep of the same samples.
usage: [torchrun --nproc-per-node N --master-addr 127.0.0.1 ...] python tools/config3.py [samples_per_rank] [steps] [n]"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import torch.distributed as dist
import burgers_b200
from burgers_b200 import pod, sweep

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 500
n = int(sys.argv[3]) if len(sys.argv) > 3 else 28
rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
g1, g2 = sweep.parameter_grid(64, 64)
total = min(B * world, g1.size)
lo, hi = sweep.shard_range(total, rank, world)
mu1 = torch.from_numpy(g1[lo:hi].copy()).to(dev); mu2 = torch.from_numpy(g2[lo:hi].copy()).to(dev)
X = np.linspace(0, 100, 512)
fem = burgers_b200.FEMBurgers(X)
u0 = torch.ones(512, dtype=torch.float64, device=dev)
U = torch.empty((hi - lo, steps + 1, 512), dtype=torch.float64, device=dev)
it = torch.empty((hi - lo, steps), dtype=torch.int32, device=dev)
acc = pod.GramAccumulator(512)

def ev():
    e = torch.cuda.Event(enable_timing=True); e.record(); return e

res = {}
for rep in range(3):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0 = ev()
    fem.fom_burgers_batch(0.05, steps, u0, mu1, 0.0, mu2, out=U, iters_out=it)
    e1 = ev()
    acc.G.zero_(); acc.add(U)
    e2 = ev()
    pod.allreduce_gram(acc.G)
    e3 = ev()
    Phi, s = pod.basis_from_gram(acc.G, n)
    e4 = ev()
    fom_its = int(it.sum())
    fem.pod_prom_burgers_batch(0.05, steps, u0, mu1, 0.0, mu2, Phi, "LSPG", out=U, iters_out=it)
    e5 = ev()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1), e1.elapsed_time(e2), e2.elapsed_time(e3), e3.elapsed_time(e4),
                      e4.elapsed_time(e5), e0.elapsed_time(e5)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res = dict(zip(["fom_ms", "gram_ms", "allreduce_ms", "eigh_ms", "prom_lspg_ms", "total_ms"], [round(float(x), 3) for x in t]))
    res.update(n_gpus=world, samples_per_rank=hi - lo, steps=steps, modes=n, fom_iterations_rank0=fom_its,
               prom_iterations_rank0=int(it.sum()), sigma_ratio=float(s[n - 1] / s[0]),
               phi_checksum=float(Phi.abs().sum()))
    if rank == 0:
        print(json.dumps(res), flush=True)
# every rank must hold the same basis
if world > 1:
    chk = torch.tensor([float(Phi.abs().sum())], dtype=torch.float64, device=dev)
    lo_, hi_ = chk.clone(), chk.clone()
    dist.all_reduce(lo_, op=dist.ReduceOp.MIN); dist.all_reduce(hi_, op=dist.ReduceOp.MAX)
    if rank == 0:
        print("basis identical on all ranks:", bool(lo_.item() == hi_.item()))
    dist.destroy_process_group()
