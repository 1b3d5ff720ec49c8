"""BASELINE config 4 end to end on N GPUs (one process per GPU, torchrun or plain python for N = 1):
  FOM sweep of this rank's block of the 64x64 grid -> snapshot Gram + all-reduce -> Phi (n modes) ->
  quadratic-manifold H on the device (tall-skinny QR per rank, all-gather of the (R, W) pairs) ->
  POD-Galerkin quadratic-manifold PROM sweep of the same samples.
usage: [torchrun --nproc-per-node N --master-addr 127.0.0.1 ...] python tools/config4.py [samples_per_rank] [steps] [n] [proj]"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import torch.distributed as dist
import burgers_b200
from burgers_b200 import pod, sweep, quadratic

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 500
n = int(sys.argv[3]) if len(sys.argv) > 3 else 28
proj = sys.argv[4] if len(sys.argv) > 4 else "Galerkin"
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
fem = burgers_b200.FEMBurgers(np.linspace(0, 100, 512))
u0 = torch.ones(512, dtype=torch.float64, device=dev)
U = torch.empty((hi - lo, steps + 1, 512), dtype=torch.float64, device=dev)
it = torch.empty((hi - lo, steps), dtype=torch.int32, device=dev)
acc = pod.GramAccumulator(512)

def ev():
    e = torch.cuda.Event(enable_timing=True); e.record(); return e

for rep in range(2):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0 = ev()
    fem.fom_burgers_batch(0.05, steps, u0, mu1, 0.0, mu2, out=U, iters_out=it)
    e1 = ev()
    acc.G.zero_(); acc.add(U); pod.allreduce_gram(acc.G)
    Phi, s = pod.basis_from_gram(acc.G, n)
    e2 = ev()
    H = quadratic.build_manifold(U, Phi, alpha=1e-2)
    e3 = ev()
    # training reconstruction error of the manifold on this rank's snapshots (build_quadratic_manifold.py:48)
    S2 = U.reshape(-1, 512)[:: max(1, U.shape[0] * U.shape[1] // 50000)]
    q = S2 @ Phi
    rec = q @ Phi.T + quadratic.sym_monomials_rows(q) @ H.T
    rel_qm = float((S2 - rec).norm() / S2.norm()); rel_pod = float((S2 - q @ Phi.T).norm() / S2.norm())
    fom_its = int(it.sum())
    e4 = ev()
    fem.pod_quadratic_manifold_batch(0.05, steps, u0, mu1, 0.0, mu2, Phi, H, proj, out=U, iters_out=it)
    e5 = ev()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1), e1.elapsed_time(e2), e2.elapsed_time(e3), e4.elapsed_time(e5)],
                     dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    its = int(it.sum())
    res = dict(zip(["fom_ms", "pod_ms", "build_H_ms", "qm_prom_ms"], [round(float(x), 3) for x in t]))
    res.update(n_gpus=world, samples_per_rank=hi - lo, steps=steps, modes=n, monomials=n * (n + 1) // 2, projection=proj,
               fom_iterations_rank0=fom_its, qm_iterations_rank0=its, qm_max_iters_per_step=int(it.max()),
               qm_iterations_per_s_rank0=round(its / float(t[3]) * 1e3), finite=bool(torch.isfinite(U).all()),
               train_rel_err_pod=rel_pod, train_rel_err_qm=rel_qm, H_checksum=float(H.abs().sum()))
    if rank == 0:
        print(json.dumps(res), flush=True)
if world > 1:
    chk = torch.tensor([float(H.abs().sum())], dtype=torch.float64, device=dev)
    lo_, hi_ = chk.clone(), chk.clone()
    dist.all_reduce(lo_, op=dist.ReduceOp.MIN); dist.all_reduce(hi_, op=dist.ReduceOp.MAX)
    if rank == 0:
        print("H identical on all ranks:", bool(lo_.item() == hi_.item()))
    dist.destroy_process_group()
