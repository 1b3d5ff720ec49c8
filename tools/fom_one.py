"""One FOM sweep configuration, timed: python tools/fom_one.py B nsteps kernel W [reps]
(kernel 0 throughput / 1 latency / -1 automatic; W warps per trajectory, -1 automatic).  Used under ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import burgers_b200

B, nsteps, kern, W = (int(x) for x in sys.argv[1:5])
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
fem = burgers_b200.FEMBurgers(np.linspace(0, 100, 512))
u0 = torch.ones(512, dtype=torch.float64, device="cuda")
g1 = np.repeat(np.linspace(4.25, 5.5, 64), 64); g2 = np.tile(np.linspace(0.015, 0.03, 64), 64)
sel = np.linspace(0, 4095, B).astype(int)
mu1 = torch.from_numpy(g1[sel].copy()).cuda(); mu2 = torch.from_numpy(g2[sel].copy()).cuda()
U = torch.empty((B, nsteps + 1, 512), dtype=torch.float64, device="cuda")
it = torch.empty((B, nsteps), dtype=torch.int32, device="cuda")
burgers_b200.set_tuning("fom_warps_per_trajectory", W)
burgers_b200.set_tuning("fom_kernel", kern)
fem.fom_burgers_batch(0.05, nsteps, u0, mu1, 0.0, mu2, out=U, iters_out=it)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(reps):
    fem.fom_burgers_batch(0.05, nsteps, u0, mu1, 0.0, mu2, out=U, iters_out=it)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
tot = int(it.sum())
print(f"B={B} nsteps={nsteps} kernel={kern} W={W}: {ms:.3f} ms, {tot} iterations, {ms*1e3/ (tot/B):.3f} us per iteration of one trajectory")
