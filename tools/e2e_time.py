"""Host-streaming FOM sweep (bench.py's e2e leg) against the raw pinned D2H rate of the box, over chunk sizes:
python tools/e2e_time.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import __graft_entry__  # noqa: F401  (registers the burgers_b200 alias)
import burgers_b200
from burgers_b200 import sweep as bsweep

N, nT, B = 512, 500, 4096
X = np.linspace(0.0, 100.0, N)
fem = burgers_b200.FEMBurgers(X)
m1, m2 = bsweep.parameter_grid(64, 64)
h1 = torch.from_numpy(np.ascontiguousarray(m1)).pin_memory(); h2 = torch.from_numpy(np.ascontiguousarray(m2)).pin_memory()
u0 = np.ones(N)

# raw D2H of one 2.1 GB chunk
d = torch.empty((1024, nT + 1, N), dtype=torch.float64, device="cuda")
h = torch.empty((1024, nT + 1, N), dtype=torch.float64).pin_memory()
for _ in range(2):
    h.copy_(d, non_blocking=True)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(4):
    h.copy_(d, non_blocking=True)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 4
print(f"raw pinned D2H: {d.numel() * 8 / dt / 1e9:.1f} GB/s ({dt * 1e3:.1f} ms per 2.1 GB chunk); "
      f"4096 x 501 x 512 doubles = {4 * dt * 1e3:.1f} ms floor", flush=True)
del d, h

for chunk in (256, 512, 1024, 2048):
    for ramp in (False, True):
        bufs = bsweep.HostSweepBuffers(fem, chunk, nT)
        for _ in range(2):
            bsweep.fom_sweep_to_host(fem, 0.05, nT, u0, h1, h2, 0.0, chunk, None, bufs, ramp=ramp)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            bsweep.fom_sweep_to_host(fem, 0.05, nT, u0, h1, h2, 0.0, chunk, None, bufs, ramp=ramp)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / 3
        print(f"chunk {chunk:5d} ramp {int(ramp)}: {dt * 1e3:7.1f} ms  {B * nT / dt / 1e6:6.2f} M time-steps/s", flush=True)
        del bufs
