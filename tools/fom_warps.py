"""FOM sweep time against warps per trajectory (1 / 2 / 4) and batch size: the input of the automatic choice in
csrc/bx_fom.cu (fom_warps_per_trajectory).  usage: python tools/fom_warps.py [nsteps] [B ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import burgers_b200
if os.environ.get("BX_VARIANT_LIB"):   # experiments: a library built by tools/build_variant.sh
    burgers_b200._lib.LIB_PATH = os.path.abspath(os.environ["BX_VARIANT_LIB"])

nsteps = int(sys.argv[1]) if len(sys.argv) > 1 else 500
Bs = [int(x) for x in sys.argv[2:]] or [148, 296, 512, 888, 1024, 1776, 2048, 4096]
fem = burgers_b200.FEMBurgers(np.linspace(0, 100, 512))
u0 = torch.ones(512, dtype=torch.float64, device="cuda")
g1 = np.repeat(np.linspace(4.25, 5.5, 64), 64); g2 = np.tile(np.linspace(0.015, 0.03, 64), 64)
CASES = [(0, 1), (1, 1), (1, 2), (1, 4)]   # (kernel: 0 throughput / 1 latency, warps per trajectory)
print(f"{'B':>6} " + " ".join(f"{'k%d W=%d ms' % c:>10}" for c in CASES) + "   its")
for B in Bs:
    sel = np.linspace(0, 4095, B).astype(int)
    mu1 = torch.from_numpy(g1[sel].copy()).cuda(); mu2 = torch.from_numpy(g2[sel].copy()).cuda()
    U = torch.empty((B, nsteps + 1, 512), dtype=torch.float64, device="cuda")
    it = torch.empty((B, nsteps), dtype=torch.int32, device="cuda")
    row, ref = [], None
    for kern, W in CASES:
        burgers_b200.set_tuning("fom_warps_per_trajectory", W)
        burgers_b200.set_tuning("fom_kernel", kern)
        for _ in range(2):
            fem.fom_burgers_batch(0.05, nsteps, u0, mu1, 0.0, mu2, out=U, iters_out=it)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(3):
            fem.fom_burgers_batch(0.05, nsteps, u0, mu1, 0.0, mu2, out=U, iters_out=it)
        e1.record(); torch.cuda.synchronize()
        row.append(e0.elapsed_time(e1) / 3)
        tot = int(it.sum())
        ref = tot if ref is None else ref
        assert tot == ref, (W, tot, ref)
    burgers_b200.set_tuning("fom_warps_per_trajectory", -1)
    burgers_b200.set_tuning("fom_kernel", -1)
    print(f"{B:>6} " + " ".join(f"{x:>10.3f}" for x in row) + f"   {ref}", flush=True)
    del U, it
