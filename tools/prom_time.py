"""Time the fused PROM sweep (config 3/4 shape) on cuda:0: python tools/prom_time.py [B] [n] [proj] [steps]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import burgers_b200
if os.environ.get("BX_VARIANT_LIB"):   # experiments: a library built by tools/build_variant.sh
    burgers_b200._lib.LIB_PATH = os.path.abspath(os.environ["BX_VARIANT_LIB"])
from burgers_b200 import sweep

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
n = int(sys.argv[2]) if len(sys.argv) > 2 else 28
proj = sys.argv[3] if len(sys.argv) > 3 else "LSPG"
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 500
X = np.linspace(0, 100, 512)
fem = burgers_b200.FEMBurgers(X)
_g = os.path.join(os.path.dirname(__file__), "..", "tests", "golden")
if n <= 96:   # the reference's stored POD modes (POD/modes/U_modes_tol_*.npy: 40 and 96 columns)
    Phi = np.ascontiguousarray(np.load(os.path.join(_g, "Phi_tol_1e-03.npy" if n <= 40 else "Phi_tol_1e-04.npy"))[:, :n])
else:         # larger bases: POD modes of a few oracle-free GPU FOM trajectories
    Us, _ = fem.fom_burgers_batch(0.05, 80, np.ones(512), np.linspace(4.3, 5.4, 6), 0.0, np.linspace(0.016, 0.029, 6))
    Phi = np.ascontiguousarray(np.linalg.svd(Us.reshape(-1, 512).T.cpu().numpy(), full_matrices=False)[0][:, :n])
g = int(round(B ** 0.5))
mu1, mu2 = sweep.parameter_grid(g, B // g)
mu1 = torch.from_numpy(mu1).cuda(); mu2 = torch.from_numpy(mu2).cuda()
U = torch.empty((mu1.numel(), steps + 1, 512), dtype=torch.float64, device="cuda")
it = torch.empty((mu1.numel(), steps), dtype=torch.int32, device="cuda")
u0 = torch.ones(512, dtype=torch.float64, device="cuda")
Phid = torch.from_numpy(Phi).cuda()
for r in range(3):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    fem.pod_prom_burgers_batch(0.05, steps, u0, mu1, 0.0, mu2, Phid, proj, out=U, iters_out=it)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    tot = int(it.sum())
    flop = tot * (2 * 512 * n * n + 12 * 512 * n + 2 * n ** 3 / 3)
    print(f"B={mu1.numel()} n={n} {proj} steps={steps}: {ms:.2f} ms, {tot} iterations, "
          f"{mu1.numel() * steps / ms * 1e3:.3e} time-steps/s, {tot / ms * 1e3:.3e} it/s, {flop / ms / 1e9:.2f} TFLOP/s (algorithmic)")
