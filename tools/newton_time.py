"""Time the Newton FOM kernel: python tools/newton_time.py [N] [B] [steps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import burgers_b200
if os.environ.get("BX_VARIANT_LIB"):
    burgers_b200._lib.LIB_PATH = os.path.abspath(os.environ["BX_VARIANT_LIB"])
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
B = int(sys.argv[2]) if len(sys.argv) > 2 else 592
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 100
X = np.linspace(0, 100, N); At = 0.05 * 511.0 / (N - 1)
fem = burgers_b200.FEMBurgers(X)
rng = np.random.default_rng(0)
mu1 = rng.uniform(4.25, 5.5, B); mu2 = rng.uniform(0.015, 0.03, B)
for r in range(3):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    U, it = fem.fom_burgers_newton_batch(At, steps, np.ones(N), mu1, 0.0, mu2)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1); tot = int(it.clamp(min=0).sum())
    print(f"newton N={N} B={B} steps={steps}: {ms:.2f} ms, {tot} Newton iterations, {tot / ms / 1e3:.2f} M it/s, finite={bool(torch.isfinite(U).all())}")
