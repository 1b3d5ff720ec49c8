"""Time the fine-mesh FOM trajectory (BASELINE config 5): python tools/fine_time.py [log2N] [steps] [cfl|dt]
cfl: At = 0.05*511/(N-1) (5a, converging); dt: At = 0.05 (5b, every step runs into the 20-iteration cap)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import burgers_b200

lg = int(sys.argv[1]) if len(sys.argv) > 1 else 20
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 500
mode = sys.argv[3] if len(sys.argv) > 3 else "cfl"
N = 1 << lg
X = np.linspace(0, 100, N)
At = 0.05 * 511.0 / (N - 1) if mode == "cfl" else 0.05
fem = burgers_b200.FEMBurgers(X)
U = torch.empty((1, steps + 1, N), dtype=torch.float64, device="cuda")
it = torch.empty((1, steps), dtype=torch.int32, device="cuda")
u0 = torch.ones(N, dtype=torch.float64, device="cuda")
mu1 = torch.tensor([4.75], dtype=torch.float64, device="cuda"); mu2 = torch.tensor([0.02], dtype=torch.float64, device="cuda")
for r in range(3):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    fem.fom_burgers_batch(At, steps, u0, mu1, 0.0, mu2, out=U, iters_out=it)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    tot = int(it.sum())
    print(f"fine mesh N=2^{lg} {mode} steps={steps}: {ms:.2f} ms, {tot} Picard iterations, {steps / ms * 1e3:.1f} time-steps/s, "
          f"{ms / tot * 1e3:.2f} us/iteration, algorithmic {88.0 * N * tot / ms / 1e6:.0f} GB/s (88 N B/iteration model), finite={bool(torch.isfinite(U[0, -1]).all())}")
