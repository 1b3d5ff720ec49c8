"""Time the quadratic-manifold PROM sweep on cuda:0: python tools/qm_time.py [B] [proj] [steps] [n] [schedule]
n = 21 uses the reference's stored Phi/H; other n truncate Phi and ridge-fit H from FOM snapshots."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import burgers_b200
from burgers_b200 import sweep

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
proj = sys.argv[2] if len(sys.argv) > 2 else "Galerkin"
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 500
n = int(sys.argv[4]) if len(sys.argv) > 4 else 21
if os.environ.get("QM_GROUP"):
    burgers_b200.set_tuning("qm_group", int(os.environ["QM_GROUP"]))
if len(sys.argv) > 5:
    burgers_b200.set_tuning("qm_schedule", int(sys.argv[5]))   # 0 v2 / 1 shared tile / 2 chunked global
G = os.path.join(os.path.dirname(__file__), "..", "tests", "golden")
X = np.linspace(0, 100, 512)
fem = burgers_b200.FEMBurgers(X)
if n == 21:
    Phi = np.load(os.path.join(G, "qm_Phi.npy")); H = np.load(os.path.join(G, "qm_H.npy"))
else:
    # offline stage as Quadratic_manifold/build_quadratic_manifold.py does it (alpha = 1e-2), from a 3x3 FOM sweep
    m1, m2 = sweep.parameter_grid(3, 3)
    Us, _ = fem.fom_burgers_batch(0.05, 500, np.ones(512), m1, 0.0, m2)
    S = Us.reshape(-1, 512).T.cpu().numpy()
    Phi = np.linalg.svd(S, full_matrices=False)[0][:, :n].copy()
    q = Phi.T @ S
    iu, ju = np.triu_indices(n)
    Q = q[iu] * q[ju]
    Uq, s, VqT = np.linalg.svd(Q, full_matrices=False)
    f = s ** 2 / (s ** 2 + 1e-2 ** 2)
    H = ((Uq * f) @ ((VqT @ (S - Phi @ q).T) / s[:, None])).T.copy()
g = int(round(B ** 0.5))
mu1, mu2 = sweep.parameter_grid(g, B // g)
mu1 = torch.from_numpy(mu1).cuda(); mu2 = torch.from_numpy(mu2).cuda()
U = torch.empty((mu1.numel(), steps + 1, 512), dtype=torch.float64, device="cuda")
it = torch.empty((mu1.numel(), steps), dtype=torch.int32, device="cuda")
u0 = torch.ones(512, dtype=torch.float64, device="cuda")
Phid = torch.from_numpy(Phi).cuda(); Hd = torch.from_numpy(H).cuda()
for r in range(3):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    fem.pod_quadratic_manifold_batch(0.05, steps, u0, mu1, 0.0, mu2, Phid, Hd, proj, out=U, iters_out=it)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    tot = int(it.sum())
    k = n * (n + 1) // 2
    flop = tot * (2 * 512 * n * n + 12 * 512 * n + 2 * n ** 3 / 3 + 4 * 512 * k + 2 * 512 * (n + k))
    print(f"QM B={mu1.numel()} n={n} {proj} steps={steps}: {ms:.2f} ms, {tot} iterations (max {int(it.max())}/step), "
          f"{mu1.numel() * steps / ms * 1e3:.3e} time-steps/s, {tot / ms * 1e3:.3e} it/s, {flop / ms / 1e9:.2f} TFLOP/s (algorithmic)")
