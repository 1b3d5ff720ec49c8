"""Small runs of every trajectory kernel family (a quick smoke driver; compute-sanitizer is closed on this pool, so
races are hunted with the bitwise run-to-run determinism tests in tests/test_gpu_determinism.py instead):
python tools/sanitize_small.py [fom|prom|promcta|qm|all]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import burgers_b200

which = sys.argv[1] if len(sys.argv) > 1 else "all"
rng = np.random.default_rng(0)


def basis(N, n):
    return np.ascontiguousarray(np.linalg.qr(rng.standard_normal((N, n)))[0])


def smooth_basis(X, n):
    k = np.arange(n)
    P = np.cos(np.outer(X / X[-1], k) * np.pi)
    return np.ascontiguousarray(np.linalg.qr(P)[0])


if which in ("all", "fom"):
    fem = burgers_b200.FEMBurgers(np.linspace(0, 100, 512))
    U, it = fem.fom_burgers_batch(0.05, 3, np.ones(512), [4.3, 5.0, 5.4], 0.0, [0.016, 0.02, 0.029])
    torch.cuda.synchronize(); print("fom", it.cpu().numpy().tolist())
if which in ("all", "promcta"):
    for N, n, proj in ((96, 33, "LSPG"), (96, 40, "Galerkin"), (512, 150, "LSPG")):
        X = np.linspace(0, 100.0 * (N - 1) / 511, N)
        fem = burgers_b200.FEMBurgers(X)
        U, it = fem.pod_prom_burgers_batch(0.05, 2, np.ones(N), [4.3, 5.0, 5.4], 0.0, [0.016, 0.02, 0.029],
                                           smooth_basis(X, n), proj)
        torch.cuda.synchronize(); print("promcta", N, n, proj, it.cpu().numpy().tolist(), bool(torch.isfinite(U).all()))
if which in ("all", "prom"):
    X = np.linspace(0, 100, 512); fem = burgers_b200.FEMBurgers(X)
    U, it = fem.pod_prom_burgers_batch(0.05, 2, np.ones(512), [4.3, 5.0, 5.4], 0.0, [0.016, 0.02, 0.029], smooth_basis(X, 28), "LSPG")
    torch.cuda.synchronize(); print("prom", it.cpu().numpy().tolist())
if which in ("all", "qm"):
    G = os.path.join(os.path.dirname(__file__), "..", "tests", "golden")
    Phi = np.load(os.path.join(G, "qm_Phi.npy")); H = np.load(os.path.join(G, "qm_H.npy"))
    X = np.linspace(0, 100, 512); fem = burgers_b200.FEMBurgers(X)
    for mode in ("1", "0", "2"):
        os.environ["BX_QM_TS"] = mode
        U, it = fem.pod_quadratic_manifold_batch(0.05, 2, np.ones(512), np.linspace(4.3, 5.4, 11), 0.0,
                                                 np.linspace(0.016, 0.029, 11), Phi, H, "LSPG")
        torch.cuda.synchronize(); print("qm mode", mode, it.cpu().numpy()[:3].tolist(), bool(torch.isfinite(U).all()))
