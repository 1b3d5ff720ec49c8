"""GPU parity of the fused fom_burgers_newton kernel (as coded in the reference: tridiagonal plus a
rank-one broadcast term, solved by Sherman-Morrison) against reference runs and the oracle.
Tolerance: per-time-step relative L2 <= 1e-10 in FP64, identical iteration counts."""
import numpy as np
import pytest

from conftest import rel_l2_per_step

pytestmark = pytest.mark.gpu
TOL = 1e-10


def test_legacy_newton_vs_reference_run(golden, oracle):
    import burgers_b200
    g = golden("newton_short.npz")
    fem = burgers_b200.FEMBurgers(g["X_64"])
    U = fem.fom_burgers_newton(0.05, 3, np.ones(64), 4.75, 0.0, 0.02)
    assert U.shape == (64, 4) and np.array_equal(fem.last_iters, g["it_64"])
    assert rel_l2_per_step(U, g["U_64"]) <= TOL
    X, T = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X, T)
    U = fem.fom_burgers_newton(0.05, 2, np.ones(512), 4.75, 0.0, 0.02)
    assert np.array_equal(fem.last_iters, g["it_512"]) and rel_l2_per_step(U, g["U_512"]) <= TOL


@pytest.mark.parametrize("N", [2, 3, 33, 64, 100, 256, 257, 512, 700, 1024])
def test_sizes_vs_oracle(N, oracle):
    import burgers_b200
    X = np.linspace(0.0, 100.0 * (N - 1) / 511.0, N)
    fem = burgers_b200.FEMBurgers(X)
    mu1 = np.array([4.3, 5.1, 4.75]); mu2 = np.array([0.016, 0.029, 0.02])
    U, it = fem.fom_burgers_newton_batch(0.05, 3, np.ones(N), mu1, 0.0, mu2)
    U = U.cpu().numpy(); it = it.cpu().numpy()
    for b in range(3):
        Uo, ito = oracle.fom_burgers_newton(X, 0.05, 3, np.ones(N), mu1[b], 0.0, mu2[b], dense=False)
        assert np.array_equal(it[b], ito), (b, it[b], ito)
        assert rel_l2_per_step(U[b].T, Uo) <= TOL, b


def test_nonuniform_mesh_diffusion_lambda_and_dense_oracle(oracle):
    import burgers_b200
    rng = np.random.default_rng(5)
    N = 90
    X = np.cumsum(0.1 + 0.2 * rng.random(N))
    u0 = 1.0 + 0.3 * rng.random(N)
    fem = burgers_b200.FEMBurgers(X)
    # lambda = 1 (undamped) converges in a handful of iterations: exercises the stopping test
    U, it = fem.fom_burgers_newton_batch(0.02, 4, u0, [2.0, 1.5], 0.01, [0.04, 0.03], lambda_=1.0, kmax=30)
    U = U.cpu().numpy(); it = it.cpu().numpy()
    for b, (m1, m2) in enumerate([(2.0, 0.04), (1.5, 0.03)]):
        Uo, ito = oracle.fom_burgers_newton(X, 0.02, 4, u0, m1, 0.01, m2, lambda_=1.0, kmax=30, dense=True)
        assert np.array_equal(it[b], ito), (it[b], ito)
        assert rel_l2_per_step(U[b].T, Uo) <= TOL
    assert it.max() < 30
