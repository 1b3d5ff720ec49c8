"""GPU parity of the fine-mesh FOM kernel (bx_fom_picard_finemesh: persistent cooperative kernel, three-level
partition solve with grid barriers) against the oracle at mesh sizes the oracle finishes in seconds, plus
size-independent properties at BASELINE config 5's full size (2^20 nodes).
The reference's scheme is only stable at constant CFL: At = 0.05 * 511 / (N - 1).
Tolerance (north_star): per-time-step relative L2 <= 1e-10 in FP64, identical iteration counts."""
import numpy as np
import pytest

from conftest import rel_l2_per_step

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _run(oracle, N, nsteps, mu1=4.75, mu2=0.02, E=0.0, X=None, u0=None, tol=1e-6, kmax=20, At=None):
    import burgers_b200
    X = np.linspace(0.0, 100.0, N) if X is None else X
    At = 0.05 * 511.0 / (N - 1) if At is None else At
    u0 = np.ones(N) if u0 is None else u0
    fem = burgers_b200.FEMBurgers(X)
    U, it = fem.fom_burgers_batch(At, nsteps, u0, [mu1], E, [mu2], tol=tol, kmax=kmax)
    Uo, ito = oracle.fom_burgers(X, At, nsteps, u0, mu1, E, mu2, tol=tol, kmax=kmax)
    return U[0].cpu().numpy().T, it[0].cpu().numpy(), Uo, ito


@pytest.mark.parametrize("rows", [-1, 16, 8, 4])
@pytest.mark.parametrize("N", [1025, 8192, 8193, 20000, 65536, 131072 + 7])
def test_finemesh_vs_oracle(N, rows, oracle):
    """rows: mesh rows per thread of the kernel instance (BX_TUNE_FINE_ROWS; -1 = automatic)."""
    import burgers_b200
    burgers_b200.set_tuning("fine_rows", rows)
    try:
        U, it, Uo, ito = _run(oracle, N, 12)
    finally:
        burgers_b200.set_tuning("fine_rows", -1)
    assert np.array_equal(it, ito), (it, ito)
    assert rel_l2_per_step(U, Uo) <= TOL


def test_finemesh_nonuniform_mesh_diffusion_batched_u0(oracle):
    import burgers_b200
    rng = np.random.default_rng(11)
    N = 30011
    X = np.cumsum(0.5 + rng.random(N)); X = 100.0 * (X - X[0]) / (X[-1] - X[0])
    At = 0.05 * 511.0 / (N - 1)
    mu1 = np.array([4.3, 5.4]); mu2 = np.array([0.017, 0.028])
    u0 = np.stack([np.ones(N), 1.0 + 0.2 * np.sin(X / 7.0)])
    fem = burgers_b200.FEMBurgers(X)
    U, it = fem.fom_burgers_batch(At, 8, u0, mu1, 1e-4, mu2)
    U = U.cpu().numpy(); it = it.cpu().numpy()
    for b in range(2):
        Uo, ito = oracle.fom_burgers(X, At, 8, u0[b], mu1[b], 1e-4, mu2[b])
        assert np.array_equal(it[b], ito)
        assert rel_l2_per_step(U[b].T, Uo) <= TOL


def test_finemesh_iteration_cap_and_legacy_surface(oracle):
    """dt = 0.05 on a 4097-node mesh does not converge (SURVEY: diverges for N >= 768): the step stops at the
    iteration cap exactly as the reference's while-loop does.  The diverging Picard map is chaotic: a 1e-15
    relative perturbation of u0 changes the ORACLE's own 20th iterate by O(1) (1e-12 after 5 iterations, 1e-7 after
    10, 8e-2 after 15 -- measured with oracle.fom_burgers), so values are compared at a cap of 5, where rounding
    differences are still below the 1e-10 bar, and the reference's cap of 20 is checked for the count only
    (no parity claim there)."""
    import burgers_b200
    N = 4097
    X = np.linspace(0.0, 100.0, N)
    fem = burgers_b200.FEMBurgers(X)
    U, it = fem.fom_burgers_batch(0.05, 1, np.ones(N), [4.75], 0.0, [0.02], kmax=5)
    Uo, ito = oracle.fom_burgers(X, 0.05, 1, np.ones(N), 4.75, 0.0, 0.02, kmax=5)
    assert np.array_equal(it[0].cpu().numpy(), ito) and np.all(ito == 5)
    assert rel_l2_per_step(U[0].cpu().numpy().T, Uo) <= TOL
    U = fem.fom_burgers(0.05, 1, np.ones(N), 4.75, 0.0, 0.02)
    assert U.shape == (N, 2) and U.dtype == np.float64
    assert np.all(fem.last_iters == 20) and np.all(np.isfinite(U)) and U[0, 1] == 4.75


def test_finemesh_full_size_properties(oracle):
    """2^20 nodes: (i) while the inflow front has not reached a probe window, the solution there matches the
    same window of a 2^16-node run scaled to the same CFL?  Not size independent -- instead check:
    (a) Dirichlet node equals mu1 after every step, (b) first steps agree with the oracle (3 steps, seconds on CPU),
    (c) determinism: two runs are bit-identical."""
    import torch
    import burgers_b200
    N = 1 << 20
    X = np.linspace(0.0, 100.0, N)
    At = 0.05 * 511.0 / (N - 1)
    fem = burgers_b200.FEMBurgers(X)
    U1, it1 = fem.fom_burgers_batch(At, 3, np.ones(N), [4.75], 0.0, [0.02])
    U2, it2 = fem.fom_burgers_batch(At, 3, np.ones(N), [4.75], 0.0, [0.02])
    assert torch.equal(U1, U2) and torch.equal(it1, it2)
    U = U1[0].cpu().numpy()
    assert np.all(U[1:, 0] == 4.75)
    Uo, ito = oracle.fom_burgers(X, At, 3, np.ones(N), 4.75, 0.0, 0.02)
    assert np.array_equal(it1[0].cpu().numpy(), ito)
    assert rel_l2_per_step(U.T, Uo) <= TOL
