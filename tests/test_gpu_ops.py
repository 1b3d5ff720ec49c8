"""GPU parity of the L1 operators and of the banded Picard assemble / solve kernels (K1, K2)
against the reference outputs in tests/golden/ops.npz and the oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
OPS_CASES = ["u512_rand", "u512_shock", "u512_signed", "u37_nonuniform", "u37_zero_velocity"]


def bands_of_csc(A):
    A = A.toarray(); N = A.shape[0]
    dl = np.zeros(N); du = np.zeros(N)
    dl[1:] = np.diag(A, -1); du[:-1] = np.diag(A, 1)
    return np.stack([dl, np.diag(A), du])


@pytest.mark.parametrize("case", OPS_CASES)
def test_legacy_operators_vs_reference(case, golden):
    import burgers_b200
    g = golden("ops.npz")
    X, U, mu2 = g[case + "_X"], g[case + "_U"], float(g[case + "_mu2"])
    fem = burgers_b200.FEMBurgers(X)
    scale = lambda a: max(1.0, np.abs(a).max())
    for name, A in (("M", fem.compute_mass_matrix()), ("K", fem.compute_diffusion_matrix()),
                    ("C", fem.compute_convection_matrix(U)), ("dC", fem.compute_convection_matrix_derivative(U))):
        assert A.format == "csc" and A.shape == (X.size, X.size)
        ref = g[f"{case}_{name}"]
        assert np.abs(bands_of_csc(A) - ref).max() <= 4e-15 * scale(ref), name
    F = fem.compute_forcing_vector(mu2); S = fem.compute_supg_term(U, mu2)
    assert np.abs(F - g[case + "_F"]).max() <= 4e-15 * scale(g[case + "_F"])
    assert np.abs(S - g[case + "_S"]).max() <= 1e-14 * scale(g[case + "_S"])


@pytest.mark.parametrize("N,B", [(512, 7), (37, 3), (64, 1), (1000, 2), (2, 2)])
def test_assemble_and_solve_vs_oracle(N, B, oracle):
    import burgers_b200
    rng = np.random.default_rng(N)
    X = np.concatenate([[0.0], np.cumsum(0.1 + 0.2 * rng.random(N - 1))])
    fem = burgers_b200.FEMBurgers(X)
    Uk = 1.0 + 4.0 * rng.random((B, N)); Un = 1.0 + 4.0 * rng.random((B, N))
    mu1 = 4.0 + rng.random(B); mu2 = 0.015 + 0.02 * rng.random(B)
    for supg, E in ((True, 0.0), (False, 0.0), (True, 0.02)):
        dl, d, du, b, R = fem.assemble_picard(Uk, Un, mu1, mu2, 0.05, E, supg)
        odl, od, odu, ob, oR = oracle.picard_system(X, Uk, Un, mu1, mu2, 0.05, E, supg)
        for got, ref in ((dl, odl), (d, od), (du, odu), (b, ob), (R, oR)):
            assert np.abs(got.cpu().numpy() - ref).max() <= 1e-13 * max(1.0, np.abs(ref).max())
        U1, err = fem.tridiag_solve_update(dl, d, du, R, Uk)
        for i in range(B):
            delta = oracle.tridiag_solve(odl[i], od[i], odu[i], -oR[i])
            ref = Uk[i] + delta
            assert np.linalg.norm(U1[i].cpu().numpy() - ref) <= 1e-12 * np.linalg.norm(ref)
            e = np.linalg.norm(delta) / np.linalg.norm(ref)
            assert abs(err[i].item() - e) <= 1e-10 * e


def test_banded_picard_loop_equals_fused_kernel(oracle):
    """Iterating K1 + K2 from the host reproduces the fused trajectory kernel (same algorithm,
    bands through HBM instead of registers)."""
    import torch
    import burgers_b200
    X, _ = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X)
    mu1 = np.array([4.75, 5.5]); mu2 = np.array([0.02, 0.015])
    Uf, itf = fem.fom_burgers_batch(0.05, 3, np.ones(512), mu1, 0.0, mu2)
    Un = torch.ones(2, 512, dtype=torch.float64, device="cuda")
    for n in range(3):
        Uk = Un.clone()
        for b in range(2):
            k = 0; err = 1.0
            ub = Uk[b:b + 1].clone()
            while err > 1e-6 and k < 20:
                dl, d, du, _, R = fem.assemble_picard(ub, Un[b:b + 1], mu1[b:b + 1], mu2[b:b + 1], 0.05, 0.0, True, False)
                ub, e = fem.tridiag_solve_update(dl, d, du, R, ub)
                err = e.item(); k += 1
            assert k == itf[b, n].item()
            Uk[b] = ub[0]
        Un = Uk
        assert (Un - Uf[:, n + 1]).norm() <= 1e-12 * Un.norm()
