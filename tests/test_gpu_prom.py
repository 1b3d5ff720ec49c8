"""GPU parity of the fused POD-PROM trajectory kernel (through the C ABI) against the oracle, the
reference-run golden vectors and the reference's stored PROM fixtures.
Tolerance (north_star): per-time-step relative L2 <= 1e-10 in FP64, identical iteration counts."""
import numpy as np
import pytest

from conftest import rel_l2_per_step, GOLDEN

pytestmark = pytest.mark.gpu
TOL = 1e-10


@pytest.fixture(scope="module")
def fem512(oracle):
    import burgers_b200
    X, T = oracle.mesh(512)
    return burgers_b200.FEMBurgers(X, T)


def _phi(tol, n=None):
    Phi = np.load(f"{GOLDEN}/Phi_tol_{tol}.npy")
    return np.ascontiguousarray(Phi if n is None else Phi[:, :n])


@pytest.mark.parametrize("n,proj", [(28, "LSPG"), (28, "Galerkin"), (9, "LSPG"), (40, "LSPG"), (40, "Galerkin")])
def test_legacy_prom_vs_reference_run(n, proj, golden, fem512):
    g = golden("prom_short.npz")
    Phi = _phi("1e-03", n) if n != 9 else _phi("1e-02")
    mu1, mu2 = (4.75, 0.02) if n != 9 else (5.19, 0.026)
    U = fem512.pod_prom_burgers(0.05, 8, np.ones(512), mu1, 0.0, mu2, Phi, proj)
    assert U.shape == (512, 9) and U.dtype == np.float64
    assert np.array_equal(fem512.last_iters, g[f"it_{n}_{proj}"])
    assert rel_l2_per_step(U, g[f"U_{n}_{proj}"]) <= TOL


def test_alias_and_bad_projection(fem512):
    assert fem512.pod_prom_newton.__func__ is fem512.pod_prom_burgers.__func__
    Phi = _phi("1e-02")
    for bad in ("lspg", "galerkin", "POD", None):
        with pytest.raises(ValueError, match="is not available"):
            fem512.pod_prom_burgers(0.05, 1, np.ones(512), 4.75, 0.0, 0.02, Phi, bad)


@pytest.mark.parametrize("name", ["U_PROM_tol_1e-02_mu1_4.750_mu2_0.0200_lspg",
                                  "U_PROM_tol_1e-02_mu1_5.190_mu2_0.0260_galerkin",
                                  "U_PROM_tol_4e-03_mu1_4.750_mu2_0.0200_galerkin",
                                  "U_PROM_tol_4e-03_mu1_4.560_mu2_0.0190_lspg",
                                  # 40 and 96 modes: the one-CTA-per-trajectory kernel (bx_prom_cta.cu);
                                  # tol 1e-03 / LSPG / (4.75, 0.02) is what POD/Results_thesis/prom_pod.py runs as shipped
                                  "U_PROM_tol_1e-03_mu1_4.750_mu2_0.0200_lspg",
                                  "U_PROM_tol_1e-03_mu1_5.190_mu2_0.0260_galerkin",
                                  "U_PROM_tol_1e-04_mu1_4.560_mu2_0.0190_lspg",
                                  "U_PROM_tol_1e-04_mu1_4.750_mu2_0.0200_galerkin"])
def test_stored_fixture_full_500_steps(name, golden, fem512):
    g = golden("prom_fixtures.npz")
    parts = name.split("_")
    tol, mu1, mu2, proj = parts[3], float(parts[5]), float(parts[7]), parts[8]
    U = fem512.pod_prom_burgers(0.05, 500, np.ones(512), mu1, 0.0, mu2, _phi(tol),
                                "LSPG" if proj == "lspg" else "Galerkin")
    assert rel_l2_per_step(U[:, g["steps"]], g[name]) <= TOL


def _basis(oracle, X, n):
    """POD basis of a few oracle FOM trajectories on the mesh X (so the PROM stays well posed)."""
    N = X.size
    mu1 = np.array([4.3, 5.4]); mu2 = np.array([0.016, 0.029])
    U, _ = oracle.fom_burgers_batch(X, 0.05, 40, np.ones(N), mu1, 0.0, mu2)
    S = U.reshape(-1, N).T
    Us = np.linalg.svd(S, full_matrices=False)[0]
    return np.ascontiguousarray(Us[:, :n])


@pytest.mark.parametrize("N,n", [(64, 1), (64, 5), (100, 8), (129, 15), (257, 16), (512, 23), (512, 24),
                                 (512, 31), (700, 12), (1024, 28)])
@pytest.mark.parametrize("proj", ["LSPG", "Galerkin"])
def test_sizes_vs_oracle(N, n, proj, oracle):
    import burgers_b200
    X = np.linspace(0.0, 100.0 * (N - 1) / 511.0, N)
    fem = burgers_b200.FEMBurgers(X)
    Phi = _basis(oracle, X, n)
    mu1 = np.array([4.4, 5.2, 4.75]); mu2 = np.array([0.017, 0.028, 0.02])
    U, it = fem.pod_prom_burgers_batch(0.05, 6, np.ones(N), mu1, 0.0, mu2, Phi, proj)
    U = U.cpu().numpy(); it = it.cpu().numpy()
    for b in range(3):
        Uo, ito = oracle.pod_prom_burgers(X, 0.05, 6, np.ones(N), mu1[b], 0.0, mu2[b], Phi, proj)
        assert np.array_equal(it[b], ito), (b, it[b], ito)
        assert rel_l2_per_step(U[b].T, Uo) <= TOL, b


def _big_basis(oracle, X, n):
    """n POD modes of six short oracle FOM trajectories (486 snapshots) on the mesh X."""
    N = X.size
    mu1 = np.linspace(4.3, 5.4, 6); mu2 = np.linspace(0.016, 0.029, 6)
    U, _ = oracle.fom_burgers_batch(X, 0.05, 80, np.ones(N), mu1, 0.0, mu2)
    Us = np.linalg.svd(U.reshape(-1, N).T, full_matrices=False)[0]
    return np.ascontiguousarray(Us[:, :n])


@pytest.mark.parametrize("N,n", [(40, 40), (64, 33), (512, 32), (300, 47), (512, 64), (1024, 72), (512, 129), (512, 160),
                                 (512, 227), (512, 240)])
@pytest.mark.parametrize("proj", ["LSPG", "Galerkin"])
def test_large_basis_vs_oracle(N, n, proj, oracle):
    """n >= 32 modes: one CTA per trajectory, Phi streamed from L2, blocked LU (reduced system in shared memory
    up to n ~ 140, in an L2-resident scratch beyond); several trajectories so the sample queue is exercised."""
    import burgers_b200
    X = np.linspace(0.0, 100.0 * (N - 1) / 511.0, N)
    fem = burgers_b200.FEMBurgers(X)
    Phi = _big_basis(oracle, X, n)
    mu1 = np.array([4.4, 5.2, 4.75]); mu2 = np.array([0.017, 0.028, 0.02])
    U, it = fem.pod_prom_burgers_batch(0.05, 5, np.ones(N), mu1, 0.0, mu2, Phi, proj)
    U = U.cpu().numpy(); it = it.cpu().numpy()
    for b in range(3):
        Uo, ito = oracle.pod_prom_burgers(X, 0.05, 5, np.ones(N), mu1[b], 0.0, mu2[b], Phi, proj)
        assert np.array_equal(it[b], ito), (b, it[b], ito)
        assert rel_l2_per_step(U[b].T, Uo) <= TOL, b


@pytest.mark.parametrize("kernel", ["register", "cta"])
@pytest.mark.parametrize("N,n", [(512, 32), (512, 33), (512, 39), (512, 40), (512, 41), (300, 47), (512, 48), (64, 40), (1000, 40)])
@pytest.mark.parametrize("proj", ["LSPG", "Galerkin"])
def test_32_to_48_modes_both_kernels_vs_oracle(N, n, proj, kernel, oracle):
    """32 <= n <= 48: the one-warp-per-trajectory register kernel (csrc/bx_prom_wide.cu; NB = 5 / 6 column blocks, R beside
    the tiles at n = 40 and 48, reduced solve in shared memory with dgetf2's pivot order) and the CTA kernel forced through
    the tuning knob, more samples than one CTA holds, a batched u0."""
    import burgers_b200
    if n > N:
        pytest.skip("n <= N")
    X = np.linspace(0.0, 100.0 * (N - 1) / 511.0, N)
    fem = burgers_b200.FEMBurgers(X)
    Phi = _big_basis(oracle, X, n)
    rng = np.random.default_rng(n)
    B = 11
    mu1 = 4.3 + 1.1 * rng.random(B); mu2 = 0.016 + 0.013 * rng.random(B)
    u0 = 1.0 + 0.02 * rng.random((B, N))
    burgers_b200.set_tuning("prom_warps_per_trajectory", 8 if kernel == "cta" else -1)
    try:
        U, it = fem.pod_prom_burgers_batch(0.05, 4, u0, mu1, 0.0, mu2, Phi, proj)
    finally:
        burgers_b200.set_tuning("prom_warps_per_trajectory", -1)
    U = U.cpu().numpy(); it = it.cpu().numpy()
    for b in (0, 5, 10):
        Uo, ito = oracle.pod_prom_burgers(X, 0.05, 4, u0[b], mu1[b], 0.0, mu2[b], Phi, proj)
        assert np.array_equal(it[b], ito), (b, it[b], ito)
        assert rel_l2_per_step(U[b].T, Uo) <= TOL, b


def test_too_many_modes_is_an_error(oracle, fem512):
    import burgers_b200
    X, _ = oracle.mesh(512)
    Phi = np.linalg.qr(np.random.default_rng(0).standard_normal((512, 241)))[0]
    with pytest.raises(burgers_b200.BurgersB200Error):
        fem512.pod_prom_burgers_batch(0.05, 1, np.ones(512), [4.75], 0.0, [0.02], Phi, "LSPG")


def test_ragged_batch_batched_u0_and_empty(oracle, fem512):
    import torch
    X, _ = oracle.mesh(512)
    Phi = _phi("1e-03", 28)
    rng = np.random.default_rng(3)
    B = 11   # not a multiple of the groups per CTA
    mu1 = 4.25 + 1.25 * rng.random(B); mu2 = 0.015 + 0.015 * rng.random(B)
    u0 = 1.0 + 0.1 * rng.random((B, 512))
    U, it = fem512.pod_prom_burgers_batch(0.05, 4, u0, mu1, 0.0, mu2, Phi, "LSPG")
    U = U.cpu().numpy(); it = it.cpu().numpy()
    for b in (0, 5, 10):
        Uo, ito = oracle.pod_prom_burgers(X, 0.05, 4, u0[b], mu1[b], 0.0, mu2[b], Phi, "LSPG")
        assert np.array_equal(it[b], ito) and rel_l2_per_step(U[b].T, Uo) <= TOL
    U, it = fem512.pod_prom_burgers_batch(0.05, 4, np.ones(512), np.zeros(0), 0.0, np.zeros(0), Phi, "LSPG")
    assert U.shape == (0, 5, 512) and it.shape == (0, 4)
    U, it = fem512.pod_prom_burgers_batch(0.05, 0, np.ones(512), [4.75], 0.0, [0.02], Phi, "Galerkin")
    torch.cuda.synchronize()
    assert U.shape == (1, 1, 512) and np.array_equal(U.cpu().numpy()[0, 0], np.ones(512))


def test_prom_sweep_properties(oracle, fem512):
    """Config 3 at reduced width: 256 samples x 500 steps LSPG-28; spot-check against the oracle and
    check size-independent properties (every snapshot lies in span(Phi) after step 0, counts in range)."""
    X, _ = oracle.mesh(512)
    Phi = _phi("1e-03", 28)
    mu1, mu2 = oracle.parameter_grid(16, 16)
    U, it = fem512.pod_prom_burgers_batch(0.05, 500, np.ones(512), mu1, 0.0, mu2, Phi, "LSPG")
    it = it.cpu().numpy()
    assert it.min() >= 1 and it.max() <= 20
    import torch
    Phid = torch.from_numpy(Phi).to(U.device)
    Ui = U[:, 1:, :]
    res = Ui - (Ui @ Phid) @ Phid.T
    assert float(res.norm(dim=2).max() / Ui.norm(dim=2).min()) <= 1e-12
    for b in (0, 100, 255):
        Uo, ito = oracle.pod_prom_burgers(X, 0.05, 60, np.ones(512), mu1[b], 0.0, mu2[b], Phi, "LSPG")
        assert np.array_equal(it[b, :60], ito)
        assert rel_l2_per_step(U[b, :61].cpu().numpy().T, Uo) <= TOL
