"""bx_reduced_system (the reduced system of one nonlinear iteration for a caller-supplied tangent) and its
first consumer, FEMBurgers.local_prom_burgers (FEM/fem_burgers.py:979-1078), through the C ABI on the GPU:
  * against Ar, br computed HERE-in-the-build-container with the reference's own operators (tests/golden/local_prom.npz);
  * against the oracle over sizes: n = 1 .. 240 (one pass and several tile passes), N = 2 .. 2048, shared and per-sample
    tangents, E != 0, supg on / off, non-uniform mesh;
  * local_prom_burgers against the reference's runs with the artefacts it ships (10 clusters, switches of basis)."""
import numpy as np
import pytest
import torch

from conftest import rel_l2_per_step

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def local(golden):
    g = golden("local_prom.npz")
    return g, [g[f"basis_{c}"] for c in range(10)]


def _close(A, Aref, rtol):
    return np.abs(A - Aref).max() <= rtol * max(np.abs(Aref).max(), 1e-300)


@pytest.mark.parametrize("proj", ["LSPG", "Galerkin"])
def test_reduced_system_vs_reference_operators(proj, oracle, local):
    import burgers_b200
    g, bases = local
    X, _ = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X)
    for j in range(3):
        Phi = bases[int(g["rs_basis_ids"][j])]
        Ar, br = fem.reduced_system_batch(float(g["rs_At"]), float(g["rs_E"]), g["rs_Uk"][j], g["rs_Un"][j],
                                          [g["rs_mu"][j, 0]], [g["rs_mu"][j, 1]], Phi, proj)
        assert _close(Ar[0].cpu().numpy(), g[f"rs_Ar_{proj}_{j}"], 1e-13)
        assert _close(br[0].cpu().numpy(), g[f"rs_br_{proj}_{j}"], 1e-11)
    # the three samples in ONE call with a per-sample tangent (zero-padded to the widest basis)
    nmax = max(bases[int(c)].shape[1] for c in g["rs_basis_ids"])
    T = np.zeros((3, 512, nmax))
    for j, c in enumerate(g["rs_basis_ids"]):
        T[j, :, :bases[int(c)].shape[1]] = bases[int(c)]
    Ar, br = fem.reduced_system_batch(float(g["rs_At"]), float(g["rs_E"]), g["rs_Uk"], g["rs_Un"], g["rs_mu"][:, 0],
                                      g["rs_mu"][:, 1], T, proj)
    for j, c in enumerate(g["rs_basis_ids"]):
        n = bases[int(c)].shape[1]
        A = Ar[j].cpu().numpy()
        assert _close(A[:n, :n], g[f"rs_Ar_{proj}_{j}"], 1e-13) and np.abs(A[n:]).max(initial=0.0) == 0.0 \
            and np.abs(A[:, n:]).max(initial=0.0) == 0.0
        assert _close(br[j].cpu().numpy()[:n], g[f"rs_br_{proj}_{j}"], 1e-11)


@pytest.mark.parametrize("N,n", [(2, 1), (9, 3), (64, 7), (300, 8), (512, 28), (512, 31), (512, 40), (700, 96), (1024, 127),
                                 (512, 128), (2048, 160), (512, 240)])
@pytest.mark.parametrize("proj", ["LSPG", "Galerkin"])
def test_reduced_system_sizes_vs_oracle(N, n, proj, oracle):
    import burgers_b200
    rng = np.random.default_rng(N + n)
    X = np.sort(np.concatenate([[0.0, 100.0], 100.0 * rng.random(N - 2)])) if N > 2 else np.array([0.0, 1.0])
    fem = burgers_b200.FEMBurgers(X)
    B = 3
    T = np.linalg.qr(rng.standard_normal((N, n)))[0] if n <= N else None
    Tb = np.stack([T, T[::-1].copy(), np.ascontiguousarray(T * (1.0 + 0.1 * rng.random((N, n))))])
    Un = 1.0 + 4.0 * rng.random((B, N)); Uk = Un * (1.0 + 0.05 * rng.random((B, N)))
    mu1 = 4.0 + rng.random(B); mu2 = 0.015 + 0.02 * rng.random(B)
    for supg, E, tang in ((True, 0.0, T), (False, 0.02, Tb)):
        Ar, br = fem.reduced_system_batch(0.05, E, Uk, Un, mu1, mu2, tang, proj, supg=supg)
        Ar = Ar.cpu().numpy(); br = br.cpu().numpy()
        for b in range(B):
            dl, d, du, _, R = oracle.picard_system(X, Uk[b], Un[b], mu1[b], mu2[b], 0.05, E, supg)
            Ao, bo = oracle.reduced_system(dl, d, du, R, tang if tang.ndim == 2 else tang[b], proj)
            assert _close(Ar[b], Ao, 1e-12), (supg, b)
            assert _close(br[b], bo, 1e-10), (supg, b)


def test_reduced_system_error_surface(oracle):
    import burgers_b200
    X, _ = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X)
    U = np.ones((2, 512))
    with pytest.raises(ValueError, match="Projection method 'Petrov' is not available"):
        fem.reduced_system_batch(0.05, 0.0, U, U, [4.5, 4.6], [0.02, 0.02], np.zeros((512, 4)), "Petrov")
    with pytest.raises(ValueError):
        fem.reduced_system_batch(0.05, 0.0, U, U, [4.5, 4.6], [0.02, 0.02], np.zeros((511, 4)), "LSPG")
    with pytest.raises(burgers_b200.BurgersB200Error):
        fem.reduced_system_batch(0.05, 0.0, U, U, [4.5, 4.6], [0.02, 0.02], np.zeros((512, 241)), "LSPG")


@pytest.mark.parametrize("proj", ["LSPG", "Galerkin"])
def test_local_prom_vs_reference(proj, oracle, local):
    import burgers_b200
    g, bases = local
    X, T = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X, T)

    class KMeansLike:          # what the drop-in reads of the sklearn model the reference passes
        cluster_centers_ = g["centers"]

    for pi, (mu1, mu2) in enumerate(g["points"]):
        U = fem.local_prom_burgers(0.05, int(g["nT"]), np.ones(512), mu1, 0.0, mu2, KMeansLike(), dict(enumerate(bases)),
                                   g["U_global"], int(g["g"]), projection=proj)
        assert U.shape == (512, int(g["nT"]) + 1)
        assert np.array_equal(fem.last_iters, g[f"it_{pi}_{proj}"]) and np.array_equal(fem.last_clusters, g[f"cl_{pi}_{proj}"])
        assert rel_l2_per_step(U, g[f"U_{pi}_{proj}"]) <= 1e-10
    # batched: both switch runs and a fresh start in one lock-step call (per-sample basis, different widths)
    u0 = np.stack([g["sw_u0_0"], g["sw_u0_1"], np.ones(512)])
    Ub, itb, clb = fem.local_prom_burgers_batch(0.05, 8, u0, [4.56, 4.56, 4.56], 0.0, [0.019, 0.019, 0.019], g["centers"],
                                                bases, g["U_global"], int(g["g"]), proj)
    Ub = Ub.cpu().numpy(); itb = itb.cpu().numpy(); clb = clb.cpu().numpy()
    for si in range(2):
        assert np.array_equal(itb[si], g[f"sw_it_{si}_{proj}"]) and np.array_equal(clb[si], g[f"sw_cl_{si}_{proj}"])
        assert rel_l2_per_step(Ub[si].T, g[f"sw_U_{si}_{proj}"]) <= 1e-10
    assert np.array_equal(itb[2], g[f"it_0_{proj}"][:8]) and rel_l2_per_step(Ub[2].T, g[f"U_0_{proj}"][:, :9]) <= 1e-10
