"""Round-robin segment scheduling of the FOM throughput kernel (csrc/bx_fom.cu): a sweep cut into work units of a
few time steps, each trajectory resumed from its own last snapshot by whichever warp draws the unit, must give
bit-identical snapshots and counts to the same kernel integrating whole trajectories — for every segment length,
including lengths that do not divide the step count, and with a trajectory that stops being finite mid-sweep."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture
def tuning():
    import burgers_b200
    yield burgers_b200.set_tuning
    for k in ("fom_kernel", "fom_segment_steps", "fom_warps_per_trajectory"):
        burgers_b200.set_tuning(k, -1)


def _sweep(fem, B, nT, mu1, mu2, N=512):
    U, it = fem.fom_burgers_batch(0.05, nT, np.ones(N), mu1[:B], 0.0, mu2[:B])
    torch.cuda.synchronize()
    return U, it


def test_segmented_sweep_is_bit_identical_to_whole_trajectories(oracle, tuning):
    import burgers_b200
    X, _ = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X)
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    B = 12 * sms + 301            # more trajectories than resident warps (12 per SM), ragged
    rng = np.random.default_rng(5)
    mu1 = 4.25 + 1.25 * rng.random(B); mu2 = 0.015 + 0.015 * rng.random(B)
    nT = 23
    tuning("fom_kernel", 0)
    tuning("fom_segment_steps", 0)           # whole trajectories from the queue
    U0, it0 = _sweep(fem, B, nT, mu1, mu2)
    for seg in (1, 4, 7, 23, 40, -1):        # -1: the default length
        tuning("fom_segment_steps", seg)
        U, it = _sweep(fem, B, nT, mu1, mu2)
        assert torch.equal(it, it0), seg
        assert torch.equal(U, U0), seg
    # and against the oracle on a strided sample
    sel = np.arange(0, B, 97)
    Uo, ito = oracle.fom_burgers_batch(X, 0.05, nT, np.ones(512), mu1[sel], 0.0, mu2[sel])
    assert np.array_equal(it0[sel].cpu().numpy(), ito)
    assert np.max(np.linalg.norm(U0[sel].cpu().numpy() - Uo, axis=2) / np.linalg.norm(Uo, axis=2)) <= 1e-10


def test_segmented_sweep_nonfinite_trajectory_keeps_its_flag(oracle, tuning):
    """A trajectory that blows up in an early segment: flag -1 and NaN snapshots from that step on in every later
    segment, the other trajectories of the batch untouched."""
    import burgers_b200
    X, _ = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X)
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    B = 12 * sms + 40
    mu1 = np.full(B, 4.75); mu2 = np.full(B, 0.02)
    bad = [3, B - 2]
    mu2[bad] = 40.0                         # exp(mu2 x) overflows the forcing: non-finite within the first steps
    nT = 12
    tuning("fom_kernel", 0)
    tuning("fom_segment_steps", 3)
    U, it = _sweep(fem, B, nT, mu1, mu2)
    it = it.cpu().numpy(); Uc = U.cpu().numpy()
    for b in bad:
        first = int(np.argmax(it[b] == -1))
        assert it[b, first] == -1 and (it[b, first:] == -1).all() and (it[b, :first] > 0).all()
        assert np.isnan(Uc[b, first + 2:]).all()
    good = np.setdiff1d(np.arange(B), bad)
    assert (it[good] > 0).all() and np.isfinite(Uc[good]).all()
    assert (it[good] == it[good[0]]).all() and (Uc[good] == Uc[good[0]]).all()


def _phi(n):
    from conftest import GOLDEN
    return np.ascontiguousarray(np.load(f"{GOLDEN}/Phi_tol_1e-04.npy")[:, :n])


@pytest.mark.parametrize("nmodes", [28, 40])
@pytest.mark.parametrize("proj", ["LSPG", "Galerkin"])
def test_prom_segmented_sweep_matches_whole_trajectories(proj, nmodes, oracle, tuning):
    """POD-PROM warp kernels (csrc/bx_prom.cu at 28 modes, csrc/bx_prom_wide.cu at 40) with more trajectories than warp slots: round-robin segments against whole
    trajectories from the queue.  A resumed trajectory projects its snapshot explicitly (q0 = Phi^T U, as the reference
    does at :770) where the running one uses (Phi^T Phi) q, so the two agree to rounding, not bit for bit: 1e-12 per
    step and identical counts; both against the oracle on a strided sample."""
    import burgers_b200
    X, _ = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X)
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    B = 8 * sms + 150
    rng = np.random.default_rng(7)
    mu1 = 4.25 + 1.25 * rng.random(B); mu2 = 0.015 + 0.015 * rng.random(B)
    Phi = _phi(nmodes)
    nT = 11
    tuning("fom_segment_steps", 0)
    U0, it0 = fem.pod_prom_burgers_batch(0.05, nT, np.ones(512), mu1, 0.0, mu2, Phi, proj)
    for seg in (1, 4, -1):
        tuning("fom_segment_steps", seg)
        U, it = fem.pod_prom_burgers_batch(0.05, nT, np.ones(512), mu1, 0.0, mu2, Phi, proj)
        assert torch.equal(it, it0), seg
        rel = (torch.linalg.norm(U - U0, dim=2) / torch.linalg.norm(U0, dim=2)).max().item()
        assert rel <= 1e-12, (seg, rel)
    sel = np.arange(0, B, 211)
    Us = U.cpu().numpy()[sel]; its = it.cpu().numpy()[sel]
    for j, b in enumerate(sel):
        Uo, ito = oracle.pod_prom_burgers(X, 0.05, nT, np.ones(512), mu1[b], 0.0, mu2[b], Phi, proj)
        assert np.array_equal(its[j], ito), b
        assert np.max(np.linalg.norm(Us[j].T - Uo, axis=0) / np.linalg.norm(Uo, axis=0)) <= 1e-10, b


@pytest.mark.parametrize("n", [21, 28])
def test_qm_segmented_sweep_is_bit_identical(n, oracle, tuning):
    """Quadratic-manifold kernel (csrc/bx_qm.cu, both tangent-row schedules by n) with more lock-step groups than SMs:
    every step starts from q = Phi^T U^m of the stored snapshot (:1129), so a resumed group reproduces the running one
    bit for bit."""
    import burgers_b200
    from conftest import GOLDEN
    X, _ = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X)
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    B = 8 * sms + 77
    rng = np.random.default_rng(9)
    mu1 = 4.25 + 1.25 * rng.random(B); mu2 = 0.015 + 0.015 * rng.random(B)
    if n == 21:
        Phi = np.load(f"{GOLDEN}/qm_Phi.npy"); H = np.load(f"{GOLDEN}/qm_H.npy")
    else:
        m1, m2 = oracle.parameter_grid(3, 3)
        Ut, _ = fem.fom_burgers_batch(0.05, 500, np.ones(512), m1, 0.0, m2)
        S = np.ascontiguousarray(Ut.cpu().numpy().reshape(-1, 512).T)
        Phi = _phi(n); q = Phi.T @ S
        H = np.ascontiguousarray(oracle.compute_H(oracle.build_Q(q), oracle.build_E(S, Phi, q), 1e-2))
    nT = 7
    tuning("fom_segment_steps", 0)
    U0, it0 = fem.pod_quadratic_manifold_batch(0.05, nT, np.ones(512), mu1, 0.0, mu2, Phi, H, "LSPG")
    for seg in (1, 3, -1):
        tuning("fom_segment_steps", seg)
        U, it = fem.pod_quadratic_manifold_batch(0.05, nT, np.ones(512), mu1, 0.0, mu2, Phi, H, "LSPG")
        assert torch.equal(it, it0), seg
        assert torch.equal(U, U0), seg
    b = B - 3
    Uo, ito = oracle.pod_quadratic_manifold(X, 0.05, nT, np.ones(512), mu1[b], 0.0, mu2[b], Phi, H, "LSPG")
    assert np.array_equal(it0[b].cpu().numpy(), ito)
    assert np.max(np.linalg.norm(U0[b].cpu().numpy().T - Uo, axis=0) / np.linalg.norm(Uo, axis=0)) <= 1e-10
