"""GPU: error surface and host-side behaviour of the drop-in class: non-finite trajectories are
flagged (-1) and never abort a batch, the single-sample drop-ins keep the reference's exceptions, verbose=True
reproduces the reference's progress lines (golden: the unmodified reference's own stdout,
tests/golden/make_verbose_golden.py), output buffers are validated, and an object bound to a device / used from
several streams runs where it should."""
import contextlib
import io
import re
import warnings

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


def _verbose_lines(fem, fn):
    buf = io.StringIO()
    fem.verbose = True
    try:
        with contextlib.redirect_stdout(buf):
            fn()
    finally:
        fem.verbose = False
    return buf.getvalue().splitlines()


def _compare_lines(got, want, rel):
    assert len(got) == len(want)
    num = re.compile(r"[-+]?\d+\.?\d*(?:[eE][-+]?\d+)?")
    for a, b in zip(got, want):
        ta, tb = num.sub("#", a), num.sub("#", b)
        assert ta == tb, (a, b)
        va, vb = [float(x) for x in num.findall(a)], [float(x) for x in num.findall(b)]
        assert np.allclose(va, vb, rtol=rel, atol=0), (a, b)


def test_verbose_fom_prints_the_reference_lines(fem512):
    want = open(f"{GOLDEN}/verbose_fom.txt").read().splitlines()
    got = _verbose_lines(fem512, lambda: fem512.fom_burgers(0.05, 3, np.ones(512), 4.75, 0.0, 0.02))
    _compare_lines(got, want, 1e-8)     # the printed error is |delta|/|U1| to 16 digits; ours agrees to rounding


def test_verbose_prom_and_qm_print_the_reference_lines(fem512):
    Phi28 = np.ascontiguousarray(np.load(f"{GOLDEN}/Phi_tol_1e-04.npy")[:, :28])
    want = open(f"{GOLDEN}/verbose_prom.txt").read().splitlines()
    got = _verbose_lines(fem512, lambda: fem512.pod_prom_burgers(0.05, 3, np.ones(512), 4.75, 0.0, 0.02, Phi28, "LSPG"))
    _compare_lines(got, want, 1e-7)
    Phi, H = np.load(f"{GOLDEN}/qm_Phi.npy"), np.load(f"{GOLDEN}/qm_H.npy")
    want = open(f"{GOLDEN}/verbose_qm.txt").read().splitlines()
    got = _verbose_lines(fem512, lambda: fem512.pod_quadratic_manifold(0.05, 3, np.ones(512), 4.75, 0.0, 0.02, Phi, H, "LSPG"))
    _compare_lines(got, want, 2e-3)     # three printed digits


def test_verbose_off_prints_nothing(fem512):
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        fem512.fom_burgers(0.05, 2, np.ones(512), 4.75, 0.0, 0.02)
    assert buf.getvalue() == ""


@pytest.mark.parametrize("W", [-1, 2, 4])
def test_nonfinite_sample_is_flagged_and_the_batch_goes_on(W, oracle, fem512):
    """One sample overflows (left boundary value 1e200: |U|^2 is inf, so the reference's error ratio is nan from the
    first iteration on); its steps carry -1 from the first step whose norm is not finite and NaN snapshots after it,
    the other samples of the batch equal the oracle."""
    import burgers_b200
    X, _ = oracle.mesh(512)
    mu1 = np.array([4.3, 1e200, 5.1, 4.75, 4.9]); mu2 = np.array([0.016, 0.02, 0.029, 0.02, 0.024])
    burgers_b200.set_tuning("fom_warps_per_trajectory", W)
    try:
        U, it = fem512.fom_burgers_batch(0.05, 6, np.ones(512), mu1, 0.0, mu2)
    finally:
        burgers_b200.set_tuning("fom_warps_per_trajectory", -1)
    U = U.cpu().numpy(); it = it.cpu().numpy()
    first = int(np.argmax(it[1] < 0))
    assert it[1, first] == -1 and (it[1, first:] == -1).all() and (it[1, :first] >= 1).all()
    with np.errstate(over="ignore"):
        assert np.isnan(U[1, first + 2:]).all() and not np.isfinite(np.sum(U[1, first + 1] ** 2))
    for b in (0, 2, 3, 4):
        Uo, ito = oracle.fom_burgers(X, 0.05, 6, np.ones(512), mu1[b], 0.0, mu2[b])
        assert np.array_equal(it[b], ito) and rel_l2_per_step(U[b].T, Uo) <= TOL
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        fem512.fom_burgers(0.05, 3, np.ones(512), 1e200, 0.0, 0.02)
    assert any(issubclass(x.category, RuntimeWarning) for x in w) and (fem512.last_iters < 0).any()


@pytest.mark.parametrize("n", [9, 40])
def test_singular_reduced_system(n, oracle, fem512):
    """Two identical basis columns make Phi^T A Phi exactly singular: np.linalg.solve raises LinAlgError in the
    reference (FEM/fem_burgers.py:767); the batched surface flags the sample, the drop-in raises."""
    X, _ = oracle.mesh(512)
    Phi = np.ascontiguousarray(np.load(f"{GOLDEN}/Phi_tol_1e-03.npy")[:, :n])
    bad = Phi.copy(); bad[:, 3] = bad[:, 2]
    with pytest.raises(np.linalg.LinAlgError):
        oracle.pod_prom_burgers(X, 0.05, 2, np.ones(512), 4.75, 0.0, 0.02, bad, "Galerkin")
    U, it = fem512.pod_prom_burgers_batch(0.05, 4, np.ones(512), [4.75, 5.0], 0.0, [0.02, 0.025], bad, "Galerkin")
    assert (it.cpu().numpy() == -1).all() and np.isnan(U.cpu().numpy()[:, 2:]).all()
    with pytest.raises(np.linalg.LinAlgError):
        fem512.pod_prom_burgers(0.05, 2, np.ones(512), 4.75, 0.0, 0.02, bad, "Galerkin")
    # and a healthy basis in the same process still runs
    Uo, ito = oracle.pod_prom_burgers(X, 0.05, 3, np.ones(512), 4.75, 0.0, 0.02, Phi, "Galerkin")
    Ug = fem512.pod_prom_burgers(0.05, 3, np.ones(512), 4.75, 0.0, 0.02, Phi, "Galerkin")
    assert np.array_equal(fem512.last_iters, ito) and rel_l2_per_step(Ug, Uo) <= TOL


def test_qm_singular_manifold_is_flagged(fem512):
    Phi, H = np.load(f"{GOLDEN}/qm_Phi.npy").copy(), np.load(f"{GOLDEN}/qm_H.npy") * 0.0
    Phi[:, 5] = Phi[:, 4]
    U, it = fem512.pod_quadratic_manifold_batch(0.05, 3, np.ones(512), [4.75, 4.9, 5.2], 0.0, [0.02, 0.02, 0.03], Phi, H,
                                                "Galerkin")
    assert (it.cpu().numpy() == -1).all()
    with pytest.raises(np.linalg.LinAlgError):
        fem512.pod_quadratic_manifold(0.05, 2, np.ones(512), 4.75, 0.0, 0.02, Phi, H, "Galerkin")


def test_output_buffers_are_validated(fem512):
    import torch
    B, nT = 3, 4
    mu1, mu2 = np.full(B, 4.75), np.full(B, 0.02)
    dev = fem512.device
    good_U = torch.empty((B, nT + 1, 512), dtype=torch.float64, device=dev)
    with pytest.raises(ValueError, match="iters_out"):
        fem512.fom_burgers_batch(0.05, nT, np.ones(512), mu1, 0.0, mu2, out=good_U,
                                 iters_out=torch.empty((B, nT), dtype=torch.int64, device=dev))
    with pytest.raises(ValueError, match="iters_out"):
        fem512.fom_burgers_batch(0.05, nT, np.ones(512), mu1, 0.0, mu2, out=good_U,
                                 iters_out=torch.empty((B, nT - 1), dtype=torch.int32, device=dev))
    with pytest.raises(ValueError, match="iters_out"):
        fem512.fom_burgers_batch(0.05, nT, np.ones(512), mu1, 0.0, mu2, out=good_U,
                                 iters_out=torch.empty((B, nT), dtype=torch.int32))            # host tensor
    with pytest.raises(ValueError, match="out must"):
        fem512.fom_burgers_batch(0.05, nT, np.ones(512), mu1, 0.0, mu2, out=torch.empty((B, nT + 1, 512), dtype=torch.float64))
    with pytest.raises(ValueError, match="err_hist"):
        fem512.fom_burgers_batch(0.05, nT, np.ones(512), mu1, 0.0, mu2,
                                 err_hist=torch.empty((B, nT, 19), dtype=torch.float64, device=dev))


def test_two_streams_share_one_object(oracle, fem512):
    """Concurrent launches of the same FEMBurgers from two streams: per-stream workspaces / queue counters."""
    import torch
    X, _ = oracle.mesh(512)
    mu1, mu2 = oracle.parameter_grid(20, 20)
    Phi = np.ascontiguousarray(np.load(f"{GOLDEN}/Phi_tol_1e-03.npy")[:, :28])
    U0, it0 = fem512.fom_burgers_batch(0.05, 10, np.ones(512), mu1, 0.0, mu2)
    P0, pit0 = fem512.pod_prom_burgers_batch(0.05, 10, np.ones(512), mu1, 0.0, mu2, Phi, "LSPG")
    torch.cuda.synchronize()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(3):
        with torch.cuda.stream(s1):
            U1, it1 = fem512.fom_burgers_batch(0.05, 10, np.ones(512), mu1, 0.0, mu2)
            P1, pit1 = fem512.pod_prom_burgers_batch(0.05, 10, np.ones(512), mu1, 0.0, mu2, Phi, "LSPG")
        with torch.cuda.stream(s2):
            U2, it2 = fem512.fom_burgers_batch(0.05, 10, np.ones(512), mu1, 0.0, mu2)
            P2, pit2 = fem512.pod_prom_burgers_batch(0.05, 10, np.ones(512), mu1, 0.0, mu2, Phi, "LSPG")
        torch.cuda.synchronize()
        for U, it in ((U1, it1), (U2, it2)):
            assert torch.equal(U, U0) and torch.equal(it, it0)
        for P, pit in ((P1, pit1), (P2, pit2)):
            assert torch.equal(P, P0) and torch.equal(pit, pit0)


def test_object_bound_to_another_device(oracle):
    import torch
    import burgers_b200
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    X, _ = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X, device="cuda:1")
    with torch.cuda.device(0):
        U, it = fem.fom_burgers_batch(0.05, 4, np.ones(512), [4.75], 0.0, [0.02])
    assert U.device == torch.device("cuda:1")
    Uo, ito = oracle.fom_burgers(X, 0.05, 4, np.ones(512), 4.75, 0.0, 0.02)
    assert np.array_equal(it[0].cpu().numpy(), ito) and rel_l2_per_step(U[0].cpu().numpy().T, Uo) <= TOL
