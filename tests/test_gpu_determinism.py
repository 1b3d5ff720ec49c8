"""GPU: run-to-run bitwise determinism of the trajectory kernels whose CTAs synchronise internally (barriers,
mbarrier rings, named barriers, cross-warp shared-memory writes, atomic work queues).  Every reduction in them has
a fixed order, so two runs of the same sweep must agree bit for bit whatever CTA picks up which sample; a data race
would show up here as a difference.  (compute-sanitizer is not available on the GPU pool.)"""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _twice(fn):
    import torch
    U1, it1 = fn()
    U1 = U1.clone(); it1 = it1.clone()
    U2, it2 = fn()
    torch.cuda.synchronize()
    assert torch.equal(it1, it2)
    assert torch.equal(U1, U2)
    assert bool(torch.isfinite(U1).all())
    return it1


@pytest.mark.parametrize("n,proj", [(40, "LSPG"), (40, "Galerkin"), (96, "Galerkin"), (28, "LSPG")])
def test_prom_sweeps_are_bitwise_reproducible(n, proj, oracle):
    import burgers_b200
    X, _ = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X)
    Phi = np.ascontiguousarray(np.load(f"{GOLDEN}/Phi_tol_{'1e-03' if n <= 40 else '1e-04'}.npy")[:, :n])
    mu1, mu2 = oracle.parameter_grid(32, 31)     # 992 samples: more than two per resident CTA (at most 3 x 148), ragged tail
    it = _twice(lambda: fem.pod_prom_burgers_batch(0.05, 8, np.ones(512), mu1, 0.0, mu2, Phi, proj))
    assert int(it.min()) >= 1


@pytest.fixture
def qm_schedule():
    """Force a tangent-row schedule of csrc/bx_qm.cu (v2 / shared tile / chunked global) for one test."""
    import burgers_b200
    yield lambda mode: burgers_b200.set_tuning("qm_schedule", mode)
    burgers_b200.set_tuning("qm_schedule", -1)


@pytest.mark.parametrize("mode", [-1, 0, 1, 2])
def test_qm_sweeps_are_bitwise_reproducible(mode, oracle, qm_schedule):
    import burgers_b200
    qm_schedule(mode)
    X, _ = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X)
    Phi = np.load(f"{GOLDEN}/qm_Phi.npy"); H = np.load(f"{GOLDEN}/qm_H.npy")
    mu1, mu2 = oracle.parameter_grid(25, 13)     # 325 samples, last group of eight is ragged
    _twice(lambda: fem.pod_quadratic_manifold_batch(0.05, 8, np.ones(512), mu1, 0.0, mu2, Phi, H, "LSPG"))


@pytest.mark.parametrize("W", [2, 4])
def test_multiwarp_fom_sweeps_are_bitwise_reproducible(W, oracle):
    """The 2- and 4-warps-per-trajectory FOM instances synchronise through named barriers and shared memory."""
    import burgers_b200
    X, _ = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X)
    mu1, mu2 = oracle.parameter_grid(37, 41)     # 1517 samples: several per resident group
    burgers_b200.set_tuning("fom_warps_per_trajectory", W)
    try:
        _twice(lambda: fem.fom_burgers_batch(0.05, 12, np.ones(512), mu1, 0.0, mu2))
    finally:
        burgers_b200.set_tuning("fom_warps_per_trajectory", -1)


def test_qm_modes_agree_with_each_other(oracle, qm_schedule):
    """The three tangent-row paths of the QM kernel are different schedules of the same arithmetic up to the
    summation order of u; they must agree to rounding and give the same iteration counts."""
    import torch
    import burgers_b200
    X, _ = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X)
    Phi = np.load(f"{GOLDEN}/qm_Phi.npy"); H = np.load(f"{GOLDEN}/qm_H.npy")
    mu1, mu2 = oracle.parameter_grid(5, 5)
    res = {}
    for mode in (0, 1, 2):
        qm_schedule(mode)
        U, it = fem.pod_quadratic_manifold_batch(0.05, 10, np.ones(512), mu1, 0.0, mu2, Phi, H, "Galerkin")
        res[mode] = (U.clone(), it.clone())
    for mode in (1, 2):
        assert torch.equal(res[mode][1], res[0][1])
        rel = ((res[mode][0] - res[0][0]).norm(dim=2) / res[0][0].norm(dim=2)).max()
        assert float(rel) <= 1e-10
