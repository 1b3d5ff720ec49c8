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
