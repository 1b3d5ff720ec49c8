"""bx_fom_picard_finemesh_ranks (the fine-mesh trajectory partitioned over R ranks, fourth partition level through
peer memory) on ONE device: the R rank kernels run side by side on R streams of this process and
exchange their interface rows through R exchange buffers in the same device memory — the same kernel, exchange
protocol and host calls as under torchrun with one process per GPU (tools/finemesh_ranks_check.py is that run).
Checked against the oracle (FEM/fem_burgers.py:646-707) and the single-kernel path (bx_fom_picard_finemesh).
Tolerance (north_star): per-time-step relative L2 <= 1e-10 in FP64, identical iteration counts.

Collected near the end: when the driver does not schedule the R cooperative kernels side by side they give up waiting
for each other after 10 s (nothing hangs) and the test is skipped, not failed."""
import numpy as np
import pytest

from conftest import rel_l2_per_step

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _run_ranks(X, At, nsteps, u0, mu1, E, mu2, R, launches=1):
    """R rank launches on R streams of cuda:0 (`launches` times on the same exchange buffers: the epochs carry over);
    returns the assembled (N, nsteps + 1) trajectory and the counts."""
    import torch
    from burgers_b200 import _lib
    from burgers_b200.finemesh_dist import PeerExchange, node_partition
    lib = _lib.load()
    dev = torch.device("cuda:0")
    N = X.size
    parts = node_partition(N, R)
    ex = PeerExchange.local(R)
    try:
        Xd = torch.from_numpy(X).to(dev)
        m1 = torch.tensor(np.atleast_1d(mu1), dtype=torch.float64, device=dev)
        m2 = torch.tensor(np.atleast_1d(mu2), dtype=torch.float64, device=dev)
        B = int(m1.numel())
        batched = 1 if np.ndim(u0) == 2 else 0
        streams = [torch.cuda.Stream(dev) for _ in range(R)]
        keep = []
        for r, (off, cnt) in enumerate(parts):
            u0d = torch.from_numpy(np.ascontiguousarray(u0[..., off:off + cnt])).to(dev)
            U = torch.full((B, nsteps + 1, cnt), float("nan"), dtype=torch.float64, device=dev)
            it = torch.full((B, nsteps), -7, dtype=torch.int32, device=dev)
            ws = torch.empty(int(lib.bx_fine_workspace_bytes(cnt)), dtype=torch.uint8, device=dev)
            keep.append((u0d, U, it, ws))
        torch.cuda.synchronize(dev)
        for r, (off, cnt) in [pr for _ in range(launches) for pr in enumerate(parts)]:
            u0d, U, it, ws = keep[r]
            _lib.check(lib.bx_fom_picard_finemesh_ranks(B, N, off, cnt, nsteps, Xd.data_ptr(), u0d.data_ptr(), batched,
                                                        m1.data_ptr(), m2.data_ptr(), float(At), float(E), 1e-6, 20,
                                                        U.data_ptr(), it.data_ptr(), ws.data_ptr(), R, r,
                                                        ex[r].pointer_array(), streams[r].cuda_stream),
                       "bx_fom_picard_finemesh_ranks")
        torch.cuda.synchronize(dev)
        if any(e.failed(streams[r].cuda_stream) for r, e in enumerate(ex)):
            pytest.skip("the rank kernels were not scheduled side by side on this device (gave up after 10 s)")
        Ufull = np.concatenate([k[1].cpu().numpy() for k in keep], axis=2)     # (B, nsteps + 1, N)
        its = [k[2].cpu().numpy() for k in keep]
        for i in its[1:]:
            assert np.array_equal(i, its[0])
        if B == 1 and np.ndim(mu1) == 0:
            return Ufull[0].T, its[0][0]
        return Ufull, its[0]
    finally:
        lib = _lib.load()
        for ptr in ex[0].bufs:
            lib.bx_peer_free(ptr)


@pytest.mark.parametrize("rows", [-1, 16, 8])
@pytest.mark.parametrize("N,R", [(100003, 1), (100003, 2), (100003, 3), (8192 * 5, 5), (131072 + 7, 8)])
def test_ranks_vs_oracle_and_single_kernel(N, R, rows, oracle):
    """rows: mesh rows per thread of the kernel instance (-1: automatic = 4 at these sizes)."""
    import burgers_b200
    burgers_b200.set_tuning("fine_rows", rows)
    try:
        _ranks_vs_oracle_and_single_kernel(N, R, oracle)
    finally:
        burgers_b200.set_tuning("fine_rows", -1)


def _ranks_vs_oracle_and_single_kernel(N, R, oracle):
    import burgers_b200
    X = np.linspace(0.0, 100.0, N)
    At = 0.05 * 511.0 / (N - 1)
    u0 = np.ones(N)
    U, it = _run_ranks(X, At, 12, u0, 4.75, 0.0, 0.02, R, launches=3 if R == 2 else 1)
    Uo, ito = oracle.fom_burgers(X, At, 12, u0, 4.75, 0.0, 0.02)
    assert np.array_equal(it, ito), (it, ito)
    assert rel_l2_per_step(U, Uo) <= TOL
    fem = burgers_b200.FEMBurgers(X)
    Us, its = fem.fom_burgers_batch(At, 12, u0, [4.75], 0.0, [0.02])
    assert np.array_equal(it, its[0].cpu().numpy())
    assert rel_l2_per_step(U, Us[0].cpu().numpy().T) <= TOL


def test_ranks_two_groups_of_ctas_nonuniform_mesh_diffusion(oracle):
    """70 + 69 CTAs on two ranks (more than 32 CTAs per rank: two-stage fold), non-uniform mesh, E != 0, ragged end."""
    rng = np.random.default_rng(5)
    N = 8192 * 139 - 4321
    X = np.cumsum(0.5 + rng.random(N)); X = 100.0 * (X - X[0]) / (X[-1] - X[0])
    At = 0.05 * 511.0 / (N - 1)
    u0 = 1.0 + 0.2 * np.sin(X / 7.0)
    import burgers_b200
    burgers_b200.set_tuning("fine_rows", 16)   # 8192 nodes per CTA: 70 + 69 CTAs fit side by side on one device
    try:
        U, it = _run_ranks(X, At, 3, u0, 4.3, 1e-4, 0.017, 2)
    finally:
        burgers_b200.set_tuning("fine_rows", -1)
    Uo, ito = oracle.fom_burgers(X, At, 3, u0, 4.3, 1e-4, 0.017)
    assert np.array_equal(it, ito), (it, ito)
    assert rel_l2_per_step(U, Uo) <= TOL


def test_ranks_two_samples_batched_initial_state(oracle):
    """B = 2 samples one after the other in ONE launch per rank (the epochs run on across the samples), (B, N_local) u0."""
    N, R = 65536 + 100, 3
    X = np.linspace(0.0, 100.0, N)
    At = 0.05 * 511.0 / (N - 1)
    mu1 = np.array([4.3, 5.4]); mu2 = np.array([0.017, 0.028])
    u0 = np.stack([np.ones(N), 1.0 + 0.2 * np.sin(X / 7.0)])
    U, it = _run_ranks(X, At, 8, u0, mu1, 0.0, mu2, R)
    for b in range(2):
        Uo, ito = oracle.fom_burgers(X, At, 8, u0[b], mu1[b], 0.0, mu2[b])
        assert np.array_equal(it[b], ito), (b, it[b], ito)
        assert rel_l2_per_step(U[b].T, Uo) <= TOL


def test_missing_rank_gives_up_instead_of_hanging():
    """Only rank 0 of 2 launches: its kernel waits 10 s for rank 1's interface row, raises the failure word and leaves."""
    import time
    import torch
    from burgers_b200 import _lib
    from burgers_b200.finemesh_dist import PeerExchange, node_partition
    lib = _lib.load()
    dev = torch.device("cuda:0")
    N = 40000
    X = torch.linspace(0.0, 100.0, N, dtype=torch.float64, device=dev)
    off, cnt = node_partition(N, 2)[0]
    ex = PeerExchange.local(2)
    try:
        u0 = torch.ones(cnt, dtype=torch.float64, device=dev)
        m1 = torch.tensor([4.75], dtype=torch.float64, device=dev); m2 = torch.tensor([0.02], dtype=torch.float64, device=dev)
        U = torch.zeros((1, 5, cnt), dtype=torch.float64, device=dev)
        ws = torch.empty(int(lib.bx_fine_workspace_bytes(cnt)), dtype=torch.uint8, device=dev)
        st = torch.cuda.Stream(dev)
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        _lib.check(lib.bx_fom_picard_finemesh_ranks(1, N, off, cnt, 4, X.data_ptr(), u0.data_ptr(), 0, m1.data_ptr(), m2.data_ptr(),
                                                    0.05 * 511.0 / (N - 1), 0.0, 1e-6, 20, U.data_ptr(), None, ws.data_ptr(), 2, 0,
                                                    ex[0].pointer_array(), st.cuda_stream), "bx_fom_picard_finemesh_ranks")
        assert ex[0].failed(st.cuda_stream)          # synchronises the stream
        dt = time.perf_counter() - t0
        assert 8.0 < dt < 40.0, dt
        assert ex[1].failed(st.cuda_stream)          # the failure word was raised in every rank's buffer
    finally:
        for ptr in ex[0].bufs:
            lib.bx_peer_free(ptr)


def test_ranks_argument_checks():
    from burgers_b200 import _lib
    lib = _lib.load()
    # block offsets that are not multiples of 8192, more than 8 ranks, a middle block that is not whole CTA blocks
    assert lib.bx_fom_picard_finemesh_ranks(1, 100000, 100, 1000, 1, None, None, 0, None, None, 0.05, 0.0, 1e-6, 20,
                                            None, None, None, 2, 1, None, None) == _lib.BX_ERR_INVALID
    assert lib.bx_fom_picard_finemesh_ranks(1, 100000, 0, 8192, 1, None, None, 0, None, None, 0.05, 0.0, 1e-6, 20,
                                            None, None, None, 9, 0, None, None) == _lib.BX_ERR_INVALID
    assert lib.bx_fom_picard_finemesh_ranks(1, 100000, 0, 8000, 1, None, None, 0, None, None, 0.05, 0.0, 1e-6, 20,
                                            None, None, None, 2, 0, None, None) == _lib.BX_ERR_INVALID
    assert b"8192" in lib.bx_last_error()
