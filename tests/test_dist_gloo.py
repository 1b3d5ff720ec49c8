"""world_size-2 gloo runs (CPU) of the host-side multi-rank logic: contiguous sample sharding of the sweep
grid (no data-path collective) and the path's single exchange step, the all-reduce of the snapshot Gram
followed by the replicated eigensolve.  The local Gram here comes from the oracle (test infrastructure);
on a GPU box it comes from bx_gram_accumulate."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import burgers_b200
    from burgers_b200 import pod, sweep
    from oracle import burgers_oracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    X, _ = O.mesh(64)
    mu1, mu2 = sweep.parameter_grid(3, 3)
    lo, hi = sweep.shard_range(mu1.size, rank, world)
    U, it = O.fom_burgers_batch(X, 0.05, 12, np.ones(64), mu1[lo:hi], 0.0, mu2[lo:hi])
    G = torch.from_numpy(O.snapshot_gram(U.reshape(-1, 64).T).copy())
    pod.allreduce_gram(G)
    Phi, s = pod.basis_from_gram(G, 6)
    tot = torch.tensor([float(it.sum())], dtype=torch.float64)
    dist.all_reduce(tot)
    np.savez(out, lo=lo, hi=hi, G=G.numpy(), Phi=Phi.numpy(), s=s.numpy(), tot=tot.numpy())
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_sharded_sweep_and_gram_allreduce(tmp_path, oracle):
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    outs = [str(tmp_path / f"r{r}.npz") for r in range(world)]
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_worker, args=(r, world, port, outs[r])) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
        assert p.exitcode == 0
    r0, r1 = np.load(outs[0]), np.load(outs[1])
    assert (int(r0["lo"]), int(r0["hi"]), int(r1["lo"]), int(r1["hi"])) == (0, 5, 5, 9)
    # both ranks hold the same reduced Gram and derive the same basis
    assert np.array_equal(r0["G"], r1["G"]) and np.array_equal(r0["Phi"], r1["Phi"])
    X, _ = oracle.mesh(64)
    mu1, mu2 = oracle.parameter_grid(3, 3)
    U, it = oracle.fom_burgers_batch(X, 0.05, 12, np.ones(64), mu1, 0.0, mu2)
    S = U.reshape(-1, 64).T
    assert np.allclose(r0["G"], S @ S.T, rtol=1e-13, atol=0)
    Us, ss = oracle.pod_basis_svd(S)
    assert np.allclose(r0["s"], ss[:6], rtol=1e-10)
    assert float(r0["tot"][0]) == float(it.sum())


def test_shard_range_covers_everything():
    from burgers_b200 import sweep
    for n in (0, 1, 7, 4096, 4099):
        for w in (1, 2, 3, 8):
            edges = [sweep.shard_range(n, r, w) for r in range(w)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[i][1] == edges[i + 1][0] for i in range(w - 1))
            sizes = [b - a for a, b in edges]
            assert max(sizes) - min(sizes) <= 1


def _qm_worker(rank, world, port, out):
    """Quadratic-manifold builder across ranks: each rank reduces its own snapshot columns to (R, W) — here with
    numpy (test infrastructure; on a GPU box burgers_b200.quadratic.reduce_snapshots does it on the device) — and the
    product's exchange step (all-gather + ordered merge) and compute_H_from_factors run on the gloo group."""
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import burgers_b200  # noqa: F401
    from burgers_b200 import quadratic, sweep
    from oracle import burgers_oracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(7)
    N, n, Ns = 24, 4, 90
    S = rng.standard_normal((N, Ns)); Phi = np.linalg.qr(rng.standard_normal((N, n)))[0]
    lo, hi = sweep.shard_range(Ns, rank, world)
    Sl = S[:, lo:hi]
    q = Phi.T @ Sl
    Q1, R = np.linalg.qr(O.build_Q(q).T)
    W = Q1.T @ O.build_E(Sl, Phi, q).T
    Rm, Wm = quadratic.merge_across_ranks(torch.from_numpy(R), torch.from_numpy(W))
    H = quadratic.compute_H_from_factors(Rm, Wm, 1e-2)
    np.savez(out, H=H.numpy(), S=S, Phi=Phi)
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_quadratic_builder_exchange_step(tmp_path, oracle):
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    outs = [str(tmp_path / f"q{r}.npz") for r in range(world)]
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_qm_worker, args=(r, world, port, outs[r])) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
        assert p.exitcode == 0
    r0, r1 = np.load(outs[0]), np.load(outs[1])
    assert np.array_equal(r0["H"], r1["H"])          # identical on every rank, no broadcast
    S, Phi = r0["S"], r0["Phi"]
    q = Phi.T @ S
    H = oracle.compute_H(oracle.build_Q(q), oracle.build_E(S, Phi, q), 1e-2)
    assert np.linalg.norm(r0["H"] - H) <= 1e-11 * np.linalg.norm(H)
