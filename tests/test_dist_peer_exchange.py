"""world_size-2 gloo runs (CPU) of the set-up of the partitioned fine-mesh trajectory (finemesh_dist.PeerExchange.create):
the exchange of the 64-byte IPC handles and — what matters for a launcher — that a rank whose local step fails does not
leave the other rank inside a collective: both raise.  The C library is replaced by a stub (no GPU here); peer memory
itself is exercised by tests/test_zy_finemesh_ranks.py on the GPU box."""
import os
import socket
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


class _StubLib:
    """bx_peer_* of the C ABI with the arguments ctypes passes (byref objects, string buffers)."""

    def __init__(self, rank, fail_alloc_on, fail_open_on):
        self.rank, self.fail_alloc_on, self.fail_open_on = rank, fail_alloc_on, fail_open_on
        self.calls = []

    def bx_peer_alloc(self, buf_ref, handle):
        self.calls.append("alloc")
        if self.rank == self.fail_alloc_on:
            return -2
        buf_ref._obj.value = 1000 + self.rank
        handle.raw = bytes([self.rank + 1]) * 64
        return 0

    def bx_peer_open(self, handle, ptr_ref):
        self.calls.append("open")
        if self.rank == self.fail_open_on:
            return -2
        ptr_ref._obj.value = 2000 + handle.raw[0] - 1
        return 0

    def bx_peer_close(self, ptr):
        self.calls.append(("close", ptr)); return 0

    def bx_peer_free(self, ptr):
        self.calls.append(("free", ptr)); return 0

    def bx_last_error(self):
        return b"stub"


def _worker(rank, world, port, out, fail_alloc_on, fail_open_on):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import burgers_b200
    from burgers_b200 import _lib, finemesh_dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    stub = _StubLib(rank, fail_alloc_on, fail_open_on)
    _lib.load = lambda: stub
    try:
        ex = finemesh_dist.PeerExchange.create()
        res = f"ok bufs={ex.bufs} own={ex.own} opened={ex._opened} part={finemesh_dist.node_partition(100003, world)[rank]}"
        ex.close()
        res += f" closed={[c for c in stub.calls if isinstance(c, tuple)]}"
    except burgers_b200.BurgersB200Error as e:
        res = f"error {e} cleanup={[c for c in stub.calls if isinstance(c, tuple)]}"
    with open(out, "w") as f:
        f.write(res)
    dist.barrier()
    dist.destroy_process_group()


def _run(tmp_path, fail_alloc_on=-1, fail_open_on=-1):
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    outs = [str(tmp_path / f"r{r}.txt") for r in range(world)]
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_worker, args=(r, world, port, outs[r], fail_alloc_on, fail_open_on)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0, "a rank hung or crashed"
    return [open(o).read() for o in outs]


@pytest.mark.timeout(300)
def test_handles_are_exchanged_and_mapped_in_rank_order(tmp_path):
    r0, r1 = _run(tmp_path)
    assert r0.startswith("ok bufs=[1000, 2001] own=1000 opened=[2001] part=(0, 57344)"), r0
    assert r1.startswith("ok bufs=[2000, 1001] own=1001 opened=[2000] part=(57344, 42659)"), r1
    assert "closed=[('close', 2001), ('free', 1000)]" in r0 and "closed=[('close', 2000), ('free', 1001)]" in r1


@pytest.mark.timeout(300)
def test_failed_allocation_on_one_rank_raises_on_both(tmp_path):
    r0, r1 = _run(tmp_path, fail_alloc_on=1)
    assert r0.startswith("error") and "bx_peer_alloc failed on rank(s) [1]" in r0, r0
    assert r1.startswith("error") and "bx_peer_alloc failed on rank(s) [1]" in r1, r1


@pytest.mark.timeout(300)
def test_failed_mapping_on_one_rank_raises_on_both_and_releases_the_buffers(tmp_path):
    r0, r1 = _run(tmp_path, fail_open_on=0)
    assert r0.startswith("error") and "bx_peer_open" in r0 and "rank(s) [0]" in r0, r0
    assert r1.startswith("error") and "rank(s) [0]" in r1, r1
    assert "('free', 1000)" in r0 and "('close', 2000)" in r1 and "('free', 1001)" in r1
