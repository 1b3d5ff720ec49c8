"""FEMBurgers.fom_burgers (FEM/fem_burgers.py:646-707) on a fine mesh partitioned over the ranks of a
torch.distributed job, one process per GPU: host side of bx_fom_picard_finemesh_ranks.

The reference has no distributed code; what is mirrored is fom_burgers' arguments and outputs, block by block: rank r
holds the node block `node_partition(N, R)[r]` of u0 and gets that block of every snapshot.  torch.distributed is
used at SET-UP only (the 64-byte IPC handles of the exchange buffers travel through all_gather_object, one barrier
before a launch so that no rank waits for a process that is still starting); inside a launch the kernels talk to
each other through peer memory."""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import _lib

BLOCK_NODES = 8192   # nodes per CTA of the fine-mesh kernel (csrc/bx_fine.cu: FM * FT)
MAX_RANKS = 8


def node_partition(N: int, R: int):
    """[(offset, count)] of the R contiguous node blocks: whole 8192-node CTA blocks, spread as evenly as possible,
    the last rank takes the ragged end.  Raises when a rank would stay empty."""
    N, R = int(N), int(R)
    if R < 1 or R > MAX_RANKS:
        raise ValueError(f"1 <= R <= {MAX_RANKS} ranks")
    ctas = -(-N // BLOCK_NODES)
    if ctas < R:
        raise ValueError(f"{N} nodes are {ctas} blocks of {BLOCK_NODES}: fewer than {R} ranks")
    base, extra = divmod(ctas, R)
    out, off = [], 0
    for r in range(R):
        g = base + (1 if r < extra else 0)
        cnt = min(g * BLOCK_NODES, N - off)
        out.append((off, cnt))
        off += cnt
    assert off == N
    return out


class PeerExchange:
    """The exchange buffers of a group of ranks.  `PeerExchange.create(group)` allocates this rank's buffer on the current
    device, all-gathers the IPC handles and maps the others; `PeerExchange.local(R, device)` builds R buffers in ONE
    process (R kernels on R streams of one device: tests)."""

    def __init__(self, rank, bufs, own, opened):
        self.rank, self.bufs, self.own, self._opened = rank, list(bufs), own, opened
        self.R = len(self.bufs)
        self.last_kernel_ms = None

    @classmethod
    def create(cls, group=None):
        """Collective over `group`: every rank must call it.  A rank whose allocation or mapping fails does not leave the
        others waiting: the outcome of every local step is exchanged before the next collective step, and all ranks raise."""
        import torch.distributed as dist
        lib = _lib.load()
        R, rank = dist.get_world_size(group), dist.get_rank(group)
        if R > MAX_RANKS:
            raise ValueError(f"at most {MAX_RANKS} ranks")

        def all_ok(ok, what):
            flags = [None] * R
            dist.all_gather_object(flags, bool(ok), group=group)
            if not all(flags):
                raise _lib.BurgersB200Error(f"PeerExchange.create: {what} failed on rank(s) {[r for r, f in enumerate(flags) if not f]}")

        own = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        all_ok(lib.bx_peer_alloc(ctypes.byref(own), handle) == _lib.BX_OK, "bx_peer_alloc")
        handles = [None] * R
        dist.all_gather_object(handles, bytes(handle.raw), group=group)
        bufs, opened, ok = [], [], True
        for r in range(R):
            if r == rank:
                bufs.append(own.value)
                continue
            ptr = ctypes.c_void_p()
            if lib.bx_peer_open(ctypes.create_string_buffer(handles[r], 64), ctypes.byref(ptr)) != _lib.BX_OK:
                ok = False
                break
            bufs.append(ptr.value); opened.append(ptr.value)
        try:
            all_ok(ok, "bx_peer_open (cudaIpcOpenMemHandle)")
        except _lib.BurgersB200Error:
            for ptr in opened:
                lib.bx_peer_close(ptr)
            lib.bx_peer_free(own.value)
            raise
        return cls(rank, bufs, own.value, opened)

    @classmethod
    def local(cls, R):
        """R exchange objects (one per rank) whose buffers all live on the current device of THIS process."""
        lib = _lib.load()
        bufs = []
        for _ in range(R):
            own = ctypes.c_void_p()
            handle = ctypes.create_string_buffer(64)
            _lib.check(lib.bx_peer_alloc(ctypes.byref(own), handle), "bx_peer_alloc")
            bufs.append(own.value)
        return [cls(r, bufs, bufs[r], []) for r in range(R)]

    def pointer_array(self):
        return (ctypes.c_void_p * self.R)(*self.bufs)

    def failed(self, stream=None):
        return bool(_lib.load().bx_fine_peer_failed(self.own, stream))

    def close(self):
        lib = _lib.load()
        for ptr in self._opened:
            lib.bx_peer_close(ptr)
        self._opened = []
        if self.own is not None:
            lib.bx_peer_free(self.own)
            self.own = None


def fom_burgers_partitioned(X, At, nTimeSteps, u0_local, mu1, E, mu2, exchange: PeerExchange, group=None,
                            tol=1e-6, kmax=20, out=None):
    """One trajectory (or B, when mu1 / mu2 are arrays) of fom_burgers on the mesh X, this rank integrating its block.
    X: the whole mesh (numpy or device tensor).  u0_local: this rank's block of the initial state.
    Returns (U_local (B, nTimeSteps + 1, N_local), iters (B, nTimeSteps)) as device tensors; iters is identical on
    every rank.  Raises BurgersB200Error when a rank gave up waiting for another one."""
    import torch.distributed as dist
    lib = _lib.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    Xd = X if torch.is_tensor(X) else torch.from_numpy(np.ascontiguousarray(X, dtype=np.float64))
    Xd = Xd.to(dev, torch.float64).contiguous()
    N = int(Xd.numel())
    R, rank = exchange.R, exchange.rank
    off, Nl = node_partition(N, R)[rank]
    m1 = torch.as_tensor(np.atleast_1d(np.asarray(mu1, dtype=np.float64))).to(dev)
    m2 = torch.as_tensor(np.atleast_1d(np.asarray(mu2, dtype=np.float64))).to(dev)
    B = int(m1.numel())
    u0d = torch.as_tensor(u0_local, dtype=torch.float64).to(dev).contiguous()
    batched = 1 if u0d.dim() == 2 else 0
    if u0d.shape[-1] != Nl:
        raise ValueError(f"u0_local holds {u0d.shape[-1]} nodes, this rank's block holds {Nl}")
    nT = int(nTimeSteps)
    U = out if out is not None else torch.empty((B, nT + 1, Nl), dtype=torch.float64, device=dev)
    iters = torch.empty((B, nT), dtype=torch.int32, device=dev)
    ws = torch.empty(int(lib.bx_fine_workspace_bytes(Nl)), dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev)
    torch.cuda.synchronize(dev)
    if dist.is_initialized():
        dist.barrier(group)   # every rank is about to launch: the in-kernel waits are short
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    _lib.check(lib.bx_fom_picard_finemesh_ranks(B, N, off, Nl, nT, Xd.data_ptr(), u0d.data_ptr(), batched, m1.data_ptr(),
                                                m2.data_ptr(), float(At), float(E), float(tol), int(kmax), U.data_ptr(),
                                                iters.data_ptr(), ws.data_ptr(), R, rank, exchange.pointer_array(),
                                                stream.cuda_stream),
               "bx_fom_picard_finemesh_ranks")
    e1.record(stream)
    if exchange.failed(stream.cuda_stream):
        raise _lib.BurgersB200Error("bx_fom_picard_finemesh_ranks: a rank gave up waiting for another one; outputs invalid")
    exchange.last_kernel_ms = e0.elapsed_time(e1)
    return U, iters
