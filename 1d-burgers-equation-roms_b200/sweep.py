"""Host-side sweep drivers: the batched replacement of the reference's
`for (mu1, mu2) in all_samples:` loop (FEM/paper_training_stage.py:8-10, 28-53).

  * parameter_grid      — the reference's mu1-major Cartesian grid (:8-10)
  * shard_range         — contiguous block of samples per rank (one process per GPU; samples are
                          independent, so the data path needs no collective)
  * fom_sweep_to_host   — streams a sweep whose parameters and results live in HOST memory:
                          H2D of the parameters, fused trajectory kernel, D2H of the snapshots
                          into pinned buffers, chunk by chunk, copy overlapped with compute
  * snapshot_filename   — the reference's on-disk name (:52)
  * to_reference_layout — device transpose (B, steps+1, N) -> (B, N, steps+1), the array the reference saves
  * sweep_to_npy        — the whole training/testing stage: sweep on the GPU, one
                          `fem_simulation_mu1_{:.3f}_mu2_{:.4f}.npy` of shape (N, steps+1) per sample (:52-53),
                          transposed on the device, copied out through pinned buffers, written by a small
                          thread pool while the next chunk integrates
  * prom_filename       — POD/Results_thesis/prom_pod.py:60
"""
from __future__ import annotations

import numpy as np
import torch


def parameter_grid(n1=64, n2=64, mu1_range=(4.25, 5.5), mu2_range=(0.015, 0.03)):
    """[[mu1, mu2] for mu1 in linspace(...) for mu2 in linspace(...)] — mu1-major."""
    mu1 = np.linspace(mu1_range[0], mu1_range[1], n1)
    mu2 = np.linspace(mu2_range[0], mu2_range[1], n2)
    return np.repeat(mu1, n2), np.tile(mu2, n1)


def shard_range(n_samples: int, rank: int, world: int):
    """Contiguous block [lo, hi) of rank `rank`; remainders go to the first ranks."""
    base, rem = divmod(int(n_samples), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def snapshot_filename(mu1: float, mu2: float) -> str:
    return f"fem_simulation_mu1_{mu1:.3f}_mu2_{mu2:.4f}.npy"


def prom_filename(tol: float, mu1: float, mu2: float, projection: str) -> str:
    """POD/Results_thesis/prom_pod.py:59-60."""
    tag = "lspg" if projection.lower() == "lspg" else "galerkin"
    return f"U_PROM_tol_{tol:.0e}_mu1_{mu1:.3f}_mu2_{mu2:.4f}_{tag}.npy"


def to_reference_layout(fem, U: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """(B, steps+1, N) trajectory buffer -> (B, N, steps+1): per sample the (N, nTimeSteps+1) C-order array
    FEMBurgers.fom_burgers returns and the drivers np.save (FEM/paper_training_stage.py:52-53)."""
    from . import _lib
    if not (U.is_cuda and U.dtype == torch.float64 and U.is_contiguous() and U.dim() == 3):
        raise ValueError("U must be a contiguous CUDA float64 tensor (B, steps+1, N)")
    B, rows, N = U.shape
    Ur = out if out is not None else torch.empty((B, N, rows), dtype=torch.float64, device=U.device)
    if tuple(Ur.shape) != (B, N, rows) or not Ur.is_contiguous() or Ur.dtype != torch.float64:
        raise ValueError("out must be a contiguous float64 tensor (B, N, steps+1)")
    lib = _lib.load()
    for lo in range(0, B, 65535):
        hi = min(B, lo + 65535)
        _lib.check(lib.bx_snapshots_to_reference_layout(hi - lo, rows, N, U[lo:hi].data_ptr(), Ur[lo:hi].data_ptr(),
                                                        torch.cuda.current_stream().cuda_stream),
                   "bx_snapshots_to_reference_layout")
    return Ur


def _write_npy(path: str, arr: np.ndarray):
    """np.save without the extra copy: version-1.0 header + the raw C-order bytes."""
    with open(path, "wb") as f:
        np.lib.format.write_array_header_1_0(f, {"descr": np.lib.format.dtype_to_descr(arr.dtype),
                                                 "fortran_order": False, "shape": arr.shape})
        f.write(memoryview(arr).cast("B"))


def sweep_to_npy(fem, directory, At, nTimeSteps, u0, mu1, mu2, E=0.0, chunk=256, method="fom_burgers",
                 name=snapshot_filename, writers=4, **kwargs):
    """Run `method` (a `*_batch` trajectory method of FEMBurgers, by name without the suffix) for every
    (mu1[b], mu2[b]) and write one .npy per sample in the reference's format and naming, so that POD/pod.py:80-82
    and the other offline stages consume GPU-generated data unchanged.
    Extra keyword arguments (Phi=..., projection=..., H=...) go to the trajectory method.
    Returns (paths, iters) with iters a (B, nTimeSteps) int32 numpy array."""
    import os
    from concurrent.futures import ThreadPoolExecutor
    os.makedirs(directory, exist_ok=True)
    mu1 = np.asarray(mu1, dtype=np.float64).reshape(-1)
    mu2 = np.asarray(mu2, dtype=np.float64).reshape(-1)
    B, N, nT = mu1.size, fem.n_nodes, int(nTimeSteps)
    if mu2.size != B:
        raise ValueError("mu1 and mu2 must have the same number of samples")
    run = getattr(fem, method + "_batch")
    chunk = max(1, min(int(chunk), max(B, 1)))
    dev = fem.device
    dU = torch.empty((chunk, nT + 1, N), dtype=torch.float64, device=dev)
    dIt = [torch.empty((chunk, nT), dtype=torch.int32, device=dev) for _ in range(2)]
    dUr = [torch.empty((chunk, N, nT + 1), dtype=torch.float64, device=dev) for _ in range(2)]
    hUr = [torch.empty((chunk, N, nT + 1), dtype=torch.float64).pin_memory() for _ in range(2)]
    hIt = [torch.empty((chunk, nT), dtype=torch.int32).pin_memory() for _ in range(2)]
    copy_stream = torch.cuda.Stream(device=dev)
    compute = torch.cuda.current_stream()
    iters = np.empty((B, nT), dtype=np.int32)
    paths = [os.path.join(directory, name(float(a), float(b))) for a, b in zip(mu1, mu2)]
    if len(set(paths)) != B:   # two samples that round to the same .3f / .4f name would race in the writer pool
        seen, dup = set(), None
        for q in paths:
            if q in seen:
                dup = q
                break
            seen.add(q)
        raise ValueError(f"two samples map to the same file name {os.path.basename(dup)!r}; pass a `name` function "
                         "with more digits")
    copied = [None, None]
    pending = [None, None]
    futures = [[], []]
    u0d = fem._dev(u0)

    with ThreadPoolExecutor(max_workers=max(1, int(writers))) as pool:
        def drain(i):
            if pending[i] is None:
                return
            copied[i].synchronize()
            lo, hi = pending[i]
            iters[lo:hi] = hIt[i][: hi - lo].numpy()
            view = hUr[i].numpy()
            futures[i] = [pool.submit(_write_npy, paths[lo + k], view[k]) for k in range(hi - lo)]
            pending[i] = None

        def wait_writes(i):
            for f in futures[i]:
                f.result()
            futures[i] = []

        for c, lo in enumerate(range(0, B, chunk)):
            hi = min(B, lo + chunk)
            n = hi - lo
            i = c & 1
            drain(1 - i)       # hand the previous chunk to the writers while this one integrates
            wait_writes(i)     # host buffer i is about to be overwritten
            a = torch.from_numpy(mu1[lo:hi]).to(dev, non_blocking=True)
            b = torch.from_numpy(mu2[lo:hi]).to(dev, non_blocking=True)
            if copied[i] is not None:
                compute.wait_event(copied[i])   # dUr[i] / dIt[i] still being copied out
            u0c = u0d if u0d.numel() == N else u0d.reshape(B, N)[lo:hi]
            run(At, nT, u0c, a, E, b, out=dU[:n], iters_out=dIt[i][:n], **kwargs)
            to_reference_layout(fem, dU[:n], out=dUr[i][:n])
            done = torch.cuda.Event()
            done.record(compute)
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(done)
                hUr[i][:n].copy_(dUr[i][:n], non_blocking=True)
                hIt[i][:n].copy_(dIt[i][:n], non_blocking=True)
                copied[i] = torch.cuda.Event()
                copied[i].record(copy_stream)
            pending[i] = (lo, hi)
        drain(0); drain(1)
        wait_writes(0); wait_writes(1)
    return paths, iters


class HostSweepBuffers:
    """Pinned host staging + device buffers for fom_sweep_to_host (allocate once, reuse)."""

    def __init__(self, fem, chunk: int, nTimeSteps: int):
        N = fem.n_nodes
        self.chunk, self.nT, self.N = int(chunk), int(nTimeSteps), N
        dev = fem.device
        self.dU = [torch.empty((chunk, nTimeSteps + 1, N), dtype=torch.float64, device=dev) for _ in range(2)]
        self.dIt = [torch.empty((chunk, nTimeSteps), dtype=torch.int32, device=dev) for _ in range(2)]
        self.dMu = [torch.empty((2, chunk), dtype=torch.float64, device=dev) for _ in range(2)]
        self.hU = [torch.empty((chunk, nTimeSteps + 1, N), dtype=torch.float64).pin_memory() for _ in range(2)]
        self.hIt = [torch.empty((chunk, nTimeSteps), dtype=torch.int32).pin_memory() for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.h2d_bytes = 0
        self.d2h_bytes = 0


def chunk_schedule(B, chunk, ramp=False):
    """[(lo, hi), ...] covering range(B) in pieces of at most `chunk` samples.  With ramp the first pieces are
    chunk/4 and chunk/2 long, so the device->host copy (the bound of a host-streaming sweep) starts after a
    quarter-size kernel instead of a full one.  Measured on B200 (tools/e2e_time.py): a 256-sample kernel takes
    as long as a 1024-sample one (a trajectory is a serial chain of ~6000 Picard iterations, ~14 ms), so the ramp
    only pays for chunks of 2048 and more; 512-sample chunks without ramp are the fastest (171 ms against the
    155 ms the 8.4 GB take at the box's 54 GB/s pinned D2H rate)."""
    out, lo = [], 0
    sizes = [max(1, chunk // 4), max(1, chunk // 2)] if ramp and B > chunk else []
    while lo < B:
        n = sizes.pop(0) if sizes else chunk
        hi = min(B, lo + n)
        out.append((lo, hi))
        lo = hi
    return out


def fom_sweep_to_host(fem, At, nTimeSteps, u0, mu1, mu2, E=0.0, chunk=1024, consumer=None,
                      buffers: HostSweepBuffers | None = None, tol=1e-6, kmax=20, ramp=False):
    """FOM sweep with HOST inputs and HOST outputs.

    mu1, mu2: pinned (or pageable) host float64 tensors / numpy arrays of shape (B,).
    consumer(lo, hi, U_host, iters_host) is called for each finished chunk with pinned-memory
    views U_host (hi-lo, nTimeSteps+1, N), iters_host (hi-lo, nTimeSteps); the views are reused two
    chunks later, so the consumer must copy or write out what it keeps (the reference writes
    one .npy per sample, FEM/paper_training_stage.py:52-53).
    Returns the total number of nonlinear iterations (int).
    """
    mu1 = torch.as_tensor(mu1, dtype=torch.float64)
    mu2 = torch.as_tensor(mu2, dtype=torch.float64)
    B = mu1.numel()
    buf = buffers or HostSweepBuffers(fem, min(chunk, max(B, 1)), nTimeSteps)
    chunk = buf.chunk
    u0d = fem._dev(u0)
    if u0d.numel() != fem.n_nodes:
        raise ValueError("fom_sweep_to_host takes one shared initial condition u0 (N,)")
    compute = torch.cuda.current_stream()
    done = [None, None]      # kernel finished on buffer i
    copied = [None, None]    # D2H finished on buffer i
    pending = [None, None]   # (lo, hi) waiting for the consumer
    total_iters = 0
    buf.h2d_bytes = 0
    buf.d2h_bytes = 0

    def drain(i):
        nonlocal total_iters
        if pending[i] is None:
            return
        copied[i].synchronize()
        lo, hi = pending[i]
        total_iters += int(buf.hIt[i][: hi - lo].sum())
        if consumer is not None:
            consumer(lo, hi, buf.hU[i][: hi - lo], buf.hIt[i][: hi - lo])
        pending[i] = None

    for c, (lo, hi) in enumerate(chunk_schedule(B, chunk, ramp)):
        n = hi - lo
        i = c & 1
        drain(i)  # host + device buffer i are free again
        buf.dMu[i][0, :n].copy_(mu1[lo:hi], non_blocking=True)
        buf.dMu[i][1, :n].copy_(mu2[lo:hi], non_blocking=True)
        buf.h2d_bytes += 2 * n * 8
        fem.fom_burgers_batch(At, nTimeSteps, u0d, buf.dMu[i][0, :n], E, buf.dMu[i][1, :n],
                              out=buf.dU[i][:n], iters_out=buf.dIt[i][:n], tol=tol, kmax=kmax)
        done[i] = torch.cuda.Event()
        done[i].record(compute)
        with torch.cuda.stream(buf.copy_stream):
            buf.copy_stream.wait_event(done[i])
            buf.hU[i][:n].copy_(buf.dU[i][:n], non_blocking=True)
            buf.hIt[i][:n].copy_(buf.dIt[i][:n], non_blocking=True)
            copied[i] = torch.cuda.Event()
            copied[i].record(buf.copy_stream)
        # device buffer i is next written by chunk c+2, after drain(i) has waited for this copy
        buf.d2h_bytes += n * ((nTimeSteps + 1) * fem.n_nodes * 8 + nTimeSteps * 4)
        pending[i] = (lo, hi)
    drain(0)
    drain(1)
    return total_iters
