"""POD basis from device-resident snapshots — the batched/multi-GPU replacement of POD/pod.py.

The reference stacks every `fem_simulation_*.npy` into S (N x n_traj*501) and takes
`np.linalg.svd(S, full_matrices=False)` (POD/pod.py:80-84), then truncates by the energy criterion
of `get_number_of_singular_values_for_given_tolerance` (:8-14) and saves `U_modes_tol_*.npy` (:68-78).

Here the snapshots never leave the GPUs that produced them:
  1. each rank forms G_r = S_r S_r^T (N x N) from its own trajectory buffer with the DMMA kernel
     `bx_gram_accumulate` (csrc/bx_gram.cu);
  2. the ONE exchange step of the whole path: all-reduce(sum) of the N*N doubles (2 MiB at N = 512)
     over NCCL / NVLink;
  3. every rank runs the same symmetric eigensolve (cuSOLVER syevd through torch.linalg.eigh — a
     library call on a 512 x 512 matrix, not a hot kernel): G = V diag(lam) V^T, so the left singular
     vectors are V (descending) and the singular values sqrt(lam).  Identical input on every rank
     => identical basis, no broadcast.
sigma_28 / sigma_1 = 1.1e-2 for this problem, so squaring the condition number costs
nothing for the mode counts the reference uses.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from . import _lib


def get_number_of_singular_values_for_given_tolerance(M, N, s, epsilon_squared):
    """Same signature and result as POD/pod.py:8-14 (M, N unused there as well)."""
    s_sorted = np.sort(np.asarray(s, dtype=np.float64))[::-1]
    squared_cumsum = np.cumsum(s_sorted ** 2)
    squared_relative_loss = 1.0 - (squared_cumsum / squared_cumsum[-1])
    return int(np.argmax(squared_relative_loss <= epsilon_squared) + 1)


def get_list_of_number_of_singular_values_for_list_of_tolerances(tolerances_squared, s, M, N):
    """POD/pod.py:16-17."""
    return [get_number_of_singular_values_for_given_tolerance(M, N, s, e) for e in tolerances_squared]


class GramAccumulator:
    """Per-GPU accumulator of G = sum over trajectory buffers of S S^T (one N x N FP64 matrix)."""

    def __init__(self, n_nodes: int, device=None):
        if not torch.cuda.is_available():
            raise RuntimeError("burgers_b200.pod needs a CUDA device (sm_100a); there is no CPU fallback")
        self.N = int(n_nodes)
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.G = torch.zeros((self.N, self.N), dtype=torch.float64, device=self.device)
        self.cols = 0
        self._ws = None

    def add(self, U: torch.Tensor):
        """U: CUDA float64 tensor whose last axis is the node axis, e.g. the (B, nsteps+1, N) trajectory
        buffer of a sweep; every row is one snapshot.  Asynchronous on the current stream."""
        if not (isinstance(U, torch.Tensor) and U.is_cuda and U.dtype == torch.float64):
            raise ValueError("snapshots must be a CUDA float64 tensor")
        if U.shape[-1] != self.N or not U.is_contiguous():
            raise ValueError("snapshots must be contiguous with the node axis last")
        cols = U.numel() // self.N
        if cols == 0:
            return self
        lib = _lib.load()
        need = int(lib.bx_gram_workspace_bytes(self.N, cols))
        if need < 0:
            raise _lib.BurgersB200Error("bx_gram_workspace_bytes failed: " + lib.bx_last_error().decode())
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        _lib.check(lib.bx_gram_accumulate(self.N, cols, U.data_ptr(), self.G.data_ptr(), 1,
                                          self._ws.data_ptr(), torch.cuda.current_stream().cuda_stream),
                   "bx_gram_accumulate")
        self.cols += cols
        return self


def snapshot_gram(U: torch.Tensor) -> torch.Tensor:
    """G = S S^T of one snapshot buffer (node axis last), (N, N) CUDA tensor."""
    return GramAccumulator(U.shape[-1], U.device).add(U).G


def allreduce_gram(G: torch.Tensor, group=None) -> torch.Tensor:
    """Sum the per-rank Gram matrices in place (the path's only collective).  No-op without an
    initialised process group (single GPU)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(G, op=dist.ReduceOp.SUM, group=group)
    return G


def basis_from_gram(G: torch.Tensor, n_modes: int | None = None):
    """eigh(G) -> (Phi, s): columns of Phi = left singular vectors of S in descending order,
    s = singular values (POD/pod.py:84 returns the same pair from the SVD).  Runs where G lives."""
    lam, V = torch.linalg.eigh(G)
    lam = torch.flip(lam, dims=(0,))
    V = torch.flip(V, dims=(1,))
    s = torch.sqrt(torch.clamp(lam, min=0.0))
    if n_modes is not None:
        V, s = V[:, :n_modes], s[:n_modes]
    return V.contiguous(), s


def pod_basis(snapshot_buffers, n_modes: int | None = None, group=None):
    """Local snapshot buffers (one tensor or a list) -> global POD basis (Phi (N, r), s (r,)) CUDA
    tensors; identical on every rank."""
    if isinstance(snapshot_buffers, torch.Tensor):
        snapshot_buffers = [snapshot_buffers]
    acc = GramAccumulator(snapshot_buffers[0].shape[-1], snapshot_buffers[0].device)
    for U in snapshot_buffers:
        acc.add(U)
    allreduce_gram(acc.G, group)
    return basis_from_gram(acc.G, n_modes)


def load_snapshot_directory(directory, device=None, prefix="fem_simulation_"):
    """POD/pod.py:80-82: every `fem_simulation_*.npy` (each (N, steps+1)) of a directory, as ONE CUDA
    snapshot buffer (n_files*(steps+1), N) with the node axis last — the layout GramAccumulator.add and
    pod_basis take.  Column order across files does not change S S^T."""
    files = sorted(f for f in os.listdir(directory) if f.endswith(".npy") and f.startswith(prefix))
    if not files:
        raise FileNotFoundError(f"no {prefix}*.npy under {directory}")
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    blocks = [torch.from_numpy(np.ascontiguousarray(np.load(os.path.join(directory, f)).T)).to(dev) for f in files]
    return torch.cat(blocks, dim=0).contiguous(), files


def save_modes(U, s, tolerances_squared, snapshots_shape=None, directory="modes"):
    """POD/pod.py:68-78: one `U_modes_tol_{eps2:.0e}.npy` + `Singular_values_modes_tol_*.npy` per tolerance."""
    U = U.cpu().numpy() if isinstance(U, torch.Tensor) else np.asarray(U)
    s = s.cpu().numpy() if isinstance(s, torch.Tensor) else np.asarray(s)
    M, N = snapshots_shape if snapshots_shape is not None else (U.shape[0], s.size)
    taken = get_list_of_number_of_singular_values_for_list_of_tolerances(tolerances_squared, s, M, N)
    os.makedirs(directory, exist_ok=True)
    out = []
    for eps2, k in zip(tolerances_squared, taken):
        np.save(os.path.join(directory, f"U_modes_tol_{eps2:.0e}.npy"), U[:, :k])
        np.save(os.path.join(directory, f"Singular_values_modes_tol_{eps2:.0e}.npy"), s[:k])
        out.append(k)
    return out
