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
        if U.device != self.device:
            raise ValueError(f"snapshots live on {U.device}, the accumulator on {self.device}")
        with torch.cuda.device(self.device):   # the library launches on the current device
            if self._ws is None or self._ws.numel() < need:
                self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
            _lib.check(lib.bx_gram_accumulate(self.N, cols, U.data_ptr(), self.G.data_ptr(), 1,
                                              self._ws.data_ptr(), torch.cuda.current_stream(self.device).cuda_stream),
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


GRAM_SAFETY = 64.0   # a mode is trusted while lam_k > GRAM_SAFETY * N * eps * lam_1


def gram_resolved_modes(lam_desc: torch.Tensor) -> int:
    """How many leading modes the Gram route determines: forming S S^T squares the condition number, so
    eigenvalues below ~N eps lam_1 (sigma_k / sigma_1 below ~1e-7 at N = 512) are rounding noise."""
    N = lam_desc.numel()
    floor = GRAM_SAFETY * N * torch.finfo(torch.float64).eps * float(lam_desc[0])
    return int((lam_desc > floor).sum().item())


def basis_from_gram(G: torch.Tensor, n_modes: int | None = None, strict: bool = False):
    """eigh(G) -> (Phi, s): columns of Phi = left singular vectors of S in descending order,
    s = singular values (POD/pod.py:84 returns the same pair from the SVD).  Runs where G lives.
    Limit of the Gram route: modes with sigma_k / sigma_1 below ~1e-7 are not resolved (enough for every
    truncation the reference ships, eps^2 >= 1e-6 -> <= 227 modes, sigma_227 / sigma_1 = 1.9e-4; not for the full
    498-column `U_modes_tol_0e+00`).  Asking for more modes than are resolved warns (raises with strict=True);
    use `pod_basis_svd` for deep truncations."""
    lam, V = torch.linalg.eigh(G)
    lam = torch.flip(lam, dims=(0,))
    V = torch.flip(V, dims=(1,))
    s = torch.sqrt(torch.clamp(lam, min=0.0))
    if n_modes is not None:
        ok = gram_resolved_modes(lam)
        if n_modes > ok:
            msg = (f"the Gram route resolves {ok} modes of this snapshot set (lam_k > {GRAM_SAFETY:g} N eps lam_1); "
                   f"{n_modes} were requested — modes beyond {ok} are rounding noise, use pod_basis_svd")
            if strict:
                raise ValueError(msg)
            import warnings
            warnings.warn(msg, RuntimeWarning, stacklevel=2)
        V, s = V[:, :n_modes], s[:n_modes]
    return V.contiguous(), s


def pod_basis_svd(snapshot_buffers, n_modes: int | None = None):
    """Deep truncations (sigma_k / sigma_1 < 1e-7): tall-skinny QR of the snapshot rows + SVD of the small R factor
    (cuSOLVER through torch.linalg — an offline library route, the same one burgers_b200.quadratic uses); no
    squared condition number.  Single GPU: all buffers must live on one device."""
    if isinstance(snapshot_buffers, torch.Tensor):
        snapshot_buffers = [snapshot_buffers]
    N = snapshot_buffers[0].shape[-1]
    R = None
    for U in snapshot_buffers:
        rows = U.reshape(-1, N)
        for lo in range(0, rows.shape[0], 1 << 17):
            blk = rows[lo:lo + (1 << 17)]
            stack = blk if R is None else torch.cat([R, blk], dim=0)
            R = torch.linalg.qr(stack, mode="r").R
    _, s, Vh = torch.linalg.svd(R, full_matrices=False)
    V = Vh.transpose(0, 1).contiguous()
    if n_modes is not None:
        V, s = V[:, :n_modes].contiguous(), s[:n_modes]
    return V, s


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


def save_modes(U, s, tolerances_squared, snapshots=None, directory="modes"):
    """POD/pod.py:68-78: one `U_modes_tol_{eps2:.0e}.npy` + `Singular_values_modes_tol_*.npy` per tolerance.
    `snapshots`: the reference passes the snapshot MATRIX as the 4th positional argument
    (`save_modes(U, s, eps2_tols, all_snapshots)`, POD/pod.py:90) and only reads its shape; an array, a tensor or
    a plain (M, N) tuple are all accepted."""
    U = U.cpu().numpy() if isinstance(U, torch.Tensor) else np.asarray(U)
    s = s.cpu().numpy() if isinstance(s, torch.Tensor) else np.asarray(s)
    shape = getattr(snapshots, "shape", snapshots)
    M, N = tuple(shape)[:2] if shape is not None else (U.shape[0], s.size)
    taken = get_list_of_number_of_singular_values_for_list_of_tolerances(tolerances_squared, s, M, N)
    os.makedirs(directory, exist_ok=True)
    out = []
    for eps2, k in zip(tolerances_squared, taken):
        np.save(os.path.join(directory, f"U_modes_tol_{eps2:.0e}.npy"), U[:, :k])
        np.save(os.path.join(directory, f"Singular_values_modes_tol_{eps2:.0e}.npy"), s[:k])
        out.append(k)
    return out
