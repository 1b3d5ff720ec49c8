"""Host-side mirror of the reference's `FEMBurgers` (FEM/fem_burgers.py:314-1574) for the hot
path: same constructor, method names, argument order/meaning, return types and error
behaviour; every number is produced by the sm_100a kernels behind include/burgers_b200.h.

Legacy single-sample methods take/return numpy (and scipy.sparse CSC for the matrices, as the
reference does) and synchronise before returning.  The `*_batch` / `*_bands` methods are the new
batched surface: torch CUDA tensors in and out, asynchronous on the current CUDA stream.

PyTorch is used for device memory, streams and (in sweep.py) torch.distributed only.
There is no CPU path: without a CUDA device or without the built library these raise.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib


def _require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("burgers_b200 needs a CUDA device (sm_100a); there is no CPU fallback")


def _ptr(t):
    return None if t is None else t.data_ptr()


def _on_device(method):
    """Run a FEMBurgers method with the object's device current: the C library enqueues on the stream it is
    given and queries the *current* device, so an object built for cuda:1 must not run while cuda:0 is current."""
    import functools

    @functools.wraps(method)
    def wrapper(self, *args, **kwargs):
        with torch.cuda.device(self.device):
            return method(self, *args, **kwargs)
    return wrapper


def validate_mesh(X, T):
    """The kernels assume the reference's structured connectivity T[e] = (e+1, e+2), 1-based
    (FEM/fom.py:16, element_nodes = T[elem,:] - 1 at FEM/fem_burgers.py:330)."""
    X = np.ascontiguousarray(np.asarray(X, dtype=np.float64))
    if X.ndim != 1 or X.size < 2:
        raise ValueError("X must be a 1-D array of at least 2 node coordinates")
    if not np.all(np.diff(X) > 0):
        raise ValueError("X must be strictly increasing")
    if T is not None:
        T = np.asarray(T)
        m = X.size - 1
        ok = T.shape == (m, 2) and np.array_equal(T[:, 0], np.arange(1, m + 1)) and \
            np.array_equal(T[:, 1], np.arange(2, m + 2))
        if not ok:
            raise ValueError("only the structured 2-node connectivity T[e] = (e+1, e+2) (1-based) is supported")
    return X


class FEMBurgers:
    WARP_MESH_MAX = 1024   # largest mesh the one-warp-per-trajectory kernel covers

    def __init__(self, X, T=None, device=None):
        self.X = validate_mesh(X, T)
        self.T = T if T is not None else np.array([np.arange(1, self.X.size), np.arange(2, self.X.size + 1)]).T
        # quadrature data kept as attributes for callers that read them (FEM/fem_burgers.py:318-322)
        self.ngaus = 2
        self.zgp = np.array([-np.sqrt(3) / 3, np.sqrt(3) / 3])
        self.wgp = np.array([1, 1])
        self.N = np.array([(1 - self.zgp) / 2, (1 + self.zgp) / 2]).T
        self.Nxi = np.array([[-1 / 2, 1 / 2], [-1 / 2, 1 / 2]])
        self.n_nodes = int(self.X.size)
        self._device = device
        self._Xd = None
        self._ws = {}       # per-stream work-queue workspaces: launches on different streams never share one
        self._big_ws = {}   # per-stream grow-only scratch
        self.last_iters = None  # per-step nonlinear iteration counts of the last legacy call
        self.verbose = False    # True: the legacy single-sample methods print the reference's progress lines

    # ------------------------------------------------------------------ device plumbing
    @property
    def device(self):
        if self._device is None:
            _require_cuda()
            self._device = torch.device("cuda", torch.cuda.current_device())
        return torch.device(self._device)

    @property
    def Xd(self):
        if self._Xd is None:
            _require_cuda()
            self._Xd = torch.from_numpy(self.X).to(self.device)
        return self._Xd

    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def _workspace(self):
        key = self._stream()
        ws = self._ws.get(key)
        if ws is None:
            ws = self._ws[key] = torch.zeros(int(_lib.load().bx_workspace_bytes()), dtype=torch.uint8,
                                             device=self.device)
        return ws

    def _scratch(self, nbytes):
        """Grow-only device scratch for the entry points whose workspace depends on the batch size; one per
        stream, because a workspace must not be shared by launches that may overlap (include/burgers_b200.h)."""
        key = self._stream()
        ws = self._big_ws.get(key)
        if ws is None or ws.numel() < nbytes:
            ws = self._big_ws[key] = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
        return ws

    def _dev(self, a, shape=None):
        """numpy / scalar / tensor -> contiguous float64 CUDA tensor."""
        if isinstance(a, torch.Tensor):
            t = a.to(device=self.device, dtype=torch.float64)
        else:
            t = torch.from_numpy(np.ascontiguousarray(np.asarray(a, dtype=np.float64))).to(self.device)
        if shape is not None:
            t = t.reshape(shape)
        return t.contiguous()

    def _new(self, *shape, dtype=torch.float64):
        return torch.empty(shape, dtype=dtype, device=self.device)

    # ------------------------------------------------------------------ batched operators
    @_on_device
    def mass_bands(self):
        """(dl, d, du) of M, each (N,) — compute_mass_matrix, FEM/fem_burgers.py:324."""
        N = self.n_nodes
        dl, d, du = self._new(N), self._new(N), self._new(N)
        lib = _lib.load()
        _lib.check(lib.bx_mass_bands(N, _ptr(self.Xd), _ptr(dl), _ptr(d), _ptr(du), self._stream()), "bx_mass_bands")
        return dl, d, du

    @_on_device
    def diffusion_bands(self):
        N = self.n_nodes
        dl, d, du = self._new(N), self._new(N), self._new(N)
        lib = _lib.load()
        _lib.check(lib.bx_diffusion_bands(N, _ptr(self.Xd), _ptr(dl), _ptr(d), _ptr(du), self._stream()),
                   "bx_diffusion_bands")
        return dl, d, du

    @_on_device
    def convection_bands(self, U):
        """U (B, N) -> (dl, d, du) each (B, N) — compute_convection_matrix, :389."""
        U = self._dev(U).reshape(-1, self.n_nodes)
        B, N = U.shape
        dl, d, du = self._new(B, N), self._new(B, N), self._new(B, N)
        lib = _lib.load()
        _lib.check(lib.bx_convection_bands(B, N, _ptr(self.Xd), _ptr(U), _ptr(dl), _ptr(d), _ptr(du), self._stream()),
                   "bx_convection_bands")
        return dl, d, du

    @_on_device
    def convection_derivative_bands(self):
        N = self.n_nodes
        dl, d, du = self._new(N), self._new(N), self._new(N)
        lib = _lib.load()
        _lib.check(lib.bx_convection_derivative_bands(N, _ptr(self.Xd), _ptr(dl), _ptr(d), _ptr(du), self._stream()),
                   "bx_convection_derivative_bands")
        return dl, d, du

    @_on_device
    def forcing_vector_batch(self, mu2):
        mu2 = self._dev(mu2).reshape(-1)
        B, N = mu2.numel(), self.n_nodes
        F = self._new(B, N)
        lib = _lib.load()
        _lib.check(lib.bx_forcing_vector(B, N, _ptr(self.Xd), _ptr(mu2), _ptr(F), self._stream()), "bx_forcing_vector")
        return F

    @_on_device
    def supg_term_batch(self, U, mu2):
        U = self._dev(U).reshape(-1, self.n_nodes)
        mu2 = self._dev(mu2).reshape(-1)
        B, N = U.shape
        if mu2.numel() != B:
            raise ValueError("mu2 must have one entry per state")
        S = self._new(B, N)
        lib = _lib.load()
        _lib.check(lib.bx_supg_term(B, N, _ptr(self.Xd), _ptr(U), _ptr(mu2), _ptr(S), self._stream()), "bx_supg_term")
        return S

    @_on_device
    def assemble_picard(self, Uk, Un, mu1, mu2, At, E=0.0, supg=True, want_b=True):
        """One Picard linearisation for B states (FEM/fem_burgers.py:667-689).
        Returns dl, d, du, b, R — all (B, N) CUDA tensors (b is None when want_b=False)."""
        Uk = self._dev(Uk).reshape(-1, self.n_nodes)
        Un = self._dev(Un).reshape(-1, self.n_nodes)
        B, N = Uk.shape
        mu1 = self._dev(mu1).reshape(-1)
        mu2 = self._dev(mu2).reshape(-1)
        if Un.shape != Uk.shape or mu1.numel() != B or mu2.numel() != B:
            raise ValueError("Uk, Un, mu1, mu2 batch sizes differ")
        dl, d, du, R = self._new(B, N), self._new(B, N), self._new(B, N), self._new(B, N)
        b = self._new(B, N) if want_b else None
        lib = _lib.load()
        _lib.check(lib.bx_assemble_picard(B, N, _ptr(self.Xd), _ptr(Uk), _ptr(Un), _ptr(mu1), _ptr(mu2),
                                          float(At), float(E), int(bool(supg)), _ptr(dl), _ptr(d), _ptr(du),
                                          _ptr(b), _ptr(R), self._stream()), "bx_assemble_picard")
        return dl, d, du, b, R

    @_on_device
    def tridiag_solve_update(self, dl, d, du, R, Uk, out=None):
        """delta = A^-1(-R); U1 = Uk + delta; err = |delta|/|U1| (FEM/fem_burgers.py:692-698)."""
        Uk = self._dev(Uk).reshape(-1, self.n_nodes)
        B, N = Uk.shape
        U1 = out if out is not None else self._new(B, N)
        err = self._new(B)
        lib = _lib.load()
        _lib.check(lib.bx_tridiag_solve_update(B, N, _ptr(dl), _ptr(d), _ptr(du), _ptr(R), _ptr(Uk), _ptr(U1),
                                               _ptr(err), self._stream()), "bx_tridiag_solve_update")
        return U1, err

    # ------------------------------------------------------------------ batched trajectories
    def _sweep_args(self, u0, mu1, mu2, nTimeSteps, out, iters_out):
        """Common argument handling and validation of the trajectory entry points."""
        mu1 = self._dev(mu1).reshape(-1)
        mu2 = self._dev(mu2).reshape(-1)
        B, N = mu1.numel(), self.n_nodes
        if mu2.numel() != B:
            raise ValueError("mu1 and mu2 must have the same number of samples")
        u0 = self._dev(u0)
        if u0.numel() == N:
            batched = 0
        elif u0.numel() == B * N:
            batched = 1
        else:
            raise ValueError("u0 must have shape (N,) or (B, N)")
        nT = int(nTimeSteps)
        U = out if out is not None else self._new(B, nT + 1, N)
        if not isinstance(U, torch.Tensor) or tuple(U.shape) != (B, nT + 1, N) or U.dtype != torch.float64 \
                or not U.is_contiguous() or U.device != self.device:
            raise ValueError(f"out must be a contiguous float64 tensor of shape (B, nTimeSteps+1, N) on {self.device}")
        iters = iters_out if iters_out is not None else self._new(B, nT, dtype=torch.int32)
        if not isinstance(iters, torch.Tensor) or tuple(iters.shape) != (B, nT) or iters.dtype != torch.int32 \
                or not iters.is_contiguous() or iters.device != self.device:
            raise ValueError(f"iters_out must be a contiguous int32 tensor of shape (B, nTimeSteps) on {self.device}")
        return mu1, mu2, u0, batched, B, N, nT, U, iters

    def _err_hist(self, err_hist, B, nT, kmax):
        if err_hist is None:
            return None
        if not isinstance(err_hist, torch.Tensor) or tuple(err_hist.shape) != (B, nT, int(kmax)) \
                or err_hist.dtype != torch.float64 or not err_hist.is_contiguous() or err_hist.device != self.device:
            raise ValueError(f"err_hist must be a contiguous float64 tensor (B, nTimeSteps, kmax) on {self.device}")
        return err_hist

    @_on_device
    def fom_burgers_batch(self, At, nTimeSteps, u0, mu1, E, mu2, out=None, iters_out=None,
                          tol=1e-6, kmax=20, err_hist=None):
        """fom_burgers (FEM/fem_burgers.py:646) for B samples: mu1, mu2 (B,).
        Returns U (B, nTimeSteps+1, N) and iters (B, nTimeSteps) int32, CUDA tensors.  iters[b, step] = -1 (and NaN
        snapshots) from the first step of sample b whose state is not finite.  err_hist: optional preallocated
        (B, nTimeSteps, kmax) float64 tensor that receives |delta_U| / |U1| of every executed iteration (:698)."""
        mu1, mu2, u0, batched, B, N, nT, U, iters = self._sweep_args(u0, mu1, mu2, nTimeSteps, out, iters_out)
        eh = self._err_hist(err_hist, B, nT, kmax)
        lib = _lib.load()
        if B == 0:
            return U, iters
        if N > self.WARP_MESH_MAX:
            if eh is not None:
                raise ValueError("err_hist is not available on the fine-mesh path (N > 1024)")
            # fine mesh (BASELINE config 5): persistent cooperative kernel, partition solve across CTAs
            ws = self._scratch(int(lib.bx_fine_workspace_bytes(N)))
            _lib.check(lib.bx_fom_picard_finemesh(B, N, nT, _ptr(self.Xd), _ptr(u0), batched, _ptr(mu1),
                                                  _ptr(mu2), float(At), float(E), float(tol), int(kmax),
                                                  _ptr(U), _ptr(iters), _ptr(ws), self._stream()),
                       "bx_fom_picard_finemesh")
            return U, iters
        _lib.check(lib.bx_fom_picard_trajectories(B, N, nT, _ptr(self.Xd), _ptr(u0), batched, _ptr(mu1),
                                                  _ptr(mu2), float(At), float(E), float(tol), int(kmax),
                                                  _ptr(U), _ptr(iters), _ptr(eh),
                                                  _ptr(self._scratch(int(lib.bx_fom_workspace_bytes(B)))),
                                                  self._stream()),
                   "bx_fom_picard_trajectories")
        return U, iters

    QM_FUSED_MAX_MODES = 31   # bx_qm_prom_trajectories (one fused trajectory kernel); larger bases go stepwise

    def _qm_stepwise(self, At, nT, u0, batched, uxa, E, mu2, Phi, H, projection, U, iters, tol, itmax, eh):
        """pod_quadratic_manifold (:1081-1175) for n > 31 modes (up to 128): the B samples advance in lock step, every
        Gauss-Newton iteration is bx_qm_decode (u = Phi q + H sym(q x q) and T = Phi + H dQ/dq for all samples, :1116-1118,
        :1148) + bx_reduced_system with T as a per-sample tangent and no SUPG term (:1141-1156) + one batched dense solve
        (cuSOLVER; :1161).  A sample whose system is singular or whose state stops being finite carries iters = -1 and
        NaN snapshots from that step on; the others are unaffected."""
        lib = _lib.load()
        B, N, n = U.shape[0], self.n_nodes, int(Phi.shape[1])
        if n > 128:
            raise _lib.BurgersB200Error(f"pod_quadratic_manifold: n = {n} modes > 128 is not supported by this build")
        U[:, 0] = u0 if batched else u0.reshape(1, N)
        q = self._new(B, n); u = self._new(B, N); T = self._new(B, N, n)
        Ar = self._new(B, n, n); br = self._new(B, n)
        bad = torch.zeros(B, dtype=torch.bool, device=self.device)

        def decode():
            _lib.check(lib.bx_qm_decode(B, N, n, _ptr(Phi), _ptr(H), _ptr(q), _ptr(u), _ptr(T), self._stream()),
                       "bx_qm_decode")

        for m in range(nT):
            Um = U[:, m].contiguous()
            Um_safe = torch.where(bad.unsqueeze(1), torch.zeros_like(Um), Um)
            q.copy_(Um_safe @ Phi)                                                   # :1129
            decode()
            active = ~bad
            count = torch.zeros(B, dtype=torch.int32, device=self.device)
            for it in range(itmax):
                self.reduced_system_batch(At, E, u, Um_safe, uxa, mu2, T, projection, supg=False, out=(Ar, br))
                dq, info = torch.linalg.solve_ex(Ar, -br.unsqueeze(2))               # :1161
                dq = dq.squeeze(2)
                dq = torch.where(active.unsqueeze(1), dq, torch.zeros_like(dq))
                q.add_(dq)                                                           # :1162
                decode()                                                             # :1163 and the next tangent
                rel = torch.linalg.norm(dq, dim=1) / torch.clamp(torch.linalg.norm(q, dim=1), min=1e-14)   # :1165
                newly_bad = active & ((info != 0) | ~torch.isfinite(rel) | ~torch.isfinite(u).all(dim=1))
                if eh is not None:
                    eh[:, m, it] = torch.where(active, rel, eh[:, m, it])
                count += active.to(torch.int32)
                bad |= newly_bad
                if bool(newly_bad.any()):
                    q[newly_bad] = 0.0
                active = active & ~(rel < tol) & ~bad
                if not bool(active.any()):
                    break
            if bool(bad.any()):
                u[bad] = float("nan")
                count[bad] = -1
            U[:, m + 1] = u
            iters[:, m] = count
        return U, iters

    @staticmethod
    def _projection_code(projection):
        """'Galerkin' / 'LSPG' exactly as the reference spells them (FEM/fem_burgers.py:754-764)."""
        if projection == "Galerkin":
            return _lib.BX_GALERKIN
        if projection == "LSPG":
            return _lib.BX_LSPG
        raise ValueError(f"Projection method '{projection}' is not available. Please use 'Galerkin' or 'LSPG'.")

    @_on_device
    def fom_burgers_newton_batch(self, At, nTimeSteps, u0, mu1, E, mu2, out=None, iters_out=None,
                                 lambda_=0.1, tol=1e-6, kmax=100):
        """fom_burgers_newton as coded (FEM/fem_burgers.py:584) for B samples.
        Returns U (B, nTimeSteps+1, N) and iters (B, nTimeSteps) int32, CUDA tensors."""
        mu1, mu2, u0, batched, B, N, nT, U, iters = self._sweep_args(u0, mu1, mu2, nTimeSteps, out, iters_out)
        if B == 0:
            return U, iters
        lib = _lib.load()
        _lib.check(lib.bx_fom_newton_trajectories(B, N, nT, _ptr(self.Xd), _ptr(u0), batched, _ptr(mu1),
                                                  _ptr(mu2), float(At), float(E), float(lambda_), float(tol),
                                                  int(kmax), _ptr(U), _ptr(iters), _ptr(self._workspace()),
                                                  self._stream()), "bx_fom_newton_trajectories")
        return U, iters

    @_on_device
    def pod_prom_burgers_batch(self, At, nTimeSteps, u0, mu1, E, mu2, Phi, projection="Galerkin", out=None,
                               iters_out=None, tol=1e-6, kmax=20, err_hist=None):
        """pod_prom_burgers (FEM/fem_burgers.py:709) for B samples: mu1, mu2 (B,), Phi (N, n) shared.
        Returns U (B, nTimeSteps+1, N) and iters (B, nTimeSteps) int32, CUDA tensors; iters = -1 flags a sample
        whose reduced system went singular / non-finite (the reference raises LinAlgError at :767)."""
        code = self._projection_code(projection)
        mu1, mu2, u0, batched, B, N, nT, U, iters = self._sweep_args(u0, mu1, mu2, nTimeSteps, out, iters_out)
        eh = self._err_hist(err_hist, B, nT, kmax)
        Phi = self._dev(Phi)
        if Phi.dim() != 2 or Phi.shape[0] != N:
            raise ValueError("Phi must have shape (N, n)")
        n = int(Phi.shape[1])
        if B == 0:
            return U, iters
        lib = _lib.load()
        need = int(lib.bx_prom_workspace_bytes(B, N, n))
        if need < 0:
            raise _lib.BurgersB200Error("bx_prom_workspace_bytes failed: " + lib.bx_last_error().decode())
        _lib.check(lib.bx_prom_trajectories(B, N, n, nT, _ptr(self.Xd), _ptr(u0), batched, _ptr(mu1), _ptr(mu2),
                                            float(At), float(E), _ptr(Phi), code, float(tol), int(kmax),
                                            _ptr(U), _ptr(iters), _ptr(eh), _ptr(self._scratch(need)),
                                            self._stream()),
                   "bx_prom_trajectories")
        return U, iters

    @_on_device
    def pod_quadratic_manifold_batch(self, At, nTimeSteps, u0, uxa, E, mu2, Phi, H, projection="LSPG", out=None,
                                     iters_out=None, newton_tol=1e-6, newton_itmax=25, err_hist=None):
        """pod_quadratic_manifold (FEM/fem_burgers.py:1081) for B samples: uxa, mu2 (B,), Phi (N, n) and
        H (N, n(n+1)/2) shared.  Returns U (B, nTimeSteps+1, N) and iters (B, nTimeSteps) int32, CUDA tensors."""
        pl = str(projection).lower()   # the reference compares projection.lower() (:1150-1158)
        if pl == "galerkin":
            code = _lib.BX_GALERKIN
        elif pl == "lspg":
            code = _lib.BX_LSPG
        else:
            raise ValueError("projection must be 'Galerkin' or 'LSPG'")
        uxa, mu2, u0, batched, B, N, nT, U, iters = self._sweep_args(u0, uxa, mu2, nTimeSteps, out, iters_out)
        eh = self._err_hist(err_hist, B, nT, newton_itmax)
        Phi = self._dev(Phi)
        H = self._dev(H)
        if Phi.dim() != 2 or Phi.shape[0] != N:
            raise ValueError("Phi must have shape (N, n)")
        n = int(Phi.shape[1])
        if H.dim() != 2 or H.shape[0] != N or H.shape[1] != n * (n + 1) // 2:
            raise ValueError("H must have shape (N, n(n+1)/2)")   # quadratic_prom_simulation.py:38
        if B == 0:
            return U, iters
        if n > self.QM_FUSED_MAX_MODES:
            return self._qm_stepwise(At, nT, u0, batched, uxa, E, mu2, Phi, H, "LSPG" if code == _lib.BX_LSPG else "Galerkin",
                                     U, iters, float(newton_tol), int(newton_itmax), eh)
        lib = _lib.load()
        need = int(lib.bx_qm_workspace_bytes(B, N, n))
        if need < 0:
            raise _lib.BurgersB200Error("bx_qm_workspace_bytes failed: " + lib.bx_last_error().decode())
        _lib.check(lib.bx_qm_prom_trajectories(B, N, n, nT, _ptr(self.Xd), _ptr(u0), batched, _ptr(uxa), _ptr(mu2),
                                               float(At), float(E), _ptr(Phi), _ptr(H), code, float(newton_tol),
                                               int(newton_itmax), _ptr(U), _ptr(iters), _ptr(eh),
                                               _ptr(self._scratch(need)), self._stream()),
                   "bx_qm_prom_trajectories")
        return U, iters

    # ------------------------------------------------------------------ legacy surface (numpy / scipy in-out)
    def _csc(self, bands):
        import scipy.sparse as sp  # return type of the reference (:353); not used for any arithmetic here
        dl, d, du = (t.reshape(-1).cpu().numpy() for t in bands)
        return sp.diags([dl[1:], d, du[:-1]], [-1, 0, 1], format="csc")

    def compute_mass_matrix(self):
        return self._csc(self.mass_bands())

    def compute_diffusion_matrix(self):
        return self._csc(self.diffusion_bands())

    def compute_convection_matrix(self, U_n):
        return self._csc(self.convection_bands(np.asarray(U_n, dtype=np.float64).reshape(1, -1)))

    def compute_convection_matrix_derivative(self, U_n):
        return self._csc(self.convection_derivative_bands())

    def compute_forcing_vector(self, mu2):
        return self.forcing_vector_batch([float(mu2)]).reshape(-1).cpu().numpy()

    def compute_supg_term(self, U_n, mu2):
        return self.supg_term_batch(np.asarray(U_n, dtype=np.float64).reshape(1, -1), [float(mu2)]) \
            .reshape(-1).cpu().numpy()

    # -- single-sample trajectories: the reference's signatures and return type, ndarray (N, nTimeSteps+1) --
    def _legacy(self, run, kmax, nT, first_err=1):
        """Run one trajectory; with self.verbose also fetch the per-iteration errors for the progress lines."""
        eh = None
        if self.verbose:
            eh = torch.full((1, int(nT), int(kmax)), float("nan"), dtype=torch.float64, device=self.device)
        U, it = run(eh)
        self.last_iters = it[0].cpu().numpy()
        Un = np.ascontiguousarray(U[0].cpu().numpy().T)
        return Un, (eh[0].cpu().numpy() if eh is not None else None)

    def _print_picard(self, At, errs):
        """The lines of FEM/fem_burgers.py:659, 664 (= :722, 727): the error printed before iteration k is the one
        iteration k-1 produced (1 before the first)."""
        for n, k in enumerate(self.last_iters):
            print(f"Time Step: {n}. Time: {n * At}")
            for j in range(max(int(k), 0)):
                print(f"Iteration: {j}, Error: {1 if j == 0 else errs[n, j - 1]}")

    def _check_finite(self, what, singular=False):
        """Non-finite trajectories are flagged, not raised, by the batched surface.  The single-sample drop-ins keep
        the reference's behaviour: np.linalg.solve raises LinAlgError on a singular reduced system (:767, :1161);
        the sparse FOM solve only warns (SuperLU) and goes on with nan."""
        bad = np.flatnonzero(self.last_iters < 0)
        if bad.size == 0:
            return
        if singular:
            raise np.linalg.LinAlgError(f"Singular matrix ({what}: reduced system not finite at time step {int(bad[0])})")
        import warnings
        warnings.warn(f"{what}: state not finite from time step {int(bad[0])} on (matrix is exactly singular / "
                      "iteration diverged)", RuntimeWarning, stacklevel=3)

    def fom_burgers(self, At, nTimeSteps, u0, mu1, E, mu2):
        """Drop-in for FEMBurgers.fom_burgers: returns ndarray (N, nTimeSteps+1)."""
        U, errs = self._legacy(lambda eh: self.fom_burgers_batch(At, nTimeSteps, np.asarray(u0, dtype=np.float64),
                                                                 [float(mu1)], E, [float(mu2)], err_hist=eh),
                               20, nTimeSteps)
        if self.verbose:
            self._print_picard(At, errs)
        self._check_finite("fom_burgers")
        return U

    def fom_burgers_newton(self, At, nTimeSteps, u0, mu1, E, mu2):
        """Drop-in for FEMBurgers.fom_burgers_newton (:584-643): returns ndarray (N, nTimeSteps+1)."""
        U, it = self.fom_burgers_newton_batch(At, nTimeSteps, np.asarray(u0, dtype=np.float64), [float(mu1)], E,
                                              [float(mu2)])
        self.last_iters = it[0].cpu().numpy()
        self._check_finite("fom_burgers_newton")
        return np.ascontiguousarray(U[0].cpu().numpy().T)

    def pod_prom_burgers(self, At, nTimeSteps, u0, mu1, E, mu2, Phi, projection="Galerkin"):
        """Drop-in for FEMBurgers.pod_prom_burgers (:709-785): returns ndarray (N, nTimeSteps+1)."""
        Phi = np.asarray(Phi, dtype=np.float64)
        U, errs = self._legacy(lambda eh: self.pod_prom_burgers_batch(At, nTimeSteps, np.asarray(u0, dtype=np.float64),
                                                                      [float(mu1)], E, [float(mu2)], Phi, projection,
                                                                      err_hist=eh), 20, nTimeSteps)
        if self.verbose:
            self._print_picard(At, errs)
        self._check_finite("pod_prom_burgers", singular=True)
        return U

    def pod_quadratic_manifold(self, At, nTimeSteps, u0, uxa, E, mu2, Phi, H, projection="LSPG",
                               newton_tol=1e-6, newton_itmax=25):
        """Drop-in for FEMBurgers.pod_quadratic_manifold (:1081-1175): returns ndarray (N, nTimeSteps+1)."""
        Phi = np.asarray(Phi, dtype=np.float64); H = np.asarray(H, dtype=np.float64)
        U, errs = self._legacy(lambda eh: self.pod_quadratic_manifold_batch(
            At, nTimeSteps, np.asarray(u0, dtype=np.float64), [float(uxa)], E, [float(mu2)], Phi, H, projection,
            newton_tol=newton_tol, newton_itmax=newton_itmax, err_hist=eh), newton_itmax, nTimeSteps)
        if self.verbose:   # the lines of :1127, :1166, :1171
            for m, k in enumerate(self.last_iters):
                print(f"\n=== time step {m+1}/{nTimeSteps}  (t = {(m+1)*At:.2f}) ===")
                for it in range(max(int(k), 0)):
                    print(f"  Newton {it:2d}:  |δq|/|q| = {errs[m, it]:.3e}")
                if k == newton_itmax and not (errs[m, k - 1] < newton_tol):
                    print("  Warning: Newton did not converge")
        self._check_finite("pod_quadratic_manifold", singular=True)
        return U

    # -- reduced system for a caller-supplied tangent and its first consumer, local POD ------------
    @_on_device
    def reduced_system_batch(self, At, E, Uk, Un, mu1, mu2, T, projection="Galerkin", supg=True, out=None):
        """Ar, br of ONE nonlinear iteration (FEM/fem_burgers.py:730-762 and the same block in local_prom_burgers,
        pod_ann_prom, pod_rbf_prom, lie_prom, ae_prom) for B samples: Uk, Un (B, N); mu1, mu2 (B,); T the decoder
        tangent, (N, n) shared by all samples or (B, N, n) one per sample.  Returns CUDA tensors Ar (B, n, n),
        br (B, n); the caller solves and updates with its own decoder (np.linalg.solve(Ar, -br) at :767)."""
        code = self._projection_code(projection)
        Uk = self._dev(Uk); Un = self._dev(Un)
        N = self.n_nodes
        if Uk.dim() == 1:
            Uk = Uk.reshape(1, N)
        if Un.dim() == 1:
            Un = Un.reshape(1, N)
        B = Uk.shape[0]
        if tuple(Uk.shape) != (B, N) or tuple(Un.shape) != (B, N):
            raise ValueError("Uk and Un must have shape (B, N)")
        mu1 = self._dev(np.atleast_1d(mu1) if not isinstance(mu1, torch.Tensor) else mu1).reshape(-1)
        mu2 = self._dev(np.atleast_1d(mu2) if not isinstance(mu2, torch.Tensor) else mu2).reshape(-1)
        if mu1.numel() != B or mu2.numel() != B:
            raise ValueError("mu1 and mu2 must have one entry per sample")
        T = self._dev(T)
        if T.dim() == 2 and T.shape[0] == N:
            n, stride = int(T.shape[1]), 0
        elif T.dim() == 3 and tuple(T.shape[:2]) == (B, N):
            n, stride = int(T.shape[2]), N * int(T.shape[2])
        else:
            raise ValueError("T must have shape (N, n) or (B, N, n)")
        if out is None:
            Ar = self._new(B, n, n); br = self._new(B, n)
        else:
            Ar, br = out
            if tuple(Ar.shape) != (B, n, n) or tuple(br.shape) != (B, n) or Ar.dtype != torch.float64 \
                    or br.dtype != torch.float64 or not Ar.is_contiguous() or not br.is_contiguous() \
                    or Ar.device != self.device or br.device != self.device:
                raise ValueError(f"out must be contiguous float64 tensors (B, n, n) and (B, n) on {self.device}")
        if B == 0:
            return Ar, br
        lib = _lib.load()
        _lib.check(lib.bx_reduced_system(B, N, n, _ptr(self.Xd), _ptr(Uk), _ptr(Un), _ptr(mu1), _ptr(mu2), float(At),
                                         float(E), 1 if supg else 0, _ptr(T), stride, code, _ptr(Ar), _ptr(br),
                                         self._stream()), "bx_reduced_system")
        return Ar, br

    @_on_device
    def local_prom_burgers_batch(self, At, nTimeSteps, u0, mu1, E, mu2, cluster_centers, local_bases, U_global,
                                 num_global_modes, projection="Galerkin", tol=1e-6, kmax=20):
        """local_prom_burgers (FEM/fem_burgers.py:979-1078) for B samples in lock step.  cluster_centers: (K, g) k-means
        centres (`kmeans.cluster_centers_`; `kmeans.predict` = nearest centre, :1011); local_bases: dict or sequence
        of K arrays (N, n_c); U_global: (N, >= g).  Every time step picks ONE basis per sample from the cluster of
        U_global[:, :g]^T U^n (:1010-1012); every Picard iteration is one bx_reduced_system call with a per-sample
        tangent, one batched solve (cuSOLVER) and two batched products for q = Phi^T U0 + dq, U1 = Phi q (:1061-1067).
        Bases are padded with zero columns to the widest one (their rows / columns of Ar become the identity).
        Returns U (B, nTimeSteps+1, N), iters (B, nTimeSteps) int32, clusters (B, nTimeSteps) int64: CUDA tensors."""
        self._projection_code(projection)
        N = self.n_nodes
        mu1 = self._dev(np.atleast_1d(mu1) if not isinstance(mu1, torch.Tensor) else mu1).reshape(-1)
        mu2 = self._dev(np.atleast_1d(mu2) if not isinstance(mu2, torch.Tensor) else mu2).reshape(-1)
        B = mu1.numel()
        if mu2.numel() != B:
            raise ValueError("mu1 and mu2 must have the same number of samples")
        nT = int(nTimeSteps)
        g = int(num_global_modes)
        centers = self._dev(cluster_centers)
        K = centers.shape[0]
        bases = [local_bases[c] for c in range(K)] if isinstance(local_bases, dict) else list(local_bases)
        if len(bases) != K:
            raise ValueError("one local basis per cluster centre is required")
        widths = [int(np.shape(b)[1]) for b in bases]
        nmax = max(widths)
        Phi_all = torch.zeros((K, N, nmax), dtype=torch.float64, device=self.device)
        for c, b in enumerate(bases):
            Phi_all[c, :, :widths[c]] = self._dev(b)
        pad = torch.zeros((K, nmax), dtype=torch.float64, device=self.device)
        for c in range(K):
            pad[c, widths[c]:] = 1.0
        Ug = self._dev(U_global)[:, :g].contiguous()
        if centers.shape[1] != g:
            raise ValueError("cluster centres and num_global_modes disagree")
        U = self._new(B, nT + 1, N)
        iters = self._new(B, nT, dtype=torch.int32)
        clusters = torch.empty((B, nT), dtype=torch.int64, device=self.device)
        u0 = self._dev(u0)
        U[:, 0] = u0 if u0.numel() == B * N else u0.reshape(1, N)
        Ar = self._new(B, nmax, nmax); br = self._new(B, nmax)
        for n in range(nT):
            Un = U[:, n].contiguous()
            qg = Un @ Ug                                                     # (B, g)
            cid = torch.cdist(qg, centers).argmin(dim=1)                     # nearest centre (:1011)
            clusters[:, n] = cid
            T = Phi_all[cid].contiguous()                                    # (B, N, nmax)
            eye_pad = torch.diag_embed(pad[cid])
            U0 = Un.clone()
            active = torch.ones(B, dtype=torch.bool, device=self.device)
            count = torch.zeros(B, dtype=torch.int32, device=self.device)
            for k in range(int(kmax)):
                self.reduced_system_batch(At, E, U0, Un, mu1, mu2, T, projection, supg=True, out=(Ar, br))
                dq = torch.linalg.solve(Ar + eye_pad, -br.unsqueeze(2)).squeeze(2)          # :1058
                q = torch.bmm(T.transpose(1, 2), U0.unsqueeze(2)).squeeze(2) + dq           # :1061
                U1 = torch.bmm(T, q.unsqueeze(2)).squeeze(2)                                # :1064
                err = torch.linalg.norm(dq, dim=1) / torch.linalg.norm(q, dim=1)            # :1067
                U0 = torch.where(active.unsqueeze(1), U1, U0)
                count += active.to(torch.int32)
                active = active & (err > tol)
                if not bool(active.any()):
                    break
            U[:, n + 1] = U0
            iters[:, n] = count
        return U, iters, clusters

    def local_prom_burgers(self, At, nTimeSteps, u0, mu1, E, mu2, kmeans, local_bases, U_global, num_global_modes,
                           projection="Galerkin"):
        """Drop-in for FEMBurgers.local_prom_burgers (:979-1078): `kmeans` is the fitted sklearn model the reference
        passes (only its cluster_centers_ are read; predict = nearest centre), or the (K, g) array of centres.
        Returns ndarray (N, nTimeSteps+1); per-step iteration counts in self.last_iters, clusters in self.last_clusters."""
        centers = getattr(kmeans, "cluster_centers_", kmeans)
        U, it, cl = self.local_prom_burgers_batch(At, nTimeSteps, np.asarray(u0, dtype=np.float64), [float(mu1)], E,
                                                  [float(mu2)], np.asarray(centers, dtype=np.float64), local_bases,
                                                  np.asarray(U_global, dtype=np.float64), num_global_modes, projection)
        self.last_iters = it[0].cpu().numpy()
        self.last_clusters = cl[0].cpu().numpy()
        return np.ascontiguousarray(U[0].cpu().numpy().T)

    # the north_star's name for the same method
    pod_prom_newton = pod_prom_burgers
