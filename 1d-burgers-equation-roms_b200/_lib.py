"""ctypes binding of libburgers_b200.so (include/burgers_b200.h).  No fallback: if the library is
missing or a call fails this raises — the product never computes on the CPU."""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_double, c_int, c_longlong, c_void_p

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "libburgers_b200.so")

BX_OK, BX_ERR_INVALID, BX_ERR_CUDA, BX_ERR_UNSUPPORTED, BX_ERR_PROJECTION = 0, -1, -2, -3, -4
BX_GALERKIN, BX_LSPG = 0, 1
BX_PROBE_DFMA, BX_PROBE_DMMA = 0, 1
# bx_set_tuning knobs (include/burgers_b200.h)
TUNING_KNOBS = {"fom_warps_per_trajectory": 0, "prom_cta_warps": 1, "prom_cta_chunk": 2, "qm_schedule": 3,
                "prom_warps_per_trajectory": 4, "fom_kernel": 5,
                "fom_segment_steps": 6, "qm_group": 7, "fine_rows": 8}

P = c_void_p  # device pointer / stream

SIGNATURES = {
    "bx_version": (c_int, []),
    "bx_last_error": (c_char_p, []),
    "bx_device_sm_count": (c_int, []),
    "bx_mass_bands": (c_int, [c_int, P, P, P, P, P]),
    "bx_diffusion_bands": (c_int, [c_int, P, P, P, P, P]),
    "bx_convection_bands": (c_int, [c_int, c_int, P, P, P, P, P, P]),
    "bx_convection_derivative_bands": (c_int, [c_int, P, P, P, P, P]),
    "bx_forcing_vector": (c_int, [c_int, c_int, P, P, P, P]),
    "bx_supg_term": (c_int, [c_int, c_int, P, P, P, P, P]),
    "bx_assemble_picard": (c_int, [c_int, c_int, P, P, P, P, P, c_double, c_double, c_int, P, P, P, P, P, P]),
    "bx_tridiag_solve_update": (c_int, [c_int, c_int, P, P, P, P, P, P, P, P]),
    "bx_workspace_bytes": (c_longlong, []),
    "bx_fom_workspace_bytes": (c_longlong, [c_int]),
    "bx_fom_plan": (c_int, [c_int, c_int]),
    "bx_fom_picard_trajectories": (c_int, [c_int, c_int, c_int, P, P, c_int, P, P, c_double, c_double,
                                           c_double, c_int, P, P, P, P, P]),
    "bx_fom_picard_trajectories_noalloc": (c_int, [c_int, c_int, c_int, P, P, c_int, P, P, c_double, c_double,
                                                   c_double, c_int, P, P, P, P]),
    "bx_fine_workspace_bytes": (c_longlong, [c_int]),
    "bx_fom_picard_finemesh": (c_int, [c_int, c_int, c_int, P, P, c_int, P, P, c_double, c_double,
                                       c_double, c_int, P, P, P, P]),
    "bx_fine_peer_bytes": (c_longlong, []),
    "bx_peer_alloc": (c_int, [P, P]),
    "bx_peer_open": (c_int, [P, P]),
    "bx_peer_close": (c_int, [P]),
    "bx_peer_free": (c_int, [P]),
    "bx_fom_picard_finemesh_ranks": (c_int, [c_int, c_int, c_int, c_int, c_int, P, P, c_int, P, P, c_double, c_double,
                                             c_double, c_int, P, P, P, c_int, c_int, P, P]),
    "bx_fine_peer_failed": (c_int, [P, P]),
    "bx_fom_newton_trajectories": (c_int, [c_int, c_int, c_int, P, P, c_int, P, P, c_double, c_double,
                                           c_double, c_double, c_int, P, P, P, P]),
    "bx_prom_workspace_bytes": (c_longlong, [c_int, c_int, c_int]),
    "bx_prom_trajectories": (c_int, [c_int, c_int, c_int, c_int, P, P, c_int, P, P, c_double, c_double,
                                     P, c_int, c_double, c_int, P, P, P, P, P]),
    "bx_qm_decode": (c_int, [c_int, c_int, c_int, P, P, P, P, P, P]),
    "bx_reduced_system": (c_int, [c_int, c_int, c_int, P, P, P, P, P, c_double, c_double, c_int, P, c_longlong, c_int,
                                  P, P, P]),
    "bx_qm_workspace_bytes": (c_longlong, [c_int, c_int, c_int]),
    "bx_qm_prom_trajectories": (c_int, [c_int, c_int, c_int, c_int, P, P, c_int, P, P, c_double, c_double,
                                        P, P, c_int, c_double, c_int, P, P, P, P, P]),
    "bx_qm_schedule_available": (c_int, [c_int, c_int, c_int]),
    "bx_set_tuning": (c_int, [c_int, c_int]),
    "bx_get_tuning": (c_int, [c_int]),
    "bx_probe_fp64_peak": (c_longlong, [c_int, c_int, P, P]),
    "bx_gram_workspace_bytes": (c_longlong, [c_int, c_longlong]),
    "bx_gram_accumulate": (c_int, [c_int, c_longlong, P, P, c_int, P, P]),
    "bx_snapshots_to_reference_layout": (c_int, [c_int, c_int, c_int, P, P, P]),
}

_lib = None


class BurgersB200Error(RuntimeError):
    """A C-ABI call returned a negative code (native precedent: std::runtime_error ->
    RuntimeError, PybindExample/sparse_solver_parallel.cpp:13-25)."""


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def set_tuning(knob: str, value: int):
    """Force a launch-shape choice (A/B measurements, tests that must reach every kernel instance); value < 0
    restores the automatic choice.  Process-wide.  Replaces round 1's environment switches."""
    if knob not in TUNING_KNOBS:
        raise KeyError(f"unknown tuning knob {knob!r}; known: {sorted(TUNING_KNOBS)}")
    check(load().bx_set_tuning(TUNING_KNOBS[knob], int(value)), "bx_set_tuning")


def get_tuning(knob: str) -> int:
    return int(load().bx_get_tuning(TUNING_KNOBS[knob]))


def qm_schedules(N: int, n: int):
    """Tangent-row schedules of the quadratic-manifold kernel that can run at (N, n) on the current device."""
    lib = load()
    return [s for s in (0, 1, 2) if lib.bx_qm_schedule_available(int(N), int(n), s) == 1]


def check(code: int, what: str):
    if code == BX_OK:
        return
    msg = load().bx_last_error().decode("utf-8", "replace")
    if code == BX_ERR_PROJECTION:
        raise ValueError(msg)
    raise BurgersB200Error(f"{what} failed with code {code}: {msg}")
