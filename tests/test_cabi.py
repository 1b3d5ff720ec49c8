"""CPU: the C-ABI library builds, loads, and exports every symbol include/burgers_b200.h declares;
host-side argument validation that needs no GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import burgers_b200
    burgers_b200.build_library()
    return burgers_b200._lib.load()


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "burgers_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bx_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(lib):
    import burgers_b200
    raw = ctypes.CDLL(burgers_b200.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 13
    for n in names:
        assert hasattr(raw, n), f"{n} declared in include/burgers_b200.h but not exported"
    # and every declared symbol has a ctypes signature in the host binding
    assert set(names) == set(burgers_b200._lib.SIGNATURES)


def test_version_and_workspace(lib):
    assert lib.bx_version() >= 100
    assert lib.bx_workspace_bytes() >= 4


def test_argument_validation_without_gpu(lib):
    # invalid sizes are rejected before any CUDA call is made
    assert lib.bx_fom_picard_trajectories(4, 1, 10, None, None, 0, None, None, 0.05, 0.0, 1e-6, 20,
                                          None, None, None, None, None) == -1
    assert b"2 nodes" in lib.bx_last_error()
    assert lib.bx_fom_picard_trajectories(4, 512, 10, None, None, 0, None, None, 0.05, 0.0, 1e-6, 0,
                                          None, None, None, None, None) == -1
    assert lib.bx_tridiag_solve_update(1, 4096, None, None, None, None, None, None, None, None) == -1
    assert lib.bx_mass_bands(1, None, None, None, None, None) == -1


def test_tuning_knobs_are_an_explicit_api(lib):
    """Launch-shape switches are set through bx_set_tuning, not the environment."""
    import burgers_b200
    for knob in burgers_b200._lib.TUNING_KNOBS:
        assert burgers_b200.get_tuning(knob) == -1
        burgers_b200.set_tuning(knob, 2)
        assert burgers_b200.get_tuning(knob) == 2
        burgers_b200.set_tuning(knob, -5)
        assert burgers_b200.get_tuning(knob) == -1
    assert lib.bx_set_tuning(99, 1) == -1
    with pytest.raises(KeyError):
        burgers_b200.set_tuning("no_such_knob", 1)
    src = "".join(open(os.path.join(ROOT, "1d-burgers-equation-roms_b200", "csrc", f)).read()
                  for f in os.listdir(os.path.join(ROOT, "1d-burgers-equation-roms_b200", "csrc")))
    assert "getenv" not in src


def test_mesh_validation():
    import burgers_b200
    X = np.linspace(0, 100, 16)
    T = np.array([np.arange(1, 16), np.arange(2, 17)]).T
    burgers_b200.validate_mesh(X, T)
    with pytest.raises(ValueError):
        burgers_b200.validate_mesh(X, T[::-1])
    with pytest.raises(ValueError):
        burgers_b200.validate_mesh(X[::-1], T)
    with pytest.raises(ValueError):
        burgers_b200.validate_mesh(X, T - 1)   # 0-based connectivity is not the reference's convention


def test_no_cpu_fallback():
    import torch
    import burgers_b200
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    fem = burgers_b200.FEMBurgers(np.linspace(0, 100, 64))
    with pytest.raises(RuntimeError):
        fem.fom_burgers(0.05, 2, np.ones(64), 4.75, 0.0, 0.02)


def test_reference_file_names_and_npy_writer(tmp_path):
    """Host logic of the on-disk side (no GPU): the reference's file names (FEM/paper_training_stage.py:52,
    POD/Results_thesis/prom_pod.py:59-60 — names of files the reference ships) and the header + raw-bytes writer."""
    from burgers_b200 import sweep
    assert sweep.snapshot_filename(4.25, 0.015) == "fem_simulation_mu1_4.250_mu2_0.0150.npy"
    assert sweep.snapshot_filename(4.875, 0.0225) == "fem_simulation_mu1_4.875_mu2_0.0225.npy"
    assert sweep.snapshot_filename(6.2, 0.04) == "fem_simulation_mu1_6.200_mu2_0.0400.npy"
    assert sweep.prom_filename(1e-3, 4.75, 0.02, "LSPG") == "U_PROM_tol_1e-03_mu1_4.750_mu2_0.0200_lspg.npy"
    assert sweep.prom_filename(4e-3, 5.19, 0.026, "Galerkin") == "U_PROM_tol_4e-03_mu1_5.190_mu2_0.0260_galerkin.npy"
    mu1, mu2 = sweep.parameter_grid(3, 3)      # mu1-major, as [[a, b] for a in mu1 for b in mu2]
    assert list(mu1[:4]) == [4.25, 4.25, 4.25, 4.875] and list(mu2[:4]) == [0.015, 0.0225, 0.03, 0.015]
    a = np.arange(512 * 7, dtype=np.float64).reshape(512, 7) / 3.0
    sweep._write_npy(str(tmp_path / "a.npy"), a)
    b = np.load(tmp_path / "a.npy")
    assert b.dtype == np.float64 and b.shape == (512, 7) and np.array_equal(a, b)
    it = np.arange(12, dtype=np.int32).reshape(3, 4)
    sweep._write_npy(str(tmp_path / "it.npy"), it)
    assert np.array_equal(np.load(tmp_path / "it.npy"), it)


@pytest.mark.parametrize("B,chunk", [(0, 512), (1, 512), (511, 512), (512, 512), (513, 512), (4096, 512), (4097, 1024), (5000, 2048)])
@pytest.mark.parametrize("ramp", [False, True])
def test_chunk_schedule_covers_the_sweep(B, chunk, ramp):
    """Host logic of the host-streaming sweep: the pieces tile range(B) in order, none is empty or longer than
    the buffers, and with ramp the copy starts after a quarter-size piece."""
    from burgers_b200 import sweep
    pieces = sweep.chunk_schedule(B, chunk, ramp)
    assert [lo for lo, _ in pieces] == [0] * min(B, 1) + [hi for _, hi in pieces[:-1]]
    assert all(0 < hi - lo <= chunk for lo, hi in pieces)
    assert (pieces[-1][1] if pieces else 0) == B
    if ramp and B > chunk:
        assert pieces[0] == (0, chunk // 4) and pieces[1] == (chunk // 4, chunk // 4 + chunk // 2)
    elif B:
        assert pieces[0] == (0, min(B, chunk))
