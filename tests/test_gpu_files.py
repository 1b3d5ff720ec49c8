"""GPU: the on-disk side of the sweep — device transpose to the reference's (N, steps+1)
layout, one .npy per sample with the reference's file names, and the POD stage consuming such a directory.
Bit-exact: the files hold exactly what the legacy drop-in method returns."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, rel_l2_per_step

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,rows,N", [(1, 1, 2), (3, 7, 33), (5, 501, 512), (2, 64, 1000), (0, 5, 16)])
def test_to_reference_layout_is_a_transpose(B, rows, N):
    import torch
    import burgers_b200
    from burgers_b200 import sweep
    fem = burgers_b200.FEMBurgers(np.linspace(0.0, 1.0, max(N, 2)))
    U = torch.randn((B, rows, N), dtype=torch.float64, device="cuda")
    Ur = sweep.to_reference_layout(fem, U)
    assert Ur.shape == (B, N, rows) and Ur.is_contiguous()
    assert torch.equal(Ur, U.permute(0, 2, 1))
    if B > 0 and rows > 1:   # (degenerate shapes stay contiguous under permute)
        with pytest.raises(ValueError):
            sweep.to_reference_layout(fem, U.permute(0, 2, 1))


def test_sweep_to_npy_matches_legacy_method_and_reference_names(tmp_path, oracle, golden):
    import burgers_b200
    from burgers_b200 import sweep, pod
    X, T = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X, T)
    # the reference's 3 x 3 training grid (FEM/paper_training_stage.py:8-10) + a ragged tail, chunked unevenly
    mu1, mu2 = sweep.parameter_grid(3, 3)
    mu1 = np.concatenate([mu1, [4.75, 6.2]]); mu2 = np.concatenate([mu2, [0.02, 0.04]])
    paths, iters = sweep.sweep_to_npy(fem, str(tmp_path / "fem_training_data"), 0.05, 20, np.ones(512), mu1, mu2,
                                      chunk=4, writers=2)
    names = [os.path.basename(p) for p in paths]
    assert names[0] == "fem_simulation_mu1_4.250_mu2_0.0150.npy" and names[4] == "fem_simulation_mu1_4.875_mu2_0.0225.npy"
    assert names[8] == "fem_simulation_mu1_5.500_mu2_0.0300.npy" and names[9] == "fem_simulation_mu1_4.750_mu2_0.0200.npy"
    g = golden("fom_fixtures.npz")
    for b in (0, 4, 9, 10):
        A = np.load(paths[b])
        assert A.shape == (512, 21) and A.dtype == np.float64 and A.flags.c_contiguous
        U = fem.fom_burgers(0.05, 20, np.ones(512), mu1[b], 0.0, mu2[b])
        assert np.array_equal(A, U)                       # bit-exact: same kernel, only the layout differs
        assert np.array_equal(iters[b], fem.last_iters)
    # the stored reference fixture for (4.75, 0.02): the file content is the reference's, step by step
    steps = g["steps"]; keep = steps[steps <= 20]
    for b, k in ((0, 0), (4, 4), (8, 8), (9, 10)):      # fixture index k of sample b (g["points"])
        assert tuple(g["points"][k]) == (mu1[b], mu2[b])
        assert rel_l2_per_step(np.load(paths[b])[:, keep], g[f"U_{k}"][:, : keep.size]) <= 1e-10
    # POD stage on the directory (POD/pod.py:80-90): same basis as numpy's SVD of the hstacked files
    S, files = pod.load_snapshot_directory(str(tmp_path / "fem_training_data"))
    assert S.shape == (11 * 21, 512) and len(files) == 11
    Phi, s = pod.pod_basis(S, n_modes=6)
    Sn = np.hstack([np.load(p) for p in paths])
    Us, ss, _ = np.linalg.svd(Sn, full_matrices=False)
    Phi = Phi.cpu().numpy(); s = s.cpu().numpy()
    assert np.max(np.abs(s - ss[:6]) / ss[:6]) <= 1e-9
    assert np.linalg.norm(Phi @ Phi.T - Us[:, :6] @ Us[:, :6].T, 2) <= 1e-7


def test_prom_files_with_reference_names(tmp_path, oracle):
    import burgers_b200
    from burgers_b200 import sweep
    X, T = oracle.mesh(512)
    fem = burgers_b200.FEMBurgers(X, T)
    Phi = np.load(f"{GOLDEN}/Phi_tol_1e-03.npy")
    name = lambda a, b: sweep.prom_filename(1e-3, a, b, "LSPG")
    paths, _ = sweep.sweep_to_npy(fem, str(tmp_path / "rom_solutions"), 0.05, 10, np.ones(512), [4.75, 5.19], [0.02, 0.026],
                                  method="pod_prom_burgers", name=name, Phi=Phi, projection="LSPG")
    assert os.path.basename(paths[0]) == "U_PROM_tol_1e-03_mu1_4.750_mu2_0.0200_lspg.npy"   # a file the reference ships
    U = fem.pod_prom_burgers(0.05, 10, np.ones(512), 5.19, 0.0, 0.026, Phi, "LSPG")
    assert np.array_equal(np.load(paths[1]), U)
