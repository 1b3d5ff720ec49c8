"""GPU parity of the snapshot Gram kernel (bx_gram_accumulate, DMMA) and the POD basis derived from it
against the oracle (numpy S S^T, np.linalg.svd as POD/pod.py:84 does).
Tolerances: Gram entries relative to |G|_max <= 1e-13 (FP64 sums of <= 2e6 terms in a different order);
projector Phi Phi^T vs the SVD's <= 1e-9 in 2-norm; singular values <= 1e-10 relative (leading modes)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _gram(S_rows):
    """S_rows: (cols, N) numpy -> G via the C ABI."""
    import torch
    from burgers_b200 import pod
    U = torch.from_numpy(np.ascontiguousarray(S_rows)).cuda()
    G = pod.snapshot_gram(U)
    torch.cuda.synchronize()
    return G.cpu().numpy()


@pytest.mark.parametrize("N,cols", [(512, 1503), (512, 16), (512, 1), (130, 333), (128, 4096), (77, 50),
                                    (258, 17), (1024, 200), (6, 5)])
def test_gram_matches_numpy(N, cols, oracle):
    rng = np.random.default_rng(N * 1000 + cols)
    S = 1.0 + 4.0 * rng.random((cols, N))
    G = _gram(S)
    Gref = oracle.snapshot_gram(S.T)
    assert G.shape == (N, N)
    assert np.array_equal(G, G.T)
    assert np.max(np.abs(G - Gref)) <= 1e-13 * np.max(np.abs(Gref))


def test_gram_accumulates_and_is_reproducible(oracle):
    import torch
    from burgers_b200 import pod
    rng = np.random.default_rng(5)
    A = rng.standard_normal((700, 512)); B = rng.standard_normal((3, 45, 512))
    acc = pod.GramAccumulator(512)
    acc.add(torch.from_numpy(A).cuda()).add(torch.from_numpy(B).cuda())
    G1 = acc.G.cpu().numpy()
    Gref = A.T @ A + B.reshape(-1, 512).T @ B.reshape(-1, 512)
    assert acc.cols == 700 + 135
    assert np.max(np.abs(G1 - Gref)) <= 1e-13 * np.max(np.abs(Gref))
    acc2 = pod.GramAccumulator(512)
    acc2.add(torch.from_numpy(A).cuda()).add(torch.from_numpy(B).cuda())
    assert np.array_equal(G1, acc2.G.cpu().numpy())   # fixed summation order


def test_gram_rejects_bad_input():
    import torch
    from burgers_b200 import pod
    acc = pod.GramAccumulator(64)
    with pytest.raises(ValueError):
        acc.add(torch.zeros((4, 64), dtype=torch.float32, device="cuda"))
    with pytest.raises(ValueError):
        acc.add(torch.zeros((4, 65), dtype=torch.float64, device="cuda"))
    acc.add(torch.zeros((0, 64), dtype=torch.float64, device="cuda"))
    assert acc.cols == 0


def test_pod_basis_of_fom_sweep_matches_svd(oracle):
    """config 3, first half, small: sweep on the GPU -> Gram -> eigh; compare with the SVD of the
    oracle's snapshots of the same sweep (POD/pod.py:80-84)."""
    import torch
    import burgers_b200
    from burgers_b200 import pod
    X, T = oracle.mesh(512)
    mu1, mu2 = oracle.parameter_grid(3, 3)
    fem = burgers_b200.FEMBurgers(X, T)
    U, _ = fem.fom_burgers_batch(0.05, 60, np.ones(512), mu1, 0.0, mu2)
    Phi, s = pod.pod_basis(U, n_modes=28)
    Phi = Phi.cpu().numpy(); s = s.cpu().numpy()
    Uo, _ = oracle.fom_burgers_batch(X, 0.05, 60, np.ones(512), mu1, 0.0, mu2)
    S = Uo.reshape(-1, 512).T                       # (N, cols) as the reference stacks it
    Us, ss = oracle.pod_basis_svd(S)
    assert np.max(np.abs(s - ss[:28]) / ss[:28]) <= 1e-10
    assert np.linalg.norm(Phi.T @ Phi - np.eye(28), 2) <= 1e-12
    # compare subspaces at a spectral gap (mode count from the reference's energy rule)
    k = pod.get_number_of_singular_values_for_given_tolerance(512, S.shape[1], ss, 1e-3)
    assert k == oracle.modes_for_tolerance(ss, 1e-3) and k <= 28
    P1 = Phi[:, :k] @ Phi[:, :k].T; P2 = Us[:, :k] @ Us[:, :k].T
    assert np.linalg.norm(P1 - P2, 2) <= 1e-9
