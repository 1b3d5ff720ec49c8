"""Oracle vs the unmodified reference for the local-POD PROM (FEM/fem_burgers.py:979-1078) and the single-iteration
reduced system (:730-762): tests/golden/local_prom.npz was produced by tests/golden/make_local_golden.py from the
reference and the artefacts it ships (Local_POD/clusters).  CPU only."""
import numpy as np
import pytest

from conftest import rel_l2_per_step


@pytest.fixture(scope="module")
def local(golden):
    g = golden("local_prom.npz")
    bases = [g[f"basis_{c}"] for c in range(10)]
    return g, bases


@pytest.mark.parametrize("proj", ["LSPG", "Galerkin"])
def test_oracle_local_prom_matches_reference(proj, oracle, local):
    g, bases = local
    X, _ = oracle.mesh(512)
    for pi, (mu1, mu2) in enumerate(g["points"]):
        U, it, cl = oracle.local_prom_burgers(X, 0.05, int(g["nT"]), np.ones(512), mu1, 0.0, mu2, g["centers"], bases,
                                              g["U_global"], int(g["g"]), proj)
        assert np.array_equal(it, g[f"it_{pi}_{proj}"]) and np.array_equal(cl, g[f"cl_{pi}_{proj}"])
        assert rel_l2_per_step(U, g[f"U_{pi}_{proj}"]) <= 1e-11
    for si in range(2):   # runs that switch basis mid-way
        U, it, cl = oracle.local_prom_burgers(X, 0.05, 8, g[f"sw_u0_{si}"], 4.56, 0.0, 0.019, g["centers"], bases,
                                              g["U_global"], int(g["g"]), proj)
        assert np.array_equal(it, g[f"sw_it_{si}_{proj}"]) and np.array_equal(cl, g[f"sw_cl_{si}_{proj}"])
        assert len(set(cl.tolist())) == 2
        assert rel_l2_per_step(U, g[f"sw_U_{si}_{proj}"]) <= 1e-11


@pytest.mark.parametrize("proj", ["LSPG", "Galerkin"])
def test_oracle_reduced_system_matches_reference_operators(proj, oracle, local):
    g, bases = local
    X, _ = oracle.mesh(512)
    for j in range(3):
        mu1, mu2 = g["rs_mu"][j]
        Phi = bases[int(g["rs_basis_ids"][j])]
        dl, d, du, b, R = oracle.picard_system(X, g["rs_Uk"][j], g["rs_Un"][j], mu1, mu2, float(g["rs_At"]), float(g["rs_E"]), True)
        Ar, br = oracle.reduced_system(dl, d, du, R, Phi, proj)
        Ar_ref, br_ref = g[f"rs_Ar_{proj}_{j}"], g[f"rs_br_{proj}_{j}"]
        assert np.abs(Ar - Ar_ref).max() <= 1e-13 * np.abs(Ar_ref).max()
        assert np.abs(br - br_ref).max() <= 1e-12 * max(np.abs(br_ref).max(), 1e-3)
