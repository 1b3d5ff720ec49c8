"""Golden vectors for the local-POD PROM from the UNMODIFIED reference and its shipped artefacts.

Run in the build container only (needs /root/reference, sklearn + joblib to unpickle the reference's k-means model):
    python tests/golden/make_local_golden.py
Writes tests/golden/local_prom.npz:
  * the artefacts the reference's driver loads (Local_POD/local_pod_galerkin_prom_burgers.py:46-53) for 10 clusters:
    k-means centres (10, 160), the ten local bases (512, 20..30) and U_global[:, :160];
  * FEMBurgers.local_prom_burgers(At = 0.05, u0 = 1, mu = (4.56, 0.019) as the driver, and (5.19, 0.026)) run HERE
    for 10 time steps, LSPG and Galerkin: the states, the per-step iteration counts parsed from the reference's
    stdout, and the clusters kmeans.predict chose;
  * Ar, br of single iterations computed with the reference's own operators (compute_* methods, :730-762) for a shared
    and for a per-sample tangent.
Nothing here comes from the oracle or the CUDA path."""
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from _refharness import import_reference, run_counted, mesh512, REF_ROOT  # noqa: E402

warnings.filterwarnings("ignore")
ref = import_reference()
import joblib  # noqa: E402

CL = os.path.join(REF_ROOT, "Local_POD", "clusters")
K = 10
G = 160
NT = 10
POINTS = [(4.56, 0.019), (5.19, 0.026)]


def main():
    X, T = mesh512()
    fem = ref.FEMBurgers(X, T)
    kmeans = joblib.load(os.path.join(CL, f"kmeans_model_{K}_clusters.pkl"))
    local_bases = np.load(os.path.join(CL, f"local_bases_overlap_{K}_clusters.npy"), allow_pickle=True).item()
    U_global = np.load(os.path.join(CL, "U_global_modes_tol_1e-05.npy"))
    out = {"centers": np.asarray(kmeans.cluster_centers_, dtype=np.float64), "U_global": U_global[:, :G].copy(),
           "points": np.array(POINTS), "nT": NT, "g": G}
    for c in range(K):
        out[f"basis_{c}"] = np.asarray(local_bases[c], dtype=np.float64)
    for pi, (mu1, mu2) in enumerate(POINTS):
        for proj in ("LSPG", "Galerkin"):
            U, counts = run_counted(fem.local_prom_burgers, 0.05, NT, np.ones_like(X), mu1, 0.0, mu2, kmeans, local_bases,
                                    U_global, G, projection=proj)
            cl = np.array([int(kmeans.predict((U_global[:, :G].T @ U[:, n]).reshape(1, -1))[0]) for n in range(NT)])
            out[f"U_{pi}_{proj}"] = U
            out[f"it_{pi}_{proj}"] = counts
            out[f"cl_{pi}_{proj}"] = cl
            print(pi, proj, counts, cl)
    # a start shortly before the cluster of the stored FOM trajectory (FEM/fem_testing_data, mu = (4.56, 0.019)) changes,
    # so that the run switches basis mid-way
    fom = np.load(os.path.join(REF_ROOT, "FEM", "fem_testing_data", "fem_simulation_mu1_4.560_mu2_0.0190.npy"))
    ids = kmeans.predict((U_global[:, :G].T @ fom).T)
    switches = np.flatnonzero(ids[1:] != ids[:-1]) + 1
    print("cluster along the stored trajectory switches at steps", switches[:12], "ids", ids[switches[:12]])
    starts = [int(switches[1]) - 3, int(switches[4]) - 2]
    out["switch_starts"] = np.array(starts)
    for si, n0 in enumerate(starts):
        for proj in ("LSPG", "Galerkin"):
            U, counts = run_counted(fem.local_prom_burgers, 0.05, 8, fom[:, n0].copy(), 4.56, 0.0, 0.019, kmeans, local_bases,
                                    U_global, G, projection=proj)
            cl = np.array([int(kmeans.predict((U_global[:, :G].T @ U[:, n]).reshape(1, -1))[0]) for n in range(8)])
            out[f"sw_u0_{si}"] = fom[:, n0].copy()
            out[f"sw_U_{si}_{proj}"] = U; out[f"sw_it_{si}_{proj}"] = counts; out[f"sw_cl_{si}_{proj}"] = cl
            print("switch run", si, n0, proj, counts, cl)
    # single-iteration reduced systems with the reference's operators
    rng = np.random.default_rng(3)
    M = fem.compute_mass_matrix(); Kd = fem.compute_diffusion_matrix()
    Uks, Uns, Ars, brs = [], [], {"LSPG": [], "Galerkin": []}, {"LSPG": [], "Galerkin": []}
    tang = [local_bases[2], local_bases[5], local_bases[0]]
    mus = [(4.56, 0.019), (5.19, 0.026), (4.3, 0.016)]
    for (mu1, mu2), Phi in zip(mus, tang):
        Un = out["U_0_LSPG"][:, 6] * (1.0 + 0.01 * rng.random(512))
        Uk = Un * (1.0 + 0.02 * rng.random(512))
        At, E = 0.05, 0.01
        C = fem.compute_convection_matrix(Uk); S = fem.compute_supg_term(Uk, mu2); F = fem.compute_forcing_vector(mu2)
        A = (M + At * C + At * E * Kd).tolil(); A[0, :] = 0.0; A[0, 0] = 1.0; A = A.tocsc()
        b = M @ Un + At * F - At * S; b[0] = mu1
        R = A @ Uk - b
        Uks.append(Uk); Uns.append(Un)
        Ars["Galerkin"].append(Phi.T @ A @ Phi); brs["Galerkin"].append(Phi.T @ R)
        J = A @ Phi
        Ars["LSPG"].append(J.T @ J); brs["LSPG"].append(J.T @ R)
    out["rs_Uk"] = np.array(Uks); out["rs_Un"] = np.array(Uns); out["rs_mu"] = np.array(mus)
    out["rs_basis_ids"] = np.array([2, 5, 0]); out["rs_At"] = 0.05; out["rs_E"] = 0.01
    for proj in ("LSPG", "Galerkin"):
        for j in range(3):
            out[f"rs_Ar_{proj}_{j}"] = np.asarray(Ars[proj][j]); out[f"rs_br_{proj}_{j}"] = np.asarray(brs[proj][j]).ravel()
    np.savez_compressed(os.path.join(HERE, "local_prom.npz"), **out)
    print("wrote local_prom.npz", os.path.getsize(os.path.join(HERE, "local_prom.npz")))


if __name__ == "__main__":
    main()
