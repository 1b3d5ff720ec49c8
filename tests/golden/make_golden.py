"""Generate the golden vectors under tests/golden/ from the UNMODIFIED reference.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py [section ...]
Sections: ops fom_short newton_short prom_short qm_short fom_fixtures prom_fixtures qm_fixtures pod qm_build
`fom_fixtures` additionally needs the full 500-step reference runs produced by
make_ref_full_fom.py (per-step iteration counts parsed from the reference's stdout); point
BURGERS_REF_FULL at the directory holding fom_<mu1>_<mu2>.npz (default /tmp/reffull).

Everything written here is OUTPUT OF THE REFERENCE ITSELF (run here) or a time-subsample of
the .npy fixtures the reference ships; nothing comes from the oracle or the CUDA path.
"""
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from _refharness import import_reference, run_counted, mesh512, REF_ROOT  # noqa: E402

warnings.filterwarnings("ignore")
ref = import_reference()

FOM_POINTS = [(4.25, 0.015), (4.25, 0.0225), (4.25, 0.03), (4.875, 0.015), (4.875, 0.0225),
              (4.875, 0.03), (5.5, 0.015), (5.5, 0.0225), (5.5, 0.03),
              (4.56, 0.019), (4.75, 0.02), (5.19, 0.026), (5.65, 0.033), (6.2, 0.04)]
KEEP_STEPS = np.array([0, 1, 2, 3, 5, 10, 20, 50, 100, 150, 200, 250, 300, 350, 400, 450, 500])


def bands_of(A):
    A = np.asarray(A.todense()) if hasattr(A, "todense") else np.asarray(A)
    N = A.shape[0]
    off = A.copy()
    for k in (-1, 0, 1):
        off -= np.diag(np.diag(A, k), k)
    assert np.abs(off).max() == 0.0, "reference operator is not tridiagonal"
    dl = np.zeros(N); du = np.zeros(N)
    dl[1:] = np.diag(A, -1); du[:-1] = np.diag(A, 1)
    return np.stack([dl, np.diag(A).copy(), du])


def fixture_path(mu1, mu2):
    for d in ("fem_training_data", "fem_testing_data"):
        p = os.path.join(REF_ROOT, "FEM", d, f"fem_simulation_mu1_{mu1:.3f}_mu2_{mu2:.4f}.npy")
        if os.path.exists(p):
            return p
    raise FileNotFoundError((mu1, mu2))


def nonuniform_mesh(n=37, seed=3):
    rng = np.random.default_rng(seed)
    w = 0.5 + rng.random(n - 1)
    X = np.concatenate([[0.0], np.cumsum(w)]) * (10.0 / w.sum())
    T = np.array([np.arange(1, n), np.arange(2, n + 1)]).T
    return X, T


def sec_ops():
    out = {}
    X, T = mesh512()
    shock = np.load(fixture_path(4.75, 0.02))[:, 300]
    rng = np.random.default_rng(0)
    cases = {"u512_rand": (X, T, 1.0 + 4.0 * rng.random(X.size), 0.02),
             "u512_shock": (X, T, shock, 0.02),
             "u512_signed": (X, T, rng.standard_normal(X.size), 0.033)}
    Xn, Tn = nonuniform_mesh()
    cases["u37_nonuniform"] = (Xn, Tn, 1.0 + 4.0 * rng.random(Xn.size), 0.04)
    # an element whose mean velocity is below the SUPG eps (FEM/fem_burgers.py:541)
    uz = 1.0 + rng.random(Xn.size); uz[5] = 1e-12; uz[6] = -1e-12
    cases["u37_zero_velocity"] = (Xn, Tn, uz, 0.015)
    for name, (Xc, Tc, U, mu2) in cases.items():
        fb = ref.FEMBurgers(Xc, Tc)
        out[name + "_X"] = Xc
        out[name + "_U"] = U
        out[name + "_mu2"] = mu2
        out[name + "_M"] = bands_of(fb.compute_mass_matrix())
        out[name + "_K"] = bands_of(fb.compute_diffusion_matrix())
        out[name + "_C"] = bands_of(fb.compute_convection_matrix(U))
        out[name + "_dC"] = bands_of(fb.compute_convection_matrix_derivative(U))
        out[name + "_F"] = fb.compute_forcing_vector(mu2)
        out[name + "_S"] = fb.compute_supg_term(U, mu2)
    np.savez_compressed(os.path.join(HERE, "ops.npz"), **out)


def sec_fom_short():
    out = {}
    X, T = mesh512()
    fb = ref.FEMBurgers(X, T)
    for mu1, mu2 in [(4.75, 0.02), (4.25, 0.015), (6.2, 0.04)]:
        U, it = run_counted(fb.fom_burgers, 0.05, 12, np.ones_like(X), mu1, 0.0, mu2)
        out[f"U_{mu1}_{mu2}"] = U
        out[f"it_{mu1}_{mu2}"] = it
    # non-uniform mesh, non-constant initial condition, diffusion switched on
    Xn, Tn = nonuniform_mesh()
    fbn = ref.FEMBurgers(Xn, Tn)
    u0 = 1.0 + 0.5 * np.sin(Xn)
    U, it = run_counted(fbn.fom_burgers, 0.02, 8, u0, 2.0, 0.0, 0.04)
    out["nonuni_X"] = Xn; out["nonuni_u0"] = u0; out["nonuni_U"] = U; out["nonuni_it"] = it
    U, it = run_counted(fbn.fom_burgers, 0.02, 8, u0, 2.0, 0.01, 0.04)
    out["nonuni_E_U"] = U; out["nonuni_E_it"] = it
    # small uniform mesh (N = 64 nodes on [0, 100]): used by tiny GPU smoke tests
    m = 63
    Xs = np.linspace(0, 100, m + 1); Ts = np.array([np.arange(1, m + 1), np.arange(2, m + 2)]).T
    U, it = run_counted(ref.FEMBurgers(Xs, Ts).fom_burgers, 0.05, 10, np.ones_like(Xs), 4.75, 0.0, 0.02)
    out["n64_X"] = Xs; out["n64_U"] = U; out["n64_it"] = it
    np.savez_compressed(os.path.join(HERE, "fom_short.npz"), **out)


def sec_newton_short():
    out = {}
    X, T = mesh512()
    fb = ref.FEMBurgers(X, T)
    U, it = run_counted(fb.fom_burgers_newton, 0.05, 2, np.ones_like(X), 4.75, 0.0, 0.02)
    out["U_512"] = U; out["it_512"] = it
    m = 63
    Xs = np.linspace(0, 100, m + 1); Ts = np.array([np.arange(1, m + 1), np.arange(2, m + 2)]).T
    U, it = run_counted(ref.FEMBurgers(Xs, Ts).fom_burgers_newton, 0.05, 3, np.ones_like(Xs), 4.75, 0.0, 0.02)
    out["X_64"] = Xs; out["U_64"] = U; out["it_64"] = it
    np.savez_compressed(os.path.join(HERE, "newton_short.npz"), **out)


def load_phi(tol):
    return np.load(os.path.join(REF_ROOT, "POD", "modes", f"U_modes_tol_{tol}.npy"))


def sec_prom_short():
    out = {}
    X, T = mesh512()
    fb = ref.FEMBurgers(X, T)
    Phi40 = load_phi("1e-03")
    for n in (28, 40):
        Phi = np.ascontiguousarray(Phi40[:, :n])
        for proj in ("LSPG", "Galerkin"):
            U, it = run_counted(fb.pod_prom_burgers, 0.05, 8, np.ones_like(X), 4.75, 0.0, 0.02, Phi, proj)
            out[f"U_{n}_{proj}"] = U; out[f"it_{n}_{proj}"] = it
    Phi9 = load_phi("1e-02")
    U, it = run_counted(fb.pod_prom_burgers, 0.05, 8, np.ones_like(X), 5.19, 0.0, 0.026, Phi9, "LSPG")
    out["U_9_LSPG"] = U; out["it_9_LSPG"] = it
    np.savez_compressed(os.path.join(HERE, "prom_short.npz"), **out)


def sec_qm_short():
    out = {}
    X, T = mesh512()
    fb = ref.FEMBurgers(X, T)
    Phi = np.load(os.path.join(REF_ROOT, "Quadratic_manifold", "Phi.npy"))
    H = np.load(os.path.join(REF_ROOT, "Quadratic_manifold", "H.npy"))
    np.save(os.path.join(HERE, "qm_Phi.npy"), Phi)
    np.save(os.path.join(HERE, "qm_H.npy"), H)
    for proj in ("LSPG", "Galerkin"):
        U, it = run_counted(fb.pod_quadratic_manifold, 0.05, 6, np.ones_like(X), 4.75, 0.0, 0.02, Phi, H, proj)
        out[f"U_{proj}"] = U; out[f"it_{proj}"] = it
    q = np.linspace(-1.0, 2.0, 5)
    out["q5"] = q
    out["sym5"] = ref.get_sym(q)
    out["dQ5"] = ref.get_dQ_dq(5, q)
    np.savez_compressed(os.path.join(HERE, "qm_short.npz"), **out)


def sec_fom_fixtures():
    full_dir = os.environ.get("BURGERS_REF_FULL", "/tmp/reffull")
    out = {"steps": KEEP_STEPS, "points": np.array(FOM_POINTS)}
    for i, (mu1, mu2) in enumerate(FOM_POINTS):
        fix = np.load(fixture_path(mu1, mu2))
        assert fix.shape == (512, 501)
        out[f"U_{i}"] = fix[:, KEEP_STEPS]
        p = os.path.join(full_dir, f"fom_{mu1}_{mu2}.npz")
        if os.path.exists(p):
            z = np.load(p)
            assert z["iters"].size == 500
            out[f"it_{i}"] = z["iters"].astype(np.int8)
            out[f"refrun_rel_{i}"] = z["rel_vs_fixture"]
            out[f"refrun_wall_{i}"] = z["wall_s"]
        else:
            print("WARNING: no full reference run for", mu1, mu2)
    np.savez_compressed(os.path.join(HERE, "fom_fixtures.npz"), **out)


def sec_prom_fixtures():
    steps = np.array([0, 1, 2, 5, 10, 50, 100, 200, 300, 400, 500])
    out = {"steps": steps}
    d = os.path.join(REF_ROOT, "POD", "Results_thesis", "rom_solutions")
    for tol in ("1e-02", "4e-03", "1e-03", "1e-04"):
        Phi = load_phi(tol)
        np.save(os.path.join(HERE, f"Phi_tol_{tol}.npy"), Phi)
        for f in sorted(os.listdir(d)):
            if f.startswith(f"U_PROM_tol_{tol}_") and f.endswith(".npy"):
                out[f[:-4]] = np.load(os.path.join(d, f))[:, steps]
    np.savez_compressed(os.path.join(HERE, "prom_fixtures.npz"), **out)


def sec_qm_fixtures():
    steps = np.array([0, 1, 2, 5, 10, 50, 100, 200, 300, 400, 500])
    out = {"steps": steps}
    d = os.path.join(REF_ROOT, "Quadratic_manifold", "quadratic_rom_solutions")
    for f in sorted(os.listdir(d)):
        if f.endswith(".npy"):
            out[f[:-4]] = np.load(os.path.join(d, f))[:, steps]
    np.savez_compressed(os.path.join(HERE, "qm_fixtures.npz"), **out)


def sec_pod():
    """Stored POD output of POD/pod.py on the 9 training trajectories."""
    s = np.load(os.path.join(REF_ROOT, "POD", "modes", "Singular_values_modes_tol_0e+00.npy"))
    out = {"s_full": s}
    # pod.py's own mode-count rule evaluated by the reference function on the stored spectrum.
    # (POD/pod.py cannot be imported without matplotlib's rc(); restate the 6-line function
    #  is NOT what we do here: exec the function source straight from the reference file.)
    src = open(os.path.join(REF_ROOT, "POD", "pod.py")).read()
    start = src.index("def get_number_of_singular_values_for_given_tolerance")
    end = src.index("def get_list_of_number_of_singular_values")
    ns = {"np": np}
    exec(src[start:end], ns)
    f = ns["get_number_of_singular_values_for_given_tolerance"]
    tols = np.array([1e-2, 4e-3, 3e-3, 1e-3, 1e-4, 1e-5, 1e-6])
    out["tols"] = tols
    out["counts"] = np.array([f(512, 4509, s, t) for t in tols])
    out["stored_counts"] = np.array([load_phi(t).shape[1] for t in
                                     ("1e-02", "4e-03", "3e-03", "1e-03", "1e-04", "1e-05", "1e-06")])
    np.savez_compressed(os.path.join(HERE, "pod.npz"), **out)


def sec_qm_build():
    """Quadratic_manifold/quad_utils.py run on a small seeded case (the builder's three functions)."""
    sys.path.insert(0, os.path.join(REF_ROOT, "Quadratic_manifold"))
    import quad_utils as qu
    rng = np.random.default_rng(5)
    n, Ns, N = 5, 60, 14
    q = rng.standard_normal((n, Ns)); S = rng.standard_normal((N, Ns))
    Phi = np.linalg.qr(rng.standard_normal((N, n)))[0]
    Q = qu.build_Q(q); E = qu.build_E(S, Phi, Phi.T @ S, None)
    H = qu.compute_H(qu.build_Q(Phi.T @ S), E, 1e-2)
    np.savez(os.path.join(HERE, "qm_build.npz"), q=q, S=S, Phi=Phi, Q=Q, E=E, H=H, alpha=1e-2)


SECTIONS = {k[4:]: v for k, v in list(globals().items()) if k.startswith("sec_")}

if __name__ == "__main__":
    todo = sys.argv[1:] or list(SECTIONS)
    for s in todo:
        print("section", s, flush=True)
        SECTIONS[s]()
