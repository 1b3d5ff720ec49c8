"""Run the UNMODIFIED reference fom_burgers (FEM/fem_burgers.py:646) for one (mu1, mu2) over the
full 500 steps in this container and record: per-step Picard iteration counts (parsed from its
stdout) and the agreement with the stored fixture.  ~6 CPU-minutes per point.
usage: python make_ref_full_fom.py mu1 mu2 out.npz
"""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from _refharness import import_reference, run_counted, mesh512, REF_ROOT

mu1, mu2, out = float(sys.argv[1]), float(sys.argv[2]), sys.argv[3]
nsteps = int(sys.argv[4]) if len(sys.argv) > 4 else 500
ref = import_reference()
X, T = mesh512()
fb = ref.FEMBurgers(X, T)
t0 = time.time()
U, iters = run_counted(fb.fom_burgers, 0.05, nsteps, np.ones_like(X), mu1, 0.0, mu2)
wall = time.time() - t0
fix = None
for d in ("fem_training_data", "fem_testing_data"):
    p = os.path.join(REF_ROOT, "FEM", d, f"fem_simulation_mu1_{mu1:.3f}_mu2_{mu2:.4f}.npy")
    if os.path.exists(p):
        fix = np.load(p)
rel = -1.0
if fix is not None:
    rel = float(np.max(np.linalg.norm(U - fix[:, :nsteps + 1], axis=0) / np.linalg.norm(fix[:, :nsteps + 1], axis=0)))
np.savez_compressed(out, U=U, iters=iters, wall_s=wall, rel_vs_fixture=rel, mu1=mu1, mu2=mu2)
sys.stderr.write(f"mu=({mu1},{mu2}) wall={wall:.1f}s total_iters={iters.sum()} rel_vs_fixture={rel:.3e}\n")
