"""Golden stdout of the UNMODIFIED reference (FEM/fem_burgers.py:659, 664, 722, 727, 1127, 1166) for the
`verbose=True` mode of the drop-in class: 3 time steps of fom_burgers, pod_prom_burgers (LSPG, 28 modes) and
pod_quadratic_manifold (LSPG, stored 21-mode manifold) at mu = (4.75, 0.02).  Run in the build container:
    python tests/golden/make_verbose_golden.py
Writes tests/golden/verbose_{fom,prom,qm}.txt."""
import contextlib
import io
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from _refharness import import_reference, mesh512  # noqa: E402


def main():
    ref = import_reference()
    X, T = mesh512()
    fem = ref.FEMBurgers(X, T)
    u0 = np.ones(512)
    Phi28 = np.ascontiguousarray(np.load(os.path.join(HERE, "Phi_tol_1e-04.npy"))[:, :28])
    Phi21 = np.load(os.path.join(HERE, "qm_Phi.npy")); H21 = np.load(os.path.join(HERE, "qm_H.npy"))
    runs = {
        "fom": lambda: fem.fom_burgers(0.05, 3, u0, 4.75, 0.0, 0.02),
        "prom": lambda: fem.pod_prom_burgers(0.05, 3, u0, 4.75, 0.0, 0.02, Phi28, "LSPG"),
        "qm": lambda: fem.pod_quadratic_manifold(0.05, 3, u0, 4.75, 0.0, 0.02, Phi21, H21, "LSPG"),
    }
    for name, fn in runs.items():
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            fn()
        with open(os.path.join(HERE, f"verbose_{name}.txt"), "w") as f:
            f.write(buf.getvalue())
        print(name, len(buf.getvalue().splitlines()), "lines")


if __name__ == "__main__":
    main()
