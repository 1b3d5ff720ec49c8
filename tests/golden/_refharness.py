"""Harness that imports the UNMODIFIED reference (read-only, /root/reference) in THIS
container to generate golden vectors.  Test infrastructure only: nothing under tests/
(other than the generator scripts in this directory), bench.py or the package imports it,
and it is never used on the GPU box (/root/reference does not exist there).

The reference module does `import matplotlib.pyplot` at import time
(FEM/fem_burgers.py:8); matplotlib is absent here, so stub modules are injected.
Iteration counts are not returned by the reference; they only surface on stdout
("Iteration: k, Error: e" at FEM/fem_burgers.py:664/727, "  Newton it: ..." at :1166,
"Error: e, Iteration: k" at :607), so stdout is replaced by a line counter.
"""
import sys
import types
import os

REF_ROOT = os.environ.get("BURGERS_REFERENCE_ROOT", "/root/reference")


def import_reference():
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.animation"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    fem_dir = os.path.join(REF_ROOT, "FEM")
    if fem_dir not in sys.path:
        sys.path.insert(0, fem_dir)
    import fem_burgers  # noqa: the reference module
    return fem_burgers


class IterationCounter:
    """File-like object: counts nonlinear-iteration lines per time step."""

    def __init__(self, step_prefixes=("Time Step:", "=== time step"),
                 iter_prefixes=("Iteration:", "  Newton", "Error:")):
        self.counts = []
        self._buf = ""
        self.step_prefixes = step_prefixes
        self.iter_prefixes = iter_prefixes

    def write(self, s):
        self._buf += s
        while "\n" in self._buf:
            line, self._buf = self._buf.split("\n", 1)
            if any(line.startswith(p) for p in self.step_prefixes):
                self.counts.append(0)
            elif any(line.startswith(p) for p in self.iter_prefixes) and self.counts:
                self.counts[-1] += 1
        return len(s)

    def flush(self):
        pass


def run_counted(fn, *args, **kwargs):
    """Run a reference method with stdout replaced; return (result, per-step iteration counts)."""
    import numpy as np
    counter = IterationCounter()
    old = sys.stdout
    sys.stdout = counter
    try:
        out = fn(*args, **kwargs)
    finally:
        sys.stdout = old
    return out, np.asarray(counter.counts, dtype=np.int32)


def mesh512():
    import numpy as np
    m = 511
    X = np.linspace(0, 100, m + 1)
    T = np.array([np.arange(1, m + 1), np.arange(2, m + 2)]).T
    return X, T
