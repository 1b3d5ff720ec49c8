import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    # `-m gpu` tests never run silently on a box without a GPU: they fail loudly inside the
    # product (no CPU fallback).  Nothing to skip here on purpose.
    pass


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(os.path.join(GOLDEN, name), allow_pickle=False)
    return load


@pytest.fixture(scope="session")
def oracle():
    from oracle import burgers_oracle
    return burgers_oracle


def rel_l2_per_step(U, Uref):
    """max over time steps of |U - Uref|_2 / |Uref|_2 ; both (N, steps)."""
    num = np.linalg.norm(U - Uref, axis=0)
    den = np.linalg.norm(Uref, axis=0)
    return float(np.max(num / den))
