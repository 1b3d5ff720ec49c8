"""burgers_b200 — B200-native (sm_100a) hot path of SADPR/1D-Burgers-Equation-ROMs.

The directory name `1d-burgers-equation-roms_b200` is not a Python identifier; import it through
the root-level shim `burgers_b200` (burgers_b200.py), which registers this package under that name.
"""
from . import _lib  # noqa: F401
from ._lib import BurgersB200Error, set_tuning, get_tuning, qm_schedules  # noqa: F401
from .build import build_library, LIB_PATH  # noqa: F401
from .fem_burgers import FEMBurgers, validate_mesh  # noqa: F401
from . import sweep  # noqa: F401
from . import pod  # noqa: F401
from . import quadratic  # noqa: F401

__all__ = ["FEMBurgers", "validate_mesh", "build_library", "LIB_PATH", "BurgersB200Error", "set_tuning",
           "get_tuning", "qm_schedules"]
