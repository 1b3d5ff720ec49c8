"""Import shim: exposes the package directory `1d-burgers-equation-roms_b200/` (not a valid Python
identifier) as the module `burgers_b200`.  `import burgers_b200` / `from burgers_b200 import FEMBurgers`."""
import importlib.util
import os
import sys

_PKG_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "1d-burgers-equation-roms_b200")
_spec = importlib.util.spec_from_file_location(
    "burgers_b200", os.path.join(_PKG_DIR, "__init__.py"), submodule_search_locations=[_PKG_DIR])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["burgers_b200"] = _mod  # replace this shim by the real package before executing it
_spec.loader.exec_module(_mod)
