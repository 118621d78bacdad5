"""Import shim: loads the package directory `taylorintegration.jl_b200/` (whose name, inherited from the
reference, contains a dot) under the importable name `taylorintegration_b200`."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "taylorintegration.jl_b200")
_spec = importlib.util.spec_from_file_location("taylorintegration_b200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["taylorintegration_b200"] = _mod
_spec.loader.exec_module(_mod)
