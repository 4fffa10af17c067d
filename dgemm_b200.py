"""`import dgemm_b200` -> the package in ./optimizing-dgemm-on-intel-cpus-with-avx512f_b200/ (its directory name
has hyphens, so it cannot be named in an import statement)."""
import importlib.util
import os
import sys

_pkg_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "optimizing-dgemm-on-intel-cpus-with-avx512f_b200")
_spec = importlib.util.spec_from_file_location("dgemm_b200", os.path.join(_pkg_dir, "__init__.py"),
                                               submodule_search_locations=[_pkg_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["dgemm_b200"] = _mod
_spec.loader.exec_module(_mod)
