"""Importable alias for the package directory `brpc-github-actions_b200/` (its name is
not a valid Python identifier).  `import brpc_b200` gives that package."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "brpc-github-actions_b200")
_spec = importlib.util.spec_from_file_location("brpc_b200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["brpc_b200"] = _mod
_spec.loader.exec_module(_mod)
