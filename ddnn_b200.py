"""Import shim: exposes the package directory ``docker-dist-nn_b200/`` (not a valid Python
identifier) as the module ``ddnn_b200``.  ``import ddnn_b200`` / ``from ddnn_b200 import Layer``.
"""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "docker-dist-nn_b200")
_spec = importlib.util.spec_from_file_location("ddnn_b200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["ddnn_b200"] = _mod
_spec.loader.exec_module(_mod)
