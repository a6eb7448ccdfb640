"""Importable alias for the package directory ``kubernetes-machine-learning-server_b200/``
(a hyphenated directory name cannot appear in an ``import`` statement). ``import kmls_b200``
loads that directory as the package ``kmls_b200``."""
import importlib.util as _u
import os as _os
import sys as _sys

_dir = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)),
                     "kubernetes-machine-learning-server_b200")
_spec = _u.spec_from_file_location("kmls_b200", _os.path.join(_dir, "__init__.py"),
                                   submodule_search_locations=[_dir])
_mod = _u.module_from_spec(_spec)
_sys.modules["kmls_b200"] = _mod
_spec.loader.exec_module(_mod)
