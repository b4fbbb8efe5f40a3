"""Import alias: ``import reidb200`` loads the package directory
``real-time-reid-tracking_b200/`` (whose name is not a valid Python identifier)."""
import importlib.util as _u
import os as _os
import sys as _sys

_dir = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "real-time-reid-tracking_b200")
_spec = _u.spec_from_file_location("reidb200", _os.path.join(_dir, "__init__.py"),
                                   submodule_search_locations=[_dir])
_mod = _u.module_from_spec(_spec)
_sys.modules["reidb200"] = _mod
_spec.loader.exec_module(_mod)
