"""Import shim: ``import motion_estimation_b200`` loads the package in ``motion-estimation_b200/``.

The package directory is named after the reference repository (hyphen included), which is not an
importable identifier; this module replaces itself in ``sys.modules`` with that package.
"""
import importlib.util as _u
import os as _os
import sys as _sys

_dir = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "motion-estimation_b200")
_spec = _u.spec_from_file_location(__name__, _os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = _u.module_from_spec(_spec)
_sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
