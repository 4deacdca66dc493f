"""Import shim: the product package lives in the directory ``uuo-mocap_b200/`` (the name the
build contract fixes), which is not a valid Python identifier.  ``import uuo_mocap_b200``
executes this file, which loads that directory as the package ``uuo_mocap_b200`` and
replaces itself in ``sys.modules``."""
import importlib.util as _ilu
import os as _os
import sys as _sys

_pkg_dir = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "uuo-mocap_b200")
_spec = _ilu.spec_from_file_location(
    "uuo_mocap_b200", _os.path.join(_pkg_dir, "__init__.py"), submodule_search_locations=[_pkg_dir])
_mod = _ilu.module_from_spec(_spec)
_sys.modules["uuo_mocap_b200"] = _mod
_spec.loader.exec_module(_mod)
