"""Import alias: ``import sr459`` is ``enph459-super-resolution_b200`` (whose directory
name, fixed by the project layout, is not a valid Python identifier)."""
import importlib as _importlib
import os as _os
import sys as _sys

_root = _os.path.abspath(_os.path.join(_os.path.dirname(__file__), ".."))
if _root not in _sys.path:
    _sys.path.insert(0, _root)
_real = "enph459-super-resolution_b200"
_pkg = _importlib.import_module(_real)
for _sub in ("_lib", "ops", "build", "driver", "bands", "run_sr", "ingest", "analysis", "calibration"):
    try:
        _m = _importlib.import_module(f"{_real}.{_sub}")
    except ModuleNotFoundError:
        continue
    _sys.modules[f"{__name__}.{_sub}"] = _m
    setattr(_pkg, _sub, _m)
_sys.modules[__name__] = _pkg
