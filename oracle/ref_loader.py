"""Import the UNMODIFIED reference scripts from /root/reference  --  TEST INFRASTRUCTURE ONLY.

Only usable in the build container (the GPU box has no /root/reference).  Used
by tests/golden/make_golden.py to generate committed fixtures and by the
``not gpu`` pinning tests to compare oracle/sr_oracle.py with the live
reference.  Nothing in the product, the ``-m gpu`` tests, smoke() or bench.py
reads /root/reference at run time.

The reference scripts ``import matplotlib`` at module top
(mono_cal_target/run_sr.py:39-41) and matplotlib is not installed here, so two
stub modules are placed in sys.modules first (SURVEY.md section 8c recipe).
"""
from __future__ import annotations

import contextlib
import importlib.util
import io
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("SR459_REFERENCE_ROOT", "/root/reference")
EXPERIMENTS = ("mono_cal_target", "rgb_cal_target", "mono_barcodes", "rgb_barcodes")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "mono_cal_target", "run_sr.py"))


def _stub_matplotlib() -> None:
    if "matplotlib" in sys.modules:
        return
    mpl = types.ModuleType("matplotlib")
    mpl.use = lambda *a, **k: None
    plt = types.ModuleType("matplotlib.pyplot")
    mpl.pyplot = plt
    sys.modules["matplotlib"] = mpl
    sys.modules["matplotlib.pyplot"] = plt


_cache = {}


def load(experiment: str = "mono_cal_target"):
    """Return the reference module ``<experiment>/run_sr.py`` (cached)."""
    if experiment not in EXPERIMENTS:
        raise KeyError(experiment)
    if experiment in _cache:
        return _cache[experiment]
    if not available():
        raise FileNotFoundError(f"reference tree not found at {REFERENCE_ROOT}")
    _stub_matplotlib()
    path = os.path.join(REFERENCE_ROOT, experiment, "run_sr.py")
    spec = importlib.util.spec_from_file_location(f"_ref_{experiment}", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    _cache[experiment] = mod
    return mod


@contextlib.contextmanager
def quiet():
    """The reference prints progress lines; swallow them."""
    with contextlib.redirect_stdout(io.StringIO()):
        yield
