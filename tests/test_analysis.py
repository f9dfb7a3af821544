"""oracle/analysis_oracle.py (the numpy checker of sr459.analysis) against the reference notebook's own functions.

Golden vectors: tests/golden/g_analysis.npz, produced by tests/golden/make_golden_analysis.py, which exec's the
function cells of /root/reference/mono_cal_target/analysis.ipynb unmodified.  Bars: ESF / LSF / MTF arrays to
1e-9 relative (same float64 arithmetic, scipy's two filters restated in numpy), bin structure and edge angle exact
to rounding, MTF50 / MTF10 to 1e-9.  When /root/reference and scipy are present the notebook code is also run live.
"""
import json
import os

import numpy as np
import pytest

from conftest import golden

from importlib import import_module

A = import_module("oracle.analysis_oracle")


def _rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    return float(np.max(np.abs(a - b)) / max(1e-300, float(np.max(np.abs(b)))))


def test_filters_match_scipy_when_available():
    ndi = pytest.importorskip("scipy.ndimage")
    rng = np.random.default_rng(3)
    a = rng.uniform(0, 255, (37, 53))
    assert _rel(A.gaussian_filter(a, 1.5), ndi.gaussian_filter(a, 1.5)) < 1e-13
    for ax in (0, 1):
        assert _rel(A.sobel(a, ax), ndi.sobel(a, axis=ax)) < 1e-13


def test_slanted_edge_pipeline_matches_notebook_goldens():
    g = golden("g_analysis")
    for i in range(int(g["n_cases"])):
        roi = g[f"roi{i}"].astype(np.float64)
        for side in ("left", "right"):
            ex, ey, ang = A.slanted_edge_esf(roi, side=side)
            assert ex.shape == g[f"esfx{i}{side}"].shape                      # same bins
            assert _rel(ex, g[f"esfx{i}{side}"]) < 1e-9 and _rel(ey, g[f"esfy{i}{side}"]) < 1e-9
            assert abs(ang - float(g[f"ang{i}{side}"])) < 1e-9
            f, m, l = A.esf_to_mtf(ex, ey)
            assert _rel(f, g[f"freq{i}{side}"]) < 1e-9 and _rel(m, g[f"mtf{i}{side}"]) < 1e-9 and _rel(l, g[f"lsf{i}{side}"]) < 1e-9
            v = f > 0
            for frac, key in ((0.5, "m50_"), (0.1, "m10_")):
                got, ref = A.mtf_at_fraction(f[v], m[v], frac), float(g[f"{key}{i}{side}"])
                assert (np.isnan(got) and np.isnan(ref)) or abs(got - ref) <= 1e-9 * max(1.0, abs(ref))


def test_local_contrast_matches_notebook_goldens():
    g = golden("g_analysis")
    assert _rel(A.local_contrast(g["profile"], window=16), g["contrast16"]) < 1e-12
    assert _rel(A.local_contrast(g["profile"]), g["contrast20"]) < 1e-12
    assert np.all(A.local_contrast(np.ones(5), window=20) == 0)


def test_mtf_at_fraction_edge_cases():
    f = np.linspace(0.01, 0.5, 50)
    assert np.isnan(A.mtf_at_fraction(f, np.ones(50), 0.5))                  # never falls
    assert np.isnan(A.mtf_at_fraction(f, np.zeros(50), 0.5))                 # never above
    m = np.linspace(1.0, 0.0, 50)
    assert abs(A.mtf_at_fraction(f, m, 0.5) - np.interp(0.5, m[::-1], f[::-1])) < 1e-12
    with pytest.raises(RuntimeError):
        A.slanted_edge_esf(np.zeros((8, 8)))


def test_sub_tolerance_differences_do_not_move_mtf50():
    """The SR parity tolerance (0.0255 DN) moves MTF50 / MTF10 by far less than their own resolution."""
    g = golden("g_analysis")
    roi = g["roi0"].astype(np.float64)
    rng = np.random.default_rng(9)
    full = np.full((300, 300), 200.0); full[80:80 + roi.shape[0], 100:100 + roi.shape[1]] = roi
    sl = (slice(80, 80 + roi.shape[0]), slice(100, 100 + roi.shape[1]))
    a50, a10 = A.mtf50_mtf10(full, sl)
    b50, b10 = A.mtf50_mtf10(full + rng.uniform(-0.0255, 0.0255, full.shape), sl)
    assert abs(a50 - b50) < 1e-3 * a50 and abs(a10 - b10) < 1e-3 * a10


@pytest.mark.skipif(not os.path.exists("/root/reference/mono_cal_target/analysis.ipynb"), reason="reference not present")
def test_live_notebook_code_on_fresh_input():
    pytest.importorskip("scipy.ndimage")
    nb = json.load(open("/root/reference/mono_cal_target/analysis.ipynb"))
    ns = {"np": np}
    exec("".join(nb["cells"][7]["source"]), ns)
    yy, xx = np.mgrid[:150, :150].astype(np.float64)
    d = (xx - 75) * np.cos(0.3) + (yy - 75) * np.sin(0.3)
    roi = np.where(np.abs(d) < 5, 40.0, 210.0) + np.random.default_rng(4).normal(0, 1.5, (150, 150))
    roi = ns["gaussian_filter"](roi, 1.0) if "gaussian_filter" in ns else roi
    for side in ("left", "right"):
        rx, ry, ra = ns["slanted_edge_esf"](roi, side=side)
        ex, ey, ea = A.slanted_edge_esf(roi, side=side)
        assert _rel(ex, rx) < 1e-9 and _rel(ey, ry) < 1e-9 and abs(ea - ra) < 1e-9
        assert _rel(A.esf_to_mtf(ex, ey)[1], ns["esf_to_mtf"](rx, ry)[1]) < 1e-9
