"""oracle/calibration_oracle.py (the numpy checker of sr459.calibration) against the reference's data_collection/psf_mtf_utils.py.

Golden vectors: tests/golden/g_calibration.npz from tests/golden/make_golden_calibration.py (the reference module
imported unmodified).  Bars: integer decisions (peak pixel, ROI shape, crop) exact; deterministic float64 steps
(background-subtracted ROI, centre of mass, radial profile, MTF) to 1e-9; the bounded Gaussian fit -- scipy's
curve_fit there, a projected Levenberg-Marquardt here -- to 1e-4 px in the centre, 1e-4 relative in the widths and
1e-4 of the peak in the fitted image (both minimise the same sum of squares; theta is compared modulo the
symmetries of the model and only where the spot is elliptical).
"""
from importlib import import_module

import numpy as np
import pytest

from conftest import golden

Cal = import_module("oracle.calibration_oracle")


def _rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    return float(np.max(np.abs(a - b)) / max(1e-300, float(np.max(np.abs(b)))))


@pytest.fixture(scope="module")
def g():
    return golden("g_calibration")


def test_peak_roi_and_centre_of_mass(g):
    for i in range(int(g["n_cases"])):
        img = g[f"img{i}"].astype(np.float64)
        pk = Cal.find_peak(img)
        assert tuple(int(v) for v in pk) == tuple(int(v) for v in g[f"peak{i}"])
        roi = Cal.extract_psf(img, pk, int(g["radius"]))
        assert roi.shape == g[f"roi{i}"].shape and _rel(roi, g[f"roi{i}"]) < 1e-12
        assert np.allclose(Cal.subpixel_centre(roi), g[f"com{i}"], rtol=0, atol=1e-10)
        assert np.array_equal(Cal.zoom_crop(img, pk, 6), g[f"zoom{i}"])
        assert img[tuple(int(v) for v in pk)] > 0 and np.array_equal(img, g[f"img{i}"].astype(np.float64))   # input untouched


def test_radial_average_and_mtf(g):
    for i in range(int(g["n_cases"])):
        roi = g[f"roi{i}"]
        rad, prof = Cal.radial_average(roi)
        assert np.array_equal(rad, g[f"rad{i}"]) and _rel(prof, g[f"radprof{i}"]) < 1e-12
        f, m, m2, label, nyq = Cal.compute_mtf(roi, 3.45)
        assert label == "cycles/mm" and abs(nyq - float(g[f"nyq{i}"])) < 1e-12
        assert _rel(f, g[f"mtf_f{i}"]) < 1e-12 and _rel(m, g[f"mtf_r{i}"]) < 1e-9
        got, ref = Cal.mtf_at_fraction(f, m, 0.5), float(g[f"mtf50_{i}"])
        assert (np.isnan(got) and np.isnan(ref)) or abs(got - ref) < 1e-7 * abs(ref)
        assert Cal.compute_mtf(roi)[3] == "cycles/pixel"


def test_gaussian_fit_and_pinhole_centre(g):
    for i in range(int(g["n_cases"])):
        roi, ref = g[f"roi{i}"], g[f"popt{i}"]
        popt, fit = Cal.fit_gaussian_psf(roi)
        assert popt is not None
        assert abs(popt[1] - ref[1]) < 1e-4 and abs(popt[2] - ref[2]) < 1e-4                     # centre (px)
        assert abs(popt[0] - ref[0]) < 1e-4 * ref[0] and abs(popt[6] - ref[6]) < 1e-4 * ref[0]   # amplitude, offset
        assert np.allclose(sorted(popt[3:5]), sorted(ref[3:5]), rtol=1e-4)                       # widths (axes may swap)
        assert np.max(np.abs(fit - g[f"fit{i}"])) < 1e-4 * ref[0]                                # the fitted surface
        cx, cy = Cal.find_pinhole_center(g[f"img{i}"].astype(np.float64))
        assert abs(cx - g[f"center{i}"][0]) < 1e-4 and abs(cy - g[f"center{i}"][1]) < 1e-4


def test_gauss2d_model_and_fallbacks():
    yy, xx = np.mgrid[:21, :25]
    p = (100.0, 12.3, 9.6, 2.0, 3.0, 0.4, 5.0)
    z = Cal.gauss2d((xx, yy), *p)
    assert z.shape == (21 * 25,) and abs(z.max() - 105.0) < 3.0
    popt, fit = Cal.fit_gaussian_psf(z.reshape(21, 25))
    assert np.max(np.abs(fit.ravel() - z)) < 1e-6                     # a noiseless surface is recovered
    assert Cal.fit_gaussian_psf(np.zeros((9, 9))) == (None, None)     # nothing to fit
    flat = np.zeros((40, 40)); flat[20, 17] = 0.0
    cx, cy = Cal.find_pinhole_center(flat)                            # all-zero image: falls through to the peak pixel
    assert (cx, cy) == (0.0, 0.0)
