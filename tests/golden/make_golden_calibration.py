#!/usr/bin/env python3
"""Golden vectors for sr459.calibration from the reference's own data_collection/psf_mtf_utils.py.

Run in the build container only (needs /root/reference and scipy):

    python tests/golden/make_golden_calibration.py

psf_mtf_utils.py is imported UNMODIFIED (matplotlib stubbed, as for run_sr.py) and applied to synthetic pinhole
images (rotated elliptical Gaussian spots + background + noise, stored as float32).  find_pinhole_center lives
in calibration_beam_shift/calibrate_shift_grid.py, which imports the camera / mirror drivers and OpenCV; its
result on the fit branch is `popt[1] + roi_c0, popt[2] + roi_r0` (:82-85), recorded here from the same calls.
Writes tests/golden/g_calibration.npz.
"""
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))
from oracle import ref_loader as R  # noqa: E402


def load_utils():
    R._stub_matplotlib()
    gs = types.ModuleType("matplotlib.gridspec"); gs.GridSpec = object
    sys.modules["matplotlib.gridspec"] = gs
    path = os.path.join(R.REFERENCE_ROOT, "data_collection", "psf_mtf_utils.py")
    spec = importlib.util.spec_from_file_location("_ref_psf_mtf_utils", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def spot(h, w, cy, cx, sx, sy, theta, amp, bg, noise, rng):
    yy, xx = np.mgrid[:h, :w].astype(np.float64)
    ct, st = np.cos(theta), np.sin(theta)
    u, v = (xx - cx) * ct + (yy - cy) * st, -(xx - cx) * st + (yy - cy) * ct
    img = bg + amp * np.exp(-0.5 * ((u / sx) ** 2 + (v / sy) ** 2)) + rng.normal(0, noise, (h, w))
    return np.clip(img, 0, 255).astype(np.float32)


def main():
    U = load_utils()
    rng = np.random.default_rng(459)
    cases = [(96, 128, 40.3, 70.8, 1.6, 2.4, 0.5, 180, 12, 1.0), (80, 80, 41.7, 37.2, 2.2, 1.3, -0.9, 120, 5, 0.5),
             (64, 96, 9.4, 12.6, 1.8, 1.8, 0.0, 200, 8, 0.8), (120, 100, 60.5, 49.5, 3.0, 2.0, 1.2, 90, 20, 1.5)]
    out = {"n_cases": np.array(len(cases)), "radius": np.array(15)}
    for i, c in enumerate(cases):
        img32 = spot(*c, rng)
        out[f"img{i}"] = img32
        img = img32.astype(np.float64)
        pk = U.find_peak(img)
        roi = U.extract_psf(img, pk, 15)
        popt, fit = U.fit_gaussian_psf(roi)
        out[f"peak{i}"] = np.array(pk)
        out[f"roi{i}"] = roi
        out[f"com{i}"] = np.array(U.subpixel_centre(roi))
        out[f"popt{i}"] = popt
        out[f"fit{i}"] = fit
        out[f"center{i}"] = np.array([popt[1] + max(pk[1] - 15, 0), popt[2] + max(pk[0] - 15, 0)])
        rad, prof = U.radial_average(roi)
        out[f"rad{i}"], out[f"radprof{i}"] = rad, prof
        f, m, m2, lab, nyq = U.compute_mtf(roi, 3.45)
        out[f"mtf_f{i}"], out[f"mtf_r{i}"], out[f"nyq{i}"] = f, m, np.array(nyq)
        out[f"mtf50_{i}"] = np.array(U.mtf_at_fraction(f, m, 0.5))
        out[f"zoom{i}"] = U.zoom_crop(img, pk, 6)
    np.savez_compressed(os.path.join(HERE, "g_calibration.npz"), **out)
    print("wrote g_calibration.npz:", len(out), "arrays")


if __name__ == "__main__":
    main()
