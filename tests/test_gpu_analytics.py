"""Device analytics against the reference's own code.

sr459.analysis / sr459.calibration run in libsr459 (csrc/analysis.cu, csrc/calib.cu).  They are compared with
(1) the golden vectors produced by the reference's unmodified notebook cells / psf_mtf_utils module
(tests/golden/g_analysis.npz, g_calibration.npz) and (2) the numpy restatements in oracle/ on fresh inputs.
Bars: integer decisions (peak pixel, ROI shape, bin count, orientation) exact; deterministic float64 chains to 1e-9;
the bounded Gaussian fit (scipy curve_fit upstream, projected Levenberg-Marquardt here) to 1e-4 px / 1e-4 relative.
"""
import numpy as np
import pytest

from conftest import golden
from oracle import analysis_oracle as AO
from oracle import calibration_oracle as CO

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    return float(np.max(np.abs(a - b)) / max(1e-300, float(np.max(np.abs(b)))))


def _same(a, b, tol):
    return (np.isnan(a) and np.isnan(b)) or abs(a - b) <= tol * max(1.0, abs(b))


@pytest.fixture(scope="module")
def A(sr):
    from sr459 import analysis
    return analysis


@pytest.fixture(scope="module")
def Cal(sr):
    from sr459 import calibration
    return calibration


# ------------------------------------------------------------------------------------------------ f3
def test_slanted_edge_chain_matches_notebook_goldens(A):
    g = golden("g_analysis")
    for i in range(int(g["n_cases"])):
        roi = g[f"roi{i}"]                                   # float32 as stored = what the notebook analysed
        for side in ("left", "right"):
            r = A.edge_mtf(roi, None, side, pitch_mm=1.0)
            assert r["esf_x"].shape == g[f"esfx{i}{side}"].shape                  # same bins
            assert _rel(r["esf_x"], g[f"esfx{i}{side}"]) < 1e-9 and _rel(r["esf_y"], g[f"esfy{i}{side}"]) < 1e-9
            assert abs(r["angle"] - float(g[f"ang{i}{side}"])) < 1e-9
            assert _rel(r["lsf"], g[f"lsf{i}{side}"]) < 1e-9
            assert _rel(r["freq"], g[f"freq{i}{side}"]) < 1e-9 and _rel(r["mtf"], g[f"mtf{i}{side}"]) < 1e-9
            assert _same(r["mtf50"], float(g[f"m50_{i}{side}"]), 1e-9) and _same(r["mtf10"], float(g[f"m10_{i}{side}"]), 1e-9)
            ex, ey, ang = A.slanted_edge_esf(roi, side=side)
            assert np.array_equal(ex, r["esf_x"]) and np.array_equal(ey, r["esf_y"]) and ang == r["angle"]


def test_slanted_edge_on_uint8_roi_inside_a_larger_device_image(A, sr):
    import torch
    rng = np.random.default_rng(21)
    n = 180
    yy, xx = np.mgrid[:n, :n].astype(np.float64)
    d = (xx - n / 2) * np.cos(1.35) + (yy - n / 2) * np.sin(1.35)          # a mostly horizontal line: column-as-x branch
    roi = np.clip(AO.gaussian_filter(np.where(np.abs(d) < 6, 40.0, 200.0), 1.3) + rng.normal(0, 1.2, (n, n)), 0, 255).astype(np.uint8)
    full = rng.integers(0, 255, (300, 420), dtype=np.uint8)
    full[70:70 + n, 150:150 + n] = roi
    t = torch.from_numpy(full).cuda()
    sl = (slice(70, 70 + n), slice(150, 150 + n))
    for side in ("left", "right"):
        ex, ey, ang = AO.slanted_edge_esf(roi.astype(np.float64), side=side)
        f, m, lsf = AO.esf_to_mtf(ex, ey)
        r = A.edge_mtf(t, sl, side)
        assert r["esf_x"].shape == ex.shape and _rel(r["esf_x"], ex) < 1e-9 and _rel(r["esf_y"], ey) < 1e-9
        assert abs(r["angle"] - ang) < 1e-9 and _rel(r["lsf"], lsf) < 1e-9 and _rel(r["mtf"], m) < 1e-9
        fm = f / A.HR_PITCH_MM
        v = fm > 0
        assert _same(r["mtf50"], AO.mtf_at_fraction(fm[v], m[v], 0.5), 1e-9) and _same(r["mtf10"], AO.mtf_at_fraction(fm[v], m[v], 0.1), 1e-9)
        again = A.mtf50_mtf10(t, sl, side)
        assert _same(again[0], r["mtf50"], 0.0) and _same(again[1], r["mtf10"], 0.0)


def test_analysis_errors_and_local_contrast(A):
    g = golden("g_analysis")
    assert _rel(A.local_contrast(g["profile"], window=16), g["contrast16"]) < 1e-12
    assert _rel(A.local_contrast(g["profile"]), g["contrast20"]) < 1e-12
    assert np.all(A.local_contrast(np.ones(5), window=20) == 0)
    with pytest.raises(RuntimeError, match="Too few edge pixels"):
        A.slanted_edge_esf(np.zeros((8, 8), dtype=np.float32))
    with pytest.raises(RuntimeError, match="Too few edge pixels"):
        A.slanted_edge_esf(np.zeros((4, 4), dtype=np.float32))


# ------------------------------------------------------------------------------------------------ f4
def test_peaks_rois_fits_and_centres_match_reference_goldens(Cal):
    g = golden("g_calibration")
    n = int(g["n_cases"])
    R = int(g["radius"])
    imgs = [g[f"img{i}"] for i in range(n)]
    for i in range(n):
        img = imgs[i].astype(np.float64)
        pk = Cal.find_peak(img)
        assert tuple(pk) == tuple(int(v) for v in g[f"peak{i}"])
        roi = Cal.extract_psf(img, pk, R)
        assert roi.shape == g[f"roi{i}"].shape and _rel(roi, g[f"roi{i}"]) < 1e-12
        ref = g[f"popt{i}"]
        popt, fit = Cal.fit_gaussian_psf(g[f"roi{i}"])
        assert popt is not None
        assert abs(popt[1] - ref[1]) < 1e-4 and abs(popt[2] - ref[2]) < 1e-4                     # centre (px)
        assert abs(popt[0] - ref[0]) < 1e-4 * ref[0] and abs(popt[6] - ref[6]) < 1e-4 * ref[0]   # amplitude, offset
        assert np.allclose(sorted(popt[3:5]), sorted(ref[3:5]), rtol=1e-4)                       # widths (axes may swap)
        assert np.max(np.abs(fit - g[f"fit{i}"])) < 1e-4 * ref[0]                                # the fitted surface
        o_popt, _ = CO.fit_gaussian_psf(g[f"roi{i}"])                                            # same algorithm on the host
        assert np.allclose(popt[[0, 1, 2, 6]], o_popt[[0, 1, 2, 6]], rtol=0, atol=1e-6 * ref[0])
        cx, cy = Cal.find_pinhole_center(img)
        assert abs(cx - g[f"center{i}"][0]) < 1e-4 and abs(cy - g[f"center{i}"][1]) < 1e-4
        f, m, m2, label, nyq = Cal.compute_mtf(g[f"roi{i}"], 3.45)
        assert label == "cycles/mm" and abs(nyq - float(g[f"nyq{i}"])) < 1e-12
        assert _rel(f, g[f"mtf_f{i}"]) < 1e-12 and _rel(m, g[f"mtf_r{i}"]) < 1e-9
        o = CO.compute_mtf(g[f"roi{i}"], 3.45)
        assert _rel(m2, o[2]) < 1e-9
    if len({im.shape for im in imgs}) == 1:                  # the batched form: one launch for all frames
        r = Cal.find_pinhole_centers(np.stack(imgs))
        for i in range(n):
            assert tuple(r["peak"][i]) == tuple(int(v) for v in g[f"peak{i}"]) and r["status"][i] == 0
            assert abs(r["center"][i, 0] - g[f"center{i}"][0]) < 1e-4 and abs(r["center"][i, 1] - g[f"center{i}"][1]) < 1e-4


def test_smoothed_argmax_is_index_exact_on_a_batch_of_uint8_frames(Cal):
    rng = np.random.default_rng(5)
    n, H, W = 6, 190, 230
    stack = rng.integers(0, 40, (n, H, W)).astype(np.uint8)
    yy, xx = np.mgrid[:H, :W]
    spots = [(3, 4), (0, 0), (H - 1, W - 2), (100, 229), (57, 113), (188, 1)]          # edges and corners included
    for k, (r, c) in enumerate(spots):
        stack[k] = np.clip(stack[k] + 200 * np.exp(-((yy - r) ** 2 + (xx - c) ** 2) / 8.0), 0, 255).astype(np.uint8)
    stack[4, 50:70, 100:130] = 255                                                         # a saturated plateau: first-maximum rule
    got = Cal.find_peaks(stack)
    for k in range(n):
        assert tuple(got[k]) == tuple(int(v) for v in CO.find_peak(stack[k].astype(np.float64))), k
    res = Cal.find_pinhole_centers(stack)
    for k in range(n):
        cx, cy = CO.find_pinhole_center(stack[k].astype(np.float64))
        assert abs(res["center"][k, 0] - cx) < 1e-4 and abs(res["center"][k, 1] - cy) < 1e-4, (k, res["center"][k], cx, cy, res["status"][k])


def test_fit_fallbacks(Cal):
    assert Cal.fit_gaussian_psf(np.zeros((9, 9))) == (None, None)
    yy, xx = np.mgrid[:21, :25]
    p = (100.0, 12.3, 9.6, 2.0, 3.0, 0.4, 5.0)
    z = Cal.gauss2d((xx, yy), *p).reshape(21, 25)
    popt, fit = Cal.fit_gaussian_psf(z)
    assert np.max(np.abs(fit - z)) < 1e-6                              # a noiseless surface is recovered
    flat = np.zeros((40, 40), dtype=np.float32)
    assert Cal.find_pinhole_center(flat) == (0.0, 0.0)                 # all-zero image: falls through to the peak pixel
