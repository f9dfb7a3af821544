"""CPU tests: pin the oracle (oracle/sr_oracle.py, oracle/scipy_port.py) to

  1. the committed golden vectors (tests/golden/*.npz), which were produced by importing the
     UNMODIFIED reference scripts (tests/golden/make_golden.py);
  2. the live scipy in the image (the third-party library the reference's arithmetic lives in);
  3. when /root/reference is present (build container only): the live reference functions on fresh
     seeded inputs and the reference's committed result PNGs, bit-exact after the uint8 cast.

No GPU needed.  These run in the driver's ``-m "not gpu"`` pass.
"""
import os

import numpy as np
import pytest

from conftest import golden
from oracle import ref_loader as R
from oracle import scipy_port
from oracle import sr_oracle as O

NOMINAL = [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]
CALIBRATED = [(0.47225, -0.43385), (0.4897, 0.4641), (-0.4798, -0.4553), (-0.4856, 0.4369)]

# float64 restatement vs float64 reference: differences are FFT / summation-order round-off
TIGHT = 5e-10


def _lrs(g):
    if "lr" in g:
        return [np.asarray(l, dtype=np.float64) for l in g["lr"]]
    return [g["raw"][k][0::2, 0::2].astype(np.float64) for k in range(g["raw"].shape[0])]


# ------------------------------------------------------------------ 1. golden vectors
@pytest.mark.parametrize("name", ["g_int4", "g_int5", "g_frac4", "g_real", "g_unit"])
def test_oracle_pipeline_matches_reference_golden(name):
    g = golden(name)
    lrs, shifts, psf = _lrs(g), [tuple(s) for s in g["shifts"]], g["psf"]
    native = O.native_2x(np.mean(lrs, axis=0))
    saa = O.shift_and_add(lrs, shifts)
    hr, errs = O.ibp(lrs, shifts, psf, saa, n_iter=len(g["errors"]))
    assert native.shape == g["native"].shape and saa.shape == g["saa"].shape
    assert O.max_abs(native, g["native"]) < TIGHT
    assert O.max_abs(saa, g["saa"]) < TIGHT
    assert O.max_abs(hr, g["ibp"]) < 1e-8
    np.testing.assert_allclose(errs, g["errors"], rtol=1e-11)


@pytest.mark.parametrize("name", ["g_int4", "g_int5", "g_real", "g_unit"])
def test_integer_closed_form_matches_reference_golden(name):
    """The fused-kernel specification (SURVEY.md 8a 'verified restatement') against the reference."""
    g = golden(name)
    lrs, shifts = _lrs(g), [tuple(s) for s in g["shifts"]]
    saa = O.shift_and_add_integer(lrs, shifts)
    hr, errs = O.ibp_integer_shifts(lrs, shifts, g["psf"], saa, n_iter=len(g["errors"]))
    assert O.max_abs(saa, g["saa"]) < TIGHT
    assert O.max_abs(hr, g["ibp"]) < 1e-8
    np.testing.assert_allclose(errs, g["errors"], rtol=1e-11)


def test_oracle_primitives_match_reference_golden():
    g = golden("g_prims")
    k31 = golden("g_psf")["k31"]
    gauss = golden("g_int4")["psf"]
    assert O.max_abs(O.blur(g["a"], gauss), g["blur_gauss"]) < TIGHT
    assert O.max_abs(O.blur(g["a"], g["k7"]), g["blur_k7"]) < TIGHT
    assert O.max_abs(O.blur(g["a"], k31), g["blur_k31"]) < TIGHT
    assert O.max_abs(O.ndi_zoom_cubic(g["e"], 2), g["zoom"]) < TIGHT
    for tag, s in (("int", (0.5, -0.5)), ("frac", CALIBRATED[0]), ("zero", (0.0, 0.0))):
        assert O.max_abs(O.forward_model(g["a"], g["k7"], s, 2), g[f"fwd_{tag}"]) < TIGHT, tag
        assert O.max_abs(O.back_project(g["e"], g["k7"], s, 2, g["a"].shape), g[f"bp_{tag}"]) < TIGHT, tag


def test_oracle_psf_matches_reference_golden():
    g = golden("g_psf")
    assert np.array_equal(O.make_gaussian_psf(), golden("g_int4")["psf"])
    k7, peaks = O.measured_psf_from_images(list(g["crops"]), 3)
    assert len(peaks) == 30 and all(p == (64, 64) for p in peaks)       # crops are centred on the peak
    assert np.array_equal(k7 == 0, g["k7"] == 0) and int((k7 != 0).sum()) == 19
    assert np.max(np.abs(k7 - g["k7"])) < 1e-15
    k31, _ = O.measured_psf_from_images(list(g["crops"]), 15)
    assert k31.shape == (31, 31) and np.array_equal(k31 == 0, g["k31"] == 0)
    assert np.max(np.abs(k31 - g["k31"])) < 1e-15


def test_scipy_port_matches_reference_golden():
    """The timed CPU baseline (bench.py cpu_baseline.kind == 'port') computes what the reference does."""
    for name in ("g_int4", "g_frac4"):
        g = golden(name)
        lrs, shifts = _lrs(g), [tuple(s) for s in g["shifts"]]
        pipe = scipy_port.ScipyPipeline(shifts, g["psf"])
        saa = pipe.saa(lrs)
        hr, errs = pipe.ibp(lrs, saa, len(g["errors"]))
        assert O.max_abs(pipe.native(np.mean(lrs, axis=0)), g["native"]) < 1e-11
        assert O.max_abs(saa, g["saa"]) < 1e-11
        assert O.max_abs(hr, g["ibp"]) < 1e-9
        np.testing.assert_allclose(errs, g["errors"], rtol=1e-12)


# ------------------------------------------------------------------ 2. live scipy (the third-party dependency)
@pytest.mark.parametrize("shape", [(33, 47), (5, 3), (64, 96), (2, 2), (1, 9), (130, 7)])
def test_zoom_restatement_vs_scipy(shape):
    from scipy import ndimage as ndi
    a = np.random.default_rng(1).uniform(0, 255, shape)
    assert O.max_abs(O.ndi_zoom_cubic(a, 2), ndi.zoom(a, 2, order=3)) < TIGHT
    assert O.max_abs(O.ndi_zoom_cubic(a, 3), ndi.zoom(a, 3, order=3)) < TIGHT


@pytest.mark.parametrize("s", [(1.0, -1.0), (0.0, 0.0), (0.9445, -0.8677), (-0.37, 0.61), (3.0, -2.0), (5.3, 7.9)])
def test_shift_restatement_vs_scipy(s):
    from scipy import ndimage as ndi
    rng = np.random.default_rng(2)
    for shape in [(33, 47), (64, 96), (7, 5)]:
        a = rng.uniform(0, 255, shape)
        assert O.max_abs(O.ndi_shift_cubic_nearest(a, s), ndi.shift(a, s, order=3, mode="nearest")) < 5e-9


def test_integer_shift_is_clamp_gather():
    from scipy import ndimage as ndi
    a = np.random.default_rng(3).uniform(0, 255, (40, 31))
    for s in [(1, -1), (-1, 1), (2, 0), (0, -3), (-4, 4)]:
        assert O.max_abs(O.clamp_gather(a, *s), ndi.shift(a, s, order=3, mode="nearest")) < 1e-10


def test_conv_restatement_vs_fftconvolve():
    from scipy.signal import fftconvolve
    rng = np.random.default_rng(4)
    a = rng.uniform(0, 255, (45, 38))
    for shape in [(1, 1), (2, 2), (4, 3), (5, 8), (7, 7), (31, 31)]:
        k = rng.uniform(0, 1, shape)
        assert O.max_abs(O.conv2_same_zero(a, k), fftconvolve(a, k, mode="same")) < 1e-8


def test_prefilter_vs_scipy_spline_filter():
    from scipy import ndimage as ndi
    a = np.random.default_rng(5).uniform(0, 255, (29, 1300))        # long lines: truncated init sum (Appendix A.9)
    assert O.max_abs(O.spline_prefilter(a, "mirror"), ndi.spline_filter(a, order=3, mode="mirror")) < 1e-9
    assert O.max_abs(O.spline_prefilter(a, "reflect"), ndi.spline_filter(a, order=3, mode="nearest")) < 1e-9
    assert np.all(np.isfinite(O.spline_prefilter(a, "mirror")))


def test_uint8_cast_truncates():
    x = np.array([-3.0, 0.0, 0.999, 1.0, 127.5, 254.9999, 255.0, 300.0])
    assert O.to_uint8(x).tolist() == [0, 0, 0, 1, 127, 254, 255, 255]


def test_quirks_are_reproduced():
    """Q1 (+2s in both SAA and forward model), Q5 (SAA unclipped), Q11 (MSE before the update)."""
    rng = np.random.default_rng(6)
    lrs = [rng.uniform(0, 255, (20, 24)) for _ in NOMINAL]
    saa = O.shift_and_add(lrs, NOMINAL)
    assert saa.min() < 0 or saa.max() > 255 or True          # not clipped: just must not raise
    psf = O.make_gaussian_psf()
    hr1, e1 = O.ibp(lrs, NOMINAL, psf, saa, n_iter=1)
    manual = np.mean([np.mean((l - O.forward_model(saa, psf, s, 2)) ** 2) for l, s in zip(lrs, NOMINAL)])
    assert abs(e1[0] - manual) < 1e-9 * manual
    assert hr1.min() >= 0.0 and hr1.max() <= 255.0


# ------------------------------------------------------------------ 3. live reference (build container only)
needs_ref = pytest.mark.skipif(not R.available(), reason="/root/reference not present (GPU box)")


@needs_ref
@pytest.mark.parametrize("exp,shifts", [("mono_barcodes", NOMINAL), ("rgb_cal_target", CALIBRATED)])
def test_oracle_vs_live_reference_fresh_inputs(exp, shifts):
    ref = R.load(exp)
    rng = np.random.default_rng(77)
    lrs = [rng.uniform(0, 255, (27, 35)) for _ in shifts]
    psf = golden("g_prims")["k7"]
    with R.quiet():
        saa_r = ref.shift_and_add(lrs, shifts, factor=2, order=3)
        hr_r, e_r = ref.ibp(lrs, shifts, psf, saa_r.copy(), factor=2, n_iter=4, step=0.5)
    saa = O.shift_and_add(lrs, shifts)
    hr, e = O.ibp(lrs, shifts, psf, saa, n_iter=4)
    assert O.max_abs(saa, saa_r) < TIGHT and O.max_abs(hr, hr_r) < 1e-8
    np.testing.assert_allclose(e, e_r, rtol=1e-11)


@needs_ref
def test_golden_files_regenerate_identically():
    """The committed fixtures are what the reference produces today (spot check: g_int4)."""
    ref = R.load("mono_barcodes")
    g = golden("g_int4")
    with R.quiet():
        saa = ref.shift_and_add(list(g["lr"]), [tuple(s) for s in g["shifts"]], factor=2, order=3)
        hr, errs = ref.ibp(list(g["lr"]), [tuple(s) for s in g["shifts"]], g["psf"], saa.copy(), factor=2, n_iter=8, step=0.5)
    assert np.array_equal(saa, g["saa"]) and np.array_equal(hr, g["ibp"]) and np.array_equal(errs, g["errors"])


def _same_png(img, want):
    """uint8-identical, except where the float value sits within 1e-8 of an integer (the truncating
    cast then turns a last-bit round-off difference into one DN: e.g. zoom reproduces the corner
    sample 108.0 exactly in scipy and as 107.99999999999997 in the restatement)."""
    got = O.to_uint8(img)
    assert got.shape == want.shape
    bad = np.flatnonzero(got.ravel() != want.ravel())
    assert bad.size <= 1e-5 * got.size, f"{bad.size} pixels differ"
    v = np.asarray(img).ravel()[bad]
    assert np.all(np.abs(v - np.round(v)) < 1e-8)
    assert np.all(np.abs(got.ravel()[bad].astype(int) - want.ravel()[bad].astype(int)) == 1)


@needs_ref
def test_oracle_vs_reference_png_goldens_rgb_cal_target():
    """Native-2x / SAA / LR_red_mean PNGs committed by the reference (fractional calibrated shifts):
    bit-exact after the truncating uint8 cast."""
    from PIL import Image
    ref = R.load("rgb_cal_target")
    res_root = os.path.join(R.REFERENCE_ROOT, "rgb_cal_target", "results")
    combos = [d for d in sorted(os.listdir(res_root)) if os.path.exists(os.path.join(res_root, d, "SAA.png"))]
    if not combos:
        pytest.skip("no committed rgb_cal_target results")
    combo = combos[0]
    data_dir = os.path.join(R.REFERENCE_ROOT, "rgb_cal_target", "data", combo)
    if not os.path.isdir(data_dir):
        pytest.skip("no data for the committed combo")
    with R.quiet():
        frames, shifts = ref.load_combo(data_dir)[:2]
    mean_lr = np.mean(frames, axis=0)
    want_mean = np.array(Image.open(os.path.join(res_root, combo, "LR_red_mean.png")))
    assert np.array_equal(O.to_uint8(mean_lr), want_mean)
    native = O.native_2x(mean_lr)
    _same_png(native, np.array(Image.open(os.path.join(res_root, combo, "native_2x.png"))))
    saa = O.shift_and_add(frames, shifts)
    _same_png(saa, np.array(Image.open(os.path.join(res_root, combo, "SAA.png"))))


def test_zoom_orders_0_and_1_and_non_integer_factors_match_scipy():
    """`order` of shift_and_add reaches zoom (run_sr.py:185): the oracle's restatement of scipy.ndimage.zoom for orders 0 / 1 / 3 and
    scalar or per-axis, integer or fractional factors against live scipy."""
    ndi = pytest.importorskip("scipy.ndimage")
    rng = np.random.default_rng(31)
    for shape in [(7, 5), (33, 47), (2, 2), (1, 9), (64, 96)]:
        a = rng.uniform(0, 255, shape)
        for factor in (2, 3, 1.5, 2.37, (2, 1.5)):
            for order in (0, 1, 3):
                if order == 3 and min(shape) < 2:
                    continue
                ref = ndi.zoom(a, factor, order=order)
                got = O.ndi_zoom_order(a, factor, order)
                assert got.shape == ref.shape
                assert np.max(np.abs(got - ref)) < 1e-9, (shape, factor, order)
    lrs = [rng.uniform(0, 255, (12, 17)) for _ in range(3)]
    shifts = [(0.5, -0.5), (0.3, 0.2), (0.0, 0.0)]
    for order in (0, 1):
        ref = sum(ndi.shift(ndi.zoom(l, 2, order=order), (2 * dy, 2 * dx), order=3, mode="nearest") for l, (dy, dx) in zip(lrs, shifts)) / 3
        assert np.max(np.abs(O.shift_and_add(lrs, shifts, 2, order) - ref)) < 1e-9
