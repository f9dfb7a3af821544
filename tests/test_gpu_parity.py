"""GPU parity: the CUDA path (through the C ABI) against the oracle and the committed
reference-generated golden vectors.  Tolerances are those of BASELINE.json's north_star:
max-abs <= 1e-4 on [0,1] images (0.0255 on the reference's [0,255] scale), |dPSNR| < 0.01 dB,
MSE vector rel-err <= 1e-4; integer decisions (argmax, shapes) exact."""
import numpy as np
import pytest

from conftest import TOL_ABS_255, TOL_MSE_REL, TOL_PSNR_DB, golden
from oracle import sr_oracle as O

pytestmark = pytest.mark.gpu

NOMINAL = [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]
CALIBRATED = [(0.47225, -0.43385), (0.4897, 0.4641), (-0.4798, -0.4553), (-0.4856, 0.4369)]


def _close(a, b, tol=TOL_ABS_255, what=""):
    d = O.max_abs(a, b)
    assert d <= tol, f"{what}: max abs diff {d:.3e} > {tol:.3e}"


def _errs_close(e, ref, what=""):
    e, ref = np.asarray(e), np.asarray(ref)
    assert e.shape == ref.shape
    rel = np.max(np.abs(e - ref) / np.maximum(np.abs(ref), 1e-12))
    assert rel <= TOL_MSE_REL, f"{what}: MSE rel err {rel:.3e}"


# ---------------------------------------------------------------- primitives
def test_gaussian_psf_matches_reference(sr):
    g = golden("g_int4")
    k = sr.make_gaussian_psf()
    assert k.shape == (7, 7)
    assert np.max(np.abs(k - g["psf"])) < 1e-15
    assert sr.make_gaussian_psf(6, 2.0).shape == (7, 7)   # reference builds mgrid[-hw:hw+1]
    np.testing.assert_allclose(sr.make_gaussian_psf(5, 0.7), O.make_gaussian_psf(5, 0.7), rtol=0, atol=1e-15)


def test_measured_psf_exact_decisions_and_values(sr):
    g = golden("g_psf")
    crops = g["crops"]
    k7, peaks, used = sr.measured_psf_from_images(list(crops), 3, return_peaks=True)
    ok7, opeaks = O.measured_psf_from_images(list(crops), 3)
    assert used == len(opeaks) == crops.shape[0]
    assert [tuple(p) for p in peaks.tolist()] == opeaks          # argmax coordinates exact
    assert np.array_equal(k7 == 0, g["k7"] == 0)                 # support / zero pattern exact
    assert np.max(np.abs(k7 - g["k7"])) < 1e-14
    k31 = sr.measured_psf_from_images(list(crops), 15)
    assert k31.shape == (31, 31)
    assert np.array_equal(k31 == 0, g["k31"] == 0)
    assert np.max(np.abs(k31 - g["k31"])) < 1e-14


def test_measured_psf_edge_skip(sr):
    rng = np.random.default_rng(3)
    imgs = [rng.integers(0, 50, (64, 80), dtype=np.uint8) for _ in range(4)]
    for i, (r, c) in enumerate([(30, 40), (4, 40), (30, 75), (20, 20)]):
        imgs[i][r, c] = 255
    k, peaks, used = sr.measured_psf_from_images(imgs, 3, return_peaks=True)
    ok, opeaks = O.measured_psf_from_images(imgs, 3)
    assert used == 2 and peaks.tolist() == [[30, 40], [-1, -1], [-1, -1], [20, 20]]
    assert np.max(np.abs(k - ok)) < 1e-14
    with pytest.raises(Exception):
        sr.measured_psf_from_images([imgs[1]], 3)


def test_argmax_first_maximum_tie_break(sr):
    img = np.zeros((40, 50), dtype=np.uint8)
    img[25, 30] = 200
    img[12, 33] = 200       # earlier in row-major order -> must win
    img[12, 40] = 200
    _, peaks, _ = sr.measured_psf_from_images([img], 1, return_peaks=True)
    assert peaks.tolist() == [[12, 33]]


def test_ingest_helpers(sr):
    rng = np.random.default_rng(5)
    raw = rng.integers(0, 256, (37, 52), dtype=np.uint8)
    red = sr.extract_red(raw)
    assert red.shape == (19, 26) and np.array_equal(red, raw[0::2, 0::2].astype(np.float64))
    frames = [rng.integers(0, 256, (21, 34)).astype(np.float64) for _ in range(5)]
    m = sr.mean_frames(frames)
    assert np.max(np.abs(m - np.mean(frames, axis=0))) < 1e-4
    x = rng.uniform(-20, 280, (33, 41))
    assert np.array_equal(sr.to_uint8(x), np.clip(x.astype(np.float32), 0, 255).astype(np.uint8))


@pytest.mark.parametrize("key", ["blur_gauss", "blur_k7", "blur_k31"])
def test_blur_vs_reference_golden(sr, key):
    g = golden("g_prims")
    k = {"blur_gauss": golden("g_int4")["psf"], "blur_k7": g["k7"], "blur_k31": golden("g_psf")["k31"]}[key]
    _close(sr.blur(g["a"], k), g[key], what=key)


def test_blur_even_and_rectangular_kernels(sr):
    rng = np.random.default_rng(6)
    a = rng.uniform(0, 255, (45, 38))
    for shape in [(1, 1), (2, 2), (4, 3), (5, 8), (3, 3)]:
        k = rng.uniform(0, 1, shape)
        k /= k.sum()
        _close(sr.blur(a, k), O.conv2_same_zero(a, k), what=f"blur {shape}")


def test_zoom_vs_reference_golden_and_odd_sizes(sr):
    g = golden("g_prims")
    _close(sr.ndi_zoom(g["e"], 2), g["zoom"], what="zoom golden")
    rng = np.random.default_rng(7)
    for shape in [(33, 47), (5, 3), (64, 96), (255, 257), (1, 9), (2, 2)]:
        a = rng.uniform(0, 255, shape)
        _close(sr.native_2x(a), O.ndi_zoom_cubic(a, 2), what=f"zoom {shape}")
    a = rng.uniform(0, 255, (20, 31))
    _close(sr.ndi_zoom(a, 3), O.ndi_zoom_cubic(a, 3), what="zoom x3")


@pytest.mark.parametrize("s", [(1.0, -1.0), (0.0, 0.0), (0.9445, -0.8677), (-0.37, 0.61), (3.0, -2.0), (5.3, 7.9)])
def test_shift_vs_oracle(sr, s):
    rng = np.random.default_rng(8)
    for shape in [(33, 47), (64, 96), (7, 5)]:
        a = rng.uniform(0, 255, shape)
        _close(sr.ndi_shift(a, s), O.ndi_shift_cubic_nearest(a, s), what=f"shift {s} {shape}")


@pytest.mark.parametrize("tag,s", [("int", (0.5, -0.5)), ("frac", CALIBRATED[0]), ("zero", (0.0, 0.0))])
def test_forward_and_back_project_vs_reference_golden(sr, tag, s):
    g = golden("g_prims")
    _close(sr.forward_model(g["a"], g["k7"], s, 2), g[f"fwd_{tag}"], what=f"forward {tag}")
    _close(sr.back_project(g["e"], g["k7"], s, 2, g["a"].shape), g[f"bp_{tag}"], what=f"back_project {tag}")


# ---------------------------------------------------------------- pipelines vs reference goldens
def _run_pipeline(sr, lrs, shifts, psf, n_iter, path=0):
    import torch
    n, (h, w) = len(lrs), lrs[0].shape
    plan = sr.Plan(n, h, w, shifts, psf)
    if path:
        try:
            plan.set_path(path)
        except sr.Sr459Error as exc:
            plan.close()
            if exc.code == -4:
                pytest.skip(f"path {path} not available for this geometry: {exc}")
            raise
    dl = [torch.from_numpy(np.asarray(l, dtype=np.float32)).cuda() for l in lrs]
    native = plan.native_2x(sr.mean_frames(dl)).cpu().numpy()
    saa = plan.saa(dl)
    hr = saa.clone()
    errs = plan.ibp_(dl, hr, n_iter)
    plan.close()
    return native, saa.cpu().numpy(), hr.cpu().numpy(), errs


def _check_against(g, native, saa, hr, errs, peak=255.0, scale=1.0, what=""):
    tol = TOL_ABS_255 * scale
    _close(native, g["native"], tol, what + " native")
    _close(saa, g["saa"], tol, what + " saa")
    _close(hr, g["ibp"], tol, what + " ibp")
    _errs_close(errs, g["errors"], what + " errors")
    if "truth" in g:
        d = abs(O.psnr(hr, g["truth"], peak) - O.psnr(g["ibp"], g["truth"], peak))
        assert d < TOL_PSNR_DB, f"{what}: PSNR differs by {d} dB"


@pytest.mark.parametrize("path", [0, 1, 2, 3, 4, 7])
@pytest.mark.parametrize("name", ["g_int4", "g_int5", "g_real"])
def test_integer_shift_goldens_all_paths(sr, name, path):
    g = golden(name)
    lrs = list(g["lr"]) if "lr" in g else [g["raw"][k][0::2, 0::2].astype(np.float64) for k in range(4)]
    res = _run_pipeline(sr, lrs, g["shifts"], g["psf"], len(g["errors"]), path)
    _check_against(g, *res, what=f"{name} path{path}")


def test_real_crop_uint8_outputs_exact(sr):
    g = golden("g_real")
    lrs = [sr.extract_red(g["raw"][k]) for k in range(4)]
    native, saa, hr, errs = _run_pipeline(sr, lrs, g["shifts"], g["psf"], len(g["errors"]))
    for got, ref in ((native, g["native"]), (saa, g["saa"])):
        a, b = sr.to_uint8(got), O.to_uint8(ref)
        # truncating cast: a float32 result an ulp below an integer may land one DN lower
        bad = np.flatnonzero(a.ravel() != b.ravel())
        assert bad.size <= 1e-4 * a.size
        assert np.all(np.abs(a.ravel()[bad].astype(int) - b.ravel()[bad].astype(int)) == 1)
        frac = np.abs(ref.ravel()[bad] - np.round(ref.ravel()[bad]))
        assert np.all(frac < 1e-3)


def test_fractional_shift_golden(sr):
    g = golden("g_frac4")
    res = _run_pipeline(sr, list(g["lr"]), g["shifts"], g["psf"], len(g["errors"]))
    _check_against(g, *res, what="g_frac4")


def test_unit_range_golden(sr):
    """The BASELINE tolerance statement itself: [0,1] float32 frames, max-abs <= 1e-4."""
    g = golden("g_unit")
    res = _run_pipeline(sr, list(g["lr"]), g["shifts"], g["psf"], len(g["errors"]))
    _check_against(g, *res, peak=1.0, scale=1.0 / 255.0, what="g_unit")


# ---------------------------------------------------------------- module-level drop-in surface
def test_dropin_functions_numpy_in_numpy_out(sr):
    g = golden("g_int4")
    lrs, shifts, psf = list(g["lr"]), [tuple(s) for s in g["shifts"]], g["psf"]
    saa = sr.shift_and_add(lrs, shifts, factor=2, order=3)
    assert isinstance(saa, np.ndarray) and saa.dtype == np.float64 and saa.shape == (66, 94)
    _close(saa, g["saa"], what="saa")
    init = saa.copy()
    hr, errs = sr.ibp(lrs, shifts, psf, saa, factor=2, n_iter=8, step=0.5)
    assert np.array_equal(saa, init), "hr_init must not be modified (reference copies it)"
    assert isinstance(errs, list) and len(errs) == 8 and isinstance(errs[0], float)
    _close(hr, g["ibp"], what="ibp")
    _errs_close(errs, g["errors"])


def test_dropin_functions_cuda_tensors(sr):
    import torch
    g = golden("g_int4")
    lrs = [torch.from_numpy(l.astype(np.float32)).cuda() for l in g["lr"]]
    saa = sr.shift_and_add(lrs, g["shifts"])
    assert saa.is_cuda and saa.dtype == torch.float32
    keep = saa.clone()
    hr, errs = sr.ibp(lrs, g["shifts"], g["psf"], saa, n_iter=8)
    assert torch.equal(saa, keep) and hr.data_ptr() != saa.data_ptr()
    _close(hr.cpu().numpy(), g["ibp"], what="ibp cuda")


def test_error_behaviour(sr):
    g = golden("g_int4")
    lrs = list(g["lr"])
    with pytest.raises(ValueError):
        sr.ibp(lrs, g["shifts"], g["psf"], np.zeros((10, 10)), n_iter=1)
    with pytest.raises(ValueError):
        sr.shift_and_add(lrs[:3], g["shifts"])
    with pytest.raises(ValueError):
        sr.shift_and_add([lrs[0], lrs[1][:-1]], g["shifts"][:2])
    with pytest.raises(FileNotFoundError):
        sr.load_measured_psf("/nonexistent/dir")
    hr, errs = sr.ibp(lrs, g["shifts"], g["psf"], g["saa"], n_iter=0)
    assert errs == [] and np.max(np.abs(hr - g["saa"])) < 1e-4


# ---------------------------------------------------------------- oracle-diff on seeded inputs, borders
@pytest.mark.parametrize("shape", [(33, 47), (64, 96), (20, 130), (6, 6), (2, 2)])
@pytest.mark.parametrize("path", [1, 2, 3, 4])
def test_integer_paths_vs_oracle_odd_sizes(sr, shape, path):
    rng = np.random.default_rng(459)
    lrs = [rng.uniform(0, 255, shape) for _ in NOMINAL]
    psf = O.make_gaussian_psf()
    h, w = shape
    if path in (3, 4) and (w % 2):
        pytest.skip("polyphase kernels need W % 4 == 0")
    saa_o = O.shift_and_add(lrs, NOMINAL)
    hr_o, e_o = O.ibp(lrs, NOMINAL, psf, saa_o, n_iter=5)
    native, saa, hr, errs = _run_pipeline(sr, lrs, NOMINAL, psf, 5, path)
    _close(saa, saa_o, what="saa")
    _close(hr, hr_o, what=f"ibp {shape} path{path}")
    # borders separately from the interior
    _close(hr[:8], hr_o[:8], what="top rows"); _close(hr[-8:], hr_o[-8:], what="bottom rows")
    _close(hr[:, :8], hr_o[:, :8], what="left cols"); _close(hr[:, -8:], hr_o[:, -8:], what="right cols")
    _errs_close(errs, e_o)


@pytest.mark.parametrize("shifts", [
    [(0.0, 0.0)],
    [(0.0, 0.0), (0.5, 0.0), (0.0, 0.5), (0.5, 0.5)],
    [(-0.5, -0.5), (0.5, 0.5)],
    [(1.0, -1.5), (-2.0, 0.5), (0.0, 1.0)],
    [(0.5, -0.5)] * 8,
])
def test_general_integer_shift_sets(sr, shifts):
    rng = np.random.default_rng(11)
    lrs = [rng.uniform(0, 255, (40, 52)) for _ in shifts]
    psf = O.make_gaussian_psf()
    saa_o = O.shift_and_add(lrs, shifts)
    hr_o, e_o = O.ibp(lrs, shifts, psf, saa_o, n_iter=4)
    for path in (0, 1, 2):
        native, saa, hr, errs = _run_pipeline(sr, lrs, shifts, psf, 4, path)
        _close(saa, saa_o, what=f"saa {shifts}")
        _close(hr, hr_o, what=f"ibp {shifts} path{path}")
        _errs_close(errs, e_o)


@pytest.mark.parametrize("shifts", [
    NOMINAL,
    [(0.0, 0.0)] + NOMINAL,                                   # mono_cal_target: two lattices
    [(0.0, 0.0), (0.5, 0.0), (0.0, 0.5), (0.5, 0.5)],         # all four lattices
    [(0.5, -0.5)] * 3 + [(-0.5, -0.5)] * 3,                   # six frames on one lattice
    [(0.0, 0.5), (-0.5, 0.0)],
])
def test_polyphase_kernel_interior_tiles_vs_oracle(sr, shifts):
    """HR 320x640 has tiles that run the unchecked instantiation as well as all edge kinds."""
    rng = np.random.default_rng(21)
    h, w = 160, 320
    lrs = [rng.uniform(0, 255, (h, w)) for _ in shifts]
    psf = golden("g_prims")["k7"]
    saa_o = O.shift_and_add_integer(lrs, shifts)
    hr_o, e_o = O.ibp_integer_shifts(lrs, shifts, psf, saa_o, n_iter=4)
    for path in (3, 2):
        native, saa, hr, errs = _run_pipeline(sr, lrs, shifts, psf, 4, path)
        _close(saa, saa_o, what="saa")
        _close(hr, hr_o, what=f"ibp path{path} {shifts}")
        _errs_close(errs, e_o)


@pytest.mark.parametrize("shifts", [
    NOMINAL,
    [(0.0, 0.0)] + NOMINAL,                                   # mono_cal_target: two lattices, both row parities
    [(0.0, 0.0), (0.5, 0.0), (0.0, 0.5), (0.5, 0.5)],         # all four lattices
    [(0.5, -0.5)] * 3 + [(-0.5, -0.5)] * 3,                   # six frames on one lattice
    [(0.0, 0.5), (-0.5, 0.0)],
    [(0.5, 0.5), (-0.5, -0.5)],                               # frame count on the last row/col is not separable
    [(-0.5, 0.5)],
])
@pytest.mark.parametrize("hw", [(160, 320), (100, 150), (33, 48)])
@pytest.mark.parametrize("path", [4])
def test_separable_tma_kernel_vs_oracle(sr, shifts, hw, path):
    """The TMA-staged warp-synchronous separable kernel (path 4, k_ibp_ws; what row bands of several lattices use):
    interior tiles, every edge kind, partial tiles, images smaller than one tile; Gaussian PSF (rank 1)."""
    rng = np.random.default_rng(22)
    h, w = hw
    lrs = [rng.uniform(0, 255, (h, w)) for _ in shifts]
    psf = O.make_gaussian_psf()
    saa_o = O.shift_and_add_integer(lrs, shifts)
    hr_o, e_o = O.ibp_integer_shifts(lrs, shifts, psf, saa_o, n_iter=4)
    native, saa, hr, errs = _run_pipeline(sr, lrs, shifts, psf, 4, path)
    _close(hr, hr_o, what=f"ibp path{path} {shifts} {hw}")
    _close(hr[:8], hr_o[:8], what="top rows"); _close(hr[-8:], hr_o[-8:], what="bottom rows")
    _close(hr[:, :8], hr_o[:, :8], what="left cols"); _close(hr[:, -8:], hr_o[:, -8:], what="right cols")
    _errs_close(errs, e_o)


def test_separable_kernel_small_psfs_and_anisotropic(sr):
    rng = np.random.default_rng(23)
    lrs = [rng.uniform(0, 255, (96, 160)) for _ in NOMINAL]
    saa_o = O.shift_and_add_integer(lrs, NOMINAL)
    gy = np.exp(-0.5 * (np.arange(-3, 4) / 0.7) ** 2); gx = np.exp(-0.5 * (np.arange(-3, 4) / 1.4) ** 2)
    aniso = np.outer(gy / gy.sum(), gx / gx.sum())
    for psf in (O.make_gaussian_psf(3, 0.6), O.make_gaussian_psf(5, 0.9), np.array([[1.0]]), aniso):
        hr_o, e_o = O.ibp_integer_shifts(lrs, NOMINAL, psf, saa_o, n_iter=4)
        native, saa, hr, errs = _run_pipeline(sr, lrs, NOMINAL, psf, 4, 4)
        _close(hr, hr_o, what=f"path4 psf {psf.shape}")
        _errs_close(errs, e_o)



# ---------------------------------------------------------------- the register-streaming kernel (path 7)
SINGLE_LATTICE_SETS = [
    NOMINAL,                                                  # (1,1): all four corner frames
    [(0.5, -0.5)] * 3 + [(-0.5, -0.5)] * 3,                   # six frames, only sx = -1
    [(0.5, 0.5), (-0.5, -0.5)],                               # frame count on the last row/col is not separable
    [(-0.5, 0.5)],                                            # one frame
    [(0.5, 0.5)],
    [(0.0, 0.0)],                                             # (0,0): centre frame only
    [(0.0, 0.0)] * 3,
    [(0.0, 0.5), (0.0, -0.5)],                                # (0,1): column clamp terms only
    [(0.5, 0.0), (-0.5, 0.0), (-0.5, 0.0)],                   # (1,0): row clamp terms only
    [(0.5, -0.5)] * 8,
]


@pytest.mark.parametrize("shifts", SINGLE_LATTICE_SETS)
@pytest.mark.parametrize("hw", [(160, 320), (100, 150), (33, 48), (7, 60), (2, 2), (231, 58)])
def test_streaming_kernel_vs_oracle(sr, shifts, hw):
    """k_ibp_stream (path 7): interior and general iterations, top segment, left strip, last site
    row / column, partial last segment and strip, images smaller than one strip."""
    rng = np.random.default_rng(24)
    h, w = hw
    lrs = [rng.uniform(0, 255, (h, w)) for _ in shifts]
    psf = O.make_gaussian_psf()
    saa_o = O.shift_and_add_integer(lrs, shifts)
    hr_o, e_o = O.ibp_integer_shifts(lrs, shifts, psf, saa_o, n_iter=4)
    native, saa, hr, errs = _run_pipeline(sr, lrs, shifts, psf, 4, 7)
    _close(hr, hr_o, what=f"ibp path7 {shifts} {hw}")
    _close(hr[:8], hr_o[:8], what="top rows"); _close(hr[-8:], hr_o[-8:], what="bottom rows")
    _close(hr[:, :8], hr_o[:, :8], what="left cols"); _close(hr[:, -8:], hr_o[:, -8:], what="right cols")
    _errs_close(errs, e_o)


def test_streaming_kernel_small_psfs_anisotropic_and_unit_range(sr):
    rng = np.random.default_rng(25)
    lrs = [rng.uniform(0, 255, (96, 160)) for _ in NOMINAL]
    saa_o = O.shift_and_add_integer(lrs, NOMINAL)
    gy = np.exp(-0.5 * (np.arange(-3, 4) / 0.7) ** 2); gx = np.exp(-0.5 * (np.arange(-3, 4) / 1.4) ** 2)
    aniso = np.outer(gy / gy.sum(), gx / gx.sum())
    for psf in (O.make_gaussian_psf(3, 0.6), O.make_gaussian_psf(5, 0.9), np.array([[1.0]]), aniso):
        hr_o, e_o = O.ibp_integer_shifts(lrs, NOMINAL, psf, saa_o, n_iter=4)
        native, saa, hr, errs = _run_pipeline(sr, lrs, NOMINAL, psf, 4, 7)
        _close(hr, hr_o, what=f"path7 psf {psf.shape}")
        _errs_close(errs, e_o)
    g = golden("g_unit")                                      # [0,1] images: tolerance 1e-4 absolute
    res = _run_pipeline(sr, list(g["lr"]), g["shifts"], g["psf"], len(g["errors"]), 7)
    _check_against(g, *res, peak=1.0, scale=1.0 / 255.0, what="g_unit path7")


_LAT_SHIFT = {0: [(0.0, 0.0)], 1: [(0.0, 0.5), (0.0, -0.5)], 2: [(0.5, 0.0), (-0.5, 0.0)], 3: [(0.5, 0.5), (-0.5, -0.5), (0.5, -0.5)]}
MULTI_LATTICE_SETS = [sum((_LAT_SHIFT[pq] for pq in range(4) if mask >> pq & 1), []) for mask in (3, 5, 6, 7, 9, 10, 11, 12, 13, 14, 15)]
MULTI_LATTICE_SETS += [
    [(0.0, 0.0)] + NOMINAL,                                   # mono_cal_target's own five frames (run_sr.py:60-66)
    [(0.0, 0.0), (-0.5, 0.0), (0.0, -0.5), (-0.5, -0.5)],     # all four lattices, row-0 / column-0 duplicates only
    [(0.0, 0.0), (0.5, 0.0), (0.0, 0.5), (0.5, 0.5)],         # all four lattices, MSE-only terms only
    [(0.0, 0.0)] * 2 + [(0.5, 0.5)] * 3 + [(-0.5, -0.5)] * 3,
]


@pytest.mark.parametrize("shifts", MULTI_LATTICE_SETS)
@pytest.mark.parametrize("hw", [(160, 320), (100, 150), (33, 48), (7, 60), (2, 2), (231, 58)])
def test_multi_lattice_streaming_kernel_vs_oracle(sr, shifts, hw):
    """k_ibp_stream_ml (path 7 with two to four parity lattices): every lattice mask, interior and general iterations,
    top segment, left strip, last site row / column, partial last segment and strip, images smaller than one strip."""
    if len(shifts) > 8:
        pytest.skip("more than 8 frames")
    rng = np.random.default_rng(26)
    h, w = hw
    lrs = [rng.uniform(0, 255, (h, w)) for _ in shifts]
    psf = O.make_gaussian_psf()
    saa_o = O.shift_and_add_integer(lrs, shifts)
    hr_o, e_o = O.ibp_integer_shifts(lrs, shifts, psf, saa_o, n_iter=4)
    native, saa, hr, errs = _run_pipeline(sr, lrs, shifts, psf, 4, 7)
    _close(hr, hr_o, what=f"ibp path7 multi-lattice {shifts} {hw}")
    _close(hr[:8], hr_o[:8], what="top rows"); _close(hr[-8:], hr_o[-8:], what="bottom rows")
    _close(hr[:, :8], hr_o[:, :8], what="left cols"); _close(hr[:, -8:], hr_o[:, -8:], what="right cols")
    _errs_close(errs, e_o)


def test_multi_lattice_streaming_kernel_is_the_default_and_handles_small_psfs(sr):
    rng = np.random.default_rng(27)
    shifts = [(0.0, 0.0)] + NOMINAL
    lrs = [rng.uniform(0, 255, (96, 160)) for _ in shifts]
    saa_o = O.shift_and_add_integer(lrs, shifts)
    gy = np.exp(-0.5 * (np.arange(-3, 4) / 0.7) ** 2); gx = np.exp(-0.5 * (np.arange(-3, 4) / 1.4) ** 2)
    aniso = np.outer(gy / gy.sum(), gx / gx.sum())
    for psf in (O.make_gaussian_psf(3, 0.6), O.make_gaussian_psf(5, 0.9), np.array([[1.0]]), aniso, O.make_gaussian_psf()):
        hr_o, e_o = O.ibp_integer_shifts(lrs, shifts, psf, saa_o, n_iter=5)
        for path in (0, 7):
            native, saa, hr, errs = _run_pipeline(sr, lrs, shifts, psf, 5, path)
            _close(hr, hr_o, what=f"path{path} multi-lattice psf {psf.shape}")
            _errs_close(errs, e_o)
    a = _run_pipeline(sr, lrs, shifts, O.make_gaussian_psf(), 5, 0)      # auto == path 7 bit for bit: the plan picks this kernel
    b = _run_pipeline(sr, lrs, shifts, O.make_gaussian_psf(), 5, 7)
    assert np.array_equal(a[2], b[2]) and np.array_equal(np.asarray(a[3]), np.asarray(b[3]))


def test_non_separable_psf_refuses_path4(sr):
    plan = sr.Plan(4, 64, 64, NOMINAL, golden("g_prims")["k7"])
    with pytest.raises(sr.Sr459Error) as ei:
        plan.set_path(4)
    assert ei.value.code == -4
    plan.close()


@pytest.mark.parametrize("ksize", [3, 5, 7, 9])
def test_psf_sizes_integer_shifts(sr, ksize):
    rng = np.random.default_rng(12)
    lrs = [rng.uniform(0, 255, (48, 64)) for _ in NOMINAL]
    psf = O.make_gaussian_psf(ksize, 0.8)
    saa_o = O.shift_and_add(lrs, NOMINAL)
    hr_o, e_o = O.ibp(lrs, NOMINAL, psf, saa_o, n_iter=4)
    native, saa, hr, errs = _run_pipeline(sr, lrs, NOMINAL, psf, 4)
    _close(hr, hr_o, what=f"psf {ksize}")
    _errs_close(errs, e_o)


def test_measured_psf_31_integer_and_fractional(sr):
    g = golden("g_psf")
    rng = np.random.default_rng(13)
    lrs = [rng.uniform(0, 255, (40, 56)) for _ in NOMINAL]
    for shifts in (NOMINAL, CALIBRATED):
        saa_o = O.shift_and_add(lrs, shifts)
        hr_o, e_o = O.ibp(lrs, shifts, g["k31"], saa_o, n_iter=3)
        native, saa, hr, errs = _run_pipeline(sr, lrs, shifts, g["k31"], 3)
        _close(hr, hr_o, what="k31")
        _errs_close(errs, e_o)


LATTICE_SHIFT_SETS = [
    NOMINAL,
    [(0.0, 0.0)] + NOMINAL,                                   # two lattices (mono_cal_target)
    [(0.0, 0.0), (0.5, 0.0), (0.0, 0.5), (0.5, 0.5)],         # all four lattices, no clamp terms
    [(0.0, 0.0), (-0.5, 0.0), (0.0, -0.5), (-0.5, -0.5)],     # all four lattices, row-0 / column-0 duplicates
    [(0.5, 0.5)] * 2 + [(-0.5, -0.5)] * 3,                    # MSE-only rows and duplicate rows on one lattice
    [(-0.5, 0.5), (0.5, 0.0)],
]


def _nonseparable_psf(size, seed):
    rng = np.random.default_rng(seed)
    r = size // 2
    y, x = np.mgrid[-r:r + 1, -r:r + 1].astype(np.float64)
    k = np.exp(-(x * x / (0.5 * r + 1.0) + y * y / (0.3 * r + 0.7) + 0.2 * x * y / (r + 1.0))) * (1.0 + 0.3 * rng.uniform(size=(size, size)))
    return k / k.sum()


@pytest.mark.parametrize("shifts", LATTICE_SHIFT_SETS)
@pytest.mark.parametrize("hw,ksize", [((40, 56), 31), ((70, 130), 31), ((33, 48), 19), ((36, 72), 9), ((64, 66), 7), ((2, 2), 31), ((5, 34), 15)])
def test_lattice_kernel_vs_oracle(sr, shifts, hw, ksize):
    """Path 8 (ibp_lat.cu): large / non-separable PSFs, integer shifts |2d| <= 1, every lattice mix and border kind."""
    rng = np.random.default_rng(81)
    h, w = hw
    lrs = [rng.uniform(0, 255, (h, w)) for _ in shifts]
    psf = _nonseparable_psf(ksize, ksize)
    saa_o = O.shift_and_add_integer(lrs, shifts)
    hr_o, e_o = O.ibp_integer_shifts(lrs, shifts, psf, saa_o, n_iter=4)
    native, saa, hr, errs = _run_pipeline(sr, lrs, shifts, psf, 4, 8)
    _close(hr, hr_o, what=f"lattice ibp {hw} k{ksize} {shifts}")
    _close(hr[:16], hr_o[:16], what="top rows"); _close(hr[:, :16], hr_o[:, :16], what="left cols")
    _close(hr[-16:], hr_o[-16:], what="bottom rows"); _close(hr[:, -16:], hr_o[:, -16:], what="right cols")
    _errs_close(errs, e_o)


def test_lattice_kernel_matches_dense_kernel_and_reference_golden(sr):
    import torch
    g = golden("g_psf")
    rng = np.random.default_rng(82)
    h, w = 96, 160
    lrs = [torch.from_numpy(rng.uniform(0, 255, (h, w)).astype(np.float32)).cuda() for _ in NOMINAL]
    plan = sr.Plan(4, h, w, NOMINAL, g["k31"])
    saa = plan.saa(lrs)
    a = saa.clone(); ea = plan.set_path(8).ibp_(lrs, a, 6)
    b = saa.clone(); eb = plan.set_path(2).ibp_(lrs, b, 6)
    plan.close()
    assert float((a - b).abs().max()) <= 2e-3
    _errs_close(ea, eb)
    with pytest.raises(sr.Sr459Error):
        sr.Plan(4, 32, 33, NOMINAL, g["k31"]).set_path(8)       # W % 4 != 0: dense path only


def test_clip_is_active_and_matches(sr):
    """S1 noise drives hr to both clip bounds (quirk Q5: literals 0 and 255)."""
    rng = np.random.default_rng(14)
    lrs = [rng.uniform(0, 255, (64, 96)) for _ in NOMINAL]
    psf = O.make_gaussian_psf()
    saa_o = O.shift_and_add(lrs, NOMINAL)
    hr_o, e_o = O.ibp_integer_shifts(lrs, NOMINAL, psf, saa_o, n_iter=30)
    assert (hr_o == 0).any() and (hr_o == 255).any()
    for path in (2, 3, 4, 7, 8):
        native, saa, hr, errs = _run_pipeline(sr, lrs, NOMINAL, psf, 30, path)
        _close(hr, hr_o, what=f"clip path{path}")
        _errs_close(errs, e_o)
        assert hr.min() >= 0.0 and hr.max() <= 255.0


# ---------------------------------------------------------------- full size, size-independent properties
def test_full_size_fast_equals_dense_and_properties(sr):
    """Config-1 shape (4x1536x2048 -> 3072x4096): the polyphase kernel against the literal dense
    kernel on the GPU, SAA linearity, and a few oracle-checked windows."""
    import torch
    rng = np.random.default_rng(459)
    h, w = 1536, 2048
    lrs = [torch.from_numpy(rng.uniform(0, 255, (h, w)).astype(np.float32)).cuda() for _ in NOMINAL]
    psf = O.make_gaussian_psf()
    plan = sr.Plan(4, h, w, NOMINAL, psf)
    saa = plan.saa(lrs)
    assert saa.shape == (3072, 4096)
    # linearity of SAA: saa(2a+b) == 2saa(a)+saa(b)
    other = [torch.from_numpy(rng.uniform(0, 255, (h, w)).astype(np.float32)).cuda() for _ in NOMINAL]
    mix = plan.saa([2 * a + b for a, b in zip(lrs, other)])
    assert float((mix - (2 * saa + plan.saa(other))).abs().max()) < 2e-3
    hr_fast = saa.clone(); e_fast = plan.set_path(3).ibp_(lrs, hr_fast, 6)
    hr_dense = saa.clone(); e_dense = plan.set_path(2).ibp_(lrs, hr_dense, 6)
    hr_sep = saa.clone(); e_sep = plan.set_path(4).ibp_(lrs, hr_sep, 6)
    with pytest.raises(sr.Sr459Error):                     # the superseded kernels are gone
        plan.set_path(5)
    plan.set_path(4)
    assert float((hr_fast - hr_dense).abs().max()) < 2e-3
    assert float((hr_sep - hr_dense).abs().max()) < 2e-3
    np.testing.assert_allclose(e_fast, e_dense, rtol=1e-6)
    np.testing.assert_allclose(e_sep, e_dense, rtol=1e-6)
    assert all(b < a for a, b in zip(e_fast, e_fast[1:])), "IBP residual must decrease"
    # full-size SAA against the oracle (integer shifts: zoom + clamp-gather)
    saa_o = O.shift_and_add_integer([l.cpu().numpy().astype(np.float64) for l in lrs], NOMINAL)
    _close(saa.cpu().numpy(), saa_o, what="full-size saa")
    # IBP is local (dependence radius 2r+2 = 8 px per iteration): oracle IBP on corner windows of the
    # GPU's own SAA image must agree away from the cut after 3 iterations
    hr3 = saa.clone(); plan.set_path(0).ibp_(lrs, hr3, 3)
    for (ys, xs) in [(slice(0, 96), slice(0, 128)), (slice(h - 96, h), slice(w - 128, w))]:
        crop = [l[ys, xs].cpu().numpy().astype(np.float64) for l in lrs]
        Ys = slice(2 * ys.start, 2 * ys.stop); Xs = slice(2 * xs.start, 2 * xs.stop)
        hr_c, _ = O.ibp_integer_shifts(crop, NOMINAL, psf, saa_o[Ys, Xs], n_iter=3)
        got = hr3[Ys, Xs].cpu().numpy()
        if ys.start == 0:
            _close(got[:160, :224], hr_c[:160, :224], what="ibp top-left window")
        else:
            _close(got[-160:, -224:], hr_c[-160:, -224:], what="ibp bottom-right window")
    plan.close()


# ---------------------------------------------------------------- row bands (multi-GPU path, emulated on one GPU)
@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("case", ["gauss7", "gauss5", "measured31", "two_lattices"])
def test_row_bands_emulated_vs_oracle(sr, world, case):
    """All ranks of a band decomposition in one process: band kernels + halo copies == full-image oracle.
    The halo exchange itself (torch.distributed) is covered by tests/test_bands_host.py over gloo."""
    from sr459.bands import BandPlan, run_emulated
    rng = np.random.default_rng(31)
    shifts = NOMINAL if case != "two_lattices" else [(0.0, 0.0)] + NOMINAL
    if case == "measured31":
        h, w, psf = 144, 80, golden("g_psf")["k31"]
    else:
        h, w = 150, 168
        psf = O.make_gaussian_psf(5, 0.9) if case == "gauss5" else O.make_gaussian_psf()
    lrs = [rng.uniform(0, 255, (h, w)) for _ in shifts]
    hr0 = O.shift_and_add_integer(lrs, shifts)
    hr_o, e_o = O.ibp_integer_shifts(lrs, shifts, psf, hr0, n_iter=4)
    plans = [BandPlan(len(shifts), h, w, shifts, psf, r, world) for r in range(world)]
    hr, errs = run_emulated(plans, lrs, hr0, 4)
    for p in plans:
        p.close()
    hr = hr.cpu().numpy()
    _close(hr, hr_o, what=f"bands {case} world {world}")
    for p in plans:                                            # band seams
        y = p.layout.own_y0
        if y:
            _close(hr[y - 8:y + 8], hr_o[y - 8:y + 8], what=f"seam at row {y}")
    _errs_close(errs, e_o)


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("case", ["gauss7", "measured31", "two_lattices", "fractional_k7", "measured7_int"])
def test_row_bands_peer_memory_path_vs_oracle(sr, world, case):
    """The product band path (csrc/bands.cu) with every band in this process: per-band SAA from LR rows only (e3), the
    iteration kernels on the local buffers (streaming kernel for gauss7, lattice kernels for the 31x31 PSF, the
    stage-by-stage iteration for fractional shifts) and the in-kernel halo exchange through peer pointers, against the
    full-image float64 oracle; seams checked separately."""
    from sr459.bands import BandPlan, run_emulated_peer
    rng = np.random.default_rng(32)
    g = golden("g_psf")
    shifts = {"two_lattices": [(0.0, 0.0)] + NOMINAL, "fractional_k7": CALIBRATED}.get(case, NOMINAL)
    if case == "measured31":
        h, w, psf = 144, 80, g["k31"]
    elif case == "fractional_k7":
        h, w, psf = 240, 96, g["k7"]
    else:
        h, w = 150, 168
        psf = g["k7"] if case == "measured7_int" else O.make_gaussian_psf()
    lrs = [rng.uniform(0, 255, (h, w)).astype(np.float32).astype(np.float64) for _ in shifts]
    integer = case != "fractional_k7"
    saa_o = O.shift_and_add_integer(lrs, shifts) if integer else O.shift_and_add(lrs, shifts)
    hr_o, e_o = (O.ibp_integer_shifts if integer else O.ibp)(lrs, shifts, psf, saa_o, n_iter=4)
    plans = [BandPlan(len(shifts), h, w, shifts, psf, r, world) for r in range(world)]
    assert plans[0].fractional == (not integer)
    saa, _ = run_emulated_peer(plans, lrs, 0)                       # SAA only: every band from its own LR rows
    _close(saa.cpu().numpy(), saa_o, what=f"banded SAA {case} world {world}")
    hr, errs = run_emulated_peer(plans, lrs, 4)
    seams = [p.layout.own_y0 for p in plans if p.layout.own_y0]
    for p in plans:
        p.close()
    hr = hr.cpu().numpy()
    _close(hr, hr_o, what=f"peer bands {case} world {world}")
    for y in seams:
        _close(hr[y - 8:y + 8], hr_o[y - 8:y + 8], what=f"seam at row {y}")
    _errs_close(errs, e_o)


def test_row_bands_peer_memory_graph_replay_two_gpus(sr):
    """One band per GPU inside one process (peer access instead of cudaIpc): each band's whole loop captured in a CUDA
    graph and replayed for a second job that re-uses plans, graphs and arenas.  Needs two GPUs: on ONE device a band's
    captured loop would wait for kernels of the other band that the device may only run afterwards (streams of one
    device can share a hardware queue); the single-GPU emulation interleaves iteration by iteration instead
    (run_emulated_peer), and bench.py --gpus N replays the graphs with one process per GPU."""
    import threading
    import torch
    from sr459.bands import BandPlan
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    rng = np.random.default_rng(33)
    h, w, K = 128, 96, 6
    psf = O.make_gaussian_psf()
    plans = [BandPlan(4, h, w, NOMINAL, psf, r, 2, device=r) for r in range(2)]
    plans[0].attach_local(None, plans[1]); plans[1].attach_local(plans[0], None)
    exts = None
    for job in range(2):
        lrs = [rng.uniform(0, 255, (h, w)) for _ in NOMINAL]
        saa_o = O.shift_and_add_integer(lrs, NOMINAL)
        hr_o, e_o = O.ibp_integer_shifts(lrs, NOMINAL, psf, saa_o, n_iter=K)
        if exts is None:
            exts = [[torch.from_numpy(l[p.layout.lr_ext_lo:p.layout.lr_ext_hi].astype(np.float32)).to(f"cuda:{p.device}") for l in lrs] for p in plans]
        else:
            for p, ext in zip(plans, exts):
                for t, l in zip(ext, lrs):
                    t.copy_(torch.from_numpy(l[p.layout.lr_ext_lo:p.layout.lr_ext_hi].astype(np.float32)))
        out = [None, None]

        def go(i):
            with torch.cuda.device(plans[i].device):
                plans[i].saa(exts[i])
                out[i] = plans[i].run_peer(exts[i], K, graph=(job == 1), reduce=False)
                torch.cuda.synchronize()
        ths = [threading.Thread(target=go, args=(i,)) for i in range(2)]
        [t.start() for t in ths]; [t.join() for t in ths]
        hr = np.zeros((2 * h, 2 * w), dtype=np.float32)
        for p in plans:
            lay = p.layout
            hr[lay.own_y0:lay.own_y1] = p.current()[lay.owned].cpu().numpy()
        _close(hr, hr_o, what=f"job {job}")
        _errs_close(((out[0] + out[1]) / (4.0 * h * w)).tolist(), e_o)
    for p in plans:
        p.close()


@pytest.mark.parametrize("lanes", [1, 2])
@pytest.mark.parametrize("use_graph", [False, True])
def test_burst_stream_matches_oracle(sr, use_graph, lanes):
    """Streaming burst mode (config 5): frames committed one by one, uint8 HR images out, ring re-use, one or two compute lanes."""
    from sr459.driver import BurstStream
    rng = np.random.default_rng(55)
    h, w, K = 48, 64, 5
    psf = O.make_gaussian_psf()
    sets = [[rng.integers(0, 256, (h, w), dtype=np.uint8) for _ in NOMINAL] for _ in range(5)]
    bs = BurstStream(4, h, w, NOMINAL, psf, K, depth=2, use_graph=use_graph, lanes=lanes)
    got = []
    for i, frames in enumerate(sets):
        for k, f in enumerate(frames):
            r = bs.push_frame(f)
            assert (r == i) if k == 3 else (r is None)
        if i >= 1:                                             # keep one set in flight
            idx, img, lat = bs.wait_result()
            got.append((idx, img.clone().numpy())); assert lat > 0
    idx, img, lat = bs.wait_result()
    got.append((idx, img.clone().numpy()))
    with pytest.raises(RuntimeError):
        bs.wait_result()
    bs.close()
    assert [g[0] for g in got] == list(range(5))
    for i, img in got:
        lrs = [f.astype(np.float64) for f in sets[i]]
        saa = O.shift_and_add_integer(lrs, NOMINAL)
        ref, _ = O.ibp_integer_shifts(lrs, NOMINAL, psf, saa, n_iter=K)
        ref8 = np.clip(ref, 0, 255).astype(np.uint8)
        # truncation makes a +-1 difference possible where the float value sits within the tolerance of an integer
        d = np.abs(img.astype(np.int16) - ref8.astype(np.int16))
        assert d.max() <= 1 and (d > 0).mean() < 1e-3, f"set {i}: {d.max()} / {(d > 0).mean()}"


@pytest.mark.parametrize("lanes", [1, 2])
@pytest.mark.parametrize("use_graph", [False, True])
@pytest.mark.parametrize("dt", ["f32", "u8"])
def test_frame_set_pipeline_matches_oracle(sr, dt, use_graph, lanes):
    """Batched driver (config 3): pinned host sets in, results out, slots re-used, optional CUDA-graph replay, one or two
    independent compute lanes (own plan and stream each; results arrive in submission order either way)."""
    import torch
    from sr459.driver import FrameSetPipeline, shard_indices
    rng = np.random.default_rng(56)
    h, w, K = 40, 64, 4
    psf = O.make_gaussian_psf()
    if dt == "u8":
        sets = [torch.from_numpy(rng.integers(0, 256, (4, h, w), dtype=np.uint8)).pin_memory() for _ in range(5)]
    else:
        sets = [torch.from_numpy(rng.uniform(0, 255, (4, h, w)).astype(np.float32)).pin_memory() for _ in range(5)]
    pipe = FrameSetPipeline(4, h, w, NOMINAL, psf, K, depth=2, in_dtype=dt, out_dtype=dt, use_graph=use_graph, lanes=lanes)
    res = pipe.run(sets, keep_results=True)
    again = pipe.run(sets[:2], keep_results=True)              # the ring is re-usable
    pipe.close()
    assert len(res) == 5 and len(again) == 2
    for i, r in enumerate(res):
        lrs = [sets[i][k].numpy().astype(np.float64) for k in range(4)]
        saa = O.shift_and_add_integer(lrs, NOMINAL)
        ref, _ = O.ibp_integer_shifts(lrs, NOMINAL, psf, saa, n_iter=K)
        if dt == "u8":
            d = np.abs(r.numpy().astype(np.int16) - np.clip(ref, 0, 255).astype(np.uint8).astype(np.int16))
            assert d.max() <= 1 and (d > 0).mean() < 1e-3
        else:
            _close(r.numpy(), ref, what=f"pipeline set {i}")
    assert torch.equal(again[0], res[0]) and torch.equal(again[1], res[1])
    assert shard_indices(10, 1, 4) == [1, 5, 9]


def test_mtf50_of_cuda_output_equals_mtf50_of_reference_output(sr):
    """Row f3: the slanted-edge MTF50 / MTF10 (analysis.ipynb cells 7-10) of the CUDA SAA+IBP image and of the
    float64 oracle image agree far inside the measurement's own resolution."""
    from sr459 import analysis as A
    rng = np.random.default_rng(77)
    n = 200
    yy, xx = np.mgrid[:n, :n].astype(np.float64)
    d = (xx - n / 2) * np.cos(0.25) + (yy - n / 2) * np.sin(0.25)
    truth = np.where(np.abs(d) < 9, 35.0, 215.0)
    psf = O.make_gaussian_psf()
    lrs = [np.clip(O.forward_model(truth, psf, s, 2) + rng.normal(0, 1, (n // 2, n // 2)), 0, 255) for s in NOMINAL]
    saa_o = O.shift_and_add(lrs, NOMINAL)
    hr_o, _ = O.ibp(lrs, NOMINAL, psf, saa_o, n_iter=10)
    hr, _ = sr.ibp(lrs, NOMINAL, psf, sr.shift_and_add(lrs, NOMINAL), n_iter=10)
    roi = (slice(20, 180), slice(20, 180))
    for img_a, img_b in ((hr, hr_o), (sr.shift_and_add(lrs, NOMINAL), saa_o)):
        a50, a10 = A.mtf50_mtf10(img_a, roi)
        b50, b10 = A.mtf50_mtf10(img_b, roi)
        assert abs(a50 - b50) <= 1e-3 * b50 and abs(a10 - b10) <= 1e-3 * b10, (a50, b50, a10, b10)


def test_band_plan_fractional_shifts_get_the_wide_halo(sr):
    from sr459.bands import BandPlan
    p = BandPlan(4, 256, 64, CALIBRATED, O.make_gaussian_psf(), 0, 2)
    assert p.fractional and p.layout.halo == 48          # 2 * (3 + 19) rounded up to 16
    p.close()
    with pytest.raises(sr.Sr459Error):                   # bands thinner than the halo
        BandPlan(4, 64, 64, CALIBRATED, O.make_gaussian_psf(), 0, 4)


def test_zoom_orders_and_factors_and_saa_order(sr):
    """ndi_zoom for spline orders 0 / 1 / 3 and integer, fractional and per-axis factors; shift_and_add(order != 3) (the order only
    reaches zoom, run_sr.py:185) -- against the oracle (pinned to live scipy in tests/test_oracle_pinning.py)."""
    rng = np.random.default_rng(41)
    for shape in [(33, 47), (7, 5), (64, 96), (2, 2)]:
        a = rng.uniform(0, 255, shape)
        for factor in (2, 3, 1.5, 2.37, (2, 1.5)):
            for order in (0, 1, 3):
                got = sr.ndi_zoom(a, factor, order)
                ref = O.ndi_zoom_order(a, factor, order)
                assert got.shape == ref.shape
                _close(got, ref, what=f"zoom {shape} x{factor} order {order}")
    with pytest.raises(NotImplementedError):
        sr.ndi_zoom(a, 2, order=2)
    lrs = [rng.uniform(0, 255, (40, 56)) for _ in NOMINAL]
    for order in (0, 1):
        for shifts in (NOMINAL, CALIBRATED):
            _close(sr.shift_and_add(lrs, shifts, 2, order), O.shift_and_add(lrs, shifts, 2, order), what=f"saa order {order}")


# ---------------------------------------------------------------- run_sr.py-compatible CLI on synthetic session folders
def _png(path, arr):
    from PIL import Image
    Image.fromarray(arr).save(path)


def _read_png(path):
    from PIL import Image
    return np.array(Image.open(path))


def _u8_close(got, want_f64, what):
    want = O.to_uint8(want_f64)
    assert got.shape == want.shape and got.dtype == np.uint8, what
    d = np.abs(got.astype(int) - want.astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 1e-3, f"{what}: {int((d > 0).sum())} pixels differ, max {d.max()}"


@pytest.mark.parametrize("experiment", ["mono_cal_target", "rgb_cal_target", "mono_barcodes", "rgb_barcodes"])
def test_run_sr_cli_writes_reference_outputs(sr, tmp_path, experiment):
    """Same directory walk, file names and pixel values as the reference's four run_sr.py scripts
    (main/process_session/load_session, mono_cal_target/run_sr.py:78-99, :262-366 and siblings)."""
    import json
    from sr459 import run_sr
    rng = np.random.default_rng(41)
    data, out = tmp_path / "data", tmp_path / "results"
    sess = data / "session_a"
    sess.mkdir(parents=True)
    (data / "not_a_session.txt").write_text("x")
    red = experiment.startswith("rgb")
    H0, W0 = (72, 104) if red else (36, 52)
    n_iter = 6
    jobs = []                                                 # (out_dir, frames f64, shifts, lr-mean name)
    if experiment == "mono_cal_target":
        names = ["center.png", "shift_0.png", "shift_1.png", "shift_2.png", "shift_3.png"]
        raws = [rng.integers(0, 256, (H0, W0), dtype=np.uint8) for _ in names]
        for nme, r in zip(names, raws):
            _png(sess / nme, r)
        jobs.append((out / "session_a", [r.astype(np.float64) for r in raws], [(0.0, 0.0)] + NOMINAL, "LR_mean.png"))
    elif experiment == "rgb_cal_target":
        meta = {"expected_shifts": {lab: {"dy_px": 2 * s[0], "dx_px": 2 * s[1]} for lab, s in zip(run_sr.CORNER_ORDER, CALIBRATED)}}
        (sess / "metadata.json").write_text(json.dumps(meta))
        frames = []
        for c in range(4):
            reps = [rng.integers(0, 256, (H0, W0), dtype=np.uint8) for _ in range(3)]
            for r, raw in enumerate(reps):
                _png(sess / f"corner{c}_rep{r:02d}.png", raw)
            frames.append(np.stack([O.extract_red(x.astype(np.float64)) for x in reps]).mean(axis=0))
        jobs.append((out / "session_a", frames, CALIBRATED, "LR_red_mean.png"))
    else:
        for r in range(2):
            raws = [rng.integers(0, 256, (H0, W0), dtype=np.uint8) for _ in range(4)]
            for c, raw in enumerate(raws):
                _png(sess / f"corner{c}_rep{r:02d}.png", raw)
            fr = [O.extract_red(x.astype(np.float64)) if red else x.astype(np.float64) for x in raws]
            jobs.append((out / "session_a" / f"rep{r}", fr, NOMINAL, "LR_red_mean.png" if red else "LR_mean.png"))
    argv = ["--experiment", experiment, "--data-dir", str(data), "--output-dir", str(out), "--iterations", str(n_iter)]
    assert run_sr.main(argv) == 0
    psf = O.make_gaussian_psf()
    for out_dir, frames, shifts, lr_name in jobs:
        assert (out_dir / "done.flag").exists()
        native_o = O.native_2x(np.mean(frames, axis=0))
        saa_o = O.shift_and_add(frames, shifts)
        hr_o, e_o = O.ibp(frames, shifts, psf, saa_o, n_iter=n_iter)
        _u8_close(_read_png(out_dir / "native_2x.png"), native_o, "native_2x.png")
        _u8_close(_read_png(out_dir / "SAA.png"), saa_o, "SAA.png")
        _u8_close(_read_png(out_dir / "SAA_IBP.png"), hr_o, "SAA_IBP.png")
        assert np.array_equal(_read_png(out_dir / lr_name), O.to_uint8(np.mean(frames, axis=0))), lr_name   # exact: float64 means
        _errs_close(json.loads((out_dir / "errors.json").read_text())["mse_per_iteration"], e_o)
        if experiment == "rgb_cal_target":
            sj = json.loads((out_dir / "shifts.json").read_text())
            assert sj["corner_labels"] == run_sr.CORNER_ORDER and np.allclose(sj["shifts_lr_yx"], CALIBRATED)
    # second run: everything is skipped (done.flag), nothing is rewritten
    stamp = {p: p.stat().st_mtime_ns for p in out.rglob("SAA_IBP.png")}
    assert run_sr.main(argv) == 0
    assert stamp == {p: p.stat().st_mtime_ns for p in out.rglob("SAA_IBP.png")}


def test_fused_ingest_matches_numpy_order_of_operations(sr):
    """sr459_ingest_groups == load_gray + extract_red + stack.mean(axis=0) + np.mean(frames, axis=0) in float64."""
    rng = np.random.default_rng(43)
    for (H0, W0, n, reps, red) in [(37, 53, 4, 5, True), (36, 52, 5, 1, False), (8, 8, 1, 3, True), (33, 47, 8, 4, False)]:
        groups = [[rng.integers(0, 256, (H0, W0), dtype=np.uint8) for _ in range(reps)] for _ in range(n)]
        frames, mean, mean_u8 = sr.ingest_groups(groups, red=red)
        ref = [np.stack([(O.extract_red(x.astype(np.float64)) if red else x.astype(np.float64)) for x in g]).mean(axis=0) for g in groups]
        ref_mean = np.mean(ref, axis=0)
        for f, r in zip(frames, ref):
            assert np.array_equal(f.cpu().numpy(), r.astype(np.float32))
        assert np.array_equal(mean.cpu().numpy(), ref_mean.astype(np.float32))
        assert np.array_equal(mean_u8.cpu().numpy(), O.to_uint8(ref_mean))
    # unequal repetition counts per corner and more than 32 files in one set: the reference's per-corner
    # np.stack(...).mean(axis=0) takes both (rgb_cal_target/run_sr.py:100-108)
    for counts in ([1, 2], [5, 9, 3, 7], [6] * 8):
        groups = [[rng.integers(0, 256, (20, 28), dtype=np.uint8) for _ in range(c)] for c in counts]
        frames, mean, mean_u8 = sr.ingest_groups(groups, red=True)
        ref = [np.stack([O.extract_red(x.astype(np.float64)) for x in g]).mean(axis=0) for g in groups]
        for f, r in zip(frames, ref):
            assert np.array_equal(f.cpu().numpy(), r.astype(np.float32))
        assert np.array_equal(mean_u8.cpu().numpy(), O.to_uint8(np.mean(ref, axis=0)))
    with pytest.raises(ValueError):
        sr.ingest_groups([[np.zeros((4, 4), np.uint8)], []])


def test_run_sr_cli_errors(sr, tmp_path):
    from sr459 import run_sr
    (tmp_path / "data" / "empty").mkdir(parents=True)
    with pytest.raises(FileNotFoundError):
        run_sr.main(["--experiment", "mono_cal_target", "--data-dir", str(tmp_path / "data"), "--output-dir", str(tmp_path / "o")])
    with pytest.raises(FileNotFoundError):
        run_sr.main(["--experiment", "mono_barcodes", "--data-dir", str(tmp_path / "data"), "--output-dir", str(tmp_path / "o")])


def test_fractional_shift_interior_polyphase_kernel_vs_oracle_and_spline_path(sr):
    """Fractional shifts at a size whose back-projection has interior tiles (k_fir_poly, fir.cu) as well as border
    tiles (k_fir_bp): both against the float64 oracle and against the stage-by-stage spline path (path 1)."""
    import torch
    rng = np.random.default_rng(4592)
    h, w = 176, 336
    lrs64 = [rng.uniform(0, 255, (h, w)) for _ in range(4)]
    psf = O.make_gaussian_psf()
    lrs = [torch.from_numpy(l.astype(np.float32)).cuda() for l in lrs64]
    plan = sr.Plan(4, h, w, CALIBRATED, psf)
    saa = plan.saa(lrs)
    a = saa.clone(); ea = plan.set_path(0).ibp_(lrs, a, 4)
    b = saa.clone(); eb = plan.set_path(1).ibp_(lrs, b, 4)
    plan.close()
    _close(a.cpu().numpy(), b.cpu().numpy(), what="polyphase + table FIR vs spline stages")
    np.testing.assert_allclose(ea, eb, rtol=1e-5)
    saa_o = O.shift_and_add([l.astype(np.float32).astype(np.float64) for l in lrs64], CALIBRATED)
    hr_o, errs_o = O.ibp([l.astype(np.float32).astype(np.float64) for l in lrs64], CALIBRATED, psf, saa_o, n_iter=4)
    _close(saa.cpu().numpy(), saa_o, what="saa")
    _close(a.cpu().numpy(), hr_o, what="ibp (polyphase interior) vs oracle")
    _errs_close(ea, errs_o)


def test_plans_on_two_devices_of_one_process(sr):
    """sr459_plan_create takes a device: kernel attributes (opt-in shared memory) are per device, so a second device
    in the same process must configure them again.  Integer and fractional paths, bit-identical to device 0."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    rng = np.random.default_rng(11)
    h, w = 176, 336
    lrs64 = [rng.uniform(0, 255, (h, w)).astype(np.float32) for _ in range(4)]
    out = {}
    for dev in (0, 1):
        with torch.cuda.device(dev):
            for tag, shifts in (("int", NOMINAL), ("frac", CALIBRATED)):
                lrs = [torch.from_numpy(l).cuda() for l in lrs64]
                plan = sr.Plan(4, h, w, shifts, O.make_gaussian_psf())
                hr = plan.saa(lrs)
                errs = plan.ibp_(lrs, hr, 3)
                out[(dev, tag)] = (hr.cpu().numpy(), errs)
                plan.close()
    for tag in ("int", "frac"):
        assert np.array_equal(out[(0, tag)][0], out[(1, tag)][0]), tag
        np.testing.assert_allclose(out[(0, tag)][1], out[(1, tag)][1], rtol=1e-12)
