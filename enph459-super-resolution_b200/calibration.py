"""Pinhole / PSF characterisation -- the upstream producers of `shifts` and of the measured PSF, on the device.

Same names, arguments and return values as the reference's `data_collection/psf_mtf_utils.py` (`find_peak` :48-51,
`extract_psf` :54-71, `fit_gaussian_psf` :115-131, `compute_mtf` :134-170) and `find_pinhole_center` of
`calibration_beam_shift/calibrate_shift_grid.py:66-96`, plus the batched forms a calibration run wants (30 pinhole
frames per XPR corner in one launch: `find_peaks`, `find_pinhole_centers`).  Everything is computed by libsr459
(csrc/calib.cu: float64 Gaussian smoothing + first-maximum argmax, ring percentile by bitonic sort, the bounded
7-parameter Gaussian fit as a projected Levenberg-Marquardt iteration inside one CTA per frame, zero-skipping 2-D DFT
+ ring means); there is no host fallback.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .ops import _is_tensor, _stream_ptr, _torch

PSF_CROP_RADIUS = 15            # calibrate_shift_grid.py: radius handed to extract_psf
_ROI = 31
_PD = C.POINTER(C.c_double)
_PI = C.POINTER(C.c_int)


def _stack(imgs):
    """(n, H, W) uint8 or float32 CUDA tensor from one image, a list of images or a stack."""
    torch = _torch()
    if _is_tensor(imgs):
        t = imgs if imgs.is_cuda else imgs.cuda()
    else:
        a = np.asarray(imgs)
        if a.dtype != np.uint8:
            a = a.astype(np.float32)
        t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    if t.dtype != torch.uint8:
        t = t.to(torch.float32)
    if t.dim() == 2:
        t = t[None]
    if t.dim() != 3:
        raise ValueError(f"expected an image or a stack of images, got shape {tuple(t.shape)}")
    return t.contiguous(), int(t.dtype != torch.uint8)


def find_peaks(imgs, smooth_sigma: float = 2.0) -> np.ndarray:
    """(n, 2) int array of (row, col): first maximum of each lightly smoothed image -- psf_mtf_utils.py:48-51, batched."""
    lib = _lib.load()
    t, is_f32 = _stack(imgs)
    n, H, W = t.shape
    flat = np.zeros(n, dtype=np.int32)
    _lib.check(lib.sr459_find_peaks(C.c_void_p(t.data_ptr()), is_f32, n, H, W, float(smooth_sigma), flat.ctypes.data_as(_PI), None,
                                    _stream_ptr(t.device)))
    return np.stack([flat // W, flat % W], axis=1).astype(np.int64)


def find_peak(img, smooth_sigma: float = 2.0):
    """(row, col) of the first maximum of the lightly smoothed image -- psf_mtf_utils.py:48-51."""
    r, c = find_peaks(img, smooth_sigma)[0]
    return int(r), int(c)


def _fit(imgs, peaks, radius, bg_percentile, max_iter):
    lib = _lib.load()
    t, is_f32 = _stack(imgs)
    n, H, W = t.shape
    pk = np.ascontiguousarray([int(r) * W + int(c) for r, c in np.asarray(peaks).reshape(-1, 2)], dtype=np.int32)
    if len(pk) != n:
        raise ValueError("one peak per image expected")
    out = np.zeros((n, 12)); roi = np.zeros((n, _ROI * _ROI)); dims = np.zeros((n, 4), dtype=np.int32)
    _lib.check(lib.sr459_pinhole_fit(C.c_void_p(t.data_ptr()), is_f32, n, H, W, pk.ctypes.data_as(_PI), int(radius), float(bg_percentile),
                                     int(max_iter), out.ctypes.data_as(_PD), roi.ctypes.data_as(_PD), dims.ctypes.data_as(_PI),
                                     _stream_ptr(t.device)))
    rois = [roi[i, :dims[i, 0] * dims[i, 1]].reshape(dims[i, 0], dims[i, 1]).copy() for i in range(n)]
    return out, rois, dims


def extract_psf(img, center, radius, bg_percentile: float = 50.0):
    """Background-subtracted ROI around `center`; the background is a percentile of the outer ring
    (pixels at least 0.6 * radius from the ROI centre) -- psf_mtf_utils.py:54-71."""
    _, rois, _ = _fit(img, [(int(center[0]), int(center[1]))], radius, bg_percentile, 1)
    return rois[0]


def gauss2d(xy, amp, x0, y0, sigma_x, sigma_y, theta, offset):
    """Rotated 2-D Gaussian, flattened -- psf_mtf_utils.py:105-112 (the model the device fit evaluates; host helper for plots)."""
    x, y = xy
    ct, st, s2 = np.cos(theta) ** 2, np.sin(theta) ** 2, np.sin(2 * theta)
    a = ct / (2 * sigma_x ** 2) + st / (2 * sigma_y ** 2)
    b = -s2 / (4 * sigma_x ** 2) + s2 / (4 * sigma_y ** 2)
    c = st / (2 * sigma_x ** 2) + ct / (2 * sigma_y ** 2)
    dx, dy = x - x0, y - y0
    return (offset + amp * np.exp(-(a * dx ** 2 + 2 * b * dx * dy + c * dy ** 2))).ravel()


def fit_gaussian_psf(psf, max_iter: int = 400):
    """Bounded least-squares fit of gauss2d to an (already background-subtracted) ROI of at most 31 x 31.
    Returns (params, fit_image) or (None, None) -- psf_mtf_utils.py:115-131 (same start point and bounds)."""
    lib = _lib.load()
    torch = _torch()
    psf = np.ascontiguousarray(psf, dtype=np.float64)
    h, w = psf.shape
    if h > _ROI or w > _ROI:
        raise ValueError(f"ROI {psf.shape} larger than {_ROI} x {_ROI}")
    t = torch.from_numpy(psf).cuda()                   # float64 ROI, taken as it is (bg_percentile < 0)
    pk = np.array([(h // 2) * w + w // 2], dtype=np.int32)
    out = np.zeros((1, 12))
    _lib.check(lib.sr459_pinhole_fit(C.c_void_p(t.data_ptr()), 2, 1, h, w, pk.ctypes.data_as(_PI), max(h, w) // 2, -1.0, int(max_iter),
                                     out.ctypes.data_as(_PD), None, None, _stream_ptr(t.device)))
    if out[0, 9] != 0.0:
        return None, None
    p = out[0, 2:9].copy()
    yy, xx = np.mgrid[:h, :w]
    return p, gauss2d((xx, yy), *p).reshape(h, w)


def find_pinhole_centers(imgs, crop_radius: int = PSF_CROP_RADIUS, smooth_sigma: float = 2.0, max_iter: int = 400):
    """Batched find_pinhole_center: dict with `center` (n, 2) as (cx, cy), `popt` (n, 7; NaN without a fit),
    `status` (0 Gaussian fit, 1 centre of mass, 2 peak pixel) and `peak` (n, 2) as (row, col)."""
    t, _ = _stack(imgs)
    peaks = find_peaks(t, smooth_sigma)
    out, _, _ = _fit(t, peaks, crop_radius, 50.0, max_iter)
    return {"center": out[:, 0:2].copy(), "popt": out[:, 2:9].copy(), "status": out[:, 9].astype(int), "peak": peaks}


def find_pinhole_center(img, crop_radius: int = PSF_CROP_RADIUS):
    """Sub-pixel (cx, cy) of a bright pinhole: Gaussian fit of the background-subtracted ROI around the peak,
    centre of mass above 10 % of the ROI maximum if the fit fails -- calibrate_shift_grid.py:66-96."""
    r = find_pinhole_centers(img, crop_radius)
    return float(r["center"][0, 0]), float(r["center"][0, 1])


def compute_mtf(psf, pixel_pitch_um=None):
    """MTF from the PSF: zero-pad to >= 256, 2-D DFT magnitude, radial average.
    Returns (freq, mtf_radial, mtf_2d, freq_label, nyquist) -- psf_mtf_utils.py:134-170."""
    lib = _lib.load()
    psf = np.ascontiguousarray(psf, dtype=np.float64)
    pad = max(256, psf.shape[0], psf.shape[1])
    pad += pad & 1
    mtf2d = np.zeros((pad, pad)); prof = np.zeros(pad // 2)
    _lib.check(lib.sr459_psf_mtf(psf.ctypes.data_as(_PD), psf.shape[0], psf.shape[1], pad, mtf2d.ctypes.data_as(_PD),
                                 prof.ctypes.data_as(_PD), _stream_ptr()))
    cpp = np.arange(pad // 2).astype(float) / pad
    if pixel_pitch_um is not None:
        return cpp / (pixel_pitch_um * 1e-3), prof, mtf2d, "cycles/mm", 1.0 / (2.0 * pixel_pitch_um * 1e-3)
    return cpp, prof, mtf2d, "cycles/pixel", 0.5


def zoom_crop(img, center, half_width):
    """Square crop of img around (row, col) -- psf_mtf_utils.py:189-197 (pure indexing)."""
    cy, cx = int(round(center[0])), int(round(center[1]))
    h, w = img.shape
    return img[max(cy - half_width, 0):min(cy + half_width + 1, h), max(cx - half_width, 0):min(cx + half_width + 1, w)]
