"""CPU oracle for the SAA / SAA+IBP hot path  --  TEST INFRASTRUCTURE ONLY.

This file is the parity checker for the CUDA engine in
``enph459-super-resolution_b200``.  It is NOT part of the product: only
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import it.  The product path
never routes through this module (and fails loudly if its CUDA library is
missing).

What is restated here
---------------------
The reference (benedikthoward/ENPH459-Super-Resolution) keeps its SR arithmetic
in four isomorphic scripts; citations below use ``mono_cal_target/run_sr.py``
(the same code sits at ``rgb_cal_target/run_sr.py:171-223``,
``mono_barcodes/run_sr.py:188-240``, ``rgb_barcodes/run_sr.py:201-255``).
The numerics themselves live in a third-party dependency that is NOT vendored
in the reference tree: **scipy, lock-pinned 1.17.0** (``uv.lock:994-995``;
pulled in transitively by scikit-image), specifically

* ``scipy.ndimage.zoom``    (call sites ``run_sr.py:185``, ``:279``)
* ``scipy.ndimage.shift``   (call sites ``run_sr.py:163``, ``:176``, ``:186``)
* ``scipy.signal.fftconvolve(mode='same')`` (call site ``run_sr.py:158``)

Their published algorithms are restated below in plain numpy float64:
cubic B-spline prefilter (pole sqrt(3)-2, gain 6, scipy's ``mirror`` /
``reflect`` boundary initialisations), the 4-tap cubic B-spline evaluation with
mirrored tap indices, scipy's 12-sample edge pre-padding for
``mode='nearest'``, align-corner zoom coordinates, and a zero-padded direct
convolution for ``fftconvolve(..., 'same')``.

Parity pinning (see tests/test_oracle_pinning.py and tests/golden/):
* every primitive here is checked against the live scipy in the image;
* every pipeline function is checked against vectors produced by importing the
  UNMODIFIED reference scripts from /root/reference
  (tests/golden/make_golden.py is the generator, the .npz files are committed);
* Native-2x and SAA are additionally checked bit-exactly (after the uint8
  truncating cast) against the reference's committed result PNGs when
  /root/reference is present.
The reference ships no valid stored SAA+IBP golden (all committed SAA_IBP.png
are stale), so IBP is pinned by live execution of the
reference ``ibp()`` only.
"""
from __future__ import annotations

import math
import os
from typing import Iterable, List, Sequence, Tuple

import numpy as np

# --------------------------------------------------------------------------
# constants of the cubic B-spline machinery (scipy ni_splines.c)
# --------------------------------------------------------------------------
POLE = math.sqrt(3.0) - 2.0          # single pole of the order-3 prefilter
GAIN = (1.0 - POLE) * (1.0 - 1.0 / POLE)   # == 6
NPAD_NEAREST = 12                    # scipy _prepad_for_spline_filter, mode='nearest'


# --------------------------------------------------------------------------
# 1. prefilter
# --------------------------------------------------------------------------
def spline_prefilter_axis(a: np.ndarray, axis: int, boundary: str) -> np.ndarray:
    """Order-3 B-spline prefilter along one axis (scipy ``spline_filter1d``).

    ``boundary`` selects scipy's initialisation of the causal/anticausal
    recursions: ``'mirror'`` is what scipy uses for extend modes
    mirror/constant/grid-constant/wrap (this is the ``zoom`` path, whose default
    mode is 'constant'); ``'reflect'`` is what it uses for nearest/reflect (the
    ``shift(mode='nearest')`` path).
    """
    c = np.moveaxis(np.array(a, dtype=np.float64, copy=True), axis, 0)
    n = c.shape[0]
    if n < 2:
        return np.moveaxis(c, 0, axis)
    z = POLE
    c *= GAIN
    if boundary == "mirror":
        zn = z ** (n - 1)
        head = c[0] + zn * c[n - 1]
        zi = z
        for i in range(1, n - 1):
            if zi == 0.0 or abs(zi) < 1e-320:
                break
            head = head + zi * (c[i] + zn * c[n - 1 - i])
            zi *= z
        c[0] = head / (1.0 - zn * zn)
    elif boundary == "reflect":
        zn = z ** n
        first = c[0].copy()
        head = c[0] + zn * c[n - 1]
        zi = z
        for i in range(1, n):
            if zi == 0.0 or abs(zi) < 1e-320:
                break
            head = head + zi * (c[i] + zn * c[n - 1 - i])
            zi *= z
        c[0] = head * (z / (1.0 - zn * zn)) + first
    else:
        raise ValueError(f"unknown boundary {boundary!r}")
    for i in range(1, n):
        c[i] += z * c[i - 1]
    if boundary == "mirror":
        c[n - 1] = (z * c[n - 2] + c[n - 1]) * (z / (z * z - 1.0))
    else:
        c[n - 1] *= z / (z - 1.0)
    for i in range(n - 2, -1, -1):
        c[i] = z * (c[i + 1] - c[i])
    return np.moveaxis(c, 0, axis)


def spline_prefilter(a: np.ndarray, boundary: str) -> np.ndarray:
    """N-d prefilter = successive 1-d prefilters, axis 0 first (scipy ``spline_filter``)."""
    out = np.asarray(a, dtype=np.float64)
    for ax in range(out.ndim):
        out = spline_prefilter_axis(out, ax, boundary)
    return out


# --------------------------------------------------------------------------
# 2. cubic evaluation with per-axis tap tables
# --------------------------------------------------------------------------
def cubic_weights(t: np.ndarray) -> np.ndarray:
    """Cubic B-spline weights for taps floor(x)-1 .. floor(x)+2, t = x - floor(x)."""
    t = np.asarray(t, dtype=np.float64)
    u = 1.0 - t
    w0 = u * u * u / 6.0
    w1 = (t * t * (t - 2.0) * 3.0 + 4.0) / 6.0
    w2 = (u * u * (u - 2.0) * 3.0 + 4.0) / 6.0
    w3 = 1.0 - w0 - w1 - w2
    return np.stack([w0, w1, w2, w3], axis=-1)


def mirror_index(idx: np.ndarray, n: int) -> np.ndarray:
    """Whole-sample symmetric index folding onto [0, n-1] (period 2n-2)."""
    idx = np.asarray(idx, dtype=np.int64)
    if n <= 1:
        return np.zeros_like(idx)
    p = 2 * n - 2
    m = np.mod(idx, p)
    return np.where(m >= n, p - m, m)


def axis_tables(coords: np.ndarray, n: int, valid: np.ndarray | None = None, fold: str = "mirror"):
    """Tap indices folded onto [0,n-1] and weights for real coordinates.

    ``fold='mirror'`` (zoom) mirrors out-of-range taps; ``fold='clamp'`` (shift,
    mode='nearest') clamps them.  For the shift path taps only leave the padded
    extent when |shift| > 11 samples, far outside the reference's usage
    (|shift| <= 1 HR px); there scipy is matched to ~1e-5 on [0,255] data, and
    to ~2e-13 for |shift| <= 11.
    """
    coords = np.asarray(coords, dtype=np.float64)
    base = np.floor(coords)
    w = cubic_weights(coords - base)
    start = base.astype(np.int64) - 1
    taps = start[:, None] + np.arange(4)[None, :]
    idx = mirror_index(taps, n) if fold == "mirror" else np.clip(taps, 0, n - 1)
    if valid is not None:
        w = w * valid[:, None]
    return idx, w


def apply_tables(coef: np.ndarray, ty, tx) -> np.ndarray:
    """Separable 4x4-tap gather: rows first, then columns."""
    iy, wy = ty
    ix, wx = tx
    tmp = np.zeros((iy.shape[0], coef.shape[1]), dtype=np.float64)
    for j in range(4):
        tmp += wy[:, j, None] * coef[iy[:, j], :]
    out = np.zeros((iy.shape[0], ix.shape[0]), dtype=np.float64)
    for j in range(4):
        out += wx[None, :, j] * tmp[:, ix[:, j]]
    return out


# --------------------------------------------------------------------------
# 3. scipy.ndimage.zoom / shift restated
# --------------------------------------------------------------------------
def zoom_coords(n_in: int, n_out: int):
    """Input coordinate of every output sample for ``zoom`` (grid_mode=False).

    scipy: zoom = (n_in-1)/(n_out-1); coordinate = o*zoom; mode 'constant'
    maps coordinates outside [0, n_in-1] to cval=0.
    """
    o = np.arange(n_out, dtype=np.float64)
    ratio = (n_in - 1) / (n_out - 1) if n_out > 1 else 1.0
    x = o * ratio
    valid = ((x >= 0.0) & (x <= n_in - 1)).astype(np.float64)
    return x, valid


def ndi_zoom_cubic(a: np.ndarray, factor) -> np.ndarray:
    """``scipy.ndimage.zoom(a, factor, order=3)`` with all other args default."""
    a = np.asarray(a, dtype=np.float64)
    fy, fx = (factor, factor) if np.isscalar(factor) else factor
    oh, ow = int(round(a.shape[0] * fy)), int(round(a.shape[1] * fx))
    coef = spline_prefilter(a, "mirror")
    xy, vy = zoom_coords(a.shape[0], oh)
    xx, vx = zoom_coords(a.shape[1], ow)
    return apply_tables(coef, axis_tables(xy, a.shape[0], vy), axis_tables(xx, a.shape[1], vx))


def ndi_zoom_order(a: np.ndarray, factor, order: int = 3) -> np.ndarray:
    """``scipy.ndimage.zoom(a, factor, order=order)`` for order 0 (nearest: floor(x + 0.5)), 1 (linear) and 3; no prefilter below
    order 2, samples outside the array are cval = 0 (mode='constant')."""
    if order == 3:
        return ndi_zoom_cubic(a, factor)
    if order not in (0, 1):
        raise NotImplementedError("oracle restates orders 0, 1 and 3")
    a = np.asarray(a, dtype=np.float64)
    h, w = a.shape
    fy, fx = (factor, factor) if np.isscalar(factor) else factor
    oh, ow = int(round(h * fy)), int(round(w * fx))

    def taps(n, m):
        x, valid = zoom_coords(n, m)
        if order == 0:
            return [(np.floor(x + 0.5).astype(int), valid)]
        fl = np.floor(x)
        return [(fl.astype(int), (1.0 - (x - fl)) * valid), (fl.astype(int) + 1, (x - fl) * valid)]

    out = np.zeros((oh, ow))
    for iy, wy in taps(h, oh):
        for ix, wx in taps(w, ow):
            wy_in, wx_in = wy * ((iy >= 0) & (iy < h)), wx * ((ix >= 0) & (ix < w))
            out += wy_in[:, None] * wx_in[None, :] * a[np.clip(iy, 0, h - 1)][:, np.clip(ix, 0, w - 1)]
    return out


def shift_coords(n: int, s: float, n_out: int | None = None, stride: int = 1):
    """Padded-array coordinate of output samples for ``shift(mode='nearest')``."""
    n_out = n if n_out is None else n_out
    o = np.arange(0, n_out * stride, stride, dtype=np.float64)
    x = o - s + NPAD_NEAREST
    return np.clip(x, 0.0, n + 2 * NPAD_NEAREST - 1)


def ndi_shift_cubic_nearest(a: np.ndarray, shift_yx: Sequence[float]) -> np.ndarray:
    """``scipy.ndimage.shift(a, shift_yx, order=3, mode='nearest')``."""
    a = np.asarray(a, dtype=np.float64)
    pad = np.pad(a, NPAD_NEAREST, mode="edge")
    coef = spline_prefilter(pad, "reflect")
    ty = axis_tables(shift_coords(a.shape[0], float(shift_yx[0])), pad.shape[0], fold="clamp")
    tx = axis_tables(shift_coords(a.shape[1], float(shift_yx[1])), pad.shape[1], fold="clamp")
    return apply_tables(coef, ty, tx)


# --------------------------------------------------------------------------
# 4. fftconvolve(..., 'same') restated as a direct zero-padded convolution
# --------------------------------------------------------------------------
def conv2_same_zero(img: np.ndarray, kernel: np.ndarray) -> np.ndarray:
    """True convolution, zero padded, centred crop (== fftconvolve mode='same')."""
    img = np.asarray(img, dtype=np.float64)
    k = np.asarray(kernel, dtype=np.float64)
    kh, kw = k.shape
    cy, cx = (kh - 1) // 2, (kw - 1) // 2
    H, W = img.shape
    pad = np.zeros((H + kh - 1, W + kw - 1), dtype=np.float64)
    pad[kh - 1 - cy: kh - 1 - cy + H, kw - 1 - cx: kw - 1 - cx + W] = img
    out = np.zeros((H, W), dtype=np.float64)
    # out[y,x] = sum_{u,v} k[u,v] * img[y + cy - u, x + cx - v]
    for u in range(kh):
        for v in range(kw):
            if k[u, v] == 0.0:
                continue
            out += k[u, v] * pad[kh - 1 - u: kh - 1 - u + H, kw - 1 - v: kw - 1 - v + W]
    return out


# --------------------------------------------------------------------------
# 5. the reference's SR operators
# --------------------------------------------------------------------------
def make_gaussian_psf(size: int = 7, sigma: float = 1.0) -> np.ndarray:
    """mono_cal_target/run_sr.py:104-111."""
    hw = size // 2
    ax = np.arange(-hw, hw + 1, dtype=np.float64)
    yy, xx = np.meshgrid(ax, ax, indexing="ij")
    k = np.exp(-(xx ** 2 + yy ** 2) / (2.0 * sigma ** 2))
    return k / k.sum()


def measured_psf_from_images(images: Iterable[np.ndarray], halfwidth: int = 3):
    """mono_cal_target/run_sr.py:114-152, with PNG decoding factored out.

    ``images`` are the decoded pos4_(0,0) pinhole frames in sorted directory
    order.  Returns (kernel, peaks) where peaks lists the (row, col) argmax of
    every image that was accepted (integer decisions the CUDA path must match
    exactly).
    """
    margin = halfwidth + 6
    patches, peaks = [], []
    for img in images:
        img = np.asarray(img, dtype=np.float64)
        pr, pc = np.unravel_index(int(np.argmax(img)), img.shape)
        if pr < margin or pr + margin + 1 > img.shape[0] or pc < margin or pc + margin + 1 > img.shape[1]:
            continue
        patches.append(img[pr - margin: pr + margin + 1, pc - margin: pc + margin + 1].copy())
        peaks.append((int(pr), int(pc)))
    if not patches:
        raise FileNotFoundError("no usable PSF images")
    avg = np.mean(patches, axis=0)
    k = avg[margin - halfwidth: margin + halfwidth + 1, margin - halfwidth: margin + halfwidth + 1].copy()
    corners = np.concatenate([k[:3, :3].ravel(), k[:3, -3:].ravel(), k[-3:, :3].ravel(), k[-3:, -3:].ravel()])
    k -= np.mean(corners)
    k = np.clip(k, 0, None)
    k /= k.sum()
    return k, peaks


def load_gray(path: str) -> np.ndarray:
    """mono_cal_target/run_sr.py:73-75."""
    from PIL import Image
    img = np.array(Image.open(path), dtype=np.float64)
    return img.mean(axis=2) if img.ndim == 3 else img


def load_measured_psf(psf_dir: str, halfwidth: int = 3) -> np.ndarray:
    """Directory walk of mono_cal_target/run_sr.py:118-126 + measured_psf_from_images."""
    imgs = []
    for d in sorted(os.listdir(psf_dir)):
        p = os.path.join(psf_dir, d, "pos4_(0,0).png")
        if os.path.isdir(os.path.join(psf_dir, d)) and os.path.exists(p):
            imgs.append(load_gray(p))
    return measured_psf_from_images(imgs, halfwidth)[0]


def extract_red(img: np.ndarray) -> np.ndarray:
    """rgb_cal_target/run_sr.py:73-75 (RGGB red sites)."""
    return np.asarray(img)[0::2, 0::2].copy()


def blur(img, kernel):
    """run_sr.py:157-158."""
    return conv2_same_zero(img, kernel)


def forward_model(hr, kernel, shift_yx, factor):
    """run_sr.py:161-165."""
    b = blur(hr, kernel)
    s = ndi_shift_cubic_nearest(b, (shift_yx[0] * factor, shift_yx[1] * factor))
    return s[::factor, ::factor]


def back_project(err_lr, kernel, shift_yx, factor, hr_shape):
    """run_sr.py:168-178."""
    H, W = hr_shape
    up = np.zeros((err_lr.shape[0] * factor, err_lr.shape[1] * factor), dtype=np.float64)
    up[::factor, ::factor] = err_lr
    if up.shape[0] < H or up.shape[1] < W:
        up = np.pad(up, ((0, max(0, H - up.shape[0])), (0, max(0, W - up.shape[1]))))
    up = up[:H, :W]
    s = ndi_shift_cubic_nearest(up, (-shift_yx[0] * factor, -shift_yx[1] * factor))
    return blur(s, np.asarray(kernel)[::-1, ::-1])


def native_2x(mean_lr, factor=2):
    """run_sr.py:279  (``ndi_zoom(mean_lr, f, order=3)``)."""
    return ndi_zoom_cubic(mean_lr, factor)


def shift_and_add(lr_list, shifts_yx, factor=2, order=3):
    """run_sr.py:181-187.  ``order`` only reaches zoom (quirk Q7); orders 0, 1 and 3 are restated."""
    acc = None
    for lr, (dy, dx) in zip(lr_list, shifts_yx):
        up = ndi_zoom_order(lr, factor, order)
        term = ndi_shift_cubic_nearest(up, (dy * factor, dx * factor))
        acc = term if acc is None else acc + term
    return acc / len(lr_list)


def ibp(lr_list, shifts_yx, kernel, hr_init, factor=2, n_iter=80, step=0.5,
        clip_lo=0.0, clip_hi=255.0):
    """run_sr.py:190-209.  Returns (hr, errors)."""
    hr = np.array(hr_init, dtype=np.float64, copy=True)
    n = len(lr_list)
    errors: List[float] = []
    for _ in range(n_iter):
        corr = np.zeros_like(hr)
        tot = 0.0
        for lr, s in zip(lr_list, shifts_yx):
            sim = forward_model(hr, kernel, s, factor)
            mh, mw = min(sim.shape[0], lr.shape[0]), min(sim.shape[1], lr.shape[1])
            err = np.asarray(lr, dtype=np.float64)[:mh, :mw] - sim[:mh, :mw]
            tot += float(np.mean(err ** 2))
            corr += back_project(err, kernel, s, factor, hr.shape)
        hr += step * corr / n
        hr = np.clip(hr, clip_lo, clip_hi)
        errors.append(tot / n)
    return hr, errors


def to_uint8(img) -> np.ndarray:
    """run_sr.py:303  clip then truncating cast."""
    return np.clip(img, 0, 255).astype(np.uint8)


# --------------------------------------------------------------------------
# 6. closed form for integer HR shifts (fast oracle for large shapes)
# --------------------------------------------------------------------------
def clamp_gather(a: np.ndarray, sy: int, sx: int) -> np.ndarray:
    """shift(a,(sy,sx),order=3,mode='nearest') for integer sy,sx: a[clamp(y-sy),clamp(x-sx)]."""
    H, W = a.shape
    yy = np.clip(np.arange(H) - sy, 0, H - 1)
    xx = np.clip(np.arange(W) - sx, 0, W - 1)
    return a[np.ix_(yy, xx)]


def shift_and_add_integer(lr_list, shifts_yx, factor=2):
    """SAA when every factor*shift is an integer: the order-3 shift of the zoomed frame is a
    pure clamp-gather (weights (1/6, 4/6, 1/6, 0) return the padded sample itself)."""
    acc = None
    for lr, (dy, dx) in zip(lr_list, shifts_yx):
        sy, sx = dy * factor, dx * factor
        if sy != round(sy) or sx != round(sx):
            raise ValueError("non-integer HR shift")
        term = clamp_gather(ndi_zoom_cubic(lr, factor), int(round(sy)), int(round(sx)))
        acc = term if acc is None else acc + term
    return acc / len(lr_list)


def ibp_integer_shifts(lr_list, shifts_yx, kernel, hr_init, factor=2, n_iter=80, step=0.5,
                       clip_lo=0.0, clip_hi=255.0):
    """IBP when every factor*shift is an integer (SURVEY.md section 8a 'verified restatement').

    The forward blur is shared by all frames and the transposed blur is applied
    once to the summed, clamp-shifted, zero-inserted residuals (exact by
    linearity, quirk Q6).
    """
    hr = np.array(hr_init, dtype=np.float64, copy=True)
    n = len(lr_list)
    H, W = hr.shape
    ints = []
    for dy, dx in shifts_yx:
        sy, sx = dy * factor, dx * factor
        if sy != round(sy) or sx != round(sx):
            raise ValueError("non-integer HR shift")
        ints.append((int(round(sy)), int(round(sx))))
    kflip = np.asarray(kernel, dtype=np.float64)[::-1, ::-1]
    errors = []
    for _ in range(n_iter):
        B = conv2_same_zero(hr, kernel)
        acc = np.zeros_like(hr)
        tot = 0.0
        for lr, (sy, sx) in zip(lr_list, ints):
            sim = clamp_gather(B, sy, sx)[::factor, ::factor]
            err = np.asarray(lr, dtype=np.float64) - sim
            tot += float(np.mean(err ** 2))
            up = np.zeros((H, W), dtype=np.float64)
            up[::factor, ::factor] = err
            acc += clamp_gather(up, -sy, -sx)
        hr = np.clip(hr + (step / n) * conv2_same_zero(acc, kflip), clip_lo, clip_hi)
        errors.append(tot / n)
    return hr, errors


# --------------------------------------------------------------------------
# 7. tolerance helpers shared by the tests
# --------------------------------------------------------------------------
def psnr(a: np.ndarray, ref: np.ndarray, peak: float) -> float:
    mse = float(np.mean((np.asarray(a, np.float64) - np.asarray(ref, np.float64)) ** 2))
    return float("inf") if mse == 0 else 10.0 * math.log10(peak * peak / mse)


def max_abs(a, b) -> float:
    return float(np.max(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64))))
