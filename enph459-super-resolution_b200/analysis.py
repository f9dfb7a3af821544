"""Quality analytics on HR outputs -- the reference's slanted-edge MTF and bar-contrast measurements.

Restates, with the same names, arguments and return values, the functions the reference defines in
`mono_cal_target/analysis.ipynb`: `local_contrast` (cell 4), `slanted_edge_esf` and `esf_to_mtf`
(cell 7), `mtf_at_fraction` (cell 10).  They validate that parity-level numerical differences of the SR
kernels do not move MTF50 / MTF10.  The inputs are 200 x 200 ROIs, so this
is host numpy code like the reference's; the two scipy.ndimage calls the notebook makes
(`gaussian_filter(sigma=1.5)`, `sobel`; both with scipy's default 'reflect' boundary) are restated here so
that the module needs numpy only.  tests/test_analysis.py pins every function against the notebook's own
code executed from /root/reference (golden vectors under tests/golden/g_analysis.npz).
"""
from __future__ import annotations

import numpy as np

SENSOR_PITCH_MM = 3.45e-3                 # analysis.ipynb cell 9
HR_PITCH_MM = SENSOR_PITCH_MM / 2


def _reflect_pad(a: np.ndarray, r: int, axis: int) -> np.ndarray:
    """scipy.ndimage 'reflect' (d c b a | a b c d | d c b a) = numpy 'symmetric'."""
    pad = [(0, 0)] * a.ndim
    pad[axis] = (r, r)
    return np.pad(a, pad, mode="symmetric")


def _correlate1d(a: np.ndarray, w: np.ndarray, axis: int) -> np.ndarray:
    r = len(w) // 2
    p = _reflect_pad(a, r, axis)
    out = np.zeros_like(a, dtype=np.float64)
    n = a.shape[axis]
    for k, wk in enumerate(w):
        sl = [slice(None)] * a.ndim
        sl[axis] = slice(k, k + n)
        out += wk * p[tuple(sl)]
    return out


def gaussian_filter(a, sigma: float, truncate: float = 4.0) -> np.ndarray:
    """scipy.ndimage.gaussian_filter(a, sigma) with the defaults (order 0, mode 'reflect', truncate 4)."""
    a = np.asarray(a, dtype=np.float64)
    r = int(truncate * float(sigma) + 0.5)
    x = np.arange(-r, r + 1)
    w = np.exp(-0.5 / (sigma * sigma) * x ** 2)
    w /= w.sum()
    for ax in range(a.ndim):
        a = _correlate1d(a, w, ax)
    return a


def sobel(a, axis: int) -> np.ndarray:
    """scipy.ndimage.sobel(a, axis) with mode 'reflect': derivative [-1 0 1] along axis, smoothing [1 2 1] across."""
    a = np.asarray(a, dtype=np.float64)
    out = _correlate1d(a, np.array([-1.0, 0.0, 1.0]), axis)
    for ax in range(a.ndim):
        if ax != axis:
            out = _correlate1d(out, np.array([1.0, 2.0, 1.0]), ax)
    return out


def local_contrast(profile, window: int = 20) -> np.ndarray:
    """(max - min) / (max + min) in a sliding window -- analysis.ipynb cell 4."""
    profile = np.asarray(profile, dtype=np.float64)
    n = len(profile)
    contrast = np.zeros(n)
    hw = window // 2
    if n > 2 * hw and hw > 0:
        win = np.lib.stride_tricks.sliding_window_view(profile, 2 * hw)[: n - 2 * hw]
        mn, mx = win.min(axis=1), win.max(axis=1)
        contrast[hw:n - hw] = (mx - mn) / (mx + mn + 1e-9)
    return contrast


def slanted_edge_esf(roi, side: str = "left", verbose: bool = False):
    """Oversampled ESF of ONE edge of a thick slanted line -- analysis.ipynb cell 7.
    Returns (esf_x, esf_y, edge_angle_deg)."""
    roi = np.asarray(roi)
    smooth = gaussian_filter(roi.astype(np.float64), sigma=1.5)
    gy = sobel(smooth, axis=0)
    gx = sobel(smooth, axis=1)
    mag = np.sqrt(gx ** 2 + gy ** 2)
    thresh = np.percentile(mag, 85)
    rs, cs = np.where(mag > thresh)
    if len(rs) < 20:
        raise RuntimeError("Too few edge pixels detected")
    use_row_as_x = (rs.max() - rs.min()) >= (cs.max() - cs.min())
    if use_row_as_x:
        cc_ = np.polyfit(rs, cs, 1)
        edge_dist = (cs - cc_[0] * rs - cc_[1]) / np.sqrt(1 + cc_[0] ** 2)
    else:
        cc_ = np.polyfit(cs, rs, 1)
        edge_dist = (rs - cc_[0] * cs - cc_[1]) / np.sqrt(1 + cc_[0] ** 2)
    cluster = edge_dist < 0 if side == "left" else edge_dist > 0
    rs_sel, cs_sel = rs[cluster], cs[cluster]
    if len(rs_sel) < 10:
        raise RuntimeError(f"Too few edge pixels on {side} side")
    nrows, ncols = roi.shape
    rr, cc = np.mgrid[:nrows, :ncols]
    if use_row_as_x:
        co = np.polyfit(rs_sel, cs_sel, 1)
        m = co[0]
        edge_angle = np.degrees(np.arctan2(1, m))
        dist = (cc - m * rr - co[1]) / np.sqrt(1 + m ** 2)
    else:
        co = np.polyfit(cs_sel, rs_sel, 1)
        m = co[0]
        edge_angle = np.degrees(np.arctan2(m, 1))
        dist = (rr - m * cc - co[1]) / np.sqrt(1 + m ** 2)
    if verbose:
        print(f"  Edge angle: {edge_angle:.1f} deg, {len(rs_sel)}/{len(rs)} edge pixels ({side} side)")
    flat_dist = dist.ravel()
    flat_val = roi.ravel().astype(np.float64)
    keep = (flat_dist > -8) & (flat_dist < 10)
    flat_dist, flat_val = flat_dist[keep], flat_val[keep]
    bin_width = 1.0 / 4
    bins = np.arange(flat_dist.min(), flat_dist.max() + bin_width, bin_width)
    esf_x = 0.5 * (bins[:-1] + bins[1:])
    esf_y = np.full(len(esf_x), np.nan)
    for i in range(len(esf_x)):                       # the notebook's half-open bins, literally
        mask = (flat_dist >= bins[i]) & (flat_dist < bins[i + 1])
        if mask.sum() > 0:
            esf_y[i] = flat_val[mask].mean()
    valid = ~np.isnan(esf_y)
    if valid.sum() > 2:
        esf_y = np.interp(esf_x, esf_x[valid], esf_y[valid])
    if esf_y[-1] < esf_y[0]:
        esf_x = -esf_x[::-1]
        esf_y = esf_y[::-1]
    return esf_x, esf_y, edge_angle


def esf_to_mtf(esf_x, esf_y):
    """ESF -> LSF (derivative) -> MTF (|FFT| of the Hann-windowed LSF), normalised -- analysis.ipynb cell 7."""
    lsf = np.gradient(esf_y, esf_x)
    lsf_w = lsf * np.hanning(len(lsf))
    n = len(lsf_w)
    mtf = np.abs(np.fft.fft(lsf_w))[: n // 2]
    if mtf[0] > 0:
        mtf = mtf / mtf[0]
    dx = np.mean(np.diff(esf_x))
    freq = np.fft.fftfreq(n, d=dx)[: n // 2]
    return freq, mtf, lsf


def mtf_at_fraction(freq, mtf, fraction):
    """Spatial frequency where the MTF first falls through `fraction` -- analysis.ipynb cell 10."""
    above = mtf >= fraction
    if not above.any() or above.all():
        return np.nan
    idx = np.where(np.diff(above.astype(int)) == -1)[0]
    if len(idx) == 0:
        return np.nan
    i = idx[0]
    f0, f1, m0, m1 = freq[i], freq[i + 1], mtf[i], mtf[i + 1]
    if abs(m1 - m0) < 1e-12:
        return f0
    return f0 + (fraction - m0) * (f1 - f0) / (m1 - m0)


def mtf50_mtf10(img, roi, side: str = "left", pitch_mm: float = HR_PITCH_MM):
    """MTF50 / MTF10 in cycles/mm of the slanted edge inside img[roi] -- cells 8 and 10 combined."""
    esf_x, esf_y, _ = slanted_edge_esf(np.asarray(img)[roi], side=side)
    freq, mtf, _ = esf_to_mtf(esf_x, esf_y)
    f = freq / pitch_mm
    valid = f > 0
    return mtf_at_fraction(f[valid], mtf[valid], 0.5), mtf_at_fraction(f[valid], mtf[valid], 0.1)
