"""Quality analytics on HR outputs -- the reference's slanted-edge MTF and bar-contrast measurements, on the device.

Same names, arguments and return values as the functions the reference defines in `mono_cal_target/analysis.ipynb`:
`local_contrast` (cell 4), `slanted_edge_esf` and `esf_to_mtf` (cell 7), `mtf_at_fraction` (cell 10), plus
`mtf50_mtf10` (cells 8 and 10 combined) which measures an HR image where it already lies: a CUDA tensor produced by
`sr459.ibp` / `shift_and_add` is analysed without a host round trip.  All arithmetic
runs in libsr459 (csrc/analysis.cu: Gaussian smoothing, Sobel magnitude, percentile threshold by radix select, line
fits, quarter-pixel binning, gradient, Hann window, DFT); there is no host fallback.  tests/test_gpu_analytics.py
pins the results against the notebook's own functions (goldens) and against the numpy restatement in oracle/.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .ops import _is_tensor, _stream_ptr, _torch

SENSOR_PITCH_MM = 3.45e-3                 # analysis.ipynb cell 9
HR_PITCH_MM = SENSOR_PITCH_MM / 2

_PD = C.POINTER(C.c_double)


def _dp(a):
    return a.ctypes.data_as(_PD)


def _device_image(img):
    """(tensor, is_f32): a 2-D uint8 or float32 CUDA tensor holding img.  HR images of this package are float32 and PNG
    results uint8, both analysed exactly; other dtypes are converted to float32 first."""
    torch = _torch()
    if _is_tensor(img):
        t = img if img.is_cuda else img.cuda()
        if t.dtype != torch.uint8:
            t = t.to(torch.float32)
    else:
        a = np.asarray(img)
        if a.dtype != np.uint8:
            a = a.astype(np.float32)
        t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    if t.dim() != 2:
        raise ValueError(f"image must be 2-D, got shape {tuple(t.shape)}")
    if t.stride(1) != 1:
        t = t.contiguous()
    return t, int(t.dtype != torch.uint8)


def _edge(img, roi, side, pitch_mm):
    lib = _lib.load()
    t, is_f32 = _device_image(img)
    H, W = t.shape
    if roi is None:
        r0, r1, c0, c1 = 0, H, 0, W
    else:
        (r0, r1, _), (c0, c1, _) = roi[0].indices(H), roi[1].indices(W)
    cap = lib.sr459_edge_mtf_capacity()
    ex, ey, lsf, fr, mtf = (np.zeros(cap) for _ in range(5))
    scal = np.zeros(8)
    rc = lib.sr459_edge_mtf(C.c_void_p(t.data_ptr()), is_f32, int(t.stride(0)), r0, c0, r1 - r0, c1 - c0,
                            0 if side == "left" else 1, float(pitch_mm), _dp(ex), _dp(ey), _dp(lsf), _dp(fr), _dp(mtf), _dp(scal),
                            _stream_ptr(t.device))
    if rc != 0:
        msg = lib.sr459_last_error().decode(errors="replace")
        if msg.startswith("Too few edge pixels"):
            raise RuntimeError(msg)                      # the notebook raises RuntimeError with this text
        raise _lib.Sr459Error(rc, msg)
    n = int(scal[1])
    return ex[:n], ey[:n], lsf[:n], fr[:n // 2], mtf[:n // 2], scal


def slanted_edge_esf(roi, side: str = "left", verbose: bool = False):
    """Oversampled ESF of ONE edge of a thick slanted line -- analysis.ipynb cell 7.
    Returns (esf_x, esf_y, edge_angle_deg)."""
    ex, ey, _, _, _, scal = _edge(roi, None, side, HR_PITCH_MM)
    if verbose:
        print(f"  Edge angle: {scal[2]:.1f} deg, {int(scal[6])}/{int(scal[5])} edge pixels ({side} side)")
    return ex, ey, float(scal[2])


def edge_mtf(img, roi=None, side: str = "left", pitch_mm: float = HR_PITCH_MM):
    """The whole chain in one device pass: dict with esf_x, esf_y, lsf, freq (cycles/pixel), mtf, angle, mtf50, mtf10
    (cycles/mm at `pitch_mm`)."""
    ex, ey, lsf, fr, mtf, scal = _edge(img, roi, side, pitch_mm)
    return {"esf_x": ex, "esf_y": ey, "lsf": lsf, "freq": fr, "mtf": mtf, "angle": float(scal[2]),
            "mtf50": float(scal[3]), "mtf10": float(scal[4]), "n_edge": int(scal[5]), "n_side": int(scal[6])}


def mtf50_mtf10(img, roi, side: str = "left", pitch_mm: float = HR_PITCH_MM):
    """MTF50 / MTF10 in cycles/mm of the slanted edge inside img[roi] -- cells 8 and 10 combined."""
    r = edge_mtf(img, roi, side, pitch_mm)
    return r["mtf50"], r["mtf10"]


def local_contrast(profile, window: int = 20) -> np.ndarray:
    """(max - min) / (max + min) in a sliding window -- analysis.ipynb cell 4."""
    lib = _lib.load()
    p = np.ascontiguousarray(profile, dtype=np.float64)
    out = np.zeros(len(p))
    if len(p):
        _lib.check(lib.sr459_local_contrast(_dp(p), len(p), int(window), _dp(out), _stream_ptr()))
    return out
