"""The reference's SR operator surface, backed by libsr459 (hand-written sm_100a CUDA).

Same names, positional order and defaults as the module-level functions every
``run_sr.py`` of the reference re-declares (citations: mono_cal_target/run_sr.py):

    make_gaussian_psf(size=7, sigma=1.0)                          :104-111
    load_measured_psf(psf_dir)                                    :114-152
    blur(img, kernel)                                             :157-158
    forward_model(hr, kernel, shift_yx, factor)                   :161-165
    back_project(error_lr, kernel, shift_yx, factor, hr_shape)    :168-178
    shift_and_add(lr_list, shifts_yx, factor=2, order=3)          :181-187
    ibp(lr_list, shifts_yx, kernel, hr_init, factor=2, n_iter=80, step=0.5)   :190-209
    native_2x(mean_lr)  ==  ndi_zoom(mean_lr, 2, order=3)         :279
    load_gray(path), extract_red(img)                             :73-75, rgb_cal_target/run_sr.py:73-75

Inputs may be numpy arrays (any real dtype) or CUDA torch tensors.  numpy in ->
numpy float64 out (what the reference's callers expect); CUDA tensor in -> CUDA
float32 tensor out with no host round trip.  PyTorch is used only to own device
buffers and streams; every arithmetic step runs in libsr459.  There is no CPU
fallback: without a CUDA device or without the library these functions raise.
"""
from __future__ import annotations

import ctypes as C
import os
from collections import OrderedDict
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import _lib

PSF_SIZE = 7
PSF_SIGMA = 1.0
PSF_HALFWIDTH = 3
UPSAMPLE_FACTOR = 2
IBP_ITERATIONS = 80
IBP_STEP_SIZE = 0.5
CLIP_LO, CLIP_HI = 0.0, 255.0

MAX_FRAMES = 8


def _torch():
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("sr459 needs a CUDA device (B200 / sm_100a); there is no CPU fallback")
    return torch


def _stream_ptr(device=None):
    """The current torch stream of `device` (default: the current device)."""
    torch = _torch()
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _is_tensor(x) -> bool:
    return type(x).__module__.startswith("torch")


def _dev(x, dtype=None, device=None):
    """Contiguous float32 CUDA tensor holding x (no copy if it already is one).  `device`: the device index the
    buffer must live on (a plan's device); host data is uploaded there, a CUDA tensor of another device is refused."""
    torch = _torch()
    dtype = dtype or torch.float32
    target = torch.device("cuda", torch.cuda.current_device() if device is None else int(device))
    if _is_tensor(x):
        t = x
        if not t.is_cuda:
            t = t.to(target)
        elif device is not None and t.device != target:
            raise ValueError(f"tensor lives on {t.device}, the plan on {target}")
        return t.to(dtype).contiguous()
    a = np.ascontiguousarray(x)
    if a.dtype == np.float64 and dtype == torch.float32:
        a = a.astype(np.float32)
    return torch.from_numpy(a).to(target).to(dtype).contiguous()


def _ret(t, like_numpy: bool):
    if like_numpy:
        return t.cpu().numpy().astype(np.float64)
    return t


def _psf_arg(kernel) -> Tuple[np.ndarray, "C._Pointer", int, int]:
    if _is_tensor(kernel):
        kernel = kernel.detach().cpu().numpy()
    k = np.ascontiguousarray(kernel, dtype=np.float64)
    if k.ndim != 2:
        raise ValueError(f"PSF kernel must be 2-D, got shape {k.shape}")
    return k, k.ctypes.data_as(C.POINTER(C.c_double)), k.shape[0], k.shape[1]


def _check2d(t, what):
    if t.dim() != 2:
        raise ValueError(f"{what} must be 2-D, got shape {tuple(t.shape)}")


def _ptr(t):
    return C.c_void_p(t.data_ptr())


# --------------------------------------------------------------------------------------
# PSF construction
# --------------------------------------------------------------------------------------
def make_gaussian_psf(size: int = PSF_SIZE, sigma: float = PSF_SIGMA) -> np.ndarray:
    """Normalised 2-D Gaussian PSF (float64, host) -- run_sr.py:104-111."""
    lib = _lib.load()
    hw = int(size) // 2
    out = np.empty((2 * hw + 1, 2 * hw + 1), dtype=np.float64)
    _lib.check(lib.sr459_psf_gaussian(int(size), float(sigma), out.ctypes.data_as(C.POINTER(C.c_double))))
    return out


def load_gray(path: str) -> np.ndarray:
    """PNG -> float64 array, RGB averaged over channels -- run_sr.py:73-75 (host decode)."""
    from PIL import Image
    img = np.array(Image.open(path), dtype=np.float64)
    return img.mean(axis=2) if img.ndim == 3 else img


def measured_psf_from_images(images, halfwidth: int = PSF_HALFWIDTH, return_peaks: bool = False):
    """Measured PSF from decoded pinhole images (uint8 or float), device argmax + patch mean."""
    torch = _torch()
    lib = _lib.load()
    imgs = list(images)
    if not imgs:
        raise FileNotFoundError("no PSF images given")
    first = imgs[0]
    H, W = int(first.shape[0]), int(first.shape[1])
    all_u8 = all((_is_tensor(i) and i.dtype == torch.uint8) or (not _is_tensor(i) and np.asarray(i).dtype == np.uint8) for i in imgs)
    for i in imgs:
        if tuple(i.shape) != (H, W):
            raise ValueError("all PSF images must have the same shape")
    if all_u8:
        stack = torch.stack([_dev(i, torch.uint8) for i in imgs]).contiguous()
    else:
        stack = torch.stack([_dev(i) for i in imgs]).contiguous()
    K = 2 * halfwidth + 1
    out = np.empty((K, K), dtype=np.float64)
    peaks = np.empty(2 * len(imgs), dtype=np.int32)
    used = C.c_int(0)
    _lib.check(lib.sr459_psf_measured(_ptr(stack), 0 if all_u8 else 1, len(imgs), H, W, int(halfwidth),
                                      out.ctypes.data_as(C.POINTER(C.c_double)),
                                      peaks.ctypes.data_as(C.POINTER(C.c_int)), C.byref(used), _stream_ptr()))
    if return_peaks:
        return out, peaks.reshape(-1, 2), used.value
    return out


def load_measured_psf(psf_dir: str, halfwidth: Optional[int] = None) -> np.ndarray:
    """Average the peak-aligned pos4_(0,0).png pinhole images under psf_dir -- run_sr.py:114-152."""
    hw = PSF_HALFWIDTH if halfwidth is None else int(halfwidth)
    from PIL import Image
    imgs = []
    for sweep_dir in sorted(os.listdir(psf_dir)):
        full = os.path.join(psf_dir, sweep_dir)
        p = os.path.join(full, "pos4_(0,0).png")
        if not os.path.isdir(full) or not os.path.exists(p):
            continue
        raw = np.array(Image.open(p))
        imgs.append(raw if (raw.ndim == 2 and raw.dtype == np.uint8) else load_gray(p))
    if not imgs:
        raise FileNotFoundError(f"No pos4_(0,0).png found under {psf_dir}")
    try:
        return measured_psf_from_images(imgs, hw)
    except _lib.Sr459Error as exc:
        if "no usable image" in str(exc):
            raise FileNotFoundError(f"No pos4_(0,0).png found under {psf_dir}") from exc
        raise


# --------------------------------------------------------------------------------------
# ingest helpers
# --------------------------------------------------------------------------------------
def u8_to_f32(img):
    """uint8 frame -> float32 on the device (the numeric half of load_gray, run_sr.py:73-75)."""
    torch = _torch()
    lib = _lib.load()
    t = _dev(img, torch.uint8)
    out = torch.empty(t.shape, dtype=torch.float32, device=t.device)
    _lib.check(lib.sr459_u8_to_f32(_ptr(t), t.numel(), _ptr(out), _stream_ptr()))
    return out


def ingest_groups(groups, red: bool = False):
    """Frame-set ingest in one kernel: `groups` is a list of n frames, each a list of uint8 repetitions (any
    count per frame, as the reference's per-corner np.stack(...).mean(axis=0) allows)
    images (H x W, numpy or CUDA uint8 tensors).  Returns (frames, mean_lr, mean_u8): the n float32
    LR frames (Bayer-red sites if `red`, mean over the repetitions), their mean and its uint8 cast --
    load_gray / extract_red / stack.mean(axis=0) / np.mean(lr_frames, axis=0) of the reference
    (rgb_cal_target/run_sr.py:73-75, :107-108; mono_cal_target/run_sr.py:274, :308)."""
    torch = _torch()
    lib = _lib.load()
    n = len(groups)
    reps = [len(g) for g in groups]
    if n < 1 or n > MAX_FRAMES or min(reps) < 1:
        raise ValueError(f"need 1..{MAX_FRAMES} frames with at least one image each")
    imgs = [_dev(i, torch.uint8) for g in groups for i in g]
    H, W = imgs[0].shape
    for t in imgs:
        _check2d(t, "image")
        if tuple(t.shape) != (H, W):
            raise ValueError("all images of a frame set must have the same shape")
    step = 2 if red else 1
    oh, ow = (H + step - 1) // step, (W + step - 1) // step
    frames = torch.empty((n, oh, ow), dtype=torch.float32, device=imgs[0].device)
    mean = torch.empty((oh, ow), dtype=torch.float32, device=imgs[0].device)
    mean_u8 = torch.empty((oh, ow), dtype=torch.uint8, device=imgs[0].device)
    ip = (C.c_void_p * len(imgs))(*[t.data_ptr() for t in imgs])
    fp = (C.c_void_p * n)(*[frames[k].data_ptr() for k in range(n)])
    rp = (C.c_int * n)(*reps)
    _lib.check(lib.sr459_ingest_groups_v(ip, n, rp, H, W, step, 0, 0, fp, _ptr(mean), _ptr(mean_u8), _stream_ptr()))
    return [frames[k] for k in range(n)], mean, mean_u8


def extract_red(img):
    """Red Bayer sites img[0::2, 0::2] -- rgb_cal_target/run_sr.py:73-75."""
    torch = _torch()
    lib = _lib.load()
    as_np = not _is_tensor(img)
    u8 = (np.asarray(img).dtype == np.uint8) if as_np else (img.dtype == torch.uint8)
    t = _dev(img, torch.uint8 if u8 else torch.float32)
    _check2d(t, "img")
    H, W = t.shape
    out = torch.empty(((H + 1) // 2, (W + 1) // 2), dtype=torch.float32, device=t.device)
    _lib.check(lib.sr459_extract_red(_ptr(t), 0 if u8 else 1, H, W, 0, 0, _ptr(out), _stream_ptr()))
    return _ret(out, as_np)


def mean_frames(frames):
    """np.mean(frames, axis=0) -- run_sr.py:274."""
    torch = _torch()
    lib = _lib.load()
    as_np = not _is_tensor(frames[0])
    ts = [_dev(f) for f in frames]
    out = torch.empty_like(ts[0])
    arr = (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])
    _lib.check(lib.sr459_mean_frames(arr, len(ts), ts[0].numel(), _ptr(out), _stream_ptr()))
    return _ret(out, as_np)


def to_uint8(img):
    """np.clip(img, 0, 255).astype(np.uint8) (truncation) -- run_sr.py:303."""
    torch = _torch()
    lib = _lib.load()
    as_np = not _is_tensor(img)
    t = _dev(img)
    out = torch.empty(t.shape, dtype=torch.uint8, device=t.device)
    _lib.check(lib.sr459_f32_to_u8(_ptr(t), t.numel(), _ptr(out), _stream_ptr()))
    return out.cpu().numpy() if as_np else out


# --------------------------------------------------------------------------------------
# stand-alone operators
# --------------------------------------------------------------------------------------
def blur(img, kernel):
    """fftconvolve(img, kernel, mode='same') -- run_sr.py:157-158."""
    torch = _torch()
    lib = _lib.load()
    as_np = not _is_tensor(img)
    t = _dev(img)
    _check2d(t, "img")
    k, kp, kh, kw = _psf_arg(kernel)
    out = torch.empty_like(t)
    _lib.check(lib.sr459_blur(_ptr(t), t.shape[0], t.shape[1], kp, kh, kw, _ptr(out), _stream_ptr()))
    return _ret(out, as_np)


def ndi_zoom(img, factor=2, order: int = 3):
    """scipy.ndimage.zoom(img, factor, order) with scipy's other defaults: any scalar or (fy, fx) zoom factor (output shape
    round(h * fy) x round(w * fx), as scipy rounds), spline order 0, 1 or 3 (the reference's setting)."""
    if order not in (0, 1, 3):
        raise NotImplementedError("spline orders 0, 1 and 3 are implemented (the reference uses 3)")
    fy, fx = (factor, factor) if np.isscalar(factor) else factor
    if fy <= 0 or fx <= 0:
        raise ValueError("zoom factors must be positive")
    torch = _torch()
    lib = _lib.load()
    as_np = not _is_tensor(img)
    t = _dev(img)
    _check2d(t, "img")
    H, W = int(round(t.shape[0] * fy)), int(round(t.shape[1] * fx))
    if H < 1 or W < 1:
        raise ValueError("zoom factor too small for this image")
    out = torch.empty((H, W), dtype=torch.float32, device=t.device)
    _lib.check(lib.sr459_zoom_to(_ptr(t), t.shape[0], t.shape[1], H, W, int(order), _ptr(out), _stream_ptr()))
    return _ret(out, as_np)


def native_2x(mean_lr, factor: int = UPSAMPLE_FACTOR):
    """Native-2x: ndi_zoom(mean_lr, 2, order=3) -- run_sr.py:279."""
    return ndi_zoom(mean_lr, factor, 3)


def ndi_shift(img, shift_yx, order: int = 3, mode: str = "nearest"):
    """scipy.ndimage.shift(img, shift_yx, order=3, mode='nearest')."""
    if order != 3 or mode != "nearest":
        raise NotImplementedError("only order=3, mode='nearest' (the reference's settings) are implemented")
    torch = _torch()
    lib = _lib.load()
    as_np = not _is_tensor(img)
    t = _dev(img)
    _check2d(t, "img")
    out = torch.empty_like(t)
    _lib.check(lib.sr459_shift(_ptr(t), t.shape[0], t.shape[1], float(shift_yx[0]), float(shift_yx[1]), _ptr(out), _stream_ptr()))
    return _ret(out, as_np)


def forward_model(hr, kernel, shift_yx, factor):
    """blur -> sub-pixel shift -> decimate -- run_sr.py:161-165."""
    torch = _torch()
    lib = _lib.load()
    as_np = not _is_tensor(hr)
    t = _dev(hr)
    _check2d(t, "hr")
    k, kp, kh, kw = _psf_arg(kernel)
    f = int(factor)
    H, W = t.shape
    out = torch.empty(((H + f - 1) // f, (W + f - 1) // f), dtype=torch.float32, device=t.device)
    _lib.check(lib.sr459_forward_model(_ptr(t), H, W, kp, kh, kw, float(shift_yx[0]), float(shift_yx[1]), f, _ptr(out), _stream_ptr()))
    return _ret(out, as_np)


def back_project(error_lr, kernel, shift_yx, factor, hr_shape):
    """zero-insert -> inverse shift -> blur with the flipped PSF -- run_sr.py:168-178."""
    torch = _torch()
    lib = _lib.load()
    as_np = not _is_tensor(error_lr)
    t = _dev(error_lr)
    _check2d(t, "error_lr")
    k, kp, kh, kw = _psf_arg(kernel)
    H, W = int(hr_shape[0]), int(hr_shape[1])
    out = torch.empty((H, W), dtype=torch.float32, device=t.device)
    _lib.check(lib.sr459_back_project(_ptr(t), t.shape[0], t.shape[1], kp, kh, kw, float(shift_yx[0]), float(shift_yx[1]),
                                      int(factor), H, W, _ptr(out), _stream_ptr()))
    return _ret(out, as_np)


# --------------------------------------------------------------------------------------
# plans
# --------------------------------------------------------------------------------------
class Plan:
    """A fixed SR geometry (n frames of h x w, shifts, PSF) with its device tables and scratch."""

    def __init__(self, n_frames, h, w, shifts_yx, kernel, factor=UPSAMPLE_FACTOR, step=IBP_STEP_SIZE,
                 clip_lo=CLIP_LO, clip_hi=CLIP_HI, device=None):
        torch = _torch()
        self._lib = _lib.load()
        self.device = torch.cuda.current_device() if device is None else int(device)
        self.n, self.h, self.w, self.factor = int(n_frames), int(h), int(w), int(factor)
        self.H, self.W = self.h * self.factor, self.w * self.factor
        sh = np.ascontiguousarray(np.asarray(shifts_yx, dtype=np.float64).reshape(-1))
        if sh.size != 2 * self.n:
            raise ValueError(f"need {self.n} (dy, dx) shifts, got {sh.size // 2}")
        if not (1 <= self.n <= MAX_FRAMES):
            raise ValueError(f"1..{MAX_FRAMES} frames supported, got {self.n}")
        k, kp, kh, kw = _psf_arg(kernel)
        handle = C.c_void_p()
        _lib.check(self._lib.sr459_plan_create(C.byref(handle), self.device, self.n, self.h, self.w, self.factor,
                                               sh.ctypes.data_as(C.POINTER(C.c_double)), kp, kh, kw,
                                               float(step), float(clip_lo), float(clip_hi)))
        self._h = handle

    def close(self):
        if getattr(self, "_h", None):
            self._lib.sr459_plan_destroy(self._h)
            self._h = None

    __del__ = close

    @staticmethod
    def shifts_are_integer(shifts_yx, factor=UPSAMPLE_FACTOR) -> bool:
        """True if every shift is a whole number of HR pixels (the fused integer kernels apply)."""
        s = np.asarray(shifts_yx, dtype=np.float64) * factor
        return bool(np.all(s == np.rint(s)))

    @property
    def is_integer(self) -> bool:
        return bool(self._lib.sr459_plan_is_integer(self._h))

    def set_path(self, path: int):
        """0 auto, 1 generic stage-by-stage, 2 dense fused, 3 polyphase fused, 4 TMA separable fused (warp-synchronous), 5 four-phase form, 6 persistent pipelined form."""
        _lib.check(self._lib.sr459_plan_set_path(self._h, int(path)))
        return self

    @property
    def scratch_bytes(self) -> int:
        return int(self._lib.sr459_plan_scratch_bytes(self._h))

    # -- helpers --
    def _stream(self):
        """torch's current stream ON THE PLAN'S DEVICE (which need not be the current device)."""
        return _stream_ptr(self.device)

    def _out(self, t):
        if not (t.is_cuda and t.device.index == self.device and t.dtype == _torch().float32 and t.is_contiguous()):
            raise ValueError(f"output buffer must be a contiguous float32 tensor on cuda:{self.device}")
        return t

    def _frames(self, lr_list):
        if len(lr_list) != self.n:
            raise ValueError(f"plan was built for {self.n} frames, got {len(lr_list)}")
        ts = [_dev(f, device=self.device) for f in lr_list]
        for t in ts:
            if tuple(t.shape) != (self.h, self.w):
                raise ValueError(f"frame shape {tuple(t.shape)} != plan shape {(self.h, self.w)}")
        arr = (C.c_void_p * self.n)(*[t.data_ptr() for t in ts])
        return ts, arr

    def native_2x(self, mean_lr, out=None):
        torch = _torch()
        t = _dev(mean_lr, device=self.device)
        if tuple(t.shape) != (self.h, self.w):
            raise ValueError(f"frame shape {tuple(t.shape)} != plan shape {(self.h, self.w)}")
        out = torch.empty((self.H, self.W), dtype=torch.float32, device=t.device) if out is None else self._out(out)
        _lib.check(self._lib.sr459_native2x(self._h, _ptr(t), _ptr(out), self._stream()))
        return out

    def saa(self, lr_list, out=None):
        torch = _torch()
        ts, arr = self._frames(lr_list)
        out = torch.empty((self.H, self.W), dtype=torch.float32, device=ts[0].device) if out is None else self._out(out)
        _lib.check(self._lib.sr459_saa(self._h, arr, _ptr(out), self._stream()))
        return out

    def ibp_(self, lr_list, hr_inout, n_iter, want_errors=True):
        """In-place IBP on a float32 CUDA tensor; returns the per-iteration MSE list."""
        ts, arr = self._frames(lr_list)
        if tuple(hr_inout.shape) != (self.H, self.W) or not hr_inout.is_contiguous():
            raise ValueError(f"hr shape {tuple(hr_inout.shape)} != plan HR shape {(self.H, self.W)} (contiguous)")
        self._out(hr_inout)
        errs = np.empty(max(int(n_iter), 0), dtype=np.float64)
        ep = errs.ctypes.data_as(C.POINTER(C.c_double)) if want_errors else None
        _lib.check(self._lib.sr459_ibp(self._h, arr, _ptr(hr_inout), int(n_iter), ep, self._stream()))
        return errs.tolist() if want_errors else None

    def errors(self, n_iter):
        errs = np.empty(int(n_iter), dtype=np.float64)
        _lib.check(self._lib.sr459_ibp_errors(self._h, int(n_iter), errs.ctypes.data_as(C.POINTER(C.c_double)), self._stream()))
        return errs.tolist()

    def saa_ibp(self, lr_list, n_iter, out=None, saa_out=None, want_errors=True):
        torch = _torch()
        ts, arr = self._frames(lr_list)
        out = torch.empty((self.H, self.W), dtype=torch.float32, device=ts[0].device) if out is None else self._out(out)
        errs = np.empty(max(int(n_iter), 0), dtype=np.float64)
        ep = errs.ctypes.data_as(C.POINTER(C.c_double)) if want_errors else None
        _lib.check(self._lib.sr459_saa_ibp(self._h, arr, _ptr(out), _ptr(self._out(saa_out)) if saa_out is not None else None,
                                           int(n_iter), ep, self._stream()))
        return out, (errs.tolist() if want_errors else None)


_PLAN_CACHE: "OrderedDict[tuple, Plan]" = OrderedDict()
_PLAN_CACHE_MAX = 8


def get_plan(n, h, w, shifts_yx, kernel, factor=UPSAMPLE_FACTOR, step=IBP_STEP_SIZE,
             clip_lo=CLIP_LO, clip_hi=CLIP_HI) -> Plan:
    """Cached Plan for a geometry (plans own device scratch, so re-use matters)."""
    torch = _torch()
    k, _, _, _ = _psf_arg(kernel)
    key = (torch.cuda.current_device(), int(n), int(h), int(w), int(factor),
           tuple(np.asarray(shifts_yx, dtype=np.float64).reshape(-1).tolist()), k.shape, k.tobytes(),
           float(step), float(clip_lo), float(clip_hi))
    plan = _PLAN_CACHE.get(key)
    if plan is None:
        plan = Plan(n, h, w, shifts_yx, k, factor, step, clip_lo, clip_hi)
        _PLAN_CACHE[key] = plan
        while len(_PLAN_CACHE) > _PLAN_CACHE_MAX:
            _, old = _PLAN_CACHE.popitem(last=False)
            old.close()
    else:
        _PLAN_CACHE.move_to_end(key)
    return plan


def clear_plan_cache():
    while _PLAN_CACHE:
        _, p = _PLAN_CACHE.popitem()
        p.close()


_DUMMY_PSF = np.array([[1.0]])


def shift_and_add(lr_list, shifts_yx, factor=2, order=3):
    """SAA: mean over frames of shift(zoom(lr, factor), factor*shift) -- run_sr.py:181-187."""
    if len(lr_list) != len(shifts_yx):
        # the reference zips the two lists; a silent truncation is never intended
        raise ValueError("lr_list and shifts_yx differ in length")
    as_np = not _is_tensor(lr_list[0])
    h, w = lr_list[0].shape
    if order != 3 or int(factor) != factor:
        # `order` only reaches zoom (run_sr.py:185, quirk Q7); the shift stays cubic: frame by frame through the stand-alone
        # operators (the fused plan path is built for the reference's order 3 and integer factors)
        terms = [ndi_shift(ndi_zoom(_dev(lr), factor, order), (dy * factor, dx * factor)) for lr, (dy, dx) in zip(lr_list, shifts_yx)]
        return _ret(mean_frames(terms), as_np)
    plan = get_plan(len(lr_list), h, w, shifts_yx, _DUMMY_PSF, factor)
    return _ret(plan.saa(lr_list), as_np)


def ibp(lr_list, shifts_yx, kernel, hr_init, factor=2, n_iter=IBP_ITERATIONS, step=IBP_STEP_SIZE):
    """Iterative back-projection -- run_sr.py:190-209.  Returns (hr, errors); hr_init is not modified."""
    if len(lr_list) != len(shifts_yx):
        raise ValueError("lr_list and shifts_yx differ in length")
    as_np = not _is_tensor(hr_init)
    h, w = lr_list[0].shape
    plan = get_plan(len(lr_list), h, w, shifts_yx, kernel, factor, step)
    hr = _dev(hr_init)
    if tuple(hr.shape) != (plan.H, plan.W):
        raise ValueError(f"hr_init shape {tuple(hr.shape)} != {(plan.H, plan.W)}")
    if _is_tensor(hr_init) and hr.data_ptr() == hr_init.data_ptr():
        hr = hr.clone()                       # the reference copies hr_init (:191)
    errors = plan.ibp_(lr_list, hr, n_iter)
    return _ret(hr, as_np), errors
