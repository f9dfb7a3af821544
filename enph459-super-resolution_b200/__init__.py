"""sr459: B200-native multi-frame super-resolution (Native-2x, SAA, SAA+IBP).

Drop-in for the SR core of benedikthoward/ENPH459-Super-Resolution's run_sr.py scripts;
all arithmetic runs in hand-written sm_100a CUDA kernels behind the C ABI declared in
include/sr459.h (libsr459.so, bound with ctypes).  Import as ``import sr459``.
"""
from . import _lib
from . import bands
from ._lib import Sr459Error
from .ops import (CLIP_HI, CLIP_LO, IBP_ITERATIONS, IBP_STEP_SIZE, PSF_HALFWIDTH, PSF_SIGMA, PSF_SIZE,
                  UPSAMPLE_FACTOR, Plan, back_project, blur, clear_plan_cache, extract_red, forward_model,
                  get_plan, ibp, ingest_groups, load_gray, load_measured_psf, make_gaussian_psf, mean_frames,
                  measured_psf_from_images, native_2x, ndi_shift, ndi_zoom, shift_and_add, to_uint8, u8_to_f32)

__all__ = [
    "Plan", "Sr459Error", "back_project", "blur", "clear_plan_cache", "extract_red", "forward_model", "get_plan",
    "ibp", "ingest_groups", "load_gray", "load_measured_psf", "make_gaussian_psf", "mean_frames", "measured_psf_from_images",
    "native_2x", "ndi_shift", "ndi_zoom", "shift_and_add", "to_uint8", "u8_to_f32",
]
__version__ = "0.1.0"
