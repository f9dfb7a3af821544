"""ctypes binding of libsr459.so (include/sr459.h).  No CPU fallback: if the CUDA
library cannot be built/loaded, or no CUDA device is present, the operators raise."""
from __future__ import annotations

import ctypes as C
import os
import re
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.abspath(os.path.join(HERE, "..", "include", "sr459.h"))
LIB_PATH = os.path.join(HERE, "libsr459.so")

_lock = threading.Lock()
_lib = None


class Sr459Error(RuntimeError):
    """A libsr459 call returned a non-zero status."""

    def __init__(self, code: int, msg: str):
        super().__init__(f"sr459 error {code}: {msg}")
        self.code = code


def declared_symbols(header: str = HEADER):
    """Every function name the public header declares (used by the CPU tests)."""
    txt = open(header).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(sr459_[a-z0-9_]+)\s*\(", txt)))


_P = C.c_void_p
_I = C.c_int
_D = C.c_double
_L = C.c_int64
_PD = C.POINTER(C.c_double)
_PI = C.POINTER(C.c_int)
_PP = C.POINTER(C.c_void_p)

_SIGNATURES = {
    "sr459_version": (_I, []),
    "sr459_last_error": (C.c_char_p, []),
    "sr459_launch_count": (C.c_uint64, []),
    "sr459_psf_gaussian": (_I, [_I, _D, _PD]),
    "sr459_fir_bp_axis": (_I, [_I, _I, _D, _PI, C.POINTER(C.c_float)]),
    "sr459_psf_measured": (_I, [_P, _I, _I, _I, _I, _I, _PD, _PI, _PI, _P]),
    "sr459_u8_to_f32": (_I, [_P, _L, _P, _P]),
    "sr459_extract_red": (_I, [_P, _I, _I, _I, _I, _I, _P, _P]),
    "sr459_mean_frames": (_I, [_PP, _I, _L, _P, _P]),
    "sr459_ingest_groups": (_I, [_PP, _I, _I, _I, _I, _I, _I, _I, _PP, _P, _P, _P]),
    "sr459_ingest_groups_v": (_I, [_PP, _I, _PI, _I, _I, _I, _I, _I, _PP, _P, _P, _P]),
    "sr459_source_stamp": (C.c_char_p, []),
    "sr459_debug_guard_bands": (_I, [_I]),
    "sr459_debug_check_guards": (_I, []),
    "sr459_debug_live_allocations": (_I, []),
    "sr459_debug_poke_guard": (_I, [_I]),
    "sr459_f32_to_u8": (_I, [_P, _L, _P, _P]),
    "sr459_blur": (_I, [_P, _I, _I, _PD, _I, _I, _P, _P]),
    "sr459_zoom": (_I, [_P, _I, _I, _I, _P, _P]),
    "sr459_zoom_to": (_I, [_P, _I, _I, _I, _I, _I, _P, _P]),
    "sr459_shift": (_I, [_P, _I, _I, _D, _D, _P, _P]),
    "sr459_forward_model": (_I, [_P, _I, _I, _PD, _I, _I, _D, _D, _I, _P, _P]),
    "sr459_back_project": (_I, [_P, _I, _I, _PD, _I, _I, _D, _D, _I, _I, _I, _P, _P]),
    "sr459_plan_create": (_I, [_PP, _I, _I, _I, _I, _I, _PD, _PD, _I, _I, _D, _D, _D]),
    "sr459_plan_destroy": (None, [_P]),
    "sr459_plan_is_integer": (_I, [_P]),
    "sr459_plan_set_path": (_I, [_P, _I]),
    "sr459_plan_scratch_bytes": (_L, [_P]),
    "sr459_native2x": (_I, [_P, _P, _P, _P]),
    "sr459_saa": (_I, [_P, _PP, _P, _P]),
    "sr459_ibp": (_I, [_P, _PP, _P, _I, _PD, _P]),
    "sr459_ibp_errors": (_I, [_P, _I, _PD, _P]),
    "sr459_saa_ibp": (_I, [_P, _PP, _P, _P, _I, _PD, _P]),
    "sr459_band_plan_create": (_I, [_PP, _I, _I, _I, _I, _I, _PD, _PD, _I, _I, _D, _D, _D, _I, _I]),
    "sr459_band_layout": (_I, [_P, _PI]),
    "sr459_band_layout_ext": (_I, [_P, _PI]),
    "sr459_band_buffers": (_I, [_P, _PP, _PP, C.POINTER(C.c_int64)]),
    "sr459_band_export": (_I, [_P, C.c_char_p, _PP]),
    "sr459_band_attach": (_I, [_P, C.c_char_p, _P, _I, C.c_char_p, _P, _I]),
    "sr459_band_current": (_I, [_P, _I]),
    "sr459_band_saa": (_I, [_P, _PP, _P]),
    "sr459_band_reserve": (_I, [_P, _PP, _I, _P]),
    "sr459_band_run_range": (_I, [_P, _PP, _I, _I, _I, _P]),
    "sr459_band_run_close": (_I, [_P, _P]),
    "sr459_band_run": (_I, [_P, _PP, _I, _P]),
    "sr459_band_begin": (_I, [_P, _PP, _I, _P]),
    "sr459_band_iterate": (_I, [_P, _PP, _P, _P, _I, _P]),
    "sr459_band_errors": (_I, [_P, _I, _PD, _P]),
    "sr459_find_peaks": (_I, [_P, _I, _I, _I, _I, _D, _PI, _P, _P]),
    "sr459_pinhole_fit": (_I, [_P, _I, _I, _I, _I, _PI, _I, _D, _I, _PD, _PD, _PI, _P]),
    "sr459_psf_mtf": (_I, [_PD, _I, _I, _I, _PD, _PD, _P]),
    "sr459_edge_mtf": (_I, [_P, _I, _I, _I, _I, _I, _I, _I, _D, _PD, _PD, _PD, _PD, _PD, _PD, _P]),
    "sr459_edge_mtf_capacity": (_I, []),
    "sr459_local_contrast": (_I, [_PD, _I, _I, _PD, _P]),
}


def load(build_if_missing: bool = True):
    """Load (building first if needed) the shared library and bind every declared symbol."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        from . import build as _build
        if build_if_missing and _build.needs_build():
            # sources differ from the binary (or there is none): a failed build is an error, never a silent fall back
            # to whatever stale library is lying around
            try:
                _build.build()
            except Exception as exc:
                raise RuntimeError(f"libsr459.so is missing or older than its sources and could not be rebuilt: {exc}") from exc
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not found; run enph459-super-resolution_b200/build.py")
        lib = C.CDLL(LIB_PATH)
        for name in declared_symbols():
            fn = getattr(lib, name)          # AttributeError if the .so does not export it
            if name in _SIGNATURES:
                fn.restype, fn.argtypes = _SIGNATURES[name]
        have, want = lib.sr459_source_stamp().decode(), _build.source_stamp()
        if have != want:
            raise RuntimeError(f"libsr459.so was built from other sources (stamp {have}, sources {want}); "
                               "run enph459-super-resolution_b200/build.py")
        _lib = lib
        return lib


def check(code: int):
    if code != 0:
        raise Sr459Error(code, load().sr459_last_error().decode(errors="replace"))
