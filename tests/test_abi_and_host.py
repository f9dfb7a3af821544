"""CPU tests of the drop-in boundary and the host-side logic: the C-ABI library builds, loads and
exports every symbol include/sr459.h declares; host-only entry points (PSF construction) compute;
device entry points fail loudly -- never fall back -- without a GPU; sharding / band layout logic."""
import ctypes as C
import inspect
import os
import re

import numpy as np
import pytest

from conftest import ROOT, golden
from oracle import sr_oracle as O


@pytest.fixture(scope="module")
def lib():
    import sr459
    return sr459._lib.load()


def test_library_exports_every_declared_symbol(lib):
    import sr459
    names = sr459._lib.declared_symbols()
    assert len(names) >= 24
    for n in names:
        assert hasattr(lib, n), f"libsr459.so does not export {n}"
        assert n in sr459._lib._SIGNATURES, f"no ctypes signature for {n}"
    assert lib.sr459_version() >= 100


def test_header_cites_reference_lines_for_each_operator():
    txt = open(os.path.join(ROOT, "include", "sr459.h")).read()
    for fn in ("sr459_psf_gaussian", "sr459_psf_measured", "sr459_blur", "sr459_forward_model", "sr459_back_project",
               "sr459_saa", "sr459_ibp", "sr459_native2x"):
        i = txt.index(fn + "(")
        assert re.search(r"run_sr\.py:\d+", txt[max(0, i - 900):i]), f"{fn}: no reference citation above it"


def test_only_c_types_cross_the_abi():
    txt = open(os.path.join(ROOT, "include", "sr459.h")).read()
    assert "torch" not in txt and "at::" not in txt and "std::" not in txt
    assert 'extern "C"' in txt


def test_gaussian_psf_host_entry_point(lib):
    import sr459
    k = sr459.make_gaussian_psf()
    assert k.shape == (7, 7) and np.array_equal(k, golden("g_int4")["psf"])
    assert np.max(np.abs(sr459.make_gaussian_psf(5, 0.7) - O.make_gaussian_psf(5, 0.7))) < 1e-16
    assert sr459.make_gaussian_psf(6, 2.0).shape == (7, 7)
    out = np.empty(49)
    assert lib.sr459_psf_gaussian(7, -1.0, out.ctypes.data_as(C.POINTER(C.c_double))) == -1
    assert b"psf_gaussian" in lib.sr459_last_error()


def test_python_surface_matches_reference_signatures():
    """Same names, positional order and defaults as the reference's module-level functions
    (mono_cal_target/run_sr.py:104-209)."""
    import sr459
    want = {
        "make_gaussian_psf": ["size", "sigma"],
        "load_measured_psf": ["psf_dir"],
        "blur": ["img", "kernel"],
        "forward_model": ["hr", "kernel", "shift_yx", "factor"],
        "back_project": ["error_lr", "kernel", "shift_yx", "factor", "hr_shape"],
        "shift_and_add": ["lr_list", "shifts_yx", "factor", "order"],
        "ibp": ["lr_list", "shifts_yx", "kernel", "hr_init", "factor", "n_iter", "step"],
    }
    for name, params in want.items():
        got = list(inspect.signature(getattr(sr459, name)).parameters)
        assert got[:len(params)] == params, (name, got)
    sig = inspect.signature(sr459.ibp)
    assert sig.parameters["factor"].default == 2 and sig.parameters["step"].default == 0.5
    assert inspect.signature(sr459.shift_and_add).parameters["order"].default == 3
    assert sr459.PSF_SIZE == 7 and sr459.PSF_SIGMA == 1.0 and sr459.UPSAMPLE_FACTOR == 2 and sr459.IBP_STEP_SIZE == 0.5


def test_no_cpu_fallback_without_a_gpu():
    import torch
    import sr459
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    lrs = [np.zeros((8, 8)) for _ in range(4)]
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        sr459.shift_and_add(lrs, [(0.5, -0.5)] * 4)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        sr459.ibp(lrs, [(0.5, -0.5)] * 4, sr459.make_gaussian_psf(), np.zeros((16, 16)), n_iter=1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        sr459.blur(np.zeros((8, 8)), sr459.make_gaussian_psf())


def test_plan_create_reports_cuda_error_without_a_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    h = C.c_void_p()
    sh = np.zeros(8)
    k = np.ones((1, 1))
    rc = lib.sr459_plan_create(C.byref(h), 0, 4, 8, 8, 2, sh.ctypes.data_as(C.POINTER(C.c_double)),
                               k.ctypes.data_as(C.POINTER(C.c_double)), 1, 1, 0.5, 0.0, 255.0)
    assert rc == -2 and not h.value
    assert len(lib.sr459_last_error()) > 0
    rc = lib.sr459_plan_create(C.byref(h), 0, 99, 8, 8, 2, sh.ctypes.data_as(C.POINTER(C.c_double)),
                               k.ctypes.data_as(C.POINTER(C.c_double)), 1, 1, 0.5, 0.0, 255.0)
    assert rc == -1 and b"n_frames" in lib.sr459_last_error()


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under the package or the CLI may import it."""
    pkg = os.path.join(ROOT, "enph459-super-resolution_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                assert "/root/reference" not in txt, f


def test_shard_indices_cover_every_frame_set_once():
    from sr459.driver import shard_indices
    for n_items, world in [(64, 8), (64, 3), (5, 8), (0, 2), (7, 1)]:
        seen = sorted(i for r in range(world) for i in shard_indices(n_items, r, world))
        assert seen == list(range(n_items))
    assert shard_indices(10, 1, 4) == [1, 5, 9]


def test_run_sr_cli_session_discovery(tmp_path):
    """The directory walks of the four reference main()s (mono_cal_target/run_sr.py:338-343,
    rgb_cal_target/run_sr.py:352-356, mono_barcodes/run_sr.py:375-378): pure host logic."""
    from sr459 import run_sr
    d = tmp_path / "data"
    for name in ("b_sess", "a_sess", "c_other"):
        (d / name).mkdir(parents=True)
    (d / "a_sess" / "center.png").write_bytes(b"")
    (d / "b_sess" / "metadata.json").write_text("{}")
    (d / "file.txt").write_text("x")
    base = lambda ps: [os.path.basename(p) for p in ps]
    assert base(run_sr.find_sessions("mono_cal_target", str(d))) == ["a_sess"]
    assert base(run_sr.find_sessions("rgb_cal_target", str(d))) == ["b_sess"]
    assert base(run_sr.find_sessions("mono_barcodes", str(d))) == ["a_sess", "b_sess", "c_other"]
    (tmp_path / "empty").mkdir()
    with pytest.raises(FileNotFoundError):
        run_sr.find_sessions("rgb_barcodes", str(tmp_path / "empty"))
    assert run_sr.EXPERIMENTS["rgb_cal_target"][0] == 50 and run_sr.EXPERIMENTS["mono_barcodes"][0] == 80


def test_fir_back_projection_tables_match_scipy_shift_of_zero_inserted_data(lib):
    """sr459_fir_bp_axis (host only): the banded matrix of one axis of back_project's
    shift(zero_insert(err), s, order=3, mode='nearest') -- run_sr.py:170-176 -- against the oracle's restatement of
    scipy.ndimage.shift.  float32 tables: 1e-5 on +-100 data."""
    rng = np.random.default_rng(1)
    for n_lr, s in [(40, 0.9445), (40, -0.8677), (33, 0.37), (64, -1.0), (25, 1.9), (48, 0.0), (30, -2.6), (20, 0.5)]:
        N = 2 * n_lr
        i0 = np.zeros(N, dtype=np.int32)
        w = np.zeros(N * 20, dtype=np.float32)
        rc = lib.sr459_fir_bp_axis(n_lr, 2, float(s), i0.ctypes.data_as(C.POINTER(C.c_int)), w.ctypes.data_as(C.POINTER(C.c_float)))
        assert rc == 0
        assert i0.min() >= 0 and i0.max() + 20 <= n_lr and np.all(np.diff(i0) >= 0)
        w = w.reshape(N, 20).astype(np.float64)
        err = rng.uniform(-100, 100, (n_lr, 5))
        up = np.zeros((N, 10)); up[::2, ::2] = err
        ref = O.ndi_shift_cubic_nearest(up, (s, 0.0))[:, ::2]
        got = np.stack([(w[o][:, None] * err[i0[o]:i0[o] + 20]).sum(axis=0) for o in range(N)])
        assert np.abs(got - ref).max() < 1e-4, (n_lr, s, np.abs(got - ref).max())
    small = np.zeros(20, dtype=np.int32)
    assert lib.sr459_fir_bp_axis(10, 2, 0.5, small.ctypes.data_as(C.POINTER(C.c_int)), w.ctypes.data_as(C.POINTER(C.c_float))) == -4


def test_rank_core_pinning_partitions_the_allowed_cores():
    """driver.pin_rank_to_cores: every rank of a node gets its own contiguous share of the cores the process may run on."""
    import os
    from importlib import import_module
    D = import_module("enph459-super-resolution_b200.driver")
    if not hasattr(os, "sched_getaffinity"):
        pytest.skip("no CPU affinity API")
    before = sorted(os.sched_getaffinity(0))
    try:
        assert D.pin_rank_to_cores(0, 1) is None and sorted(os.sched_getaffinity(0)) == before      # single rank: untouched
        world = 2 if len(before) >= 2 else 1
        got = []
        for r in range(world):
            os.sched_setaffinity(0, before)
            rng = D.pin_rank_to_cores(r, world)
            if world == 1:
                assert rng is None
                continue
            mine = sorted(os.sched_getaffinity(0))
            assert rng == [mine[0], mine[-1]] and len(mine) == len(before) // world
            got.append(set(mine))
        if world == 2:
            assert not (got[0] & got[1])
        os.sched_setaffinity(0, before)
        assert D.pin_rank_to_cores(0, 10 * len(before)) is None                                     # fewer cores than ranks: untouched
    finally:
        os.sched_setaffinity(0, before)
    assert D.shard_indices(10, 1, 4) == [1, 5, 9]
