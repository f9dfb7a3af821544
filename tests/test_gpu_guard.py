"""Memory-safety evidence without compute-sanitizer (the GPU pool refuses it): every device allocation of the
library gets patterned guard bands (csrc/alloc.cu), every caller buffer sits inside a patterned arena, each kernel
path runs at a ragged size, and afterwards (1) all bands and arenas are intact -- no out-of-range WRITE -- and
(2) the results still match the oracle -- an out-of-range READ that mattered would have pulled the pattern
(-391.5f / NaN) into them."""
import numpy as np
import pytest

from conftest import TOL_ABS_255, golden
from oracle import sr_oracle as O

pytestmark = pytest.mark.gpu

NOMINAL = [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]
MONOCAL = [(0.0, 0.0)] + NOMINAL
FOUR = [(0.0, 0.0), (0.5, 0.0), (0.0, 0.5), (0.5, 0.5)]
CALIBRATED = [(0.47225, -0.43385), (0.4897, 0.4641), (-0.4798, -0.4553), (-0.4856, 0.4369)]
GUARD = 1 << 16


class Arena:
    """Caller buffers carved out of one NaN-filled device tensor with GUARD floats between them."""

    def __init__(self, torch, total_floats):
        self.torch = torch
        self.buf = torch.full((total_floats,), float("nan"), dtype=torch.float32, device="cuda")
        self.used = torch.zeros((total_floats,), dtype=torch.bool, device="cuda")
        self.off = GUARD

    def take(self, shape, src=None):
        n = int(np.prod(shape))
        n_al = (n + 63) & ~63
        assert self.off + n_al + GUARD <= self.buf.numel()
        v = self.buf[self.off:self.off + n].view(*shape)
        self.used[self.off:self.off + n] = True
        self.off += n_al + GUARD
        if src is not None:
            v.copy_(self.torch.from_numpy(np.asarray(src, dtype=np.float32)))
        return v

    def intact(self):
        return bool(self.torch.isnan(self.buf[~self.used]).all())


@pytest.fixture()
def guarded(sr):
    lib = sr._lib.load()
    sr.clear_plan_cache()
    assert lib.sr459_debug_guard_bands(4096) == 0
    yield lib
    lib.sr459_debug_guard_bands(0)


CASES = [
    # (name, shifts, psf key, (h, w), path)
    ("stream", NOMINAL, "gauss", (77, 116), 7),
    ("stream-one-strip", NOMINAL, "gauss", (9, 20), 7),
    ("ws", NOMINAL, "gauss", (77, 116), 4),
    ("auto-two-lattices", MONOCAL, "gauss", (77, 116), 0),
    ("auto-four-lattices", FOUR, "gauss", (61, 100), 0),
    ("stream-ml-one-strip", MONOCAL, "gauss", (9, 20), 7),
    ("stream-ml-three-lattices", FOUR[:3], "gauss", (130, 60), 7),
    ("fast-k7", NOMINAL, "k7", (77, 116), 3),
    ("auto-k7", MONOCAL, "k7", (77, 116), 0),
    ("dense-odd-width", NOMINAL, "gauss", (33, 47), 2),
    ("lattice-k31", NOMINAL, "k31", (70, 132), 8),
    ("lattice-k9", MONOCAL, "k9", (37, 72), 8),
    ("generic", NOMINAL, "gauss", (45, 61), 1),
    ("fractional-small", CALIBRATED, "k7", (45, 61), 0),
    ("fractional-interior", CALIBRATED, "k7", (176, 336), 0),
]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_guard_bands_intact_and_parity_holds(sr, guarded, case):
    import torch
    name, shifts, key, (h, w), path = case
    g = golden("g_psf")
    psf = {"gauss": O.make_gaussian_psf(), "k7": g["k7"], "k31": g["k31"], "k9": O.make_gaussian_psf(9, 1.3)}[key]
    rng = np.random.default_rng(99)
    lrs = [rng.uniform(0, 255, (h, w)).astype(np.float32).astype(np.float64) for _ in shifts]
    n, K = len(shifts), 3
    arena = Arena(torch, n * h * w + 3 * 4 * h * w + h * w + (n + 8) * (GUARD + 64))
    dl = [arena.take((h, w), l) for l in lrs]
    mean = arena.take((h, w), np.mean(lrs, axis=0))
    native, saa, hr = (arena.take((2 * h, 2 * w)) for _ in range(3))
    plan = sr.Plan(n, h, w, shifts, psf)
    if path:
        try:
            plan.set_path(path)
        except sr.Sr459Error as exc:
            plan.close()
            if exc.code == -4:
                pytest.skip(str(exc))
            raise
    plan.native_2x(mean, out=native)
    plan.saa(dl, out=saa)
    hr.copy_(saa)
    errs = plan.ibp_(dl, hr, K)
    torch.cuda.synchronize()
    bad = guarded.sr459_debug_check_guards()
    msg = guarded.sr459_last_error().decode() if bad else ""
    assert guarded.sr459_debug_live_allocations() > 0, "guard mode is not active"
    plan.close()
    assert bad == 0, f"{name}: {bad} plan allocations were written out of range: {msg}"
    assert arena.intact(), f"{name}: a caller buffer was written out of range"
    integer = all(float(2 * a).is_integer() and float(2 * b).is_integer() for a, b in shifts)
    saa_o = O.shift_and_add_integer(lrs, shifts) if integer else O.shift_and_add(lrs, shifts)
    hr_o, e_o = (O.ibp_integer_shifts if integer else O.ibp)(lrs, shifts, psf, saa_o, n_iter=K)
    assert O.max_abs(native.cpu().numpy(), O.native_2x(np.mean(lrs, axis=0))) <= TOL_ABS_255
    assert O.max_abs(saa.cpu().numpy(), saa_o) <= TOL_ABS_255
    assert O.max_abs(hr.cpu().numpy(), hr_o) <= TOL_ABS_255
    assert np.max(np.abs(np.array(errs) / np.array(e_o) - 1)) <= 1e-4


def test_guard_band_checker_detects_an_overrun(sr, guarded):
    """The checker itself: one byte written just past the end of a plan buffer must be reported, and only then."""
    import torch
    plan = sr.Plan(4, 16, 16, NOMINAL, O.make_gaussian_psf())
    lrs = [torch.zeros((16, 16), device="cuda") for _ in range(4)]
    hr = plan.saa(lrs)
    plan.ibp_(lrs, hr, 1)
    assert guarded.sr459_debug_check_guards() == 0
    assert guarded.sr459_debug_poke_guard(0) == 0
    assert guarded.sr459_debug_check_guards() == 1
    assert b"guard band" in guarded.sr459_last_error()
    assert guarded.sr459_debug_poke_guard(1) == 0
    assert guarded.sr459_debug_check_guards() == 0
    plan.close()


def test_guard_bands_intact_for_the_analytics_and_calibration_kernels(sr, guarded):
    """csrc/analysis.cu and csrc/calib.cu allocate their scratch through the same guarded allocator: ragged ROI / image sizes,
    peaks at the image border (clipped ROIs), then every band must be intact (the entry points free their scratch, so the bands are
    checked by the allocator when they are released: sr459_debug_check_guards also reports bands damaged at free time)."""
    import torch
    from sr459 import analysis, calibration
    rng = np.random.default_rng(17)
    n = 150
    yy, xx = np.mgrid[:n, :n].astype(np.float64)
    d = (xx - n / 2) * np.cos(0.3) + (yy - n / 2) * np.sin(0.3)
    roi = (np.where(np.abs(d) < 6, 40.0, 205.0) + rng.normal(0, 1.5, (n, n))).astype(np.float32)
    full = torch.full((n + 37, n + 53), 200.0, device="cuda")
    full[11:11 + n, 29:29 + n] = torch.from_numpy(roi).cuda()
    r = analysis.edge_mtf(full, (slice(11, 11 + n), slice(29, 29 + n)), "left")
    assert r["mtf50"] > 0 and len(r["esf_x"]) > 30
    analysis.local_contrast(rng.uniform(0, 255, 333), 20)
    H, W = 97, 131
    stack = rng.integers(0, 30, (5, H, W)).astype(np.uint8)
    yy, xx = np.mgrid[:H, :W]
    for k, (pr, pc) in enumerate([(2, 3), (50, 60), (H - 1, W - 1), (0, 70), (60, 0)]):
        stack[k] = np.clip(stack[k] + 210 * np.exp(-((yy - pr) ** 2 + (xx - pc) ** 2) / 7.0), 0, 255).astype(np.uint8)
    res = calibration.find_pinhole_centers(stack)
    assert res["center"].shape == (5, 2) and np.isfinite(res["center"]).all()
    calibration.compute_mtf(calibration.extract_psf(stack[1].astype(np.float64), (50, 60), 15), 3.45)
    torch.cuda.synchronize()
    bad = guarded.sr459_debug_check_guards()
    assert bad == 0, guarded.sr459_last_error().decode()
    assert torch.equal(full[:11], torch.full((11, n + 53), 200.0, device="cuda"))          # the image around the ROI is untouched
