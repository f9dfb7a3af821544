#!/usr/bin/env python3
"""Dev helper: per-stage device timings (CUDA events) of the BASELINE configs that are not the bench line.

    python tools/time_configs.py [--path=N] [c1 c2 c2int monocal c1meas c4 ...]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
NOMINAL = [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]
CALIBRATED = [(0.47225, -0.43385), (0.4897, 0.4641), (-0.4798, -0.4553), (-0.4856, 0.4369)]


def measured7():
    """A non-separable 7x7 stand-in with the measured PSF's structure (19 non-zero taps, peak at the centre)."""
    y, x = np.mgrid[-3:4, -3:4].astype(np.float64)
    k = np.exp(-(x * x / 1.1 + y * y / 0.7 + 0.3 * x * y))
    k[k < 0.004] = 0.0
    return k / k.sum()


def measured31():
    """The non-separable 31x31 stand-in of tools/bench_bands.py (the pinhole PNGs do not travel)."""
    y, x = np.mgrid[-15:16, -15:16].astype(np.float64)
    k = np.exp(-(x * x / 18.0 + y * y / 8.0 + 0.05 * x * y)) + 0.15 * np.exp(-((x - 4) ** 2 + (y + 3) ** 2) / 6.0)
    k[np.abs(k) < 1e-4 * k.max()] = 0.0
    return k / k.sum()


CONFIGS = {
    # name: (h, w, shifts, psf, iters)
    "c1": (1536, 2048, NOMINAL, "gauss", 50),
    "c1meas": (1536, 2048, NOMINAL, "k7", 50),
    "c2": (768, 1024, CALIBRATED, "k7", 50),
    "c2gauss": (768, 1024, CALIBRATED, "gauss", 50),
    "c2int": (768, 1024, NOMINAL, "gauss", 50),
    "monocal": (1536, 2048, [(0.0, 0.0)] + NOMINAL, "gauss", 80),
    "c4": (1536, 2048, NOMINAL, "k31", 20),
}


def main():
    import torch
    import sr459
    argv = [a for a in sys.argv[1:] if not a.startswith("--path=")]
    force = [int(a.split("=")[1]) for a in sys.argv[1:] if a.startswith("--path=")]
    names = argv or ["c1", "c2", "c2int", "monocal", "c1meas", "c4"]
    rng = np.random.default_rng(459)
    for name in names:
        h, w, shifts, psfname, iters = CONFIGS[name]
        if psfname in ("k7", "k31"):                      # the committed reference-generated measured PSFs
            psf = np.load(os.path.join(ROOT, "tests", "golden", "g_psf.npz"))[psfname]
        else:
            psf = sr459.make_gaussian_psf() if psfname == "gauss" else (measured31() if psfname == "meas31" else measured7())
        n = len(shifts)
        frames = [torch.from_numpy(rng.uniform(0, 255, (h, w)).astype(np.float32)).cuda() for _ in range(n)]
        plan = sr459.Plan(n, h, w, shifts, psf)
        if force:
            plan.set_path(force[0])
        hr = torch.empty((2 * h, 2 * w), dtype=torch.float32, device="cuda")
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        best = None
        for rep in range(4):
            ev[0].record()
            nat = plan.native_2x(frames[0])
            ev[1].record()
            plan.saa(frames, out=hr)
            ev[2].record()
            plan.ibp_(frames, hr, iters, want_errors=False)
            ev[3].record()
            torch.cuda.synchronize()
            t = [ev[i].elapsed_time(ev[i + 1]) for i in range(3)]
            best = t if best is None else [min(a, b) for a, b in zip(best, t)]
        errs = plan.errors(iters)
        mpix = 4.0 * h * w / 1e6
        print(f"{name:8s} n={n} {h}x{w} psf={psfname} int={plan.is_integer}: native {best[0]*1e3:7.1f} us  saa {best[1]*1e3:7.1f} us  "
              f"ibp({iters}) {best[2]:7.3f} ms = {best[2]/iters*1e3:7.1f} us/iter  job {mpix/((best[1]+best[2])/1e3):8.0f} HR Mpix/s  mse[-1]={errs[-1]:.4f}",
              flush=True)
        plan.close()


if __name__ == "__main__":
    main()
