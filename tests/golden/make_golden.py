#!/usr/bin/env python3
"""Generate the committed golden vectors by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

Every .npz written next to this file holds inputs plus the outputs of the
reference's own functions (imported via oracle/ref_loader.py from
/root/reference/<experiment>/run_sr.py) on those inputs.  The files are small
enough to commit; the GPU box (which has no /root/reference) reads only them.

Cases
-----
g_int4      mono_barcodes : 4 frames 33x47  S1 noise, nominal +-0.5 px shifts, Gaussian PSF, K=8
g_int5      mono_cal_target: 5 frames (centre + corners) 40x56 S2-like scene, Gaussian PSF, K=8
g_frac4     rgb_cal_target: 4 frames 48x64, calibrated fractional shifts, measured 7x7 PSF, K=8
g_real      rgb_barcodes  : 96x128 red-channel crops of the real rep00 frames, Gaussian PSF, K=10
g_unit      mono_barcodes : 64x96 S3 unit-range frames in [0,1], Gaussian PSF, K=12
g_psf       real pinhole crops (128x128 around each peak, uint8) -> measured 7x7 and 31x31 PSF
g_prims     primitives: blur / forward_model / back_project on one 37x53 array
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))

from oracle import ref_loader as R  # noqa: E402

NOMINAL = [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]
CALIBRATED = [(0.47225, -0.43385), (0.4897, 0.4641), (-0.4798, -0.4553), (-0.4856, 0.4369)]
SEED = 459


def scene(h, w, rng):
    """S2-like HR truth: bars + slanted edge + blobs on [20,235]."""
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    img = np.full((h, w), 60.0)
    img += 80.0 * (((xx // 5) % 2) == 0) * (yy < h / 2)
    img += 90.0 * ((xx - w / 2) * 0.25 + (yy - h / 2) > 0)
    for _ in range(6):
        cy, cx, s = rng.uniform(0, h), rng.uniform(0, w), rng.uniform(2, 6)
        img += 60.0 * np.exp(-((yy - cy) ** 2 + (xx - cx) ** 2) / (2 * s * s))
    return np.clip(img, 20, 235)


def run_pipeline(ref, lrs, shifts, psf, n_iter):
    with R.quiet():
        mean_lr = np.mean(lrs, axis=0)
        native = ref.ndi_zoom(mean_lr, 2, order=3)
        saa = ref.shift_and_add(lrs, shifts, factor=2, order=3)
        hr, errs = ref.ibp(lrs, shifts, psf, saa.copy(), factor=2, n_iter=n_iter, step=0.5)
    return dict(native=native, saa=saa, ibp=hr, errors=np.array(errs))


def simulate(ref, truth, shifts, psf, rng, noise=1.0):
    with R.quiet():
        lrs = [np.clip(ref.forward_model(truth, psf, s, 2) + rng.normal(0, noise, (truth.shape[0] // 2, truth.shape[1] // 2)), 0, 255)
               for s in shifts]
    return lrs


def real_psf_crops():
    ref = R.load("mono_cal_target")
    psf_dir = os.path.join(R.REFERENCE_ROOT, "calibration_beam_shift", "data")
    crops, names = [], []
    for d in sorted(os.listdir(psf_dir)):
        p = os.path.join(psf_dir, d, "pos4_(0,0).png")
        if not (os.path.isdir(os.path.join(psf_dir, d)) and os.path.exists(p)):
            continue
        from PIL import Image
        img = np.array(Image.open(p))
        assert img.dtype == np.uint8 and img.ndim == 2
        pr, pc = np.unravel_index(img.argmax(), img.shape)
        y0, x0 = int(pr) - 64, int(pc) - 64
        assert y0 >= 0 and x0 >= 0 and y0 + 128 <= img.shape[0] and x0 + 128 <= img.shape[1]
        crops.append(img[y0:y0 + 128, x0:x0 + 128].copy())
        names.append(d)
    with R.quiet():
        k7 = ref.load_measured_psf(psf_dir)
        ref.PSF_HALFWIDTH = 15          # quirk Q10: module constant read at call time
        k31 = ref.load_measured_psf(psf_dir)
        ref.PSF_HALFWIDTH = 3
    return np.stack(crops), names, k7, k31


def main():
    assert R.available(), "needs /root/reference"
    rng = np.random.default_rng(SEED)
    out = {}

    crops, names, k7, k31 = real_psf_crops()
    np.savez_compressed(os.path.join(HERE, "g_psf.npz"), crops=crops, names=np.array(names), k7=k7, k31=k31)

    # g_int4 --------------------------------------------------------------
    ref = R.load("mono_barcodes")
    with R.quiet():
        gauss = ref.make_gaussian_psf()
    lrs = [rng.uniform(0, 255, (33, 47)) for _ in NOMINAL]
    res = run_pipeline(ref, lrs, NOMINAL, gauss, 8)
    np.savez_compressed(os.path.join(HERE, "g_int4.npz"), lr=np.stack(lrs), shifts=np.array(NOMINAL), psf=gauss, **res)

    # g_int5 --------------------------------------------------------------
    ref = R.load("mono_cal_target")
    shifts5 = [s for _, s in ref.IMAGE_SHIFTS]
    truth = scene(80, 112, rng)
    lrs = simulate(ref, truth, shifts5, gauss, rng)
    res = run_pipeline(ref, lrs, shifts5, gauss, 8)
    np.savez_compressed(os.path.join(HERE, "g_int5.npz"), lr=np.stack(lrs), shifts=np.array(shifts5), psf=gauss, truth=truth, **res)

    # g_frac4 -------------------------------------------------------------
    ref = R.load("rgb_cal_target")
    truth = scene(96, 128, rng)
    lrs = simulate(ref, truth, CALIBRATED, k7, rng)
    res = run_pipeline(ref, lrs, CALIBRATED, k7, 8)
    np.savez_compressed(os.path.join(HERE, "g_frac4.npz"), lr=np.stack(lrs), shifts=np.array(CALIBRATED), psf=k7, truth=truth, **res)

    # g_real --------------------------------------------------------------
    ref = R.load("rgb_barcodes")
    sess = os.path.join(R.REFERENCE_ROOT, "rgb_barcodes", "data", "2_3_5_mil_color_tilt 0.28256_settle50ms")
    from PIL import Image
    raw = [np.array(Image.open(os.path.join(sess, f"corner{c}_rep00.png"))) for c in range(4)]
    # raw Bayer crop 192x256 at an even origin -> red channel 96x128
    y0, x0 = 600, 900
    raw_crops = [r[y0:y0 + 192, x0:x0 + 256].copy() for r in raw]
    lrs = [ref.extract_red(rc.astype(np.float64)) for rc in raw_crops]
    res = run_pipeline(ref, lrs, list(ref.CORNER_SHIFTS), gauss, 10)
    np.savez_compressed(os.path.join(HERE, "g_real.npz"), raw=np.stack(raw_crops), shifts=np.array(ref.CORNER_SHIFTS), psf=gauss, **res)

    # g_unit --------------------------------------------------------------
    ref = R.load("mono_barcodes")
    truth = scene(128, 192, rng)
    lrs = [l / 255.0 for l in simulate(ref, truth, NOMINAL, gauss, rng)]
    lrs = [np.asarray(l, dtype=np.float32).astype(np.float64) for l in lrs]
    res = run_pipeline(ref, lrs, NOMINAL, gauss, 12)
    np.savez_compressed(os.path.join(HERE, "g_unit.npz"), lr=np.stack(lrs), shifts=np.array(NOMINAL), psf=gauss, truth=truth / 255.0, **res)

    # g_prims -------------------------------------------------------------
    ref = R.load("mono_cal_target")
    a = rng.uniform(0, 255, (74, 106))
    e = rng.normal(0, 10, (37, 53))
    prim = {"a": a, "e": e, "k7": k7}
    with R.quiet():
        prim["blur_gauss"] = ref.blur(a, gauss)
        prim["blur_k7"] = ref.blur(a, k7)
        prim["blur_k31"] = ref.blur(a, k31)
        for tag, s in (("int", (0.5, -0.5)), ("frac", CALIBRATED[0]), ("zero", (0.0, 0.0))):
            prim[f"fwd_{tag}"] = ref.forward_model(a, k7, s, 2)
            prim[f"bp_{tag}"] = ref.back_project(e, k7, s, 2, a.shape)
        prim["zoom"] = ref.ndi_zoom(e, 2, order=3)
    np.savez_compressed(os.path.join(HERE, "g_prims.npz"), **prim)

    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)) // 1024, "KiB")


if __name__ == "__main__":
    main()
