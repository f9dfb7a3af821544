#!/usr/bin/env python3
"""Golden vectors for sr459.analysis from the reference's OWN notebook code.

Run in the build container only (needs /root/reference and scipy):

    python tests/golden/make_golden_analysis.py

The function definitions of mono_cal_target/analysis.ipynb (cells 4, 7 and 10: local_contrast,
slanted_edge_esf, esf_to_mtf, mtf_at_fraction) are exec'd unmodified -- only the plotting statements
that follow local_contrast in cell 4 are cut off -- and applied to synthetic ROIs: a thick dark slanted
line on a bright background (the notebook's ROI 2) at several angles / blurs, and a bar profile.
Writes tests/golden/g_analysis.npz.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
NB = "/root/reference/mono_cal_target/analysis.ipynb"


def notebook_functions():
    nb = json.load(open(NB))
    src = {i: "".join(c["source"]) for i, c in enumerate(nb["cells"]) if c["cell_type"] == "code"}
    ns = {"np": np}
    cell4 = src[4].split("fig, ax = plt.subplots")[0]           # the function only, not the plot
    exec(cell4, ns)
    exec(src[7], ns)
    cell10 = src[10].split("print(f'{\"Method\"")[0]
    exec(cell10, ns)
    return ns


def slanted_line_roi(n, angle_deg, width, blur_sigma, rng, lo=30.0, hi=220.0, noise=1.0):
    from scipy.ndimage import gaussian_filter
    yy, xx = np.mgrid[:n, :n].astype(np.float64)
    t = np.radians(angle_deg)
    d = (xx - n / 2) * np.cos(t) + (yy - n / 2) * np.sin(t)      # distance from the line's axis
    img = np.where(np.abs(d) < width / 2, lo, hi)
    img = gaussian_filter(img, blur_sigma) + rng.normal(0, noise, img.shape)
    return np.clip(img, 0, 255)


def main():
    ns = notebook_functions()
    rng = np.random.default_rng(459)
    out = {}
    cases = [(140, 12.0, 9.0, 1.2), (120, -20.0, 10.0, 2.0), (120, 75.0, 9.0, 0.8), (140, 100.0, 11.0, 1.5)]
    out["n_cases"] = np.array(len(cases))
    for i, (n, ang, wd, sg) in enumerate(cases):
        roi = slanted_line_roi(n, ang, wd, sg, rng).astype(np.float32)      # stored as float32, analysed as float64
        out[f"roi{i}"] = roi
        roi = roi.astype(np.float64)
        for side in ("left", "right"):
            ex, ey, a = ns["slanted_edge_esf"](roi, side=side)
            f, m, l = ns["esf_to_mtf"](ex, ey)
            out[f"esfx{i}{side}"], out[f"esfy{i}{side}"], out[f"ang{i}{side}"] = ex, ey, np.array(a)
            out[f"freq{i}{side}"], out[f"mtf{i}{side}"], out[f"lsf{i}{side}"] = f, m, l
            v = f > 0
            out[f"m50_{i}{side}"] = np.array(ns["mtf_at_fraction"](f[v], m[v], 0.5))
            out[f"m10_{i}{side}"] = np.array(ns["mtf_at_fraction"](f[v], m[v], 0.1))
    prof = 128 + 90 * np.sign(np.sin(np.arange(320) * 2 * np.pi / np.linspace(24, 6, 320))) + rng.normal(0, 2, 320)
    out["profile"] = prof
    out["contrast16"] = ns["local_contrast"](prof, window=16)
    out["contrast20"] = ns["local_contrast"](prof)
    np.savez_compressed(os.path.join(HERE, "g_analysis.npz"), **out)
    print("wrote g_analysis.npz:", len(out), "arrays")


if __name__ == "__main__":
    sys.exit(main())
