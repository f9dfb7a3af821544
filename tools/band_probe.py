#!/usr/bin/env python3
"""Dev probe: the kernels of ONE thin row band, timed on one GPU.  Emulates `world` bands of the config-4 image in one
process (run_emulated_peer) for a few iterations; run it under `ncu --metrics gpu__time_duration.sum` to get the
per-kernel durations a rank of an N-GPU run sees (the exchange itself is local here).

    python tools/band_probe.py [world] [psf: measured31|gaussian7] [iters]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
NOMINAL = [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]


def main():
    import torch
    import sr459
    from sr459.bands import BandPlan, run_emulated_peer
    world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    which = sys.argv[2] if len(sys.argv) > 2 else "measured31"
    K = int(sys.argv[3]) if len(sys.argv) > 3 else 4
    g = np.load(os.path.join(ROOT, "tests", "golden", "g_psf.npz"))
    psf = g["k31"] if which == "measured31" else sr459.make_gaussian_psf()
    h, w = 1536, 2048
    rng = np.random.default_rng(459)
    lrs = [rng.uniform(0, 255, (h, w)).astype(np.float32) for _ in NOMINAL]
    plans = [BandPlan(4, h, w, NOMINAL, psf, r, world) for r in range(world)]
    for rep in range(2):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        hr, e = run_emulated_peer(plans, lrs, K)
        ev1.record()
        torch.cuda.synchronize()
    print(f"world={world} psf={which} K={K}: {ev0.elapsed_time(ev1) * 1e3 / K:.1f} us per iteration for ALL bands on one GPU (serial), mse[-1]={e[-1]:.4f}")
    for p in plans:
        p.close()


if __name__ == "__main__":
    main()
