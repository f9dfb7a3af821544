#!/usr/bin/env python3
"""Dev probe: FrameSetPipeline end-to-end throughput (config 1, uint8 in / uint8 out) for 1 and 2 compute lanes."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
NOMINAL = [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]
import torch, sr459
from sr459.driver import FrameSetPipeline
n, h, w, K = 4, 1536, 2048, 50
rng = np.random.default_rng(1)
sets = [torch.from_numpy(rng.integers(0, 256, (n, h, w), dtype=np.uint8)).pin_memory() for _ in range(4)]
for lanes, depth in ((1, 3), (2, 4), (3, 6), (4, 8)):
    pipe = FrameSetPipeline(n, h, w, NOMINAL, sr459.make_gaussian_psf(), K, depth=depth, in_dtype="u8", out_dtype="u8", use_graph=True, lanes=lanes)
    pipe.run([sets[i % 4] for i in range(8)])
    best = 1e9
    for rep in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        pipe.run([sets[i % 4] for i in range(60)])
        torch.cuda.synchronize(); best = min(best, (time.perf_counter() - t0) / 60)
    print(f"lanes={lanes} depth={pipe.depth}: {best*1e3:.3f} ms per job = {4*h*w/1e6/best:.0f} HR Mpix/s", flush=True)
    pipe.close()
