#!/usr/bin/env python3
"""Streaming burst mode (BASELINE config 5): continuous 4-frame sets at camera rate on one GPU.

    python tools/bench_stream.py [--sets 200] [--fps 0|F] [--iters 50] [--h 1536 --w 2048] [--graph] [--check]

A producer thread plays the camera: it writes one uint8 frame per 1/fps seconds (fps 0 = as fast as the
consumer drains) into sr459.driver.BurstStream's pinned ring (frame_buffer / commit_frame, the shape of
AlliedCamera.stream_burst's frame callback, api/allied_vision_camera.py:90-115); the main thread waits for
the HR images.  Reported: p50 / p99 latency from the commit of a set's LAST frame to its uint8 HR image
being on the host, and sustained HR frames/s.  --check compares a few results with the oracle.
Prints one JSON line.
"""
import argparse
import json
import os
import statistics
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
NOMINAL = [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sets", type=int, default=200)
    ap.add_argument("--fps", type=float, default=0.0, help="camera frame rate (frames/s); 0 = free running")
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--h", type=int, default=1536)
    ap.add_argument("--w", type=int, default=2048)
    ap.add_argument("--depth", type=int, default=3)
    ap.add_argument("--graph", action="store_true")
    ap.add_argument("--check", action="store_true")
    args = ap.parse_args()

    import torch
    import sr459
    from sr459.driver import BurstStream, play_camera
    n, h, w = 4, args.h, args.w
    psf = sr459.make_gaussian_psf()
    rng = np.random.default_rng(459)
    pool = [rng.integers(0, 256, (h, w), dtype=np.uint8) for _ in range(8)]      # the "camera" cycles through 8 frames
    bs = BurstStream(n, h, w, NOMINAL, psf, args.iters, depth=args.depth, use_graph=args.graph)
    for warm in range(2):                                # warm-up sets
        for k in range(n):
            bs.push_frame(pool[k])
        bs.wait_result()
    keep = {}

    def on_result(s, img):
        if args.check and s < 2:
            keep[s] = img.clone().numpy()

    lat, wall = play_camera(bs, pool, args.sets, args.fps, on_result)
    lat_ms = sorted(1e3 * x for x in lat)
    line = {
        "workload": f"config5: continuous {n}-frame uint8 sets {h}x{w} -> {2*h}x{2*w} uint8, Gaussian 7x7 PSF, SAA + {args.iters} IBP iterations",
        "sets": args.sets, "camera_fps": args.fps or "free-running", "cuda_graph": bool(args.graph), "depth": args.depth,
        "latency_ms_p50": statistics.median(lat_ms), "latency_ms_p99": lat_ms[min(len(lat_ms) - 1, int(0.99 * len(lat_ms)))],
        "latency_ms_max": lat_ms[-1], "hr_frames_per_s": args.sets / wall,
        "hr_mpix_per_s": args.sets * 4.0 * h * w / 1e6 / wall,
        "h2d_bytes_per_set": bs.h2d_bytes, "d2h_bytes_per_set": bs.d2h_bytes,
    }
    if args.check:
        from oracle import sr_oracle as O
        worst = 0
        for s, img in keep.items():
            lrs = [pool[(s * n + k) % len(pool)].astype(np.float64) for k in range(n)]
            saa = O.shift_and_add_integer(lrs, NOMINAL)
            ref, _ = O.ibp_integer_shifts(lrs, NOMINAL, O.make_gaussian_psf(), saa, n_iter=args.iters)
            ref8 = np.clip(ref, 0, 255).astype(np.uint8)
            worst = max(worst, int(np.abs(img.astype(np.int16) - ref8.astype(np.int16)).max()))
        line["max_abs_u8_vs_oracle"] = worst
    print(json.dumps(line), flush=True)
    bs.close()


if __name__ == "__main__":
    main()
