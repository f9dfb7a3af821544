#!/usr/bin/env python3
"""Row-band IBP of ONE 3072x4096 HR image across N GPUs (BASELINE config 4), launched by torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/bench_bands.py [--psf gaussian7|measured31|measured7] [--iters 200] [--h 1536 --w 2048] [--fractional]

Every rank owns a band of HR rows, builds its initial image from its own LR rows (banded SAA) and runs the iteration
kernels on it; the halo rows travel as peer-memory stores from k_band_push (csrc/bands.cu).  The loop is one CUDA graph.
Timing: CUDA events around the replay, barrier on both sides, per-rank values and the max.  SR459_BAND_NOPUSH=1 in the
environment skips the exchange (compute-only probe; the seams are then wrong).  The measured PSFs are the committed
reference-generated ones (tests/golden/g_psf.npz).  Prints one JSON line on rank 0.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
NOMINAL = [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]
CALIBRATED = [(0.47225, -0.43385), (0.4897, 0.4641), (-0.4798, -0.4553), (-0.4856, 0.4369)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--psf", choices=["gaussian7", "measured31", "measured7"], default="measured31")
    ap.add_argument("--iters", type=int, default=200)
    ap.add_argument("--h", type=int, default=1536)
    ap.add_argument("--w", type=int, default=2048)
    ap.add_argument("--fractional", action="store_true")
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    import sr459
    from sr459.bands import BandPlan

    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    g = np.load(os.path.join(ROOT, "tests", "golden", "g_psf.npz"))
    psf = {"gaussian7": sr459.make_gaussian_psf(), "measured31": g["k31"], "measured7": g["k7"]}[args.psf]
    shifts = CALIBRATED if args.fractional else NOMINAL
    h, w, K = args.h, args.w, args.iters
    rng = np.random.default_rng(459)                       # identical on every rank
    lrs = [torch.from_numpy(rng.uniform(0, 255, (h, w)).astype(np.float32)).cuda() for _ in shifts]

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    bp = BandPlan(4, h, w, shifts, psf, rank, world)
    if world > 1:
        bp.attach()
    lay = bp.layout
    ext = [f[lay.lr_ext_lo:lay.lr_ext_hi].contiguous() for f in lrs]
    bp.saa(ext)
    bp.run_peer(ext, K, reduce=False)                        # eager warm-up: lazy scratch sized by K
    sync()
    runs, errs = [], None
    for rep in range(args.reps + 1):
        bp.saa(ext)
        sync()
        errs = bp.run_peer(ext, K, graph=True, reduce=False)
        sync()
        if rep:
            runs.append(bp.last_loop_ms)
    t = torch.tensor([min(runs)], dtype=torch.float64, device="cuda")
    allt = [torch.zeros_like(t) for _ in range(world)]
    if world > 1:
        dist.all_gather(allt, t)
    else:
        allt = [t]
    per_rank = [float(x.item()) for x in allt]
    line = {
        "workload": f"config4: one {2*h}x{2*w} HR image, 4 frames, {args.psf} PSF, {'fractional' if args.fractional else 'integer'} shifts, "
                    f"{K} IBP iterations, {world} row band(s), halo {lay.halo} rows; {bp.exchange_kind}",
        "n_gpus": world, "iters": K, "no_push_probe": bool(os.environ.get("SR459_BAND_NOPUSH")),
        "us_per_iteration": 1e3 * max(per_rank) / K, "per_rank_us_per_iteration": [1e3 * x / K for x in per_rank],
        "hr_mpix_per_s": (4.0 * h * w / 1e6) / (max(per_rank) / 1e3),
        "halo_bytes_per_iteration_per_neighbour": lay.halo * lay.W * 4,
    }
    if rank == 0:
        print(json.dumps(line), flush=True)
    bp.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
