#!/usr/bin/env python3
"""Dev probe: FrameSetPipeline end-to-end throughput (config 1, uint8 in / uint8 out) for several lane counts; under torchrun every
rank runs its own pipeline (batch mode) and the slowest rank is reported, with and without core pinning."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
NOMINAL = [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]
import torch, sr459
import torch.distributed as dist
from sr459.driver import FrameSetPipeline, pin_rank_to_cores
world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n, h, w, K = 4, 1536, 2048, 50
rng = np.random.default_rng(1 + rank)
sets = [torch.from_numpy(rng.integers(0, 256, (n, h, w), dtype=np.uint8)).pin_memory() for _ in range(4)]
all_cores = sorted(os.sched_getaffinity(0))
variants = [("nopin", 1, 3), ("nopin", 4, 8), ("pin", 4, 8), ("pin", 2, 4)] if world > 1 else [("nopin", 1, 3), ("nopin", 2, 4), ("nopin", 4, 8)]
for pin, lanes, depth in variants:
    os.sched_setaffinity(0, all_cores)
    if pin == "pin":
        pin_rank_to_cores(local, world)
    pipe = FrameSetPipeline(n, h, w, NOMINAL, sr459.make_gaussian_psf(), K, depth=depth, in_dtype="u8", out_dtype="u8", use_graph=True, lanes=lanes)
    pipe.run([sets[i % 4] for i in range(8)])
    best = 1e9
    for rep in range(3):
        torch.cuda.synchronize()
        if world > 1: dist.barrier()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        pipe.run([sets[i % 4] for i in range(60)])
        torch.cuda.synchronize(); best = min(best, (time.perf_counter() - t0) / 60)
    t = torch.tensor([best], dtype=torch.float64, device="cuda")
    if world > 1:
        allt = [torch.zeros_like(t) for _ in range(world)]; dist.all_gather(allt, t); per = [float(x.item()) for x in allt]
    else:
        per = [best]
    if rank == 0:
        print(f"{pin} lanes={lanes} depth={pipe.depth}: slowest rank {max(per)*1e3:.3f} ms per job, per rank {[round(x*1e3,3) for x in per]} -> {world*4*h*w/1e6/max(per):.0f} HR Mpix/s", flush=True)
    pipe.close()
if world > 1:
    dist.barrier(); dist.destroy_process_group()
