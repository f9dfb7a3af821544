#!/usr/bin/env python3
"""bench.py -- SAA+IBP HR Mpix/s on B200 (BASELINE.json's metric), one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A "step" is one complete SR job of config 1 (BASELINE.json configs[0]: 4 XPR-shifted
1536x2048 mono frames -> 3072x4096, Gaussian 7x7 PSF): SAA initialisation followed by 50 IBP
iterations, on synthetic S1-noise frames (np.random.default_rng(459)).

  value    whole-job throughput with the LR frames already resident in HBM, CUDA-event timed,
           max over ranks; a ring of 4 distinct frame sets (200 MB of LR data + 150 MB of HR
           working set per job > the 126 MB L2) is cycled so no step finds its inputs cached.
  e2e      the same jobs through the public host API (FrameSetPipeline) in the formats run_sr.py moves: pinned host
           uint8 frames (what load_gray reads, run_sr.py:73-75) -> H2D -> u8->f32, SAA+IBP, clip + truncation ->
           D2H of the uint8 HR image (what run_sr.py:303 writes), copies inside the timed region.
  e2e_f32  the same with float32 frames in and the float32 HR image out (4x the PCIe bytes: 100 MB per job, which
           saturates the host side beyond 2 GPUs; reported beside e2e, not instead of it).
  roofline the fused IBP-iteration kernel: algorithmic bytes (8+n)*H*W per launch divided by its
           mean duration, measured live with CUDA events around the IBP loop of every timed step.
  cpu_baseline  the reference's CPU path (oracle/scipy_port.py: the same scipy calls the reference
           makes) on the host cores, on a bounded sample, rank 0 at N=1 only.

Multi-GPU: independent frame sets per rank (weak scaling, no data-path collective), launched by
torchrun; barrier + max-over-ranks timing.  --impl reference times the CPU port only.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

NOMINAL = [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]
WORKLOADS = {
    # name: (n, h, w, iters)
    "c1": (4, 1536, 2048, 50),
    "c2size": (4, 768, 1024, 50),
    "small": (4, 384, 512, 50),
}
METRIC = "SAA+IBP HR Mpix/s (4 frames -> 3072x4096, 50 IBP iterations)"
UNIT = "HR Mpix/s"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


_REAL_STDOUT = None


def guard_stdout():
    """stdout carries exactly ONE JSON line: anything libraries print there (NCCL's version banner comes from C)
    is sent to stderr; emit() writes the line to the real stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md: 6.65 TB/s)"


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        rows = []
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                rows.append((int(parts[0]), float(parts[1]), float(parts[2]), float(parts[3]), parts[5:9]))
            except ValueError:
                continue
        mine = [r for r in rows if r[0] == self.index] or rows
        if not mine:
            return out
        # "under load": the upper half of the power samples
        pw = sorted(r[3] for r in mine)
        thr = pw[len(pw) // 2]
        loaded = [r for r in mine if r[3] >= thr] or mine
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in loaded for i in range(4) if r[4][i].lower().startswith("active")})
        out.update(sm_mhz=statistics.median(r[1] for r in loaded), sm_max_mhz=max(r[2] for r in mine),
                   reasons=reasons, samples=len(mine), power_w_max=max(r[3] for r in mine))
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        return out


# ------------------------------------------------------------------------------------------------
# CPU baseline (reference's path on scipy, host cores)
# ------------------------------------------------------------------------------------------------
def cpu_sample(n, h, w, k_iter, cores, seed0=459):
    """Run `cores` concurrent single-threaded copies of the scipy port; returns aggregate HR Mpix/s
    extrapolated to the K=50 job (per-iteration cost is constant), plus details."""
    import multiprocessing as mp
    from oracle import scipy_port, sr_oracle
    psf = sr_oracle.make_gaussian_psf()
    args = [(seed0 + i, n, h, w, NOMINAL, psf, k_iter) for i in range(cores)]
    t0 = time.perf_counter()
    if cores == 1:
        res = [scipy_port.timed_sample(args[0])]
    else:
        ctx = mp.get_context("fork")
        with ctx.Pool(cores) as pool:
            res = pool.map(scipy_port.timed_sample, args)
    wall = time.perf_counter() - t0
    hr_px = 4.0 * h * w
    rate = 0.0
    for t_saa, t_iter in res:
        rate += hr_px / (t_saa + 50.0 * t_iter) / 1e6
    return rate, wall, res


def run_reference_arm(args):
    """--impl reference: the CPU port with all host threads, each step a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    n, h, w, iters = WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    sh, sw, k_iter = 384, 512, 2
    vals, walls = [], []
    for i in range(args.warmup + args.steps):
        rate, wall, _ = cpu_sample(n, sh, sw, k_iter, cores, seed0=459 + 100 * i)
        if i >= args.warmup:
            vals.append(rate)
            walls.append(wall)
    value = statistics.mean(vals)
    sample = (f"each step: {cores} concurrent single-threaded processes, each SAA + {k_iter} IBP iterations on "
              f"{n}x{sh}x{sw} S1-noise frames (float64, scipy); per-iteration cost extrapolated to the 50-iteration job")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * statistics.mean(walls), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"config1: {n}x{h}x{w} mono frames -> {2*h}x{2*w}, Gaussian 7x7 PSF, SAA + {iters} IBP iterations",
                   "n_frames": n, "lr_shape": [h, w], "hr_shape": [2 * h, 2 * w], "ibp_iterations": iters, "psf": "gaussian7"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------
def run_b200_arm(args):
    import torch
    import torch.distributed as dist
    import sr459
    from sr459.driver import FrameSetPipeline

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device; there is no CPU fallback (use --impl reference for the CPU port)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = sr459._lib.load()

    n, h, w, iters = WORKLOADS[args.workload]
    H, W = 2 * h, 2 * w
    psf = sr459.make_gaussian_psf()
    ring = 4
    rng = np.random.default_rng(459 + 1000 * rank)
    host_sets = [torch.from_numpy(rng.uniform(0, 255, (n, h, w)).astype(np.float32)).pin_memory() for _ in range(ring)]
    dev_sets = [hs.cuda() for hs in host_sets]
    plan = sr459.Plan(n, h, w, NOMINAL, psf)
    hr = torch.empty((H, W), dtype=torch.float32, device="cuda")

    def job_eager(i):
        frames = [dev_sets[i % ring][k] for k in range(n)]
        plan.saa(frames, out=hr)
        plan.ibp_(frames, hr, iters, want_errors=False)

    # One captured launch sequence per stage and ring slot (CUDA graphs: the 55 launches of a job replay without
    # host launch gaps); the stage events are recorded between the replays, on the same stream.
    graphs = []

    def job(i, ev=None):
        g_saa, g_ibp = graphs[i % ring]
        if ev: ev[0].record()
        g_saa.replay()
        if ev: ev[1].record()
        g_ibp.replay()
        if ev: ev[2].record()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    job_eager(0)                                     # lazy plan scratch is allocated outside any capture
    torch.cuda.synchronize()
    cap = torch.cuda.Stream()
    launches_per_job = 0
    for i in range(ring):
        frames = [dev_sets[i][k] for k in range(n)]
        c0 = lib.sr459_launch_count()
        g_saa, g_ibp = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.graph(g_saa, stream=cap):
            plan.saa(frames, out=hr)
        with torch.cuda.graph(g_ibp, stream=cap):
            plan.ibp_(frames, hr, iters, want_errors=False)
        graphs.append((g_saa, g_ibp))
        launches_per_job = lib.sr459_launch_count() - c0      # kernels captured = kernels every replay launches
    for i in range(args.warmup):
        job(i)
    barrier()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    start.record()
    for i in range(args.steps):
        job(args.warmup + i, evs[i])
    end.record()
    barrier()
    launches = launches_per_job * args.steps
    total_ms = start.elapsed_time(end)
    saa_ms = sum(e[0].elapsed_time(e[1]) for e in evs) / args.steps
    ibp_ms = sum(e[1].elapsed_time(e[2]) for e in evs) / args.steps
    errs = plan.errors(iters)

    # ---- e2e through the host API ------------------------------------------------------------
    pipe = FrameSetPipeline(n, h, w, NOMINAL, psf, iters, depth=3, in_dtype="f32", out_dtype="f32", use_graph=True)
    sets = [host_sets[i % ring] for i in range(args.steps)]
    pipe.run([host_sets[i % ring] for i in range(max(args.warmup, 3))])
    barrier()
    t0 = time.perf_counter()
    pipe.run(sets)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    lat = pipe.run_latency(sets[: min(5, len(sets))])
    barrier()
    # the headline e2e: the same jobs in run_sr.py's file formats, uint8 frames in (load_gray, :73-75), uint8 HR image out (:303)
    pipe8 = FrameSetPipeline(n, h, w, NOMINAL, psf, iters, depth=3, in_dtype="u8", out_dtype="u8", use_graph=True)
    sets8 = [torch.from_numpy(rng.integers(0, 256, (n, h, w), dtype=np.uint8)).pin_memory() for _ in range(ring)]
    pipe8.run([sets8[i % ring] for i in range(max(args.warmup, 3))])
    barrier()
    t0 = time.perf_counter()
    pipe8.run([sets8[i % ring] for i in range(args.steps)])
    torch.cuda.synchronize()
    e2e8_s = time.perf_counter() - t0
    lat8 = pipe8.run_latency([sets8[i % ring] for i in range(min(5, args.steps))])
    barrier()
    clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([total_ms, e2e_s * 1e3, ibp_ms, e2e8_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms, ibp_ms_max, e2e8_ms = [float(x) for x in t.tolist()]

    hr_mpix = H * W / 1e6
    value = world * args.steps * hr_mpix / (total_ms / 1e3)
    e2e_value = world * args.steps * hr_mpix / (e2e_ms / 1e3)

    peak, peak_src = measured_peak()
    alg_iter = (8 + n) * H * W                       # bytes: hr in 4 + hr out 4 + n LR frames at 1/4 density
    k_ms = ibp_ms / iters
    achieved = alg_iter / (k_ms * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get("k_ibp_stream_bytes_per_launch")
    except Exception:
        pass
    alg_job = ((n + 4) + iters * (8 + n)) * H * W

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"config1: {n}x{h}x{w} mono frames -> {H}x{W}, Gaussian 7x7 PSF, nominal +-0.5 px shifts, "
                               f"SAA + {iters} IBP iterations per step; one independent frame set per step per GPU; each stage replayed as a CUDA graph",
                   "n_frames": n, "lr_shape": [h, w], "hr_shape": [H, W], "ibp_iterations": iters, "psf": "gaussian7",
                   "cache": f"ring of {ring} distinct frame sets ({ring * n * h * w * 4 >> 20} MiB LR) + {3 * H * W * 4 >> 20} MiB HR "
                            "working set per job, larger than the 126 MB L2; no explicit flush",
                   "parallelism": f"batch-dp{world}"},
        "stage_ms": {"saa": saa_ms, "ibp": ibp_ms, "ibp_per_iteration": k_ms},
        "job_hbm_frac": alg_job / (total_ms / args.steps * 1e-3) / 1e9 / peak,
        "roofline": {"bound": "hbm", "kernel": "k_ibp_stream (fused IBP iteration: register-streaming warps, TMA row ring; rank-1 PSF, integer HR shifts)",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                     "algorithmic_bytes_per_launch": alg_iter, "peak_source": peak_src, "launch_ms": k_ms,
                     "dram_achieved": (traffic / (k_ms * 1e-3) / 1e9) if traffic else None,
                     "kernel_bytes_per_launch": 9 * H * W,
                     "note": "frac is SURVEY 8(d)'s algorithmic bytes, (8 + n) per HR pixel, over the measured DRAM copy peak; it can exceed 1 "
                             "because the kernel reads ONE pre-summed LR plane instead of the n frames (kernel_bytes_per_launch = 9 per HR "
                             "pixel) and part of the hr ping-pong pair (96 MiB) stays in the 126 MB L2: ncu counts `traffic` DRAM bytes per "
                             "launch (cold L2), dram_achieved = traffic / launch time"},
        "e2e": {"value": world * args.steps * hr_mpix / (e2e8_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": pipe8.h2d_bytes,
                "d2h_bytes_per_step": pipe8.d2h_bytes, "ms_per_step": e2e8_ms / args.steps, "latency_ms_p50": 1e3 * statistics.median(lat8),
                "api": "sr459.driver.FrameSetPipeline.run(in_dtype='u8', out_dtype='u8'): run_sr.py's own I/O formats -- pinned uint8 frames in "
                       "(load_gray, :73-75), uint8 HR image out (clip + truncation, :303), conversions on the device; 3 jobs in flight, "
                       "one CUDA-graph replay per job"},
        "e2e_f32": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": pipe.h2d_bytes, "d2h_bytes_per_step": pipe.d2h_bytes,
                    "ms_per_step": e2e_ms / args.steps, "latency_ms_p50": 1e3 * statistics.median(lat),
                    "api": "the same pipeline with pinned float32 frames in and the float32 HR image out (100 MB of PCIe traffic per job)"},
        "gpu_launches": int(launches),
        "final_mse": errs[-1],
        "clocks": clocks,
    }

    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        os.environ.setdefault("OMP_NUM_THREADS", "1")
        sh, sw, k_iter = 768, 1024, 2
        rate1, wall1, res1 = cpu_sample(n, sh, sw, k_iter, 1)
        rate_all, wall_all, _ = cpu_sample(n, 384, 512, k_iter, cores)
        line["cpu_baseline"] = {
            "value": rate_all, "unit": UNIT, "cores": cores, "kind": "port",
            "one_core_value": rate1,
            "sample": (f"all-core: {cores} concurrent single-threaded processes, each SAA + {k_iter} IBP iterations on {n}x384x512 "
                       f"S1-noise frames; one-core: the same on {n}x{sh}x{sw} frames ({res1[0][0]:.1f} s SAA, {res1[0][1]:.1f} s/iteration); "
                       "float64 scipy calls identical to the reference's; per-iteration cost extrapolated to the 50-iteration job"),
        }
    if rank == 0:
        emit(line)
    pipe.close()
    pipe8.close()
    plan.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="c1")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    guard_stdout()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_b200_arm(args)


if __name__ == "__main__":
    sys.exit(main())
