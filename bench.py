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
  roofline the fused IBP-iteration kernel: algorithmic bytes (8+n)*H*W per launch divided by its
           mean duration, measured live with CUDA events around the IBP loop of every timed step.
  cpu_baseline  the reference's CPU path (oracle/scipy_port.py: the same scipy calls the reference
           makes) on the host cores, rank 0 at N=1 only: one sample of the work of ONE FRAME at the
           job's FULL size (4x1536x2048 -> 3072x4096) on every core at once.
  configs  the other BASELINE.json configurations, measured in the same run (device-resident, CUDA
           graphs, CUDA events): c2 (fractional shifts + measured 7x7), monocal5, c1_measured7,
           c4 (31x31 measured PSF, 200 iterations), c5 (streaming burst); N > 1 adds `bands`
           (config 4 as row bands with the halo exchange).

Multi-GPU: independent frame sets per rank (weak scaling, no data-path collective), launched by
torchrun; barrier + max-over-ranks timing.  --impl reference times the CPU port only: every step is
one frame-iteration of the job at full size on every host core at once (see run_reference_arm).
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
# workload description shared by both arms (the driver compares the `config` of the two lines)
# ------------------------------------------------------------------------------------------------
def workload_config(n, h, w, iters, world):
    H, W = 2 * h, 2 * w
    ring = 4
    return {
        "workload": f"config1: {n}x{h}x{w} mono frames -> {H}x{W}, Gaussian 7x7 PSF, nominal +-0.5 px shifts, "
                    f"SAA + {iters} IBP iterations per job; one independent frame set per job per GPU",
        "n_frames": n, "lr_shape": [h, w], "hr_shape": [H, W], "ibp_iterations": iters, "psf": "gaussian7",
        "shifts_lr_px": [list(s) for s in NOMINAL],
        "parallelism": f"batch-dp{world}",
        "b200_arm": {
            "step": "one whole job; each stage replayed as a CUDA graph; jobs round-robin over `lanes` independent compute lanes (own plan, stream, graphs)",
            "cache": f"ring of {ring} distinct frame sets ({ring * n * h * w * 4 >> 20} MiB LR) + {3 * H * W * 4 >> 20} MiB HR "
                     "working set per job, larger than the 126 MB L2; no explicit flush",
        },
        "reference_arm": {
            "step": "bounded sample of the same job AT ITS FULL SIZE: every host core at once runs the reference's scipy "
                    f"calls for ONE frame of ONE iteration (forward_model + residual + back_project on {H}x{W} float64); "
                    "the SAA part (zoom + shift of one frame) is sampled once before the timed steps",
            "job_time_model": "t_job = n*t_saa_frame + K*(n*t_frame_iteration + t_update): identical scipy calls on identical "
                              "shapes repeated n*K times, nothing extrapolated across image sizes",
            "extrapolated_over": "frame and iteration count only",
        },
    }


# ------------------------------------------------------------------------------------------------
# CPU baseline (reference's path on scipy, host cores)
# ------------------------------------------------------------------------------------------------
def cpu_procs(cores, per_proc_gb=1.6):
    """Concurrent full-size workers the host memory allows (each holds ~12 float64 HR-sized arrays)."""
    try:
        import psutil
        avail = psutil.virtual_memory().available / 2 ** 30
        return max(1, min(cores, int(avail / per_proc_gb)))
    except Exception:
        return cores


def cpu_full_size_step(n, h, w, iters, procs, with_saa, seed0, t_saa_frame=None):
    """One bounded sample at the job's full size on `procs` cores at once (see workload_config.reference_arm).
    Returns dict(rate HR Mpix/s summed over the workers, wall, per-piece mean times)."""
    import multiprocessing as mp
    from oracle import scipy_port, sr_oracle
    psf = sr_oracle.make_gaussian_psf()
    args = [(seed0 + i, h, w, NOMINAL[i % len(NOMINAL)], psf, with_saa) for i in range(procs)]
    t0 = time.perf_counter()
    if procs == 1:
        res = [scipy_port.full_size_frame_sample(args[0])]
    else:
        with mp.get_context("fork").Pool(procs) as pool:
            res = pool.map(scipy_port.full_size_frame_sample, args)
    wall = time.perf_counter() - t0
    if with_saa:
        t_saa_frame = statistics.mean(r[0] for r in res)
    hr_mpix = 4.0 * h * w / 1e6
    rate = sum(hr_mpix / (n * t_saa_frame + iters * (n * r[1] + r[2])) for r in res)
    return {"rate": rate, "wall": wall, "t_saa_frame": t_saa_frame, "t_frame_iteration": statistics.mean(r[1] for r in res),
            "t_update": statistics.mean(r[2] for r in res), "procs": procs}


def run_reference_arm(args):
    """--impl reference: the reference's CPU path (oracle/scipy_port.py: the scipy calls of mono_cal_target/run_sr.py:157-209)
    on all host cores.  Every step runs, on every core at once, the work of one frame of one IBP iteration at the FULL
    4x1536x2048 -> 3072x4096 size; value = sum over the cores of H*W / (n*t_saa_frame + K*(n*t_frame_iteration + t_update))."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    n, h, w, iters = WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    procs = cpu_procs(cores)
    calib = cpu_full_size_step(n, h, w, iters, procs, True, 459)            # SAA piece + first warm-up
    log(f"reference arm: {procs} workers, t_saa_frame {calib['t_saa_frame']:.2f} s, t_frame_iteration {calib['t_frame_iteration']:.2f} s")
    steps = []
    for i in range(max(args.warmup - 1, 0) + args.steps):
        r = cpu_full_size_step(n, h, w, iters, procs, False, 1000 + 100 * i, calib["t_saa_frame"])
        if i >= max(args.warmup - 1, 0):
            steps.append(r)
    value = statistics.mean(r["rate"] for r in steps)
    t_fi = statistics.mean(r["t_frame_iteration"] for r in steps)
    t_up = statistics.mean(r["t_update"] for r in steps)
    sample = (f"each step: {procs} concurrent single-threaded processes (host has {cores} cores), each the scipy calls of ONE frame of ONE "
              f"IBP iteration on full-size arrays ({2*h}x{2*w} float64 HR, {h}x{w} LR): {t_fi:.2f} s per frame-iteration, "
              f"{calib['t_saa_frame']:.2f} s per SAA frame (sampled once), {t_up:.3f} s per update; job = {n}*t_saa + {iters}*({n}*t_fi + t_upd) "
              f"= {n * calib['t_saa_frame'] + iters * (n * t_fi + t_up):.0f} s per core")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * statistics.mean(r["wall"] for r in steps), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(n, h, w, iters, args.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": procs, "kind": "port", "sample": sample,
                         "one_core_job_seconds_under_load": n * calib["t_saa_frame"] + iters * (n * t_fi + t_up)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------
CALIBRATED = [(0.47225, -0.43385), (0.4897, 0.4641), (-0.4798, -0.4553), (-0.4856, 0.4369)]   # rgb_cal_target/results/*/shifts.json
MONOCAL = [(0.0, 0.0)] + NOMINAL                                                              # mono_cal_target/run_sr.py:60-66
FFMA2_PEAK_TMAC = 35.7      # tools/probes/ffma2_probe.cu on B200: packed fma.rn.f32x2 issue ceiling


def committed_psfs():
    """The reference-generated measured PSFs (load_measured_psf on the shipped pinhole crops, PSF_HALFWIDTH 3 and 15),
    committed under tests/golden by tests/golden/make_golden.py."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "g_psf.npz"))
    return g["k7"], g["k31"]


def time_config(sr459, torch, name, h, w, shifts, psf, iters, reps, peak, seed, desc):
    """SAA + K iterations of one configuration, device-resident, each stage one CUDA-graph replay, CUDA events."""
    n, H, W = len(shifts), 2 * h, 2 * w
    rng = np.random.default_rng(seed)
    frames = [torch.from_numpy(rng.uniform(0, 255, (h, w)).astype(np.float32)).cuda() for _ in range(n)]
    plan = sr459.Plan(n, h, w, shifts, psf)
    hr = torch.empty((H, W), dtype=torch.float32, device="cuda")
    lib = sr459._lib.load()
    plan.saa(frames, out=hr); plan.ibp_(frames, hr, iters, want_errors=False)      # lazy scratch (sized by K), outside capture
    torch.cuda.synchronize()
    cap = torch.cuda.Stream()
    g_saa, g_ibp = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
    c0 = lib.sr459_launch_count()
    with torch.cuda.graph(g_saa, stream=cap):
        plan.saa(frames, out=hr)
    c1 = lib.sr459_launch_count()
    with torch.cuda.graph(g_ibp, stream=cap):
        plan.ibp_(frames, hr, iters, want_errors=False)
    c2 = lib.sr459_launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    saa_ms, ibp_ms = [], []
    for r in range(reps + 1):
        ev[0].record(); g_saa.replay(); ev[1].record(); g_ibp.replay(); ev[2].record()
        torch.cuda.synchronize()
        if r:                                        # replay 0 is the warm-up
            saa_ms.append(ev[0].elapsed_time(ev[1])); ibp_ms.append(ev[1].elapsed_time(ev[2]))
    errs = plan.errors(iters)
    # the same job on `lanes` independent compute lanes at once (independent frame sets, as the reference's per-session loops):
    # the kernels of these configurations are latency- or wave-bound one job at a time, concurrent jobs fill the machine
    lanes = 4
    lane = [dict(plan=plan, hr=hr, g=(g_saa, g_ibp), st=torch.cuda.Stream())]
    for _ in range(lanes - 1):
        pl = sr459.Plan(n, h, w, shifts, psf)
        hl = torch.empty((H, W), dtype=torch.float32, device="cuda")
        pl.saa(frames, out=hl); pl.ibp_(frames, hl, iters, want_errors=False)
        torch.cuda.synchronize()
        ga, gb = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.graph(ga, stream=cap):
            pl.saa(frames, out=hl)
        with torch.cuda.graph(gb, stream=cap):
            pl.ibp_(frames, hl, iters, want_errors=False)
        lane.append(dict(plan=pl, hr=hl, g=(ga, gb), st=torch.cuda.Stream()))
    batch_ms = []
    for r in range(reps + 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for ln in lane:
            ln["st"].wait_event(e0)
            with torch.cuda.stream(ln["st"]):
                ln["g"][0].replay(); ln["g"][1].replay()
        for ln in lane:
            ej = torch.cuda.Event(); ej.record(ln["st"]); torch.cuda.current_stream().wait_event(ej)
        e1.record()
        torch.cuda.synchronize()
        if r:
            batch_ms.append(e0.elapsed_time(e1))
    t_batch = statistics.median(batch_ms)
    for ln in lane[1:]:
        ln["g"] = None
        ln["plan"].close()
    plan.close()
    t_saa, t_ibp = statistics.median(saa_ms), statistics.median(ibp_ms)
    us_iter = 1e3 * t_ibp / iters
    alg_iter = (8 + n) * H * W
    return {
        "workload": desc, "n_frames": n, "hr_shape": [H, W], "ibp_iterations": iters, "integer_hr_shifts": bool(sr459.Plan.shifts_are_integer(shifts)),
        "saa_us": 1e3 * t_saa, "us_per_iteration": us_iter, "job_ms": t_saa + t_ibp,
        "hr_mpix_per_s": H * W / 1e6 / ((t_saa + t_ibp) / 1e3),
        "batch": {"lanes": lanes, "jobs": lanes, "ms": t_batch, "hr_mpix_per_s": lanes * H * W / 1e6 / (t_batch / 1e3),
                  "speedup_vs_one_job_at_a_time": lanes * (t_saa + t_ibp) / t_batch,
                  "note": "`lanes` independent frame sets of this configuration at once (own plan and stream each)"},
        "launches_per_iteration": (c2 - c1 - 1) / iters, "launches_saa": c1 - c0,
        "roofline": {"bound": "hbm", "algorithmic_bytes_per_iteration": alg_iter, "achieved": alg_iter / (us_iter * 1e-6) / 1e9,
                     "peak": peak, "unit": "GB/s", "frac": alg_iter / (us_iter * 1e-6) / 1e9 / peak},
        "final_mse": errs[-1], "mse_monotone": all(b <= a for a, b in zip(errs, errs[1:])),
    }


def stream_record(sr459, torch, n, h, w, psf, iters):
    """Config 5: streaming burst mode through driver.BurstStream (uint8 frames in, uint8 HR image out, PCIe both ways)."""
    from sr459.driver import BurstStream, play_camera
    rng = np.random.default_rng(4595)
    pool = [rng.integers(0, 256, (h, w), dtype=np.uint8) for _ in range(8)]
    bs = BurstStream(n, h, w, NOMINAL, psf, iters, depth=3, use_graph=True)
    for _ in range(3):
        for k in range(n):
            bs.push_frame(pool[k])
        bs.wait_result()
    out = {"workload": f"config5: continuous {n}-frame uint8 sets {h}x{w} -> {2*h}x{2*w} uint8, Gaussian 7x7, SAA + {iters} IBP iterations; "
                       "latency = commit of a set's last frame -> uint8 HR image on the host",
           "h2d_bytes_per_set": bs.h2d_bytes, "d2h_bytes_per_set": bs.d2h_bytes}
    for tag, sets, fps in (("free_running", 100, 0.0), ("camera_48fps", 16, 48.0)):
        lat, wall = play_camera(bs, pool, sets, fps)
        ms = sorted(1e3 * x for x in lat)
        out[tag] = {"sets": sets, "latency_ms_p50": statistics.median(ms), "latency_ms_p99": ms[min(len(ms) - 1, int(0.99 * len(ms)))],
                    "hr_frames_per_s": sets / wall, "hr_mpix_per_s": sets * 4.0 * h * w / 1e6 / wall}
    bs.close()
    # a free-running source on one compute lane per ring slot: the sets overlap at their kernel boundaries (sustained rate up,
    # per-set latency up while several sets are in flight); the camera-rate figures above are the latency the design is for
    bs = BurstStream(n, h, w, NOMINAL, psf, iters, depth=4, use_graph=True, lanes=4)
    for _ in range(4):
        for k in range(n):
            bs.push_frame(pool[k])
        bs.wait_result()
    lat, wall = play_camera(bs, pool, 100, 0.0)
    ms = sorted(1e3 * x for x in lat)
    out["free_running_4_lanes"] = {"sets": 100, "latency_ms_p50": statistics.median(ms), "latency_ms_p99": ms[min(len(ms) - 1, 99)],
                                   "hr_frames_per_s": 100 / wall, "hr_mpix_per_s": 100 * 4.0 * h * w / 1e6 / wall}
    bs.close()
    return out


def parity_spot_check(sr459, iters):
    """Outside every timed region: the kernels the bench runs (auto path) against the float64 oracle at the benchmarked
    iteration count on a small frame set.  The oracle is the checker here, never the thing measured."""
    from oracle import sr_oracle as O
    rng = np.random.default_rng(459)
    lrs = [rng.uniform(0, 255, (96, 128)) for _ in NOMINAL]
    psf = O.make_gaussian_psf()
    saa_o = O.shift_and_add_integer(lrs, NOMINAL)
    hr_o, e_o = O.ibp_integer_shifts(lrs, NOMINAL, psf, saa_o, n_iter=iters)
    hr, errs = sr459.ibp(lrs, NOMINAL, sr459.make_gaussian_psf(), sr459.shift_and_add(lrs, NOMINAL), n_iter=iters)
    d = float(np.max(np.abs(hr - hr_o)))
    rel = float(np.max(np.abs(np.array(errs) / np.array(e_o) - 1.0)))
    return {"shape": "4x96x128 S1 noise", "iterations": iters, "max_abs_vs_oracle_dn": d, "mse_rel_vs_oracle": rel,
            "ok": bool(d <= 0.0255 and rel <= 1e-4)}


def bands_record(sr459, torch, dist, world, rank, peak, k31):
    """Config 4 under torchrun: ONE 3072x4096 HR image as `world` row bands, 200 iterations, every band starting from its
    own LR rows (banded SAA), the halo rows exchanged by a kernel over NVLink peer memory (csrc/bands.cu; no collective on
    the data path), against the same job on one GPU (every rank runs that too, to check its own band)."""
    from sr459.bands import BandPlan
    h, w, K = 1536, 2048, 200
    rng = np.random.default_rng(4594)                       # identical on every rank
    lrs = [torch.from_numpy(rng.uniform(0, 255, (h, w)).astype(np.float32)).cuda() for _ in NOMINAL]
    out = {}

    def sync():
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()

    how = {"gaussian7": "the iteration kernel (k_ibp_stream) stores its boundary rows straight into the neighbours' halo rows over NVLink peer "
                        "memory -- border segments run first, one remote arrival count per warp, the next iteration's border warps wait on the counts",
           "measured31": "after the lattice kernels (narrow 32 x 32-site tiles on thin bands) k_band_push stores the boundary rows into the neighbours' "
                         "halo rows over NVLink peer memory and waits for theirs"}
    for tag, psf in (("measured31", k31), ("gaussian7", sr459.make_gaussian_psf())):
        full = sr459.Plan(4, h, w, NOMINAL, psf)
        saa = full.saa(lrs)
        ref = saa.clone()
        full.ibp_(lrs, ref, K, want_errors=False)               # warm-up: lazy scratch sized by K
        ref.copy_(saa)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(); ref_errs = full.ibp_(lrs, ref, K); e1.record()
        torch.cuda.synchronize()
        single_ms = e0.elapsed_time(e1)
        bp = BandPlan(4, h, w, NOMINAL, psf, rank, world).attach()
        lay = bp.layout
        ext = [f[lay.lr_ext_lo:lay.lr_ext_hi].contiguous() for f in lrs]
        bp.saa(ext); bp.run_peer(ext, K)                        # eager warm-up: scratch, peer mappings touched
        sync()
        errs, ms_runs, saa_ms = None, [], 0.0
        for rep in range(3):                                    # rep 0 captures the loop; reps 1-2 are measured
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record(); bp.saa(ext); s1.record()
            sync()
            saa_ms = s0.elapsed_time(s1)
            errs = bp.run_peer(ext, K, graph=True)
            sync()
            if rep:
                ms_runs.append(bp.last_loop_ms)
        res = bp.current()
        d = float((res[lay.owned] - ref[lay.own_y0:lay.own_y1]).abs().max().item())
        t = torch.tensor([min(ms_runs), d, single_ms, saa_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, dmax, single, saa_band = [float(x) for x in t.tolist()]
        out[tag] = {
            "workload": f"config4: one {2*h}x{2*w} HR image, 4 frames, {tag} PSF, {K} IBP iterations, {world} row bands, "
                        f"halo {lay.halo} rows; {how[tag]}; no collective on the data path, the whole loop is one CUDA graph; "
                        f"every band's initial image from its own LR rows (banded SAA)",
            "us_per_iteration": 1e3 * ms / K, "single_gpu_us_per_iteration": 1e3 * single / K,
            "speedup_vs_single_gpu": single / ms, "efficiency": single / ms / world,
            "banded_saa_us": 1e3 * saa_band,
            "max_abs_vs_single_gpu": dmax, "mse_rel_vs_single_gpu": float(np.max(np.abs(np.array(errs) / np.array(ref_errs) - 1.0))),
            "halo_bytes_per_iteration_per_neighbour": lay.halo * lay.W * 4,
            "nvlink_bytes_per_iteration_per_neighbour": lay.halo * lay.W * 4 + 4 * ((lay.W + 111) // 112 if tag == "gaussian7" else 1),
            "hr_mpix_per_s": 4.0 * h * w / 1e6 / (ms / 1e3),
            "timing": "CUDA events around one graph replay of the whole loop, best of 2, max over ranks",
        }
        bp.close(); full.close()
        sync()
    return out


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
    from sr459.driver import pin_rank_to_cores
    cpu_aff = pin_rank_to_cores(local, world)          # every rank on its own share of the host cores (no-op at N = 1)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = sr459._lib.load()

    n, h, w, iters = WORKLOADS[args.workload]
    H, W = 2 * h, 2 * w
    psf = sr459.make_gaussian_psf()
    ring = 4
    rng = np.random.default_rng(459 + 1000 * rank)
    dev_sets = [torch.from_numpy(rng.uniform(0, 255, (n, h, w)).astype(np.float32)).cuda() for _ in range(ring)]
    plan = sr459.Plan(n, h, w, NOMINAL, psf)
    hr = torch.empty((H, W), dtype=torch.float32, device="cuda")

    def job_eager(i):
        frames = [dev_sets[i % ring][k] for k in range(n)]
        plan.saa(frames, out=hr)
        plan.ibp_(frames, hr, iters, want_errors=False)

    # One captured launch sequence per stage and ring slot (CUDA graphs: the launches of a job replay without
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
    single_ms = start.elapsed_time(end)
    saa_ms = sum(e[0].elapsed_time(e[1]) for e in evs) / args.steps
    ibp_ms = sum(e[1].elapsed_time(e[2]) for e in evs) / args.steps
    errs = plan.errors(iters)

    # ---- the same K jobs on `lanes` independent compute lanes (batch mode: independent frame sets): every lane has its own
    # plan (scratch), hr buffer, graphs and stream, job i runs on lane i mod lanes.  The kernels of different jobs then overlap at
    # their launch boundaries and tails (each iteration kernel is ONE wave; a lone stream leaves the machine partly idle between
    # launches).  `value` is this pass; `stage_ms` / `roofline` come from the one-job-at-a-time pass above, where a launch has the
    # GPU to itself and its duration means something.
    L = max(1, args.lanes)
    lane_objs = [dict(plan=plan, hr=hr, graphs=graphs, stream=torch.cuda.Stream())]
    for l in range(1, L):
        pl = sr459.Plan(n, h, w, NOMINAL, psf)
        hl = torch.empty((H, W), dtype=torch.float32, device="cuda")
        f0 = [dev_sets[0][k] for k in range(n)]
        pl.saa(f0, out=hl); pl.ibp_(f0, hl, iters, want_errors=False)        # lazy scratch outside any capture
        torch.cuda.synchronize()
        gl = []
        for i in range(ring):
            frames = [dev_sets[i][k] for k in range(n)]
            g_saa, g_ibp = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
            with torch.cuda.graph(g_saa, stream=cap):
                pl.saa(frames, out=hl)
            with torch.cuda.graph(g_ibp, stream=cap):
                pl.ibp_(frames, hl, iters, want_errors=False)
            gl.append((g_saa, g_ibp))
        lane_objs.append(dict(plan=pl, hr=hl, graphs=gl, stream=torch.cuda.Stream()))

    def job_lane(i):
        lane = lane_objs[i % L]
        g_saa, g_ibp = lane["graphs"][i % ring]
        with torch.cuda.stream(lane["stream"]):
            g_saa.replay(); g_ibp.replay()

    def join_lanes():
        for lane in lane_objs:
            ev = torch.cuda.Event(); ev.record(lane["stream"]); torch.cuda.current_stream().wait_event(ev)

    for i in range(args.warmup):
        job_lane(i)
    join_lanes()
    start2, end2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    start2.record()
    for lane in lane_objs:
        lane["stream"].wait_event(start2)
    for i in range(args.steps):
        job_lane(args.warmup + i)
    join_lanes()
    end2.record()
    barrier()
    total_ms = start2.elapsed_time(end2)

    # ---- e2e through the host API: run_sr.py's file formats, uint8 frames in (load_gray, :73-75), uint8 HR image out (:303) ----
    pipe8 = FrameSetPipeline(n, h, w, NOMINAL, psf, iters, depth=2 * L, in_dtype="u8", out_dtype="u8", use_graph=True, lanes=L)
    sets8 = [torch.from_numpy(rng.integers(0, 256, (n, h, w), dtype=np.uint8)).pin_memory() for _ in range(ring)]
    pipe8.run([sets8[i % ring] for i in range(max(args.warmup, 3))])
    e2e_runs = []
    e2e_jobs = max(args.steps, 40)                    # jobs per timed run: the pipeline's fill + drain (one H2D and one D2H copy that
                                                      # cannot overlap compute, about 0.5 ms) must not dominate a short run
    for rep in range(3):                              # median of three timed runs of `e2e_jobs` jobs each
        barrier()
        t0 = time.perf_counter()
        pipe8.run([sets8[i % ring] for i in range(e2e_jobs)])
        torch.cuda.synchronize()
        e2e_runs.append((time.perf_counter() - t0) * args.steps / e2e_jobs)      # scaled to `steps` jobs
    e2e8_s = statistics.median(e2e_runs)
    lat8 = pipe8.run_latency([sets8[i % ring] for i in range(min(5, args.steps))])
    barrier()
    clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([total_ms, ibp_ms, e2e8_s * 1e3, max(e2e_runs) * 1e3, single_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, ibp_ms_max, e2e8_ms, e2e8_worst_ms, single_ms = [float(x) for x in t.tolist()]
    # every rank's own figures beside the max-over-ranks ones (a box whose GPUs do not all sit behind the same PCIe path shows up here)
    mine = torch.tensor([e2e8_s * 1e3 / args.steps, start2.elapsed_time(end2) / args.steps], dtype=torch.float64, device="cuda")
    per_rank = [mine.clone() for _ in range(world)]
    if world > 1:
        dist.all_gather(per_rank, mine)
    per_rank_e2e_ms = [round(float(x[0]), 4) for x in per_rank]
    per_rank_step_ms = [round(float(x[1]), 4) for x in per_rank]

    hr_mpix = H * W / 1e6
    value = world * args.steps * hr_mpix / (total_ms / 1e3)

    peak, peak_src = measured_peak()
    alg_iter = (8 + n) * H * W                       # bytes: hr in 4 + hr out 4 + n LR frames at 1/4 density
    k_ms = ibp_ms / iters
    achieved = alg_iter / (k_ms * 1e-3) / 1e9
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f)
            traffic, traffic_src = tj.get("k_ibp_stream_bytes_per_launch"), tj.get("source")
    except Exception:
        pass
    alg_job = ((n + 4) + iters * (8 + n)) * H * W
    kernel_bytes = 9 * H * W

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": workload_config(n, h, w, iters, world),
        "lanes": L, "per_rank_ms_per_step": per_rank_step_ms,
        "single_lane": {"value": world * args.steps * hr_mpix / (single_ms / 1e3), "ms_per_step": single_ms / args.steps,
                        "note": "the same K jobs one at a time on one stream; stage_ms and roofline.launch_ms are measured in this pass"},
        "stage_ms": {"saa": saa_ms, "ibp": ibp_ms, "ibp_per_iteration": k_ms},
        "job_hbm_frac": alg_job / (total_ms / args.steps * 1e-3) / 1e9 / peak,
        "roofline": {"bound": "hbm", "kernel": "k_ibp_stream (fused IBP iteration: register-streaming warps, TMA row ring; rank-1 PSF, integer HR shifts)",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                     "algorithmic_bytes_per_launch": alg_iter, "peak_source": peak_src, "launch_ms": k_ms,
                     "dram_achieved": (traffic / (k_ms * 1e-3) / 1e9) if traffic else None,
                     "kernel_bytes_per_launch": kernel_bytes, "kernel_bytes_frac": kernel_bytes / (k_ms * 1e-3) / 1e9 / peak,
                     "note": "frac is SURVEY 8(d)'s algorithmic bytes, (8 + n) per HR pixel, over the measured DRAM copy peak; it can exceed 1 "
                             "because the kernel reads ONE pre-summed LR plane instead of the n frames (kernel_bytes_per_launch = 9 per HR "
                             "pixel; kernel_bytes_frac is the same fraction on those bytes) and part of the hr ping-pong pair (96 MiB) stays "
                             "in the 126 MB L2: ncu counts `traffic` DRAM bytes per launch (cold L2), dram_achieved = traffic / launch time"},
        "e2e": {"value": world * args.steps * hr_mpix / (e2e8_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": pipe8.h2d_bytes,
                "d2h_bytes_per_step": pipe8.d2h_bytes, "ms_per_step": e2e8_ms / args.steps, "latency_ms_p50": 1e3 * statistics.median(lat8),
                "runs": 3, "jobs_per_run": e2e_jobs, "per_rank_ms_per_step": per_rank_e2e_ms, "cpu_affinity": cpu_aff, "worst_run_value": world * args.steps * hr_mpix / (e2e8_worst_ms / 1e3),
                "api": "sr459.driver.FrameSetPipeline.run(in_dtype='u8', out_dtype='u8'): run_sr.py's own I/O formats -- pinned uint8 frames in "
                       "(load_gray, :73-75), uint8 HR image out (clip + truncation, :303), conversions on the device; 2 jobs in flight on each of `lanes` compute lanes, "
                       "one CUDA-graph replay per job; median of 3 runs of `jobs_per_run` jobs each, reported per `steps` jobs (max over ranks of each rank's median)"},
        "gpu_launches": int(launches),
        "final_mse": errs[-1],
        "mse_monotone": all(b < a for a, b in zip(errs, errs[1:])),
        "clocks": clocks,
    }
    pipe8.close()
    for lane in lane_objs[1:]:
        lane["graphs"] = None
        lane["plan"].close()
    plan.close()
    del graphs, dev_sets, sets8, hr, lane_objs
    torch.cuda.empty_cache()

    # ---- the other BASELINE configurations, same run (rank 0 measures the single-GPU ones) -------------------
    if not args.no_configs:
        k7, k31 = committed_psfs()
        cfg = {}
        if rank == 0:
            cfg["c2"] = time_config(sr459, torch, "c2", 768, 1024, CALIBRATED, k7, 50, 5, peak, 4592,
                                    "config2: rgb_cal_target red-Bayer 4x768x1024 -> 1536x2048, calibrated fractional shifts, measured 7x7 PSF "
                                    "(tests/golden/g_psf.npz k7, generated by the reference's load_measured_psf), SAA + 50 iterations")
            cfg["monocal5"] = time_config(sr459, torch, "monocal5", 1536, 2048, MONOCAL, psf, 80, 3, peak, 4596,
                                          "mono_cal_target's own geometry: centre + 4 shifted frames (two parity lattices) 5x1536x2048 -> 3072x4096, "
                                          "Gaussian 7x7, SAA + 80 iterations (run_sr.py:56-66)")
            cfg["c1_measured7"] = time_config(sr459, torch, "c1_measured7", 1536, 2048, NOMINAL, k7, 50, 3, peak, 4597,
                                              "config1 with --psf measured: 4x1536x2048 -> 3072x4096, nominal shifts, measured 7x7 PSF (19 non-zero taps), SAA + 50 iterations")
            c4 = time_config(sr459, torch, "c4", 1536, 2048, NOMINAL, k31, 200, 2, peak, 4594,
                             "config4 on ONE GPU: 4x1536x2048 -> 3072x4096, measured 31x31 PSF (tests/golden/g_psf.npz k31: 830 non-zero taps), 200 iterations")
            macs = 2.0 * 961 * (H * W / 4)                 # forward blur at the sampled sites + transposed blur of the site plane
            c4["fma"] = {"bound": "fp32-fma", "useful_mac_per_iteration": macs, "achieved_tmac_s": macs / (c4["us_per_iteration"] * 1e-6) / 1e12,
                         "peak_tmac_s": FFMA2_PEAK_TMAC, "frac": macs / (c4["us_per_iteration"] * 1e-6) / 1e12 / FFMA2_PEAK_TMAC,
                         "peak_source": "tools/probes/ffma2_probe.cu (packed fma.rn.f32x2 issue ceiling measured on B200)"}
            cfg["c4"] = c4
            cfg["c5"] = stream_record(sr459, torch, n, h, w, psf, iters)
            cfg["c3"] = {"workload": "config3: 64 independent config-1 frame sets, data-parallel: this line's own `value` / `e2e` "
                                     "(one set per step per GPU, no data-path collective)", "see": "value, e2e, n_gpus"}
        if world > 1:
            barrier()
            b = bands_record(sr459, torch, dist, world, rank, peak, k31)
            if rank == 0:
                cfg["bands"] = b
        line["configs"] = cfg

    if rank == 0 and not args.no_cpu:
        line["parity"] = parity_spot_check(sr459, iters)
        line["parity_checked"] = bool(line["parity"]["ok"] and line["mse_monotone"])
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        os.environ.setdefault("OMP_NUM_THREADS", "1")
        one = cpu_full_size_step(n, h, w, iters, 1, True, 459)
        procs = cpu_procs(cores)
        allc = cpu_full_size_step(n, h, w, iters, procs, True, 4590)
        line["cpu_baseline"] = {
            "value": allc["rate"], "unit": UNIT, "cores": procs, "kind": "port",
            "one_core_value": one["rate"],
            "one_core_job_seconds": n * one["t_saa_frame"] + iters * (n * one["t_frame_iteration"] + one["t_update"]),
            "sample": (f"FULL-SIZE sample ({n}x{h}x{w} -> {H}x{W}, float64, the reference's scipy calls): the work of ONE frame -- zoom + shift "
                       f"(SAA) and forward_model + residual + back_project (one IBP iteration) -- on 1 core alone ({one['t_saa_frame']:.2f} s + "
                       f"{one['t_frame_iteration']:.2f} s) and on {procs} cores at once ({allc['t_saa_frame']:.2f} s + {allc['t_frame_iteration']:.2f} s each; "
                       f"host has {cores}); job = {n}*t_saa + {iters}*({n}*t_iter + t_update), nothing extrapolated across image sizes"),
        }
    if rank == 0:
        emit(line)
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="c1")
    ap.add_argument("--lanes", type=int, default=4, help="independent compute lanes per GPU (own plan, stream and graphs): job i runs on lane i mod lanes")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline and oracle spot-check legs")
    ap.add_argument("--no-configs", action="store_true", help="skip the `configs` object (other BASELINE configurations)")
    args = ap.parse_args()
    guard_stdout()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_b200_arm(args)


if __name__ == "__main__":
    sys.exit(main())
