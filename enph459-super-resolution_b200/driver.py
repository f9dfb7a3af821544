"""Batched / streaming driver: independent frame sets through SAA + IBP on one GPU.

The reference processes sessions and reps in plain ``for`` loops
(mono_cal_target/run_sr.py:358-360, mono_barcodes/run_sr.py:301, :393-395); every
iteration is an independent SR job.  On a B200 one job is a few milliseconds, so
the PCIe copies either side of it matter as much as the kernels.  FrameSetPipeline
keeps ``depth`` jobs in flight on three CUDA streams (H2D, compute, D2H) with
pinned host staging, so the copy of job i+1 and the read-back of job i-1 overlap
the kernels of job i.  Multi-GPU batch mode (one process per GPU) simply gives
every rank its own pipeline and its own share of the frame sets: no collective is
needed on the data path.
"""
from __future__ import annotations

import time
from typing import Iterable, List, Optional, Sequence

import numpy as np

from . import ops


class FrameSetPipeline:
    """SAA+IBP over a stream of frame sets with overlapped host<->device traffic.

    in_dtype : 'f32' (synthetic / pre-converted frames) or 'u8' (camera / PNG frames,
               converted on the device -- load_gray, run_sr.py:73-75)
    out_dtype: 'f32' (the HR float image) or 'u8' (clip + truncation, run_sr.py:303)
    """

    def __init__(self, n_frames, h, w, shifts_yx, kernel, n_iter, factor=2, step=ops.IBP_STEP_SIZE,
                 depth=3, in_dtype="f32", out_dtype="f32", device=None, use_graph=False, lanes=1):
        """lanes: independent compute lanes (each its own plan = its own scratch, and its own compute stream); job i runs on lane
        i mod lanes, so that the kernels of consecutive jobs overlap at their launch boundaries and tails."""
        torch = ops._torch()
        self.torch = torch
        self.use_graph = bool(use_graph)
        if device is not None:
            torch.cuda.set_device(device)
        self.dev = torch.device("cuda", torch.cuda.current_device())
        self.lanes = max(1, int(lanes))
        self.plans = [ops.Plan(n_frames, h, w, shifts_yx, kernel, factor, step) for _ in range(self.lanes)]
        self.plan = self.plans[0]
        self.n, self.h, self.w, self.n_iter = int(n_frames), int(h), int(w), int(n_iter)
        self.H, self.W = self.plan.H, self.plan.W
        self.depth = int(depth)
        self.in_dtype, self.out_dtype = in_dtype, out_dtype
        tin = torch.uint8 if in_dtype == "u8" else torch.float32
        tout = torch.uint8 if out_dtype == "u8" else torch.float32
        self.s_h2d, self.s_cmp, self.s_d2h = (torch.cuda.Stream() for _ in range(3))
        self.s_lane = [self.s_cmp] + [torch.cuda.Stream() for _ in range(self.lanes - 1)]
        if self.depth % self.lanes:
            self.depth += self.lanes - self.depth % self.lanes          # slot i always runs on lane i mod lanes
        self.slots = []
        for i in range(self.depth):
            self.slots.append(dict(
                lane=i % self.lanes,
                raw=torch.empty((self.n, self.h, self.w), dtype=tin, device=self.dev),
                lr=torch.empty((self.n, self.h, self.w), dtype=torch.float32, device=self.dev) if in_dtype == "u8" else None,
                hr=torch.empty((self.H, self.W), dtype=torch.float32, device=self.dev),
                out=torch.empty((self.H, self.W), dtype=torch.uint8, device=self.dev) if out_dtype == "u8" else None,
                host_out=torch.empty((self.H, self.W), dtype=tout).pin_memory(),
                ev_in=torch.cuda.Event(), ev_cmp=torch.cuda.Event(), ev_out=torch.cuda.Event(),
                busy=False, errors=None, graph=None,
            ))
        self.h2d_bytes = self.n * self.h * self.w * (1 if in_dtype == "u8" else 4)
        self.d2h_bytes = self.H * self.W * (1 if out_dtype == "u8" else 4)
        self._next = 0
        if self.use_graph:
            # one captured launch sequence per slot (slot buffers and the plan's scratch are baked in); the
            # un-captured warm-up run allocates the plan's lazy scratch first
            for lane in range(self.lanes):
                with torch.cuda.stream(self.s_lane[lane]):
                    self._compute(self.slots[lane])
            torch.cuda.synchronize()
            for slot in self.slots:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=self.s_lane[slot["lane"]]):
                    self._compute(slot)
                slot["graph"] = g

    def close(self):
        self.torch.cuda.synchronize()
        for slot in self.slots:
            slot["graph"] = None
        for p in self.plans:
            p.close()

    def _compute(self, slot):
        """(u8 -> f32,) SAA + IBP (, clip + truncation to u8) of one slot on the current stream."""
        import ctypes as C
        lib = ops._lib.load()
        st = C.c_void_p(self.torch.cuda.current_stream().cuda_stream)
        if self.in_dtype == "u8":
            ops._lib.check(lib.sr459_u8_to_f32(ops._ptr(slot["raw"]), slot["raw"].numel(), ops._ptr(slot["lr"]), st))
            frames = slot["lr"]
        else:
            frames = slot["raw"]
        self.plans[slot["lane"]].saa_ibp([frames[k] for k in range(self.n)], self.n_iter, out=slot["hr"], want_errors=False)
        if self.out_dtype == "u8":
            ops._lib.check(lib.sr459_f32_to_u8(ops._ptr(slot["hr"]), slot["hr"].numel(), ops._ptr(slot["out"]), st))

    # -- one job ------------------------------------------------------------------------
    def _submit(self, slot, host_frames):
        torch = self.torch
        lib = ops._lib.load()
        import ctypes as C
        if slot["busy"]:
            slot["ev_out"].synchronize()            # the slot's previous result has landed
        with torch.cuda.stream(self.s_h2d):
            self.s_h2d.wait_event(slot["ev_cmp"])   # previous compute on this slot is done with raw/lr
            slot["raw"].copy_(host_frames, non_blocking=True)
            slot["ev_in"].record(self.s_h2d)
        s_cmp = self.s_lane[slot["lane"]]
        with torch.cuda.stream(s_cmp):
            s_cmp.wait_event(slot["ev_in"])
            s_cmp.wait_event(slot["ev_out"])        # previous D2H of this slot finished reading hr/out
            if slot["graph"] is not None:
                slot["graph"].replay()
            else:
                self._compute(slot)
            slot["ev_cmp"].record(s_cmp)
        with torch.cuda.stream(self.s_d2h):
            self.s_d2h.wait_event(slot["ev_cmp"])
            src = slot["out"] if self.out_dtype == "u8" else slot["hr"]
            slot["host_out"].copy_(src, non_blocking=True)
            slot["ev_out"].record(self.s_d2h)
        slot["busy"] = True

    def run(self, host_sets: Sequence, on_result=None, keep_results=False):
        """Process pinned host tensors of shape (n, h, w); returns results (copies) if keep_results.

        on_result(index, pinned_host_tensor) is called when a job's output has landed; the tensor is
        only valid during the callback (its slot is re-used)."""
        results = [] if keep_results else None
        pending: List[tuple] = []
        for idx, hs in enumerate(host_sets):
            slot = self.slots[self._next]
            self._next = (self._next + 1) % self.depth
            if slot["busy"] and pending and pending[0][1] is slot:
                i0, s0 = pending.pop(0)
                s0["ev_out"].synchronize()
                if on_result: on_result(i0, s0["host_out"])
                if keep_results: results.append(s0["host_out"].clone())
                s0["busy"] = False
            self._submit(slot, hs)
            pending.append((idx, slot))
        for i0, s0 in pending:
            s0["ev_out"].synchronize()
            if on_result: on_result(i0, s0["host_out"])
            if keep_results: results.append(s0["host_out"].clone())
            s0["busy"] = False
        return results

    def run_latency(self, host_sets: Sequence):
        """Unpipelined: per-job wall latency (submit -> result landed), seconds."""
        lat = []
        for hs in host_sets:
            slot = self.slots[0]
            t0 = time.perf_counter()
            self._submit(slot, hs)
            slot["ev_out"].synchronize()
            slot["busy"] = False
            lat.append(time.perf_counter() - t0)
        return lat


def pin_rank_to_cores(local_rank: int, local_world: int):
    """One process per GPU: give every rank of a node its own contiguous share of the cores this process may run on, so that
    the submitting threads of the ranks do not migrate onto each other (the e2e path is host-submission bound at 8 GPUs).
    Returns the core list it set, or None (single rank, no affinity support, fewer cores than ranks)."""
    import os
    if local_world <= 1 or not hasattr(os, "sched_getaffinity"):
        return None
    try:
        cores = sorted(os.sched_getaffinity(0))
        per = len(cores) // local_world
        if per < 1:
            return None
        mine = cores[local_rank * per:(local_rank + 1) * per]
        os.sched_setaffinity(0, mine)
        return [mine[0], mine[-1]]
    except OSError:
        return None


def shard_indices(n_items: int, rank: int, world: int) -> List[int]:
    """Item i goes to rank i mod world (batch mode)."""
    return list(range(rank, n_items, world))


class BurstStream:
    """Continuous frame sets at camera rate (BASELINE config 5): one HR image per n-frame XPR cycle.

    The reference acquires one frame per mirror corner (settle, trigger, capture --
    data_collection/collect_hw_triggered.py:183-188) or streams bursts through a frame callback
    that copies every frame into a list (api/allied_vision_camera.py:90-115) and only then runs
    run_sr.py offline.  Here the callback side is `frame_buffer()` / `commit_frame()` (or
    `push_frame()`): every uint8 frame goes into a pinned ring slot and its H2D copy is enqueued
    at once, so when the LAST frame of a set arrives only that frame's copy (3 MB) is still
    ahead of the kernels.  The job (u8 -> f32, SAA, K IBP iterations, clip + truncation to u8)
    runs on its own stream -- optionally as one captured CUDA graph launch -- and the uint8 HR
    image is copied back to pinned host memory.  `wait_result()` returns it with the latency
    from the last frame's commit to the result being on the host.
    """

    def __init__(self, n_frames, h, w, shifts_yx, kernel, n_iter, factor=2, step=ops.IBP_STEP_SIZE,
                 depth=3, use_graph=False, device=None, lanes=1):
        """lanes = 1: every job on one compute stream (lowest latency per set: a job has the GPU to itself).
        lanes > 1: slot i computes on stream i mod lanes, so that the jobs of a free-running source overlap at their kernel
        boundaries (higher sustained rate, longer latency per set while several sets are in flight)."""
        torch = ops._torch()
        self.torch = torch
        if device is not None:
            torch.cuda.set_device(device)
        self.dev = torch.device("cuda", torch.cuda.current_device())
        self.n, self.h, self.w, self.n_iter = int(n_frames), int(h), int(w), int(n_iter)
        self.depth = int(depth)
        self.use_graph = bool(use_graph)
        self.s_h2d, self.s_cmp, self.s_d2h = (torch.cuda.Stream() for _ in range(3))
        self.lanes = max(1, min(int(lanes), self.depth))
        self.s_lane = [self.s_cmp] + [torch.cuda.Stream() for _ in range(self.lanes - 1)]
        self.slots = []
        for _ in range(self.depth):
            plan = ops.Plan(n_frames, h, w, shifts_yx, kernel, factor, step)     # one plan per slot: graphs bake pointers
            H, W = plan.H, plan.W
            self.slots.append(dict(
                plan=plan,
                host_in=torch.empty((self.n, self.h, self.w), dtype=torch.uint8).pin_memory(),
                raw=torch.empty((self.n, self.h, self.w), dtype=torch.uint8, device=self.dev),
                lr=torch.empty((self.n, self.h, self.w), dtype=torch.float32, device=self.dev),
                hr=torch.empty((H, W), dtype=torch.float32, device=self.dev),
                out=torch.empty((H, W), dtype=torch.uint8, device=self.dev),
                host_out=torch.empty((H, W), dtype=torch.uint8).pin_memory(),
                ev_in=torch.cuda.Event(), ev_cmp=torch.cuda.Event(), ev_out=torch.cuda.Event(),
                graph=None, set_index=-1, t_commit=0.0,
            ))
        self.H, self.W = self.slots[0]["plan"].H, self.slots[0]["plan"].W
        self.frames_in = 0                       # frames committed so far
        self.sets_out = 0                        # results handed out so far
        self.sets_committed = 0                  # sets whose job AND read-back have been enqueued (set last in commit_frame)
        self.h2d_bytes = self.n * self.h * self.w
        self.d2h_bytes = self.H * self.W
        for slot in self.slots:                  # warm-up: lazy plan scratch is allocated outside any capture
            self._job(slot)
        torch.cuda.synchronize()
        if self.use_graph:
            for slot in self.slots:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=self.s_lane[self.slots.index(slot) % self.lanes]):
                    self._job(slot)
                slot["graph"] = g

    def close(self):
        self.torch.cuda.synchronize()
        for slot in self.slots:
            slot["graph"] = None
            slot["plan"].close()

    def _job(self, slot):
        """u8 -> f32, SAA + IBP, clip + truncation, on the current stream (run_sr.py:73-75, :181-209, :303)."""
        import ctypes as C
        lib = ops._lib.load()
        st = C.c_void_p(self.torch.cuda.current_stream().cuda_stream)
        ops._lib.check(lib.sr459_u8_to_f32(ops._ptr(slot["raw"]), slot["raw"].numel(), ops._ptr(slot["lr"]), st))
        slot["plan"].saa_ibp([slot["lr"][k] for k in range(self.n)], self.n_iter, out=slot["hr"], want_errors=False)
        ops._lib.check(lib.sr459_f32_to_u8(ops._ptr(slot["hr"]), slot["hr"].numel(), ops._ptr(slot["out"]), st))

    # -- producer side (camera callback) -------------------------------------------------------------
    def frame_buffer(self):
        """Pinned uint8 view (h, w) the next frame should be written into (zero-copy hand-over)."""
        set_idx, k = divmod(self.frames_in, self.n)
        slot = self.slots[set_idx % self.depth]
        if k == 0 and slot["set_index"] >= 0 and slot["set_index"] >= self.sets_out:
            raise RuntimeError("BurstStream ring is full: wait_result() must drain older sets first")
        return slot["host_in"][k]

    def commit_frame(self):
        """The frame written into frame_buffer() is complete: enqueue its H2D copy; launch the job on the last one."""
        torch = self.torch
        set_idx, k = divmod(self.frames_in, self.n)
        slot = self.slots[set_idx % self.depth]
        with torch.cuda.stream(self.s_h2d):
            if k == 0:
                self.s_h2d.wait_event(slot["ev_cmp"])        # the slot's previous job has consumed raw
            slot["raw"][k].copy_(slot["host_in"][k], non_blocking=True)
            if k == self.n - 1:
                slot["ev_in"].record(self.s_h2d)
        self.frames_in += 1
        if k != self.n - 1:
            return None
        slot["t_commit"] = time.perf_counter()
        slot["set_index"] = set_idx
        s_cmp = self.s_lane[(set_idx % self.depth) % self.lanes]
        with torch.cuda.stream(s_cmp):
            s_cmp.wait_event(slot["ev_in"])
            s_cmp.wait_event(slot["ev_out"])                 # previous read-back of this slot's out buffer
            if slot["graph"] is not None:
                slot["graph"].replay()
            else:
                self._job(slot)
            slot["ev_cmp"].record(s_cmp)
        with torch.cuda.stream(self.s_d2h):
            self.s_d2h.wait_event(slot["ev_cmp"])
            slot["host_out"].copy_(slot["out"], non_blocking=True)
            slot["ev_out"].record(self.s_d2h)
        self.sets_committed = set_idx + 1                    # only now may a consumer thread wait on ev_out
        return set_idx

    def push_frame(self, frame):
        """Copy a uint8 frame (numpy or torch, h x w) into the ring and commit it."""
        buf = self.frame_buffer()
        if isinstance(frame, np.ndarray):
            buf.numpy()[...] = frame
        else:
            buf.copy_(frame)
        return self.commit_frame()

    # -- consumer side -----------------------------------------------------------------------------------
    def wait_result(self):
        """Block until the oldest outstanding HR image is on the host.  Returns (set_index, pinned uint8 H x W
        tensor -- valid until `depth` further sets have been committed --, latency in seconds)."""
        if self.sets_out >= self.sets_committed:
            raise RuntimeError("no complete frame set is outstanding")
        slot = self.slots[self.sets_out % self.depth]
        slot["ev_out"].synchronize()
        lat = time.perf_counter() - slot["t_commit"]
        idx = self.sets_out
        self.sets_out += 1
        return idx, slot["host_out"], lat


def play_camera(bs: "BurstStream", frame_pool: Sequence[np.ndarray], n_sets: int, fps: float = 0.0, on_result=None):
    """Drive a BurstStream like a camera would (api/allied_vision_camera.py:90-115): a producer thread writes one uint8
    frame per 1/fps seconds (fps 0 = as fast as the consumer drains) into the pinned ring and commits it; the calling
    thread collects the HR images.  Returns (latencies in seconds -- commit of a set's LAST frame to its uint8 HR image
    being on the host --, wall seconds for all sets)."""
    import threading
    n = bs.n
    sem = threading.Semaphore(bs.depth - 1)              # the producer may run depth-1 sets ahead of the consumer
    period = 1.0 / fps if fps and fps > 0 else 0.0

    def producer():
        t_next = time.perf_counter()
        for f in range(n_sets * n):
            if f % n == 0:
                sem.acquire()
            if period:
                t_next += period
                while time.perf_counter() < t_next:
                    time.sleep(0.0002)
            bs.frame_buffer().numpy()[...] = frame_pool[f % len(frame_pool)]   # the camera's DMA into the pinned buffer
            bs.commit_frame()

    base = bs.sets_out
    lat = []
    th = threading.Thread(target=producer, daemon=True)
    t0 = time.perf_counter()
    th.start()
    for s in range(n_sets):
        while bs.sets_out >= bs.sets_committed:          # the set has not been committed yet
            time.sleep(0.0001)
        idx, img, l = bs.wait_result()
        lat.append(l)
        if on_result:
            on_result(idx - base, img)
        sem.release()
    wall = time.perf_counter() - t0
    th.join()
    return lat, wall
