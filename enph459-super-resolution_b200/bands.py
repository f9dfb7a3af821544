"""Row-band IBP of ONE large HR image across the GPUs of a node (BASELINE config 4).

The reference has no multi-device path: `ibp()` (mono_cal_target/run_sr.py:190-209) walks one
image on one CPU thread.  Here rank g owns the HR rows [2a_g, 2a_{g+1}) (LR rows [a_g, a_{g+1}) =
an even split of h) and keeps `halo = 2*reach + 2` further hr rows on each side.  One iteration
= the fused kernel over the owned rows (sr459_band_iterate) followed by ONE neighbour exchange
of the freshly written boundary rows; the per-iteration squared-error sums are all-reduced once
per call.  Image borders keep the reference's boundary rules on the first / last band only.

Two exchange mechanisms:

* `BandPlan.attach()` + `saa()` + `run_peer()` -- the product path.  Each band's hr buffers live in
  a device arena the neighbours map (cudaIpc handles gathered once through torch.distributed); a
  kernel stores the boundary rows straight into the neighbours' halo rows over NVLink and waits on
  their arrival counters (csrc/bands.cu).  No collective on the data path, the whole K-iteration
  loop is one CUDA graph, SAA is computed per band from LR rows only, fractional shifts included.
* `BandPlan.run()` -- caller-owned buffers and torch.distributed send/recv after every iteration
  (NCCL over NVLink on the GPU box, gloo in the CPU tests): the portable fallback and the
  comparison baseline for the numbers in DESIGN.md.

`BandLayout`, `exchange_halos` and `band_loop` are device-agnostic host logic (the world_size-2
gloo tests run them on CPU tensors); `BandPlan` binds the CUDA side through the C ABI.
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, List, Optional, Sequence

import numpy as np

from . import _lib


LR_MARGIN = 32          # extra LR rows either side for zoom's column prefilter (csrc/bands.cu kLrMargin)


def band_halo(reach: int, integer_shifts: bool) -> int:
    """Halo rows of a band: 2*reach+2 for integer HR shifts, 2*(reach+19) rounded up to 16 otherwise."""
    return 2 * reach + 2 if integer_shifts else (2 * (reach + 19) + 15) // 16 * 16


def exchange_handles(handle: bytes, rank: int, world: int, group=None):
    """One all-gather of the 64-byte arena handles; returns (handle of rank-1 or None, handle of rank+1 or None).
    Device-agnostic (gloo in the CPU tests, NCCL on the GPU box)."""
    import torch.distributed as dist
    if world == 1:
        return None, None
    got = [None] * world
    dist.all_gather_object(got, bytes(handle), group=group)
    return (got[rank - 1] if rank > 0 else None), (got[rank + 1] if rank < world - 1 else None)


class _DeviceView:
    """A raw device pointer as a torch-importable array (torch.as_tensor reads __cuda_array_interface__)."""

    def __init__(self, ptr: int, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


class BandLayout:
    """Row ranges of one rank; the same arithmetic as sr459_band_plan_create (checked in the tests)."""

    def __init__(self, h: int, w: int, reach: int, rank: int, world: int, factor: int = 2, halo: Optional[int] = None):
        if factor != 2:
            raise ValueError("row bands are implemented for factor 2")
        if not (0 <= rank < world):
            raise ValueError(f"bad rank {rank} of {world}")
        if h < world:
            raise ValueError(f"{h} LR rows cannot be split into {world} bands")
        self.h, self.w, self.H, self.W = h, w, 2 * h, 2 * w
        self.rank, self.world, self.reach = rank, world, reach
        self.halo = 2 * reach + 2 if halo is None else int(halo)
        self.own_i0 = (rank * h) // world
        self.own_i1 = ((rank + 1) * h) // world
        self.own_y0, self.own_y1 = 2 * self.own_i0, 2 * self.own_i1
        self.hr_lo = max(0, self.own_y0 - self.halo)
        self.hr_hi = min(self.H, self.own_y1 + self.halo)
        self.lr_lo = self.hr_lo // 2
        self.lr_hi = min(h, self.hr_hi // 2 + 1)
        self.lr_ext_lo = max(0, self.hr_lo // 2 - LR_MARGIN)      # rows the peer-memory path wants (SAA margin)
        self.lr_ext_hi = min(h, self.hr_hi // 2 + LR_MARGIN)
        if world > 1 and min(((r + 1) * h) // world - (r * h) // world for r in range(world)) * 2 < self.halo:
            raise ValueError(f"bands of {2 * (h // world)} HR rows are thinner than the {self.halo}-row halo; use fewer ranks")

    def as_tuple(self):
        return (self.own_y0, self.own_y1, self.hr_lo, self.hr_hi, self.lr_lo, self.lr_hi, self.halo, self.world)

    # local row slices of the band's hr buffer (rows hr_lo .. hr_hi)
    @property
    def owned(self) -> slice:
        return slice(self.own_y0 - self.hr_lo, self.own_y1 - self.hr_lo)

    def split_hr(self, hr_full):
        return hr_full[self.hr_lo:self.hr_hi]

    def split_frames(self, lr_full: Sequence):
        return [f[self.lr_lo:self.lr_hi] for f in lr_full]


def exchange_halos(hr_local, lay: BandLayout, group=None) -> None:
    """Refresh the halo rows of hr_local from the neighbours' owned boundary rows (in place).
    Works on CPU tensors (gloo) and CUDA tensors (NCCL send/recv over NVLink, stream-ordered)."""
    import torch.distributed as dist
    if lay.world == 1:
        return
    R, o0, o1 = lay.halo, lay.own_y0 - lay.hr_lo, lay.own_y1 - lay.hr_lo
    ops = []
    if lay.rank > 0:                                   # upper neighbour: my first R owned rows <-> my upper halo
        ops.append(dist.P2POp(dist.isend, hr_local[o0:o0 + R], lay.rank - 1, group))
        ops.append(dist.P2POp(dist.irecv, hr_local[o0 - R:o0], lay.rank - 1, group))
    if lay.rank < lay.world - 1:                       # lower neighbour
        ops.append(dist.P2POp(dist.isend, hr_local[o1 - R:o1], lay.rank + 1, group))
        ops.append(dist.P2POp(dist.irecv, hr_local[o1:o1 + R], lay.rank + 1, group))
    for req in dist.batch_isend_irecv(ops):
        req.wait()


def band_loop(lay: BandLayout, hr_local, scratch, n_iter: int, iterate: Callable, exchange: Callable = exchange_halos,
              group=None):
    """The K-iteration loop of ibp() (run_sr.py:197-208) on one band: iterate(cur, nxt, it) updates
    nxt's owned rows from cur; the exchange refreshes nxt's halo rows; buffers swap.  Returns the
    buffer that holds the result (hr_local or scratch)."""
    cur, nxt = hr_local, scratch
    for it in range(n_iter):
        iterate(cur, nxt, it)
        exchange(nxt, lay, group)
        cur, nxt = nxt, cur
    return cur


def reduce_errors(local_sums, n_frames: int, lay: BandLayout, group=None, device=None) -> List[float]:
    """errors[it] = (1/n) sum_k mean(err_k^2) (run_sr.py:202,206) from the per-rank squared-error sums."""
    import torch
    import torch.distributed as dist
    t = torch.as_tensor(np.asarray(local_sums, dtype=np.float64), device=device)
    if lay.world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return (t / float(n_frames * lay.h * lay.w)).tolist()


class BandPlan:
    """CUDA side of one rank's band (sr459_band_* in include/sr459.h)."""

    def __init__(self, n_frames, h, w, shifts_yx, kernel, rank, world, factor=2, step=0.5, clip_lo=0.0, clip_hi=255.0,
                 device=None):
        from . import ops
        torch = ops._torch()
        self._lib = _lib.load()
        self.device = torch.cuda.current_device() if device is None else int(device)
        self.n, self.h, self.w = int(n_frames), int(h), int(w)
        sh = np.ascontiguousarray(np.asarray(shifts_yx, dtype=np.float64).reshape(-1))
        if sh.size != 2 * self.n:
            raise ValueError(f"need {self.n} (dy, dx) shifts, got {sh.size // 2}")
        k, kp, kh, kw = ops._psf_arg(kernel)
        handle = C.c_void_p()
        _lib.check(self._lib.sr459_band_plan_create(C.byref(handle), self.device, self.n, self.h, self.w, int(factor),
                                                    sh.ctypes.data_as(C.POINTER(C.c_double)), kp, kh, kw,
                                                    float(step), float(clip_lo), float(clip_hi), int(rank), int(world)))
        self._h = handle
        self.exchange_kind = ("boundary rows stored into the neighbours' halo rows over NVLink peer memory (by the streaming iteration kernel itself, "
                              "else by k_band_push after the iteration's kernels); no collective; one CUDA graph")
        out = (C.c_int * 8)()
        _lib.check(self._lib.sr459_band_layout(self._h, out))
        ext = (C.c_int * 4)()
        _lib.check(self._lib.sr459_band_layout_ext(self._h, ext))
        self.fractional = bool(ext[2])
        reach = max(k.shape) // 2
        self.layout = BandLayout(self.h, self.w, reach, int(rank), int(world), int(factor), halo=int(out[6]))
        if tuple(out) != self.layout.as_tuple() or (int(ext[0]), int(ext[1])) != (self.layout.lr_ext_lo, self.layout.lr_ext_hi):
            raise RuntimeError(f"band layout mismatch: library {tuple(out)} {tuple(ext)} vs host {self.layout.as_tuple()}")
        self.rank, self.world = int(rank), int(world)
        self._attached = world == 1
        self._graph = None

    def close(self):
        if getattr(self, "_h", None):
            self._lib.sr459_plan_destroy(self._h)
            self._h = None

    __del__ = close

    def set_path(self, path: int):
        _lib.check(self._lib.sr459_plan_set_path(self._h, int(path)))
        return self

    # -- peer-memory path (csrc/bands.cu) ------------------------------------------------------------------
    def current(self):
        """The band's current image buffer as a float32 CUDA tensor of HR rows [hr_lo, hr_hi) (a view of the arena:
        ask again after run_peer -- the two buffers swap with every iteration)."""
        from . import ops
        torch = ops._torch()
        cur = C.c_void_p()
        _lib.check(self._lib.sr459_band_buffers(self._h, C.byref(cur), None, None))
        lay = self.layout
        with torch.cuda.device(self.device):
            return torch.as_tensor(_DeviceView(cur.value, (lay.hr_hi - lay.hr_lo, lay.W)), device=torch.device("cuda", self.device))

    def attach(self, group=None):
        """Map the neighbouring ranks' arenas (one all-gather of 64-byte cudaIpc handles)."""
        handle = C.create_string_buffer(64)
        _lib.check(self._lib.sr459_band_export(self._h, handle, None))
        up, down = exchange_handles(handle.raw, self.rank, self.world, group)
        _lib.check(self._lib.sr459_band_attach(self._h, up, None, 0, down, None, 0))
        self._attached = True
        return self

    def attach_local(self, up: Optional["BandPlan"], down: Optional["BandPlan"]):
        """Neighbours that live in THIS process (tests; one process driving several GPUs): raw pointers."""
        def raw(p):
            if p is None:
                return None, 0
            ptr = C.c_void_p()
            _lib.check(p._lib.sr459_band_export(p._h, None, C.byref(ptr)))
            return ptr, p.device
        (pu, du), (pd, dd) = raw(up), raw(down)
        _lib.check(self._lib.sr459_band_attach(self._h, None, pu, du, None, pd, dd))
        self._attached = True
        return self

    def _ext_frames(self, lr_ext: Sequence):
        from . import ops
        lay = self.layout
        ts = [ops._dev(f, device=self.device) for f in lr_ext]
        for t in ts:
            if tuple(t.shape) != (lay.lr_ext_hi - lay.lr_ext_lo, self.w):
                raise ValueError(f"local frame shape {tuple(t.shape)} != {(lay.lr_ext_hi - lay.lr_ext_lo, self.w)} "
                                 f"(LR rows {lay.lr_ext_lo}..{lay.lr_ext_hi})")
        return ts, (C.c_void_p * self.n)(*[t.data_ptr() for t in ts])

    def saa(self, lr_ext: Sequence):
        """shift_and_add of this band (halo rows included) from its LR rows only -> the current buffer."""
        from . import ops
        ts, arr = self._ext_frames(lr_ext)
        _lib.check(self._lib.sr459_band_saa(self._h, arr, ops._stream_ptr(self.device)))
        return self.current()

    def run_peer(self, lr_ext: Sequence, n_iter: int, group=None, graph: bool = False, reduce: bool = True):
        """n_iter IBP iterations on the current buffer with the in-kernel halo exchange.  Returns the error list
        (all-reduced over the ranks unless reduce=False, which returns this band's raw sums).
        graph=True replays the loop as one CUDA graph (captured on first use for this n_iter)."""
        from . import ops
        torch = ops._torch()
        if not self._attached:
            raise RuntimeError("call attach() / attach_local() first")
        ts, arr = self._ext_frames(lr_ext)
        n_iter = int(n_iter)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if graph and n_iter > 0:
            key = (n_iter, tuple(t.data_ptr() for t in ts))
            if self._graph is None or self._graph[0] != key:
                start = self._lib.sr459_band_current(self._h, -1)
                cap = torch.cuda.Stream(device=self.device)
                cap.wait_stream(torch.cuda.current_stream(self.device))
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=cap, capture_error_mode="thread_local"):
                    _lib.check(self._lib.sr459_band_run(self._h, arr, n_iter, ops._stream_ptr(self.device)))
                self._graph = (key, g, start, self._lib.sr459_band_current(self._h, -1))
                self._lib.sr459_band_current(self._h, start)      # nothing has run yet
            _, g, start, end = self._graph
            if self._lib.sr459_band_current(self._h, -1) != start:
                raise RuntimeError("the captured loop starts from the other buffer; run an even number of iterations or re-capture")
            e0.record()
            g.replay()
            e1.record()
            self._lib.sr459_band_current(self._h, end)
        else:
            e0.record()
            _lib.check(self._lib.sr459_band_run(self._h, arr, n_iter, ops._stream_ptr(self.device)))
            e1.record()
        sums = np.zeros(max(n_iter, 0), dtype=np.float64)
        if n_iter > 0:
            _lib.check(self._lib.sr459_band_errors(self._h, n_iter, sums.ctypes.data_as(C.POINTER(C.c_double)),
                                                   ops._stream_ptr(self.device)))
            self.last_loop_ms = e0.elapsed_time(e1)
        else:
            self.last_loop_ms = 0.0
        if not reduce:
            return sums
        return reduce_errors(sums, self.n, self.layout, group, device=torch.device("cuda", self.device))

    def run(self, lr_local: Sequence, hr_local, n_iter: int, group=None, exchange: Optional[Callable] = exchange_halos,
            graph: bool = False):
        """IBP on this rank's band.  lr_local: the n frames' rows [lr_lo, lr_hi); hr_local: float32 CUDA
        tensor with HR rows [hr_lo, hr_hi) (halos filled).  Returns (hr_local_result, errors).

        graph=True captures the whole K-iteration loop -- kernels AND the NCCL send/recv of every
        iteration -- into one CUDA graph and replays it: the loop is launch-bound otherwise (one
        ctypes call + one batched P2P group per iteration is ~60 us of host time against 20-200 us of
        kernels).  Every rank must pass the same flag."""
        from . import ops
        torch = ops._torch()
        lay = self.layout
        ts = [ops._dev(f) for f in lr_local]
        for t in ts:
            if tuple(t.shape) != (lay.lr_hi - lay.lr_lo, self.w):
                raise ValueError(f"local frame shape {tuple(t.shape)} != {(lay.lr_hi - lay.lr_lo, self.w)}")
        if tuple(hr_local.shape) != (lay.hr_hi - lay.hr_lo, lay.W) or not hr_local.is_contiguous():
            raise ValueError(f"local hr shape {tuple(hr_local.shape)} != {(lay.hr_hi - lay.hr_lo, lay.W)} (contiguous)")
        arr = (C.c_void_p * self.n)(*[t.data_ptr() for t in ts])
        _lib.check(self._lib.sr459_band_begin(self._h, arr, int(n_iter), ops._stream_ptr()))
        scratch = hr_local.clone()

        def iterate(cur, nxt, it):
            _lib.check(self._lib.sr459_band_iterate(self._h, arr, ops._ptr(cur), ops._ptr(nxt), int(it), ops._stream_ptr()))

        exch = exchange or (lambda *a: None)
        if graph and n_iter > 0:
            hr0 = hr_local.clone()
            # un-captured warm-up: lazy plan scratch gets allocated, NCCL P2P channels get connected.
            # It only writes `scratch` (rewritten by iteration 0) and partial-sum slot 0 (ditto).
            iterate(hr_local, scratch, 0)
            exch(scratch, lay, group)
            torch.cuda.synchronize()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side, capture_error_mode="thread_local"):
                res = band_loop(lay, hr_local, scratch, int(n_iter), iterate, exch, group)
            hr_local.copy_(hr0)                              # the warm-up left hr_local untouched, the capture too
            if lay.world > 1:
                import torch.distributed as dist
                dist.barrier(group)                          # replays start together (capture time differs per rank)
                torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            g.replay()
            e1.record()
            self._graph = g                                  # keeps the captured launches alive until the next run
        else:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            res = band_loop(lay, hr_local, scratch, int(n_iter), iterate, exch, group)
            e1.record()
        sums = np.zeros(max(int(n_iter), 0), dtype=np.float64)
        if n_iter > 0:
            _lib.check(self._lib.sr459_band_errors(self._h, int(n_iter), sums.ctypes.data_as(C.POINTER(C.c_double)),
                                                   ops._stream_ptr()))
        errors = reduce_errors(sums, self.n, lay, group, device=hr_local.device)
        self.last_loop_ms = e0.elapsed_time(e1) if n_iter > 0 else 0.0      # device time of the K-iteration loop
        return res, errors


def run_emulated(plans: Sequence[BandPlan], lr_full: Sequence, hr_full, n_iter: int):
    """All ranks of a band decomposition inside ONE process on one GPU (tests; a single-GPU check of
    the multi-GPU path): the bands iterate in lock step and the halo exchange is a row copy."""
    from . import ops
    torch = ops._torch()
    world = len(plans)
    lays = [p.layout for p in plans]
    hr_full = ops._dev(hr_full)
    frames = [ops._dev(f) for f in lr_full]
    cur = [lay.split_hr(hr_full).clone() for lay in lays]
    nxt = [c.clone() for c in cur]
    arrs, keep = [], []
    for p, lay in zip(plans, lays):
        ts = [f[lay.lr_lo:lay.lr_hi].contiguous() for f in frames]
        keep.append(ts)
        arr = (C.c_void_p * p.n)(*[t.data_ptr() for t in ts])
        arrs.append(arr)
        _lib.check(p._lib.sr459_band_begin(p._h, arr, int(n_iter), ops._stream_ptr()))
    for it in range(n_iter):
        for g, p in enumerate(plans):
            _lib.check(p._lib.sr459_band_iterate(p._h, arrs[g], ops._ptr(cur[g]), ops._ptr(nxt[g]), it, ops._stream_ptr()))
        for g, lay in enumerate(lays):                  # what exchange_halos does, as copies
            R, o0, o1 = lay.halo, lay.own_y0 - lay.hr_lo, lay.own_y1 - lay.hr_lo
            if g > 0:
                up = lays[g - 1]
                nxt[g][o0 - R:o0].copy_(nxt[g - 1][up.own_y1 - up.hr_lo - R:up.own_y1 - up.hr_lo])
            if g < world - 1:
                dn = lays[g + 1]
                nxt[g][o1:o1 + R].copy_(nxt[g + 1][dn.own_y0 - dn.hr_lo:dn.own_y0 - dn.hr_lo + R])
        cur, nxt = nxt, cur
    out = torch.empty_like(hr_full)
    total = np.zeros(max(n_iter, 0), dtype=np.float64)
    for g, (p, lay) in enumerate(zip(plans, lays)):
        out[lay.own_y0:lay.own_y1] = cur[g][lay.owned]
        if n_iter > 0:
            sums = np.zeros(n_iter, dtype=np.float64)
            _lib.check(p._lib.sr459_band_errors(p._h, int(n_iter), sums.ctypes.data_as(C.POINTER(C.c_double)), ops._stream_ptr()))
            total += sums
    errors = (total / float(plans[0].n * lays[0].h * lays[0].w)).tolist()
    return out, errors


def run_emulated_peer(plans: Sequence[BandPlan], lr_full: Sequence, n_iter: int, hr_init=None):
    """All bands of one image inside ONE process (one GPU, or one band per visible GPU): the product path itself --
    per-band SAA from LR rows, the in-kernel halo exchange through plain peer pointers -- with every band on a stream
    of its own so the exchange kernels of neighbouring bands can meet.  Returns (hr image on the first band's device,
    errors)."""
    from . import ops
    torch = ops._torch()
    world = len(plans)
    for g, p in enumerate(plans):
        p.attach_local(plans[g - 1] if g > 0 else None, plans[g + 1] if g < world - 1 else None)
    streams, exts = [], []
    for p in plans:
        lay = p.layout
        with torch.cuda.device(p.device):
            streams.append(torch.cuda.Stream(device=p.device))
            exts.append([ops._dev(np.ascontiguousarray(np.asarray(f)[lay.lr_ext_lo:lay.lr_ext_hi]) if not ops._is_tensor(f)
                                  else f[lay.lr_ext_lo:lay.lr_ext_hi].to(torch.device("cuda", p.device)).contiguous(), device=p.device)
                         for f in lr_full])
    torch.cuda.synchronize()
    for p, st, ext in zip(plans, streams, exts):
        with torch.cuda.device(p.device), torch.cuda.stream(st):
            if hr_init is None:
                p.saa(ext)
            else:
                lay = p.layout
                src = hr_init[lay.hr_lo:lay.hr_hi]
                p.current().copy_(src if ops._is_tensor(src) else torch.from_numpy(np.ascontiguousarray(src, dtype=np.float32)))
    for p in plans:
        with torch.cuda.device(p.device):
            torch.cuda.synchronize()
    arrs = []
    for p, st, ext in zip(plans, streams, exts):          # allocations first: growing a buffer synchronises the device
        arr = (C.c_void_p * p.n)(*[t.data_ptr() for t in ext])
        arrs.append(arr)
        with torch.cuda.device(p.device):
            _lib.check(p._lib.sr459_band_reserve(p._h, arr, int(n_iter), C.c_void_p(st.cuda_stream)))
    for p in plans:
        with torch.cuda.device(p.device):
            torch.cuda.synchronize()
    # iteration by iteration over the bands: a kernel only ever waits for kernels enqueued BEFORE it, so the emulation
    # cannot deadlock even if the device runs the streams one after the other (they may share a hardware queue)
    for it in range(int(n_iter)):
        for p, st, arr in zip(plans, streams, arrs):
            with torch.cuda.device(p.device):
                _lib.check(p._lib.sr459_band_run_range(p._h, arr, it, it + 1, int(n_iter), C.c_void_p(st.cuda_stream)))
    if n_iter > 0:
        for p, st in zip(plans, streams):
            with torch.cuda.device(p.device):
                _lib.check(p._lib.sr459_band_run_close(p._h, C.c_void_p(st.cuda_stream)))
    total = np.zeros(max(int(n_iter), 0), dtype=np.float64)
    lay0 = plans[0].layout
    out = torch.empty((lay0.H, lay0.W), dtype=torch.float32, device=torch.device("cuda", plans[0].device))
    for p, st in zip(plans, streams):
        lay = p.layout
        with torch.cuda.device(p.device):
            if n_iter > 0:
                sums = np.zeros(int(n_iter), dtype=np.float64)
                _lib.check(p._lib.sr459_band_errors(p._h, int(n_iter), sums.ctypes.data_as(C.POINTER(C.c_double)), C.c_void_p(st.cuda_stream)))
                total += sums
            st.synchronize()
            out[lay.own_y0:lay.own_y1] = p.current()[lay.owned].to(out.device)
    errors = (total / float(plans[0].n * lay0.h * lay0.w)).tolist()
    return out, errors
