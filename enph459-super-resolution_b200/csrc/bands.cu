// bands.cu -- ONE large HR image as row bands across the GPUs of a node (BASELINE config 4).
//
// The reference has no multi-device path: ibp() (mono_cal_target/run_sr.py:190-209) walks one image on one CPU thread.
// Rank g owns the HR rows [2a_g, 2a_{g+1}) and keeps `halo` further rows on each side (2 reach + 2 for integer HR
// shifts, 2 (reach + 19) for fractional ones: blur reach + the spline prefilter's 16-sample tail + 2 taps + |shift|).
//
//   arena      every band's two hr buffers (ping-pong) and its flag words live in ONE plain cudaMalloc block, so that
//              the neighbours can map it -- cudaIpc handles between the processes of a torchrun job, plain pointers
//              (+ cudaDeviceEnablePeerAccess) inside one process.
//   exchange   after iteration `it` has written the owned rows of the next buffer, k_band_push copies the band's first
//              and last `halo` owned rows STRAIGHT INTO the neighbours' halo rows over NVLink (peer stores, float4),
//              fences, bumps the neighbours' arrival counters and waits for its own two counters: ONE small kernel per
//              iteration, no NCCL launch, no host involvement, capturable in a CUDA graph (the epoch lives in device
//              memory).  Write-after-read on the neighbour's halo is covered by the same wait chain (see band_run).
//   SAA        every band builds its own initial image from LR rows only (zoom's column prefilter is local up to its
//              16-sample tail: the caller provides `lr margin` extra LR rows), so a job starts from LR bands.
//   fractional the generic stage-by-stage iteration runs on the band's local buffer as if it were an image of its own
//              ("child" plan): what the artificial cuts get wrong stays inside the halo, which is never stored and is
//              refreshed by the neighbours.
#include "plan.h"

#include <math.h>
#include <stdlib.h>
#include <new>

using namespace sr459;

namespace {
constexpr int kLrMargin   = 32;          // extra LR rows beyond hr_lo/2 and hr_hi/2 the caller provides (SAA)
constexpr int kFlagWords  = 64;          // 256 B at the head of the arena
constexpr int F_FROM_UP = BF_FROM_UP, F_FROM_DOWN = BF_FROM_DOWN, F_ERROR = BF_ERROR, F_BASE = BF_BASE, F_TICKET = BF_TICKET;
constexpr int kPushCtas   = 64;          // CTAs per direction (512 KB per direction for a 31x31 PSF: 2 float4 per thread)
}

struct BandState {
    unsigned char* arena = nullptr; size_t arena_bytes = 0; size_t buf_bytes = 0;
    float* hr[2] = {nullptr, nullptr};
    unsigned* flags = nullptr;
    int cur = 0;                                          // index of the buffer holding the current image
    unsigned char* peer[2] = {nullptr, nullptr};          // mapped arenas of the upper / lower neighbour
    bool peer_ipc[2] = {false, false};
    int peer_rows[2] = {0, 0}, peer_hr_lo[2] = {0, 0};    // the neighbours' local buffers
    int lr_ext_lo = 0, lr_ext_hi = 0;                     // LR rows the caller's local frames hold
    sr459_plan* child = nullptr;                          // fractional shifts: generic iteration on the local image
    // SAA scratch / tables
    AxisTap *d_ty = nullptr, *d_tx = nullptr, *d_zy = nullptr, *d_zx = nullptr;
    float *d_coef = nullptr, *d_zoomed = nullptr, *d_pad = nullptr, *d_tmp = nullptr;
    int e_lo = 0, e_hi = 0;                               // fractional SAA: rows of the zoomed extent
    bool recipe_valid = false; int iters = 0;
};

void band_state_free(BandState* bs) {
    if (!bs) return;
    for (int d = 0; d < 2; ++d)
        if (bs->peer[d] && bs->peer_ipc[d]) cudaIpcCloseMemHandle(bs->peer[d]);
    if (bs->arena) cudaFree(bs->arena);
    if (bs->child) plan_free(bs->child);
    void* ptrs[] = {bs->d_ty, bs->d_tx, bs->d_zy, bs->d_zx, bs->d_coef, bs->d_zoomed, bs->d_pad, bs->d_tmp};
    for (void* q : ptrs) if (q) gfree(q);
    delete bs;
}

static size_t arena_buf_bytes(int rows, int W) { return (((size_t)rows * W * sizeof(float)) + 255) & ~(size_t)255; }
static size_t arena_off(int which, int rows, int W) { return 256 + (size_t)which * arena_buf_bytes(rows, W); }

// the deterministic layout every rank can compute for every other rank
struct Layout { int own_i0, own_i1, hr_lo, hr_hi; };
static Layout layout_of(int h, int halo, int rank, int nranks) {
    Layout l;
    l.own_i0 = (int)(((long long)rank * h) / nranks);
    l.own_i1 = (int)(((long long)(rank + 1) * h) / nranks);
    l.hr_lo = std::max(0, 2 * l.own_i0 - halo);
    l.hr_hi = std::min(2 * h, 2 * l.own_i1 + halo);
    return l;
}

// ---- the exchange kernel -----------------------------------------------------------------------------------------------
struct PushArgs {
    const float4* src[2];          // my first / last `halo` owned rows of the buffer just written
    float4* dst[2];                // the upper neighbour's LOWER halo rows / the lower neighbour's UPPER halo rows (peer memory)
    unsigned* peer_flag[2];        // the counter I bump in the upper / lower neighbour's arena
    unsigned* flags;               // my own flag words
    long long n4;                  // float4s per direction
    int has[2];
    int it;                        // iteration within the run: publishes BASE + it + 1 and waits for the same from both neighbours
    unsigned long long timeout_ns;
};

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// grid = 2 * kPushCtas CTAs: the first half feeds the upper neighbour, the second half the lower one.  The last CTA to
// finish (ticket) publishes the new epoch to both neighbours and waits until both have published theirs to me.
__global__ void __launch_bounds__(256) k_band_push(const __grid_constant__ PushArgs a) {
    const int dir = blockIdx.x / kPushCtas, part = blockIdx.x % kPushCtas;
    if (a.has[dir]) {
        const float4* __restrict__ s = a.src[dir];
        float4* __restrict__ d = a.dst[dir];
        for (long long i = (long long)part * 256 + threadIdx.x; i < a.n4; i += (long long)kPushCtas * 256) d[i] = s[i];
    }
    __threadfence_system();                               // my peer stores are ordered before the ticket / the flag
    __syncthreads();
    if (threadIdx.x != 0) return;
    const unsigned t = atomicAdd(a.flags + F_TICKET, 1u);
    if (t != gridDim.x - 1) return;
    a.flags[F_TICKET] = 0;                                // re-armed for the next launch (stream order)
    __threadfence_system();
    const unsigned epoch = a.flags[F_BASE] + (unsigned)a.it + 1u;
    if (a.has[0]) st_release_sys(a.peer_flag[0], epoch);
    if (a.has[1]) st_release_sys(a.peer_flag[1], epoch);
    const unsigned long long t0 = globaltimer_ns();
    for (int dirw = 0; dirw < 2; ++dirw) {
        if (!a.has[dirw]) continue;
        const unsigned* f = a.flags + (dirw == 0 ? F_FROM_UP : F_FROM_DOWN);
        while ((int)(ld_acquire_sys(f) - epoch) < 0) {
            if (globaltimer_ns() - t0 > a.timeout_ns) { a.flags[F_ERROR] = 1u; __threadfence_system(); return; }
            __nanosleep(100);
        }
    }
}

// end of a run of K iterations: wait until both neighbours' rows of the last iteration have arrived (so the halos are
// valid and nobody still writes into this arena when the caller moves on), then advance the base of the arrival counts
__global__ void k_band_advance(unsigned* flags, int K, int unit, int word_up, int has_up, int has_down, unsigned long long timeout_ns) {
    const unsigned base = flags[F_BASE] + (unsigned)K;
    const unsigned want = base * (unsigned)unit;          // arrivals are counted per pushing warp (in-kernel) or per launch (k_band_push)
    const unsigned long long t0 = globaltimer_ns();
    for (int d = 0; d < 2; ++d) {
        if (!(d == 0 ? has_up : has_down)) continue;
        const unsigned* f = flags + word_up + d;
        while ((int)(ld_acquire_sys(f) - want) < 0) {
            if (globaltimer_ns() - t0 > timeout_ns) { flags[F_ERROR] = 1u; break; }
            __nanosleep(100);
        }
    }
    flags[F_BASE] = base;
}

// what an iteration kernel needs to do the exchange itself (ibp_stream.cu); buf = the buffer the iteration writes
static int band_fill_peer(sr459_plan* p, int buf, int it, const FusedRecipe& r, BandPeer* bp) {
    BandState* bs = p->bs;
    memset(bp, 0, sizeof(*bp));
    const int W = p->W, R = p->halo;
    const int o0 = 2 * p->own_i0 - p->hr_lo, o1 = 2 * p->own_i1 - p->hr_lo;
    bp->flags = bs->flags; bp->it = it; bp->timeout_ns = 4000000000ull;
    { static const int probe = getenv("SR459_BAND_PROBE") ? atoi(getenv("SR459_BAND_PROBE")) : 0; bp->probe = probe; }
    for (int d = 0; d < 2; ++d) {
        const bool has = d == 0 ? p->rank > 0 : p->rank < p->nranks - 1;
        bp->has[d] = has ? 1 : 0;
        bp->push_y0[d] = d == 0 ? o0 : o1 - R; bp->push_y1[d] = d == 0 ? o0 + R : o1;
        if (!has) continue;
        if (!bs->peer[d]) { set_error("band_run: neighbour %s is not attached (sr459_band_attach)", d == 0 ? "above" : "below"); return SR459_EINVAL; }
        float* peer_buf = (float*)(bs->peer[d] + arena_off(buf, bs->peer_rows[d], W));
        bp->peer_out[d] = peer_buf + ((ptrdiff_t)p->hr_lo - bs->peer_hr_lo[d]) * W;
        bp->peer_flag[d] = (unsigned*)bs->peer[d] + (d == 0 ? BF_ARR_DOWN : BF_ARR_UP);
        int segs = 0;                                     // segments of the streaming kernel with rows in the pushed range
        if (r.st_bnit > 0) segs = 1;                      // the border segment (8 b_nit - 6 >= halo rows)
        else
            for (int sgm = 0; sgm < r.st_nseg; ++sgm) {
                const int ya = o0 + sgm * r.st_seg, yb = std::min(ya + r.st_seg, o1);
                if (ya < bp->push_y1[d] && yb > bp->push_y0[d]) ++segs;
            }
        bp->push_warps[d] = segs * r.st_nstrip;
        bp->unit = r.st_nstrip;
    }
    return SR459_OK;
}

static int band_push(sr459_plan* p, int buf, int it, cudaStream_t st) {
    BandState* bs = p->bs;
    if (p->nranks == 1) return SR459_OK;
    // development probe (tools/bench_bands.py --no-push): compute only, results are WRONG near the seams
    { static const bool skip = getenv("SR459_BAND_NOPUSH") != nullptr; if (skip) return SR459_OK; }
    const int W = p->W, R = p->halo;
    const int rows = p->hr_hi - p->hr_lo;
    PushArgs a{};
    a.flags = bs->flags;
    a.n4 = (long long)R * W / 4;
    a.timeout_ns = 4000000000ull;
    a.it = it;
    const int o0 = 2 * p->own_i0 - p->hr_lo, o1 = 2 * p->own_i1 - p->hr_lo;
    (void)rows;
    for (int d = 0; d < 2; ++d) {
        const bool has = d == 0 ? p->rank > 0 : p->rank < p->nranks - 1;
        a.has[d] = has ? 1 : 0;
        if (!has) continue;
        if (!bs->peer[d]) { set_error("band_run: neighbour %s is not attached (sr459_band_attach)", d == 0 ? "above" : "below"); return SR459_EINVAL; }
        const float* src = bs->hr[buf] + (size_t)(d == 0 ? o0 : o1 - R) * W;
        // upper neighbour: its owned rows end at my own_y0; its lower halo = global rows [own_y0, own_y0 + R)
        // lower neighbour: its upper halo = global rows [own_y1 - R, own_y1)
        const int g0 = d == 0 ? 2 * p->own_i0 : 2 * p->own_i1 - R;
        float* peer_buf = (float*)(bs->peer[d] + arena_off(buf, bs->peer_rows[d], W));
        a.src[d] = (const float4*)src;
        a.dst[d] = (float4*)(peer_buf + (size_t)(g0 - bs->peer_hr_lo[d]) * W);
        a.peer_flag[d] = (unsigned*)bs->peer[d] + (d == 0 ? F_FROM_DOWN : F_FROM_UP);
    }
    k_band_push<<<2 * kPushCtas, 256, 0, st>>>(a);
    SR_LAUNCH_CHECK("k_band_push");
    return SR459_OK;
}

// ---- plan construction -------------------------------------------------------------------------------------------------
static int band_halo_for(const sr459_plan* p) {
    const int reach = std::max(p->taps_f.reach, p->taps_b.reach);
    if (fused_supported(p)) return 2 * reach + 2;
    return ((2 * (reach + 19)) + 15) & ~15;               // fractional: a multiple of 16 keeps LR offsets multiples of 8
}

static int band_arena_alloc(sr459_plan* p) {
    BandState* bs = p->bs;
    if (bs->arena) return SR459_OK;
    const int rows = p->hr_hi - p->hr_lo;
    bs->buf_bytes = arena_buf_bytes(rows, p->W);
    bs->arena_bytes = 256 + 2 * bs->buf_bytes;
    cudaError_t e = cudaMalloc((void**)&bs->arena, bs->arena_bytes);      // plain cudaMalloc: cudaIpc needs it
    if (e != cudaSuccess) { set_error("band arena of %zu bytes: %s", bs->arena_bytes, cudaGetErrorString(e)); bs->arena = nullptr; return SR459_ENOMEM; }
    SR_CUDA(cudaMemset(bs->arena, 0, 256));
    bs->flags = (unsigned*)bs->arena;
    bs->hr[0] = (float*)(bs->arena + arena_off(0, rows, p->W));
    bs->hr[1] = (float*)(bs->arena + arena_off(1, rows, p->W));
    p->scratch += (int64_t)bs->arena_bytes;
    return SR459_OK;
}

extern "C" {

int sr459_band_plan_create(sr459_plan** out, int device, int n_frames, int h, int w, int factor, const double* h_shifts_yx,
                           const double* h_psf, int kh, int kw, double step, double clip_lo, double clip_hi,
                           int rank, int nranks) {
    if (!out) { set_error("band_plan_create: null output"); return SR459_EINVAL; }
    *out = nullptr;
    if (nranks < 1 || rank < 0 || rank >= nranks) { set_error("band_plan_create: bad rank %d of %d", rank, nranks); return SR459_EINVAL; }
    SR_TRY(plan_args_check(n_frames, h, w, factor, h_shifts_yx, h_psf, kh, kw));
    if (h < nranks) { set_error("band_plan_create: %d LR rows cannot be split into %d bands", h, nranks); return SR459_EINVAL; }
    if (factor != 2) { set_error("band plans are implemented for factor 2"); return SR459_EUNSUPPORTED; }
    SR_DEVICE(device);
    sr459_plan* p = new (std::nothrow) sr459_plan();
    if (!p) { set_error("out of host memory"); return SR459_ENOMEM; }
    int rc = plan_common_init(p, device, n_frames, h, w, factor, h_shifts_yx, h_psf, kh, kw, step, clip_lo, clip_hi);
    if (rc != SR459_OK) { plan_free(p); return rc; }
    p->bs = new (std::nothrow) BandState();
    if (!p->bs) { plan_free(p); set_error("out of host memory"); return SR459_ENOMEM; }
    p->band = true;
    p->rank = rank; p->nranks = nranks;
    p->halo = band_halo_for(p);
    const Layout l = layout_of(h, p->halo, rank, nranks);
    for (int r = 0; r < nranks && nranks > 1; ++r) {
        const Layout o = layout_of(h, p->halo, r, nranks);
        if (2 * (o.own_i1 - o.own_i0) < p->halo) {
            plan_free(p);
            set_error("band_plan_create: bands of %d HR rows are thinner than the %d-row halo; use fewer ranks", 2 * (o.own_i1 - o.own_i0), p->halo);
            return SR459_EINVAL;
        }
    }
    p->own_i0 = l.own_i0; p->own_i1 = l.own_i1; p->hr_lo = l.hr_lo; p->hr_hi = l.hr_hi;
    p->lr_lo = p->hr_lo / 2;
    p->lr_hi = std::min(h, p->hr_hi / 2 + 1);
    BandState* bs = p->bs;
    bs->lr_ext_lo = std::max(0, p->hr_lo / 2 - kLrMargin);
    bs->lr_ext_hi = std::min(h, p->hr_hi / 2 + kLrMargin);
    for (int d = 0; d < 2; ++d) {
        const int r = rank + (d == 0 ? -1 : 1);
        if (r < 0 || r >= nranks) continue;
        const Layout o = layout_of(h, p->halo, r, nranks);
        bs->peer_rows[d] = o.hr_hi - o.hr_lo; bs->peer_hr_lo[d] = o.hr_lo;
    }
    if (!fused_supported(p)) {
        // fractional shifts (rgb_cal_target/run_sr.py:88-96): the generic iteration on the local image
        const int h_loc = (p->hr_hi - p->hr_lo) / 2;
        sr459_plan* c = new (std::nothrow) sr459_plan();
        if (!c) { plan_free(p); set_error("out of host memory"); return SR459_ENOMEM; }
        rc = plan_common_init(c, device, n_frames, h_loc, w, factor, h_shifts_yx, h_psf, kh, kw, step, clip_lo, clip_hi);
        if (rc != SR459_OK) { plan_free(c); plan_free(p); return rc; }
        c->store_y0 = 2 * p->own_i0 - p->hr_lo; c->store_y1 = 2 * p->own_i1 - p->hr_lo;
        c->mse_i0 = p->own_i0 - p->lr_lo;       c->mse_i1 = p->own_i1 - p->lr_lo;
        bs->child = c;
    }
    *out = p;
    return SR459_OK;
}

int sr459_band_layout(const sr459_plan* plan, int* h_out8) {
    if (!plan || !h_out8 || !plan->band) { set_error("band_layout: not a band plan"); return SR459_EINVAL; }
    h_out8[0] = 2 * plan->own_i0; h_out8[1] = 2 * plan->own_i1;
    h_out8[2] = plan->hr_lo; h_out8[3] = plan->hr_hi;
    h_out8[4] = plan->lr_lo; h_out8[5] = plan->lr_hi;
    h_out8[6] = plan->halo;  h_out8[7] = plan->nranks;
    return SR459_OK;
}

int sr459_band_layout_ext(const sr459_plan* plan, int* h_out4) {
    if (!plan || !h_out4 || !plan->band) { set_error("band_layout_ext: not a band plan"); return SR459_EINVAL; }
    h_out4[0] = plan->bs->lr_ext_lo; h_out4[1] = plan->bs->lr_ext_hi;
    h_out4[2] = plan->bs->child ? 1 : 0; h_out4[3] = kLrMargin;
    return SR459_OK;
}

int sr459_band_buffers(sr459_plan* plan, void** d_cur, void** d_other, int64_t* arena_bytes) {
    if (!plan || !plan->band) { set_error("band_buffers: not a band plan"); return SR459_EINVAL; }
    SR_DEVICE(plan->device);
    SR_TRY(band_arena_alloc(plan));
    BandState* bs = plan->bs;
    if (d_cur) *d_cur = bs->hr[bs->cur];
    if (d_other) *d_other = bs->hr[bs->cur ^ 1];
    if (arena_bytes) *arena_bytes = (int64_t)bs->arena_bytes;
    return SR459_OK;
}

// which of the two buffers holds the current image (0 / 1); set_to >= 0 changes it (a captured run is replayed from the
// state it was captured in)
int sr459_band_current(sr459_plan* plan, int set_to) {
    if (!plan || !plan->band || set_to > 1) { set_error("band_current: bad arguments"); return SR459_EINVAL; }
    if (set_to >= 0) plan->bs->cur = set_to;
    return plan->bs->cur;
}

// 64-byte cudaIpcMemHandle_t of the arena, for a neighbour in ANOTHER process; h_arena_ptr (may be NULL) receives the
// raw device pointer for a neighbour in the SAME process
int sr459_band_export(sr459_plan* plan, unsigned char* h_handle64, void** h_arena_ptr) {
    if (!plan || !plan->band) { set_error("band_export: not a band plan"); return SR459_EINVAL; }
    SR_DEVICE(plan->device);
    SR_TRY(band_arena_alloc(plan));
    if (h_handle64) {
        static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
        cudaIpcMemHandle_t hnd;
        SR_CUDA(cudaIpcGetMemHandle(&hnd, plan->bs->arena));
        memcpy(h_handle64, &hnd, 64);
    }
    if (h_arena_ptr) *h_arena_ptr = plan->bs->arena;
    return SR459_OK;
}

// Map the neighbours' arenas.  For each direction (0 = the band above, 1 = the band below) pass EITHER the 64-byte handle
// another process exported OR the raw arena pointer of a plan in this process (peer_device = its device); NULL / NULL
// for a direction without a neighbour.
int sr459_band_attach(sr459_plan* plan, const unsigned char* up_handle64, void* up_arena, int up_device,
                      const unsigned char* down_handle64, void* down_arena, int down_device) {
    if (!plan || !plan->band) { set_error("band_attach: not a band plan"); return SR459_EINVAL; }
    SR_DEVICE(plan->device);
    SR_TRY(band_arena_alloc(plan));
    BandState* bs = plan->bs;
    const unsigned char* hnd[2] = {up_handle64, down_handle64};
    void* raw[2] = {up_arena, down_arena};
    const int dev[2] = {up_device, down_device};
    for (int d = 0; d < 2; ++d) {
        const bool has = d == 0 ? plan->rank > 0 : plan->rank < plan->nranks - 1;
        if (!has) continue;
        if (bs->peer[d] && bs->peer_ipc[d]) { cudaIpcCloseMemHandle(bs->peer[d]); }
        bs->peer[d] = nullptr; bs->peer_ipc[d] = false;
        if (hnd[d]) {
            cudaIpcMemHandle_t h;
            memcpy(&h, hnd[d], 64);
            void* ptr = nullptr;
            cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) { set_error("band_attach: cudaIpcOpenMemHandle failed: %s", cudaGetErrorString(e)); return SR459_ECUDA; }
            bs->peer[d] = (unsigned char*)ptr; bs->peer_ipc[d] = true;
        } else if (raw[d]) {
            if (dev[d] != plan->device) {
                cudaError_t e = cudaDeviceEnablePeerAccess(dev[d], 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
                if (e != cudaSuccess) { set_error("band_attach: no peer access from device %d to %d: %s", plan->device, dev[d], cudaGetErrorString(e)); return SR459_ECUDA; }
            }
            bs->peer[d] = (unsigned char*)raw[d];
        } else {
            set_error("band_attach: rank %d needs the band %s", plan->rank, d == 0 ? "above" : "below");
            return SR459_EINVAL;
        }
    }
    return SR459_OK;
}

}  // extern "C"

// ---- SAA on a band (e3) ------------------------------------------------------------------------------------------------
static int upload_tbl(const std::vector<AxisTap>& v, AxisTap** d) {
    if (*d) { gfree(*d); *d = nullptr; }
    return upload_taps_tbl(v, d);
}

static int band_saa_tables(sr459_plan* p) {
    BandState* bs = p->bs;
    if (bs->d_ty) return SR459_OK;
    const int H_loc = p->hr_hi - p->hr_lo, rows_ext = bs->lr_ext_hi - bs->lr_ext_lo;
    std::vector<AxisTap> zy, zx;
    build_zoom_axis(p->h, p->H, zy);
    build_zoom_axis(p->w, p->W, zx);
    auto local = [&](AxisTap t) {                         // global LR row -> row of the caller's local frames
        for (int j = 0; j < 4; ++j) t.idx[j] = std::min(std::max(t.idx[j] - bs->lr_ext_lo, 0), rows_ext - 1);
        return t;
    };
    if (p->is_int) {
        // shift(zoom(lr), s)[y] = zoom(lr)[clamp(y - s)] (run_sr.py:185-186 with integer HR shifts): re-indexed zoom taps
        std::vector<AxisTap> ty((size_t)p->n * H_loc), tx((size_t)p->n * p->W);
        for (int k = 0; k < p->n; ++k) {
            for (int y = 0; y < H_loc; ++y) ty[(size_t)k * H_loc + y] = local(zy[std::min(std::max(p->hr_lo + y - p->isy[k], 0), p->H - 1)]);
            for (int x = 0; x < p->W; ++x) tx[(size_t)k * p->W + x] = zx[std::min(std::max(x - p->isx[k], 0), p->W - 1)];
        }
        SR_TRY(upload_tbl(ty, &bs->d_ty));
        SR_TRY(upload_tbl(tx, &bs->d_tx));
        return SR459_OK;
    }
    // fractional: zoom the rows [e_lo, e_hi) of every frame, then shift that extent as an image of its own
    bs->e_lo = std::max(0, p->hr_lo - 24); bs->e_hi = std::min(p->H, p->hr_hi + 24);
    const int E = bs->e_hi - bs->e_lo;
    std::vector<AxisTap> zyl(E), one;
    for (int y = 0; y < E; ++y) zyl[y] = local(zy[bs->e_lo + y]);
    SR_TRY(upload_tbl(zyl, &bs->d_zy));
    SR_TRY(upload_tbl(zx, &bs->d_zx));
    std::vector<AxisTap> ty((size_t)p->n * E), tx((size_t)p->n * p->W);
    for (int k = 0; k < p->n; ++k) {
        build_shift_axis(E, p->shifts[2 * k] * p->factor, E, 1, one);
        std::copy(one.begin(), one.end(), ty.begin() + (size_t)k * E);
        build_shift_axis(p->W, p->shifts[2 * k + 1] * p->factor, p->W, 1, one);
        std::copy(one.begin(), one.end(), tx.begin() + (size_t)k * p->W);
    }
    SR_TRY(upload_tbl(ty, &bs->d_ty));
    SR_TRY(upload_tbl(tx, &bs->d_tx));
    return SR459_OK;
}

static int band_saa(sr459_plan* p, const float* const* lr_ext, cudaStream_t st) {
    BandState* bs = p->bs;
    SR_TRY(band_arena_alloc(p));
    SR_TRY(band_saa_tables(p));
    const int H_loc = p->hr_hi - p->hr_lo, rows_ext = bs->lr_ext_hi - bs->lr_ext_lo;
    const size_t lpx = (size_t)rows_ext * p->w;
    if (!bs->d_coef) { SR_TRY(dev_alloc(&bs->d_coef, (size_t)p->n * lpx)); p->scratch += (int64_t)(p->n * lpx * 4); }
    PrefilterJob jobs[kMaxFrames];
    for (int k = 0; k < p->n; ++k)
        jobs[k] = PrefilterJob{lr_ext[k], bs->d_coef + k * lpx, rows_ext, p->w, rows_ext, p->w, rows_ext, p->w, SRC_PLAIN, 1};
    SR_TRY(launch_prefilter2d(jobs, p->n, st));
    float* out = bs->hr[bs->cur];
    if (p->is_int) {
        ResampleSumArgs a{};
        a.n = p->n; a.Hout = H_loc; a.Wout = p->W; a.out = out; a.scale = 1.0f / (float)p->n;
        a.ty = bs->d_ty; a.tx = bs->d_tx;
        for (int k = 0; k < p->n; ++k) { a.coef[k] = bs->d_coef + k * lpx; a.cpitch[k] = p->w; }
        return launch_resample_tile(a, st);
    }
    const int E = bs->e_hi - bs->e_lo, PH = E + 2 * kPad, PW = p->W + 2 * kPad;
    const size_t epx = (size_t)E * p->W;
    if (!bs->d_zoomed) { SR_TRY(dev_alloc(&bs->d_zoomed, (size_t)p->n * epx)); p->scratch += (int64_t)(p->n * epx * 4); }
    if (!bs->d_pad) { SR_TRY(dev_alloc(&bs->d_pad, (size_t)p->n * PH * PW)); p->scratch += (int64_t)((size_t)p->n * PH * PW * 4); }
    if (!bs->d_tmp) { SR_TRY(dev_alloc(&bs->d_tmp, epx)); p->scratch += (int64_t)(epx * 4); }
    for (int k = 0; k < p->n; ++k) {
        ResampleSumArgs z{};
        z.coef[0] = bs->d_coef + k * lpx; z.cpitch[0] = p->w; z.n = 1;
        z.ty = bs->d_zy; z.tx = bs->d_zx; z.Hout = E; z.Wout = p->W; z.scale = 1.0f;
        z.out = bs->d_zoomed + k * epx;
        SR_TRY(launch_resample_sum(z, st));
        jobs[k] = PrefilterJob{bs->d_zoomed + k * epx, bs->d_pad + (size_t)k * PH * PW, E, p->W, E, p->W, PH, PW, SRC_EDGEPAD, 1};
    }
    SR_TRY(launch_prefilter2d(jobs, p->n, st));
    ResampleSumArgs a{};
    a.n = p->n; a.Hout = E; a.Wout = p->W; a.out = bs->d_tmp; a.scale = 1.0f / (float)p->n;
    a.ty = bs->d_ty; a.tx = bs->d_tx;
    for (int k = 0; k < p->n; ++k) { a.coef[k] = bs->d_pad + (size_t)k * PH * PW; a.cpitch[k] = PW; }
    SR_TRY(launch_resample_sum(a, st));
    SR_CUDA(cudaMemcpyAsync(out, bs->d_tmp + (size_t)(p->hr_lo - bs->e_lo) * p->W, (size_t)H_loc * p->W * sizeof(float),
                            cudaMemcpyDeviceToDevice, st));
    return SR459_OK;
}

// ---- the K-iteration loop with the in-kernel halo exchange ---------------------------------------------------------------
static int band_prepare(sr459_plan* p, const float* const* lr_ibp, int n_iter, cudaStream_t st) {
    BandState* bs = p->bs;
    bs->recipe_valid = false;
    p->band_recipe_valid = false;
    if (bs->child) {
        sr459_plan* c = bs->child;
        c->per_iter = generic_per_iter(c);
        SR_TRY(plan_ensure_errors(c, n_iter));
        SR_TRY(plan_ensure_partial(c, (size_t)n_iter * c->per_iter));
    } else {
        FusedRecipe r = fused_recipe(p, 2 * p->own_i0, 2 * p->own_i1);
        p->per_iter = r.per_iter;
        SR_TRY(plan_ensure_errors(p, n_iter));
        SR_TRY(plan_ensure_partial(p, (size_t)n_iter * r.per_iter));
        p->band_uses_vsum = false;
        p->lat_s1 = false;
        if (r.use_sep || r.use_stream || (r.use_lat && p->nslot > 0 && !getenv("SR459_LAT_NO_S1"))) {
            SR_TRY(sep_prepare(p, lr_ibp, r.use_stream, st)); p->band_uses_vsum = true; p->lat_s1 = r.use_lat;
        }
    }
    bs->recipe_valid = true; bs->iters = n_iter;
    p->band_recipe_valid = true; p->band_iters = n_iter;
    return SR459_OK;
}

static int band_iterate(sr459_plan* p, const float* const* lr_ibp, const float* cur, float* nxt, int it, cudaStream_t st) {
    BandState* bs = p->bs;
    if (bs && bs->child) return iterate_generic(bs->child, lr_ibp, cur, nxt, it, st);
    const int uy0 = 2 * p->own_i0, uy1 = 2 * p->own_i1;
    FusedRecipe r = fused_recipe(p, uy0, uy1);
    return iterate_fused(p, r, lr_ibp, cur, nxt, p->d_partial + (size_t)it * r.per_iter, uy0, uy1, st, it);
}

extern "C" {

// SAA of this band from its LR rows [lr_ext_lo, lr_ext_hi) (sr459_band_layout_ext) into the band's current buffer:
// shift_and_add (run_sr.py:181-187) restricted to the HR rows [hr_lo, hr_hi), halos included, no exchange needed.
int sr459_band_saa(sr459_plan* plan, const float* const* h_d_lr_ext, void* stream) {
    SR_TRY(check_lr(plan, h_d_lr_ext));
    if (!plan->band) { set_error("band_saa: not a band plan"); return SR459_EINVAL; }
    SR_DEVICE(plan->device);
    return band_saa(plan, h_d_lr_ext, (cudaStream_t)stream);
}

// Allocate (or grow) everything sr459_band_run(n_iter) needs.  Optional: band_run does it itself, but growing frees
// device memory, which synchronises the whole device -- one process driving several bands of ONE GPU (tests) must not do
// that while another band's loop is already waiting for this one.
int sr459_band_reserve(sr459_plan* plan, const float* const* h_d_lr_ext, int n_iter, void* stream) {
    SR_TRY(check_lr(plan, h_d_lr_ext));
    if (!plan->band || n_iter < 0) { set_error("band_reserve: not a band plan or n_iter < 0"); return SR459_EINVAL; }
    SR_DEVICE(plan->device);
    SR_TRY(band_arena_alloc(plan));
    if (n_iter == 0) return SR459_OK;
    const float* lr_ibp[kMaxFrames];
    for (int k = 0; k < plan->n; ++k) lr_ibp[k] = h_d_lr_ext[k] + (size_t)(plan->lr_lo - plan->bs->lr_ext_lo) * plan->w;
    return band_prepare(plan, lr_ibp, n_iter, (cudaStream_t)stream);
}

// n_iter IBP iterations (run_sr.py:197-208) on the band's current buffer, the halo exchange after every iteration done
// by k_band_push over peer memory.  Every rank of the image must make the same call.  Enqueues only; capturable.
int sr459_band_run_range(sr459_plan* plan, const float* const* h_d_lr_ext, int it0, int it1, int n_iter, void* stream) {
    SR_TRY(check_lr(plan, h_d_lr_ext));
    if (!plan->band || n_iter < 0 || it0 < 0 || it1 < it0 || it1 > n_iter) { set_error("band_run: not a band plan or a bad iteration range"); return SR459_EINVAL; }
    SR_DEVICE(plan->device);
    cudaStream_t st = (cudaStream_t)stream;
    BandState* bs = plan->bs;
    SR_TRY(band_arena_alloc(plan));
    if (n_iter == 0 || it0 == it1) return SR459_OK;
    const float* lr_ibp[kMaxFrames];
    for (int k = 0; k < plan->n; ++k) lr_ibp[k] = h_d_lr_ext[k] + (size_t)(plan->lr_lo - bs->lr_ext_lo) * plan->w;
    if (it0 == 0) SR_TRY(band_prepare(plan, lr_ibp, n_iter, st));
    else if (!bs->recipe_valid || bs->iters != n_iter) { set_error("band_run: iterations must start at 0"); return SR459_EINVAL; }
    FusedRecipe rec{};
    if (!bs->child) rec = fused_recipe(plan, 2 * plan->own_i0, 2 * plan->own_i1);
    // the streaming kernel (with its border segments) stores its boundary rows into the neighbours' buffers itself and
    // its border warps wait for the neighbours' arrival counts; every other iteration kernel is followed by k_band_push
    const bool probe_nopush = getenv("SR459_BAND_NOPUSH") != nullptr;
    const bool in_kernel = !bs->child && rec.use_stream && rec.st_bnit > 0 && plan->nranks > 1 && !getenv("SR459_BAND_PUSH_KERNEL") && !probe_nopush;
    for (int it = it0; it < it1; ++it) {
        // iteration `it` reads buffer cur (owned rows + halos) and stores the owned rows of buffer cur^1; the boundary rows
        // also go into the neighbours' halo rows of THEIR buffer cur^1, and iteration it+1 starts on the halo-reading part
        // only once both neighbours' rows have arrived.  Write-after-read: a neighbour last READ its cur^1 halos in
        // iteration it-1, and its arrivals for iteration it-1 were counted only after the warps that read them had
        // finished; my halo-writing warps of iteration `it` waited for exactly that count before they started.
        BandPeer bp;
        if (in_kernel) { SR_TRY(band_fill_peer(plan, bs->cur ^ 1, it, rec, &bp)); plan->kpush = &bp; }
        int rc = band_iterate(plan, lr_ibp, bs->hr[bs->cur], bs->hr[bs->cur ^ 1], it, st);
        plan->kpush = nullptr;
        SR_TRY(rc);
        if (!in_kernel) SR_TRY(band_push(plan, bs->cur ^ 1, it, st));
        bs->cur ^= 1;
    }
    return SR459_OK;
}

// end of a run: wait for the neighbours' rows of the last iteration, advance the base of the arrival counts
int sr459_band_run_close(sr459_plan* plan, void* stream) {
    if (!plan || !plan->band || !plan->bs->recipe_valid) { set_error("band_run_close: no run is open"); return SR459_EINVAL; }
    SR_DEVICE(plan->device);
    BandState* bs = plan->bs;
    if (plan->nranks == 1 || bs->iters == 0 || getenv("SR459_BAND_NOPUSH") || getenv("SR459_BAND_PROBE")) return SR459_OK;
    FusedRecipe rec{};
    if (!bs->child) rec = fused_recipe(plan, 2 * plan->own_i0, 2 * plan->own_i1);
    const bool in_kernel = !bs->child && rec.use_stream && rec.st_bnit > 0 && !getenv("SR459_BAND_PUSH_KERNEL");
    k_band_advance<<<1, 1, 0, (cudaStream_t)stream>>>(bs->flags, bs->iters, in_kernel ? rec.st_nstrip : 1, in_kernel ? BF_ARR_UP : BF_FROM_UP,
                                                      plan->rank > 0, plan->rank < plan->nranks - 1, 4000000000ull);
    SR_LAUNCH_CHECK("k_band_advance");
    return SR459_OK;
}

int sr459_band_run(sr459_plan* plan, const float* const* h_d_lr_ext, int n_iter, void* stream) {
    SR_TRY(sr459_band_run_range(plan, h_d_lr_ext, 0, n_iter, n_iter, stream));
    return n_iter > 0 ? sr459_band_run_close(plan, stream) : SR459_OK;
}

// the old two-call form (caller-owned buffers, exchange done by the host layer) stays available
int sr459_band_begin(sr459_plan* plan, const float* const* h_d_lr_local, int n_iter, void* stream) {
    SR_TRY(check_lr(plan, h_d_lr_local));
    if (!plan->band || n_iter < 0) { set_error("band_begin: not a band plan or n_iter < 0"); return SR459_EINVAL; }
    SR_DEVICE(plan->device);
    plan->band_recipe_valid = false;
    if (n_iter == 0) return SR459_OK;
    return band_prepare(plan, h_d_lr_local, n_iter, (cudaStream_t)stream);
}

int sr459_band_iterate(sr459_plan* plan, const float* const* h_d_lr_local, const float* d_cur, float* d_nxt, int it, void* stream) {
    SR_TRY(check_lr(plan, h_d_lr_local));
    if (!plan->band || !plan->band_recipe_valid || !d_cur || !d_nxt || it < 0 || it >= plan->band_iters) {
        set_error("band_iterate: call sr459_band_begin first; it must be in [0, n_iter)"); return SR459_EINVAL;
    }
    SR_DEVICE(plan->device);
    return band_iterate(plan, h_d_lr_local, d_cur, d_nxt, it, (cudaStream_t)stream);
}

// per-iteration squared-error sums over the band's owned sites (the host layer all-reduces them and divides by
// n*h*w, run_sr.py:202,206); also reports an exchange that timed out
int sr459_band_errors(sr459_plan* plan, int n_iter, double* h_sums, void* stream) {
    if (!plan || !plan->band || !h_sums || n_iter < 0) { set_error("band_errors: bad arguments"); return SR459_EINVAL; }
    sr459_plan* q = (plan->bs && plan->bs->child) ? plan->bs->child : plan;
    if (n_iter > q->errors_cap) { set_error("band_errors: only %d iterations were run", q->errors_cap); return SR459_EINVAL; }
    if (n_iter == 0) return SR459_OK;
    cudaStream_t st = (cudaStream_t)stream;
    SR_DEVICE(plan->device);
    SR_TRY(launch_reduce_partials(q->d_partial, q->per_iter, n_iter, 1.0, (q == plan && plan->band_uses_vsum) ? plan->d_vsum : nullptr,
                                  q->d_errors, st));
    SR_CUDA(cudaMemcpyAsync(h_sums, q->d_errors, sizeof(double) * n_iter, cudaMemcpyDeviceToHost, st));
    unsigned flags[kFlagWords] = {0};
    if (plan->bs && plan->bs->arena) SR_CUDA(cudaMemcpyAsync(flags, plan->bs->flags, sizeof(flags), cudaMemcpyDeviceToHost, st));
    SR_CUDA(cudaStreamSynchronize(st));
    if (flags[F_ERROR]) { set_error("band halo exchange timed out (a neighbour never published its rows)"); return SR459_ECUDA; }
    return SR459_OK;
}

}  // extern "C"
