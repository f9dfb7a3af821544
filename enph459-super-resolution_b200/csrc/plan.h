// plan.h -- the plan object behind the C ABI and the helpers api.cu and bands.cu share (not part of the public ABI).
#pragma once
#include "internal.h"

#include <string.h>
#include <algorithm>
#include <vector>

namespace sr459 {

// The caller's current device is restored when an entry point returns: a plan may live on any device of the
// process, and plan_free is reached from Python's garbage collector at arbitrary times.
struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev == dev) prev = -1;                       // nothing to switch, nothing to restore
        else ok = (cudaSetDevice(dev) == cudaSuccess);
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    DeviceGuard(const DeviceGuard&) = delete;
    DeviceGuard& operator=(const DeviceGuard&) = delete;
};
#define SR_DEVICE(dev)                                                                      \
    sr459::DeviceGuard _device_guard(dev);                                                  \
    if (!_device_guard.ok) { sr459::set_error("cudaSetDevice(%d) failed", (int)(dev)); return SR459_ECUDA; }

template <typename T>
inline int dev_alloc(T** p, size_t count) {
    if (*p) return SR459_OK;
    cudaError_t e = gmalloc((void**)p, std::max<size_t>(count, 1) * sizeof(T));
    if (e != cudaSuccess) {
        set_error("cudaMalloc of %zu bytes failed: %s", count * sizeof(T), cudaGetErrorString(e));
        *p = nullptr;
        return SR459_ENOMEM;
    }
    return SR459_OK;
}

inline int upload_taps_tbl(const std::vector<AxisTap>& v, AxisTap** d) {
    SR_TRY(dev_alloc(d, v.size()));
    SR_CUDA(cudaMemcpy(*d, v.data(), v.size() * sizeof(AxisTap), cudaMemcpyHostToDevice));
    return SR459_OK;
}

}  // namespace sr459

using namespace sr459;                    // internal header: only api.cu and bands.cu include it

struct BandState;                         // bands.cu: arena, peers, flags of a row-band plan

struct sr459_plan {
    int device = 0, n = 0, h = 0, w = 0, factor = 2, H = 0, W = 0;
    double shifts[2 * kMaxFrames] = {0};
    std::vector<double> psf; int kh = 0, kw = 0;
    double step = 0.5, lo = 0.0, hi = 255.0;
    bool is_int = false; int isy[kMaxFrames] = {0}, isx[kMaxFrames] = {0}, smax = 0;
    int path = 0;
    int64_t scratch = 0;

    // band geometry (whole image when not a band plan)
    bool band = false;
    int own_i0 = 0, own_i1 = 0;       // owned LR rows
    int hr_lo = 0, hr_hi = 0, lr_lo = 0, lr_hi = 0, halo = 0;
    int rank = 0, nranks = 1;
    bool band_recipe_valid = false, band_uses_vsum = false; int band_iters = 0;

    // tables
    AxisTap *d_zoom_y = nullptr, *d_zoom_x = nullptr;
    AxisTap *d_saa_ty = nullptr, *d_saa_tx = nullptr;
    AxisTap *d_fwd_ty = nullptr, *d_fwd_tx = nullptr, *d_bwd_ty = nullptr, *d_bwd_tx = nullptr;
    int *d_fir_iy0 = nullptr, *d_fir_ix0 = nullptr; float *d_fir_wy = nullptr, *d_fir_wx = nullptr; bool fir_ok = false;   // fir.cu
    FirPolyArgs fir_poly{}; bool fir_poly_ok = false;                      // interior polyphase form of the same tables
    cudaStream_t side = nullptr; cudaEvent_t ev_fork = nullptr, ev_join = nullptr;   // k_fir_poly runs beside k_fir_bp's border tiles
    TapList taps_f, taps_b;
    float psf7[49] = {0};
    bool  separable = false;          // psf7[u][v] == acol[u] * brow[v]
    float acol[7] = {0}, brow[7] = {0};

    // separable fused path: pre-summed frames per parity lattice (ibp_sep.cu)
    int   nslot = 0, slot_pq[kMaxSlots] = {0};
    int   slot_n[kMaxSlots] = {0}, slot_k[kMaxSlots][kMaxFrames] = {{0}};
    float mtab[kMaxSlots][4] = {{0}};
    float* d_s1[kMaxSlots] = {nullptr, nullptr, nullptr, nullptr};
    int   s1_rows = 0, s1_pitch = 0;
    CUtensorMap tm_s1w[kMaxSlots];    // box of ibp_ws.cu
    double* d_vsum = nullptr;
    bool lat_s1 = false;              // the lattice kernels of the current call use the S1 planes (elementwise residual pass)
    float* d_edge = nullptr;          // ibp_stream.cu edge vectors: colS1[h], leftS1[h], rowS1[w], topS1[w], scal[4]
    CUtensorMap tm_s1s;               // box of ibp_stream.cu (single lattice)
    CUtensorMap tm_s1ml[sr459::kMaxSlots];   // same box, one per lattice (ibp_stream_ml.cu)
    float* d_edge_ml[sr459::kMaxSlots] = {};  // per-lattice edge vectors of ibp_stream_ml.cu
    LatPlan* lat = nullptr;           // ibp_lat.cu: large / non-separable PSFs
    int   n_sm = 0;

    // scratch
    float *d_coef_lr = nullptr, *d_zoomed = nullptr, *d_coef_pad = nullptr;
    float *d_B = nullptr, *d_acc = nullptr, *d_err = nullptr, *d_hr2 = nullptr;
    double *d_partial = nullptr; size_t partial_cap = 0;
    double *d_errors = nullptr; int errors_cap = 0;

    // per-iteration partial-sum bookkeeping of the last ibp call
    int per_iter = 0;

    // rows of the iteration's output the generic path may store, LR rows whose residuals it counts (row-band children;
    // -1 = everything)
    int store_y0 = -1, store_y1 = -1, mse_i0 = -1, mse_i1 = -1;
    BandState* bs = nullptr;              // row-band plans (bands.cu)
    const BandPeer* kpush = nullptr;      // set by band_run around one iteration: the kernel exchanges the halo rows itself
};

template <typename T>
inline int plan_alloc(sr459_plan* p, T** ptr, size_t count) {
    if (*ptr) return SR459_OK;
    SR_TRY(dev_alloc(ptr, count));
    p->scratch += (int64_t)(count * sizeof(T));
    return SR459_OK;
}


// ---- api.cu internals used by bands.cu --------------------------------------------------------------------------
struct DensePass { sr459::Rect rects[4]; int nrect; int tile_h, tile_w; };
struct FusedRecipe {
    bool use_fast;
    bool use_sep;
    bool use_ws;
    bool use_stream;
    bool use_stream_ml;                   // 2..4 lattices (ibp_stream_ml.cu); st_nit .. st_nstrip are its geometry
    bool use_lat;
    int st_nit, st_seg, st_nseg, st_nstrip;
    int st_bnit, st_btop, st_bbot;        // row bands: border segments of the streaming kernel (internal.h StreamArgs)
    int fast_tx, fast_ty;
    DensePass all;
    int n_fast, n_all, per_iter;
};
int  plan_args_check(int n_frames, int h, int w, int factor, const double* shifts, const double* psf, int kh, int kw);
int  plan_common_init(sr459_plan* p, int device, int n_frames, int h, int w, int factor, const double* shifts,
                      const double* psf, int kh, int kw, double step, double lo, double hi);
void plan_free(sr459_plan* p);
int  plan_ensure_partial(sr459_plan* p, size_t count);
int  plan_ensure_errors(sr459_plan* p, int n_iter);
bool fused_supported(const sr459_plan* p);
bool use_fused(const sr459_plan* p);
FusedRecipe fused_recipe(const sr459_plan* p, int uy0, int uy1);
int  sep_prepare(sr459_plan* p, const float* const* lr, bool stream_kernel, cudaStream_t st, bool stream_ml = false);
int  iterate_fused(sr459_plan* p, const FusedRecipe& r, const float* const* lr, const float* cur, float* nxt,
                   double* partial, int uy0, int uy1, cudaStream_t st, int it = -1);
int  iterate_generic(sr459_plan* p, const float* const* lr, const float* cur, float* nxt, int it, cudaStream_t st);
int  generic_per_iter(const sr459_plan* p);
int  check_lr(const sr459_plan* p, const float* const* lr);
void band_state_free(BandState* bs);      // bands.cu
