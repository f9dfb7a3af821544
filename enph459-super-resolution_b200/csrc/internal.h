// internal.h -- shared declarations of the sr459 CUDA engine (not part of the public ABI).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <atomic>
#include <vector>
#include "../../include/sr459.h"

namespace sr459 {

// ---- error plumbing ---------------------------------------------------------
void set_error(const char* fmt, ...);
extern std::atomic<uint64_t> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

// numpy's pairwise summation of a contiguous float64 run (api.cu)
double np_pairwise_sum(const double* a, size_t n);

// every device allocation of the library goes through these (alloc.cu: optional guard bands)
cudaError_t gmalloc(void** p, size_t bytes);
void gfree(void* p);

#define SR_CUDA(call)                                                                       \
    do {                                                                                    \
        cudaError_t _e = (call);                                                            \
        if (_e != cudaSuccess) {                                                            \
            sr459::set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__,           \
                             cudaGetErrorString(_e));                                       \
            return SR459_ECUDA;                                                             \
        }                                                                                   \
    } while (0)

#define SR_LAUNCH_CHECK(name)                                                               \
    do {                                                                                    \
        cudaError_t _e = cudaGetLastError();                                                \
        if (_e != cudaSuccess) {                                                            \
            sr459::set_error("launch of %s failed at %s:%d: %s", name, __FILE__, __LINE__,  \
                             cudaGetErrorString(_e));                                       \
            return SR459_ECUDA;                                                             \
        }                                                                                   \
        sr459::count_launch();                                                              \
    } while (0)

#define SR_TRY(call)                                                                        \
    do {                                                                                    \
        int _r = (call);                                                                    \
        if (_r != SR459_OK) return _r;                                                      \
    } while (0)

// ---- cubic B-spline machinery (scipy.ndimage restated; see spline.cu) ---------
constexpr int   kPad       = 12;      // scipy _prepad_for_spline_filter(mode='nearest')
constexpr float kPoleF     = -0.26794919243112270647f;   // sqrt(3) - 2
constexpr float kSqrt3F    = 1.73205080756887729353f;

struct AxisTap {            // one output sample along one axis: 4 taps into the coefficient line
    int   idx[4];
    float w[4];
};

// how the prefilter kernel reads its (virtual) input image
enum SrcMode : int {
    SRC_PLAIN      = 0,     // the array itself, mirror boundary            (zoom path)
    SRC_EDGEPAD    = 1,     // np.pad(a, 12, 'edge'), reflect boundary      (shift path)
    SRC_ZI_EDGEPAD = 2      // zero-inserted LR residual -> HR, then as 1   (back_project path)
};

struct PrefilterJob {
    const float* src;       // source image
    float*       dst;       // coefficient image, ny x nx, pitch nx
    int sh, sw;             // source dims
    int vh, vw;             // virtual (unpadded) dims: == sh,sw except SRC_ZI_EDGEPAD (HR dims)
    int ny, nx;             // coefficient dims (vh+24, vw+24 for the padded modes)
    int mode;
    int factor;             // zero-insert stride for SRC_ZI_EDGEPAD
};

int launch_prefilter2d(const PrefilterJob* h_jobs, int njobs, cudaStream_t st);

// host-side table builders (double precision, converted to float weights)
void build_zoom_axis(int n_in, int n_out, std::vector<AxisTap>& out, int order = 3);   // order 0, 1 or 3
// shift tables: output o = 0..n_out-1 at stride `stride` (decimation), coordinate o*stride - s + 12
void build_shift_axis(int n, double s, int n_out, int stride, std::vector<AxisTap>& out);

// out[y,x] = scale * sum_k R_k(y,x);   tables stacked per frame: ty[k*Hout + y], tx[k*Wout + x]
struct ResampleSumArgs {
    const float* coef[8];
    int   cpitch[8];
    int   n;
    const AxisTap* ty;
    const AxisTap* tx;
    int   Hout, Wout;
    float scale;
    float* out;
};
int launch_resample_sum(const ResampleSumArgs& a, cudaStream_t st);
// same result for upsampling tables (factor >= 2: zoom, zoom re-indexed by integer shifts): tiled, separable
int launch_resample_tile(const ResampleSumArgs& a, cudaStream_t st);

// per frame k: sim = R_k(i,j); if lr: err = lr_k - sim stored to out_k and sum(err^2) -> partial sums
struct ResampleErrArgs {
    const float* coef;      // shared coefficient image (prefiltered padded blur)
    int   cpitch;
    int   n;
    const AxisTap* ty;      // [k*h + i]
    const AxisTap* tx;      // [k*w + j]
    int   h, w;
    const float* lr[8];     // may be all null -> store sim
    float* out[8];
    double* partial;        // [gridDim.x*gridDim.y*n] or null
    int mse_i0, mse_i1;     // LR rows whose squared residuals are counted (row bands: the owned ones)
};
int launch_resample_err(const ResampleErrArgs& a, cudaStream_t st, int* nblocks_out);
// one CTA per kRsTileH x kRsTileW LR tile and all frames, coefficient region staged in shared memory; ONE partial per CTA
constexpr int kRsTileH = 16, kRsTileW = 32;
int launch_resample_err_tile(const ResampleErrArgs& a, cudaStream_t st);

// ---- FIR form of the fractional-shift back-projection (fir.cu) ---------------------
constexpr int kFirTaps = 20;            // LR taps per axis: (2*16 + 4) HR samples / factor 2, rounded up
// cudaFuncSetAttribute is per DEVICE: a plan may live on any device of the process (sr459_plan_create's `device`), so
// the "already configured" memo of a launch function is kept per device.
struct PerDeviceOnce {
    bool done[64] = {};
    bool need() {
        int d = 0;
        if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= 64) return true;
        if (done[d]) return false;
        done[d] = true;
        return true;
    }
};

struct PerDeviceMax {                  // same, for launchers whose dynamic shared-memory size varies: true if `s` exceeds the device's memo
    size_t v[64] = {};
    bool raise(size_t s) {
        int d = 0;
        if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= 64) return true;
        if (s <= v[d]) return false;
        v[d] = s;
        return true;
    }
};

// one axis of  shift(zero_insert(err, f), s, order=3, mode='nearest'): out[o] = sum_a w[o][a] * err[i0[o] + a]
bool build_fir_bp_axis(int n_lr, int f, double s, std::vector<int>& i0, std::vector<float>& w);
struct FirBpArgs {
    const float* err[8];                // LR residuals, h x w, pitch w
    int n, h, w, H, W;
    const int* iy0; const int* ix0;     // [k * H + Y], [k * W + X]
    const float* wy; const float* wx;   // [(k * H + Y) * kFirTaps + a], [(k * W + X) * kFirTaps + b]
    float* out;                         // sum over frames, H x W
    int skip_y0, skip_y1, skip_x0, skip_x1;   // HR-pixel rectangle (tile aligned) left to launch_fir_poly; empty if y1 <= y0
};
int launch_fir_bp(const FirBpArgs& a, cudaStream_t st);
// Polyphase fast path of the same map (factor 2): away from the borders the tables are 2-periodic -- output o of parity
// p reads LR samples (o - p) / 2 + c + d[p] + tap with ONE weight set per parity -- so interior tiles run a register-
// window FIR with the weights in registers; launch_fir_bp then skips the tiles of the region [y0, y1) x [x0, x1).
struct FirPolyFrame { int cy, cx, dy[2], dx[2]; };
struct FirPolyArgs {
    const float* err[8];
    int n, h, w, H, W;
    float* out;
    int y0, y1, x0, x1;                 // region in HR pixels: rows a multiple of 32, columns a multiple of 128
    FirPolyFrame fr[8];
    float wy[8][2][kFirTaps], wx[8][2][kFirTaps];
};
// host analysis of the per-frame tables (layout of FirBpArgs); false if no interior region exists
bool fir_poly_analyse(int n, int H, int W, const std::vector<int>& iy0, const std::vector<int>& ix0,
                      const std::vector<float>& wy, const std::vector<float>& wx, FirPolyArgs& out);
int launch_fir_poly(const FirPolyArgs& a, cudaStream_t st);

// ---- correlation / convolution ------------------------------------------------
// out[y,x] = sum_t w[t] * in0[y+dy[t], x+dx[t]]   (zero outside), taps with w==0 dropped
struct TapList {
    int    ntaps = 0;
    int    reach = 0;       // max |dy|,|dx|
    float* d_w   = nullptr; // device
    int*   d_dy  = nullptr;
    int*   d_dx  = nullptr;
    std::vector<float> w; std::vector<int> dy, dx;   // host copies
};
// true convolution 'same' with k (flip=false) or with k[::-1, ::-1] (flip=true)
void build_taps_host(const double* k, int kh, int kw, bool flip, TapList& t);
int  upload_taps(TapList& t);
void free_taps(TapList& t);

struct CorrArgs {
    const float* in; int H, W;
    const TapList* taps;
    float* out;
    // epilogue: mode 0 -> out = v ; mode 1 -> out = clamp(base + alpha*v, lo, hi)
    int mode; const float* base; float alpha, lo, hi;
    int y0 = 0, y1 = -1;    // rows stored (y1 < 0: all); row bands keep their halo rows untouched
};
int launch_corr2d(const CorrArgs& a, cudaStream_t st);

// ---- fused integer-shift IBP iteration ------------------------------------------
constexpr int kMaxFrames = 8;
constexpr int kMaxIntShift = 4;

struct IbpGeom {
    int H, W, h, w;                 // global HR / LR dims (factor 2)
    int n;
    int sy[kMaxFrames], sx[kMaxFrames];   // integer HR shifts
    int smax;
    float alpha, lo, hi;            // step/n, clip bounds
    // local buffers (row bands): local hr row 0 == global row hr_lo, rows [hr_lo, hr_hi) present
    int hr_lo, hr_hi;
    int lr_lo, lr_hi;               // LR rows present in the local LR buffers
    // rows this launch may update / LR rows whose residuals it counts
    int upd_y0, upd_y1;             // HR rows (even bounds)
    int own_i0, own_i1;             // LR rows
};

struct Rect { int y0, x0, y1, x1; };

struct DenseArgs {
    IbpGeom g;
    const float* lr[kMaxFrames];
    const float* hr_in;
    float* hr_out;
    const float* wf; const int* fdy; const int* fdx; int nf;   // forward taps (device)
    const float* wb; const int* bdy; const int* bdx; int nb;   // backward taps
    int R;                          // tap reach
    Rect rect[4]; int nrect;        // regions (even bounds) tiled by tile_h x tile_w
    int tile_h, tile_w;
    int tiles_before[5];            // prefix sums of tiles per rect
    int tiles_x[4];
    double* partial;                // one slot per CTA (this launch's slice)
};
int launch_ibp_dense(const DenseArgs& a, cudaStream_t st);
size_t ibp_dense_smem(int R, int smax, int tile_h, int tile_w, int nf, int nb);

struct FastArgs {
    IbpGeom g;                      // upd_y0/upd_y1: HR rows this launch updates and whose sites it owns
    const float* lr[kMaxFrames];
    const float* hr_in;
    float* hr_out;
    float psf[49];                  // 7x7 row-major k[u][v]; smaller odd kernels centred & zero padded
    int ph_n[4];                    // frames on lattice (P,Q), index P*2+Q
    int ph_k[4][kMaxFrames];        // their frame ids
    int tiles_x, tiles_y;
    int ty_origin;                  // global HR row of tile row 0 (even) == g.upd_y0
    double* partial;                // one slot per CTA
};
int  launch_ibp_fast(const FastArgs& a, cudaStream_t st);
bool ibp_fast_supported(int kh, int kw, int smax);
void ibp_fast_tile(int* th, int* tw);

// ---- separable-PSF fused iteration (lattice.cu, ibp_ws.cu): TMA-staged tiles, pre-summed co-sited frames -------
constexpr int kMaxSlots = 4;            // active parity lattices

struct LatticeArgs {                    // k_lattice_sums: S1 planes + sum of V, once per ibp() call
    const float* lr[kMaxFrames];
    int h, w;                           // global LR dims
    int lr_lo, lr_hi;                   // LR rows present in the local LR buffers
    int nslot;
    int slot_n[kMaxSlots];
    int slot_k[kMaxSlots][kMaxFrames];  // frame ids on the lattice
    int slot_oy[kMaxSlots][kMaxFrames]; // LR sample of site (a, b) is lr[a + oy][b + ox]
    int slot_ox[kMaxSlots][kMaxFrames];
    float* s1[kMaxSlots];               // planes: rows x pitch floats
    int rows, row0, pitch;              // plane row r holds site row a = row0 + r
    int own_a0, own_a1;                 // site rows whose V this launch counts
    double* vsum;
};
int launch_lattice_sums(const LatticeArgs& a, cudaStream_t st);

struct alignas(64) SepArgs {
    CUtensorMap tm_hr;                  // hr_in as a {W, local rows} tensor, box 144 x 80
    CUtensorMap tm_s1[kMaxSlots];       // S1 planes, box 72 x 36
    IbpGeom g;
    const float* lr[kMaxFrames];        // raw frames (edge fix-ups only)
    float* hr_out;
    float psf[49];                      // 7x7 row-major (fix-ups)
    float acol[7], brow[7];             // psf[u][v] == acol[u] * brow[v]
    int nph;
    int ph_pq[kMaxSlots];               // P*2+Q of every active lattice, ascending
    float mtab[kMaxSlots][4];           // frames with a sample at a site: [2*(last site row) + (last site col)]
    int s1_row0;                        // site row held by plane row 0
    int tiles_x, tiles_y;
    int ty_origin;                      // global HR row of tile row 0 (even)
    double* partial;                    // one slot per tile
    int n_sm;                           // SMs of the device (grid of the pipelined kernel)
    int dbg;                            // development probes (SR459_DBG): 1 = no arithmetic, 2 = no TMA loads
};
bool   ibp_sep_supported(int kh, int kw, int smax, bool separable, int W);
// warp-synchronous form of the same iteration (ibp_ws.cu): tile 64 x 112, own TMA box shapes
int    launch_ibp_ws(const SepArgs& a, cudaStream_t st);
void   ibp_ws_tile(int* th, int* tw);
int    ws_make_hr_map(CUtensorMap* out, const float* base, int W, int rows);
int    ws_make_s1_map(CUtensorMap* out, const float* base, int w, int rows, int pitch);

// ---- row bands: flag words at the head of a band arena, and what an iteration kernel needs to exchange halos itself (bands.cu)
constexpr int BF_FROM_UP = 0, BF_FROM_DOWN = 1, BF_ERROR = 2, BF_BASE = 3, BF_TICKET = 4, BF_ARR_UP = 5, BF_ARR_DOWN = 6;
struct BandPeer {
    unsigned* flags;                    // my flag words (null: no in-kernel exchange)
    unsigned* peer_flag[2];             // the arrival counter (BF_ARR_*) of the band above / below: one remote add per pushing warp
    float* peer_out[2];                 // the neighbour's copy of the buffer this launch writes, biased so that
                                        // [local row * W + x] is the same global pixel as in my buffer
    int push_y0[2], push_y1[2];         // local rows that are ALSO stored to the band above / below
    int push_warps[2];                  // warps with a share in those rows
    int unit;                           // arrivals per iteration and direction (= strips): wait for unit * (BASE + it)
    int has[2];
    int it;                             // iteration within the run: wait for BASE + it, publish BASE + it + 1
    unsigned long long timeout_ns;
    int probe;                          // development timing probes (SR459_BAND_PROBE): 1 no waits, 2 no peer stores, 4 no fence / publish
};

// ---- register-streaming fused iteration (ibp_stream.cu): one lattice, rank-1 PSF, whole image ---------
struct alignas(64) StreamArgs {
    CUtensorMap tm_hr;                  // hr_in as a {W, H} tensor, box 128 x 8
    CUtensorMap tm_s1;                  // the lattice's S1 plane {w, h}, box 64 x 4
    float* hr_out;
    int H, W, h, w;                     // dims of the (local) image behind tm_hr / tm_s1
    int y_base, y_end;                  // rows this launch updates, partitioned into segments (whole image: 0, H)
    int nit;                            // loop iterations (4 lattice rows each) per warp
    int seg_rows;                       // HR rows a warp updates = 8 nit - 6
    int nseg, nstrip;                   // interior warps = nseg * nstrip
    int b_nit, b_top, b_bot;            // row bands: short border segments (8 b_nit - 6 rows) at the band's top / bottom edge, mapped to
                                        // the FIRST warps of the grid so that their rows reach the neighbours while the interior still runs
    float alpha, lo, hi;
    float acol[7], brow[7];
    float mtab[4];                      // frames with a sample at a site: [2*(last site row) + (last site col)]
    int has_left;                       // lattice column parity 1: column-0 clamp terms exist
    const float* colS1; const float* leftS1; const float* rowS1; const float* topS1; const float* scal;
    float mc[2], ml[2];                 // frames with sx = -1 / +1 valid at a site row: [last site row]
    float mr[2], mt[2];                 // frames with sy = -1 / +1 valid at a site col: [last site col]
    float m_dupc, m_msec;               // frames (-1,-1) / (+1,+1)
    double* partial;                    // one slot per warp
    BandPeer bp;                        // row bands: halo rows stored straight into the neighbours' buffers
};
struct EdgeSumArgs;
int  launch_ibp_stream(const StreamArgs& a, int P, int Q, cudaStream_t st);
bool ibp_stream_supported(int kh, int kw, int smax, bool separable, int W, int nslot);
int  stream_make_hr_map(CUtensorMap* out, const float* base, int W, int rows);
int  stream_make_s1_map(CUtensorMap* out, const float* base, int w, int rows, int pitch);
void stream_geometry(int H, int W, int n_sm, int* nit, int* seg_rows, int* nseg, int* nstrip);
constexpr int kStreamBorderNit = 2;     // border segments of row bands: 10 rows (>= the 8 rows a neighbour needs)
// edge vectors + their variances (added to *vsum); buf holds 2h + 2w + 4 floats: colS1, leftS1, rowS1, topS1, scal
int  launch_stream_edge_sums(const float* const* lr, int n, int h, int w, const int* sy, const int* sx, int P, int Q,
                             float* buf, double* vsum, int own_a0, int own_a1, int top_owned, cudaStream_t st);

// ---- register-streaming fused iteration for 2..4 lattices (ibp_stream_ml.cu): rank-1 PSF, whole image -----------
struct alignas(64) StreamMLArgs {
    CUtensorMap tm_hr;                  // hr_in as a {W, H} tensor, box 128 x 8
    CUtensorMap tm_s1[kMaxSlots];       // S1 planes {w, h} in slot order (ascending P*2+Q), box 64 x 4
    float* hr_out;
    int H, W, h, w;
    int nit;                            // loop iterations (4 pairs of HR rows each) per warp
    int seg_rows;                       // HR rows a warp updates = 8 nit - 8
    int nseg, nstrip;
    float alpha, lo, hi;
    float acol[7], brow[7];
    float mtab[kMaxSlots][4];           // frames with a sample at a site: [2*(last site row) + (last site col)]
    const float* edge[kMaxSlots];       // per lattice: colS1[h], leftS1[h], rowS1[w], topS1[w], scal[4] (k_edge_sums)
    float mc[kMaxSlots][2], ml[kMaxSlots][2];   // frames with sx = -1 / +1 valid at a site row: [last site row]
    float mr[kMaxSlots][2], mt[kMaxSlots][2];   // frames with sy = -1 / +1 valid at a site col: [last site col]
    float m_dupc[kMaxSlots], m_msec[kMaxSlots]; // frames (-1,-1) / (+1,+1)
    double* partial;                    // one slot per warp
};
int  launch_ibp_stream_ml(const StreamMLArgs& a, int mask, cudaStream_t st);
bool ibp_stream_ml_supported(int kh, int kw, int smax, bool separable, int W, int nslot);
int  stream_ml_make_hr_map(CUtensorMap* out, const float* base, int W, int rows);
void stream_ml_geometry(int H, int W, int n_sm, int* nit, int* seg_rows, int* nseg, int* nstrip);

// ---- lattice form for large / non-separable PSFs (ibp_lat.cu): reach <= 15, |integer HR shift| <= 1 ---------
struct LatPlan;                         // tables + site planes, owned by the plan
struct LatCall {
    int H, W, h, w, n;
    int sy[kMaxFrames], sx[kMaxFrames];
    int hr_lo, hr_hi, lr_lo, lr_hi, own_i0, own_i1, upd_y0, upd_y1;
    float alpha, lo, hi;
    const float* lr[kMaxFrames];
    const float* hr_in; float* hr_out;
    double* partial;                    // lat_per_iter() slots
    // optional pre-summed S1 planes (k_lattice_sums, once per ibp() call): the site residuals become A = S1 - m B, one elementwise
    // pass instead of a pass over the n frames; the caller then adds the planes' sum of V to the squared-error sums
    int s1_n;                           // 0: not available
    int s1_pq[kMaxSlots];               // P*2+Q of plane t
    const float* s1[kMaxSlots];         // rows x pitch floats, pitch = (w + 3) & ~3; plane row r = site row s1_row0 + r
    int s1_row0, s1_rows;
    cudaStream_t side; cudaEvent_t ev_fork, ev_join;   // optional: k_lat_edge_b runs beside k_lat_fwd (null: in line)
};
bool    lat_supported(int kh, int kw, int smax, int W);
int     lat_create(LatPlan** out, const double* psf, int kh, int kw);
void    lat_destroy(LatPlan* L);
int     lat_per_iter(const LatPlan* L, const LatCall& c);
int     lat_iterate(LatPlan* L, const LatCall& c, cudaStream_t st);
int64_t lat_scratch_bytes(const LatPlan* L);

// errors[it] = (sum(partials of iteration it) + *d_add) * scale     (d_add may be null)
int launch_reduce_partials(const double* partial, int per_iter, int n_iter, double scale,
                           const double* d_add, double* d_errors, cudaStream_t st);

// ---- misc kernels (misc.cu) -------------------------------------------------------
int launch_u8_to_f32(const uint8_t* in, int64_t n, float* out, cudaStream_t st);
int launch_f32_to_u8(const float* in, int64_t n, uint8_t* out, cudaStream_t st);
int launch_extract_red(const void* in, int is_f32, int H, int W, int ro, int co, float* out, cudaStream_t st);
int launch_ingest_groups(const uint8_t* const* h_imgs, int n_groups, const int* h_reps, int H, int W, int step, int ro, int co,
                         float* const* h_frames, float* mean_f32, uint8_t* mean_u8, cudaStream_t st);
int launch_mean_frames(const float* const* h_frames, int n, int64_t npx, float* out, cudaStream_t st);
int launch_argmax(const void* imgs, int is_f32, int n_img, int64_t npx, long long* d_out /*n_img packed*/, cudaStream_t st);
int launch_patch_sum(const void* imgs, int is_f32, int n_img, int H, int W, const int* d_peaks, int margin,
                     double* d_patch /* (2m+1)^2 */, cudaStream_t st);

}  // namespace sr459
