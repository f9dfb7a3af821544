// spline.cu -- cubic B-spline prefilter + table-driven resampling.
//
// Replaces the arithmetic of scipy.ndimage.zoom / scipy.ndimage.shift as the
// reference calls them (mono_cal_target/run_sr.py:163, :176, :185-186, :279):
//   * prefilter: pole z = sqrt(3)-2, gain 6.  scipy runs a causal then an
//     anticausal recursion over the whole line with a boundary-dependent
//     initialisation ('mirror' for zoom's default mode, 'reflect' for
//     mode='nearest').  Both are exactly c = h * x~ with h[k] = sqrt(3) z^|k|
//     and x~ the infinite mirror / reflect extension of the line, so a tile
//     with a 16-sample warm-up halo on every side reproduces them to
//     z^16 = 7e-10 relative -- far below float32 resolution.  The kernel does
//     both axes of one 64x64 tile in shared memory (column pass on 96 columns,
//     row pass on 64 rows), so a 2-D prefilter is ONE launch and one HBM round
//     trip instead of scipy's two full passes.
//   * mode='nearest' first pads the array by 12 edge samples
//     (_prepad_for_spline_filter); the kernel reads that padded array
//     virtually (clamped source index), and can also read a zero-inserted
//     low-resolution residual (back_project, run_sr.py:170-176) virtually.
//   * evaluation: per-axis tables of 4 tap indices + 4 weights, computed on the
//     host in double precision exactly as NI_ZoomShift does, then gathered.
#include "internal.h"
#include <math.h>
#include <algorithm>

namespace sr459 {

// ------------------------------------------------------------------------------------
// host tables
// ------------------------------------------------------------------------------------
static inline void cubic_w(double t, double w[4]) {
    double u = 1.0 - t;
    w[0] = u * u * u / 6.0;
    w[1] = (t * t * (t - 2.0) * 3.0 + 4.0) / 6.0;
    w[2] = (u * u * (u - 2.0) * 3.0 + 4.0) / 6.0;
    w[3] = 1.0 - w[0] - w[1] - w[2];
}

static inline int mirror_idx(long long i, int n) {
    if (n <= 1) return 0;
    long long p = 2LL * n - 2;
    long long m = i % p;
    if (m < 0) m += p;
    return (int)(m >= n ? p - m : m);
}

void build_zoom_axis(int n_in, int n_out, std::vector<AxisTap>& out, int order) {
    out.resize(n_out);
    double ratio = (n_out > 1) ? (double)(n_in - 1) / (double)(n_out - 1) : 1.0;
    for (int o = 0; o < n_out; ++o) {
        double x = (double)o * ratio;
        bool valid = (x >= 0.0) && (x <= (double)(n_in - 1));   // mode='constant', cval=0
        if (order < 2) {                                        // nearest / linear: no prefilter, 1 or 2 taps, samples outside are 0
            const double fl = order == 0 ? floor(x + 0.5) : floor(x), t = x - fl;
            const long long i0 = (long long)fl;
            const double wv[2] = {order == 0 ? 1.0 : 1.0 - t, order == 0 ? 0.0 : t};
            for (int j = 0; j < 4; ++j) {
                const long long i = i0 + (j < 2 ? j : 0);
                const bool in = i >= 0 && i < n_in;
                out[o].idx[j] = (int)std::min<long long>(std::max<long long>(i, 0), n_in - 1);
                out[o].w[j] = (j < 2 && valid && in) ? (float)wv[j] : 0.0f;
            }
            continue;
        }
        double fl = floor(x);
        double w[4];
        cubic_w(x - fl, w);
        long long start = (long long)fl - 1;
        for (int j = 0; j < 4; ++j) {
            out[o].idx[j] = mirror_idx(start + j, n_in);
            out[o].w[j]   = valid ? (float)w[j] : 0.0f;
        }
    }
}

void build_shift_axis(int n, double s, int n_out, int stride, std::vector<AxisTap>& out) {
    out.resize(n_out);
    const int np = n + 2 * kPad;
    for (int o = 0; o < n_out; ++o) {
        double x = (double)(o * (long long)stride) - s + (double)kPad;
        x = std::min(std::max(x, 0.0), (double)(np - 1));       // mode='nearest'
        double fl = floor(x);
        double w[4];
        cubic_w(x - fl, w);
        long long start = (long long)fl - 1;
        for (int j = 0; j < 4; ++j) {
            long long t = start + j;
            out[o].idx[j] = (int)std::min<long long>(std::max<long long>(t, 0), np - 1);
            out[o].w[j]   = (float)w[j];
        }
    }
}

// ------------------------------------------------------------------------------------
// prefilter kernel
// ------------------------------------------------------------------------------------
constexpr int PF_T    = 64;              // tile edge (outputs)
constexpr int PF_WARM = 16;              // warm-up halo
constexpr int PF_S    = PF_T + 2 * PF_WARM;   // 96
constexpr int PF_P    = PF_S + 1;        // smem pitch (odd: conflict-free row walks)

__device__ __forceinline__ int fold_mirror(int i, int n) {      // period 2n-2
    if (n <= 1) return 0;
    int p = 2 * n - 2;
    int m = i % p;
    if (m < 0) m += p;
    return m >= n ? p - m : m;
}
__device__ __forceinline__ int fold_reflect(int i, int n) {     // period 2n
    int p = 2 * n;
    int m = i % p;
    if (m < 0) m += p;
    return m >= n ? p - 1 - m : m;
}

__device__ __forceinline__ float pf_fetch(const PrefilterJob& j, int Y, int X) {
    // fold only when the (virtual) coordinate is outside the coefficient array: integer modulo is
    // ~20 instructions and almost every fetch is inside
    if (j.mode == SRC_PLAIN) {
        if ((unsigned)Y >= (unsigned)j.ny) Y = fold_mirror(Y, j.ny);
        if ((unsigned)X >= (unsigned)j.nx) X = fold_mirror(X, j.nx);
        return __ldg(j.src + (size_t)Y * j.sw + X);
    }
    if ((unsigned)Y >= (unsigned)j.ny) Y = fold_reflect(Y, j.ny);
    if ((unsigned)X >= (unsigned)j.nx) X = fold_reflect(X, j.nx);
    const int yy = min(max(Y - kPad, 0), j.vh - 1);
    const int xx = min(max(X - kPad, 0), j.vw - 1);
    if (j.mode == SRC_EDGEPAD) return __ldg(j.src + (size_t)yy * j.sw + xx);
    // zero-inserted residual (run_sr.py:170-175): up[::f, ::f] = err, zero elsewhere / beyond
    if (j.factor == 2) {
        if ((yy | xx) & 1) return 0.0f;
        const int i = yy >> 1, k = xx >> 1;
        if (i >= j.sh || k >= j.sw) return 0.0f;
        return __ldg(j.src + (size_t)i * j.sw + k);
    }
    const int f = j.factor;
    if ((yy % f) | (xx % f)) return 0.0f;
    const int i = yy / f, k = xx / f;
    if (i >= j.sh || k >= j.sw) return 0.0f;
    return __ldg(j.src + (size_t)i * j.sw + k);
}

// Lane-dependent carry factors of the scans (16 lanes per line, 6 samples per lane): z^6, z^12 where the neighbour
// 1 / 2 lanes away exists on that side of the line, 0 otherwise (a multiply by zero instead of a select per step),
// and the 0/1 gates of the carry-in.
struct PfLane { float u1, u2, ug, d1, d2, dg; };
__device__ __forceinline__ PfLane pf_lane(int lane) {
    const float z = kPoleF, z3 = z * z * z, z6 = z3 * z3, z12 = z6 * z6;
    const int l = lane & 15;
    PfLane f;
    f.u1 = l >= 1 ? z6 : 0.f;   f.u2 = l >= 2 ? z12 : 0.f;   f.ug = l >= 1 ? 1.f : 0.f;
    f.d1 = l <= 14 ? z6 : 0.f;  f.d2 = l <= 13 ? z12 : 0.f;  f.dg = l <= 14 ? 1.f : 0.f;
    // a shuffle from the own lane: opaque to ptxas, which otherwise re-materialises the compare + select pairs for
    // every line instead of keeping the values in registers
    f.u1 = __shfl_sync(0xffffffffu, f.u1, lane); f.u2 = __shfl_sync(0xffffffffu, f.u2, lane); f.ug = __shfl_sync(0xffffffffu, f.ug, lane);
    f.d1 = __shfl_sync(0xffffffffu, f.d1, lane); f.d2 = __shfl_sync(0xffffffffu, f.d2, lane); f.dg = __shfl_sync(0xffffffffu, f.dg, lane);
    return f;
}

// Two 96-sample lines per warp (one per half warp), in place: q points at this lane's 6 consecutive samples
// q[0], q[stride], ...  Symmetric exponential filter c = P + Q - x with P[i] = sum_k z^k x[i-k], Q[i] = sum_k z^k x[i+k]:
// serial inside the lane, two shuffle steps (carry factors z^6, z^12) across the 16 lanes of the line, i.e. 24 samples
// of carry -- the same z^16-or-better truncation as the 16-sample warm-up halo.  The result is multiplied by `gain`
// (1 after the first axis, 3 = sqrt(3)^2 after the second).
__device__ __forceinline__ void pf_scan_line6(float* q, int stride, const PfLane& f, float gain) {
    const float z = kPoleF, z2 = z * z, z3 = z2 * z, z4 = z2 * z2, z5 = z4 * z, z6 = z3 * z3;
    const float zp[7] = {1.f, z, z2, z3, z4, z5, z6};
    float x[6], a[6], b[6];
    #pragma unroll
    for (int i = 0; i < 6; ++i) x[i] = q[i * stride];
    a[0] = x[0];
    #pragma unroll
    for (int i = 1; i < 6; ++i) a[i] = fmaf(z, a[i - 1], x[i]);
    b[5] = x[5];
    #pragma unroll
    for (int i = 4; i >= 0; --i) b[i] = fmaf(z, b[i + 1], x[i]);
    float C = a[5], D = b[0];
    C = fmaf(f.u1, __shfl_up_sync(0xffffffffu, C, 1, 16), C);
    D = fmaf(f.d1, __shfl_down_sync(0xffffffffu, D, 1, 16), D);
    C = fmaf(f.u2, __shfl_up_sync(0xffffffffu, C, 2, 16), C);
    D = fmaf(f.d2, __shfl_down_sync(0xffffffffu, D, 2, 16), D);
    const float cin = f.ug * __shfl_up_sync(0xffffffffu, C, 1, 16);
    const float din = f.dg * __shfl_down_sync(0xffffffffu, D, 1, 16);
    // P[i] + Q[i] - x[i] = a[i] + z b[i+1] + z^(i+1) cin + z^(6-i) din      (b[6] = 0)
    #pragma unroll
    for (int i = 0; i < 6; ++i) {
        float v = fmaf(zp[i + 1], cin, a[i]);
        v = fmaf(zp[6 - i], din, v);
        if (i < 5) v = fmaf(z, b[i + 1], v);
        q[i * stride] = gain * v;
    }
}

struct PrefilterJobs { PrefilterJob job[8]; };

// 256 threads = 8 warps.  Load / store: warp w takes tile rows w, w+8, ..., lanes run along x (coalesced).
// Passes: one thread per line (96 columns, then the 64 interior rows); 6 CTAs are resident per SM, so
// the serial recursions of one tile overlap the loads and stores of the others.
__global__ void __launch_bounds__(256) k_prefilter2d(const __grid_constant__ PrefilterJobs jobs) {
    extern __shared__ float s[];
    const PrefilterJob& j = jobs.job[blockIdx.z];
    const int ty0 = blockIdx.y * PF_T, tx0 = blockIdx.x * PF_T;
    if (ty0 >= j.ny || tx0 >= j.nx) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const PfLane pl = pf_lane(lane);
    const int half = lane >> 4, l16 = lane & 15;

    // region [ty0-16, ty0+80) x [tx0-16, tx0+80) of the (virtual) coefficient array
    const int off = (j.mode == SRC_PLAIN) ? 0 : kPad;
    const bool fast = (j.mode != SRC_ZI_EDGEPAD) &&
                      (ty0 - PF_WARM - off >= 0) && (ty0 + PF_T + PF_WARM - off <= j.sh) &&
                      (tx0 - PF_WARM - off >= 0) && (tx0 + PF_T + PF_WARM - off <= j.sw);
    if (fast) {
        const float* src = j.src + (size_t)(ty0 - PF_WARM - off) * j.sw + (tx0 - PF_WARM - off);
        #pragma unroll 4
        for (int ly = warp; ly < PF_S; ly += 8) {
            const float* row = src + (size_t)ly * j.sw;
            float* d = s + ly * PF_P;
            d[lane] = __ldg(row + lane); d[lane + 32] = __ldg(row + lane + 32); d[lane + 64] = __ldg(row + lane + 64);
        }
    } else {
        for (int ly = warp; ly < PF_S; ly += 8) {
            float* d = s + ly * PF_P;
            #pragma unroll
            for (int q = 0; q < 3; ++q) d[lane + 32 * q] = pf_fetch(j, ty0 - PF_WARM + ly, tx0 - PF_WARM + lane + 32 * q);
        }
    }
    __syncthreads();
    // Both passes as half-warp scans (pf_scan_line6): a line of 96 samples = 6 consecutive samples per lane, two lines per
    // warp.  No serial recursion over the line, no idle threads, 3 shuffles per direction for two lines.
    // axis 0 (down the columns), all 96 columns: warp w takes the column pairs w, w+8, ...; lane l holds rows 6l..6l+5
    // (pitch 97: the 16 lanes of a column hit 16 distinct even banks, the neighbouring column the odd ones)
    for (int cp = warp; cp < PF_S / 2; cp += 8) pf_scan_line6(s + 2 * cp + half + 6 * l16 * PF_P, PF_P, pl, 1.0f);
    __syncthreads();
    // axis 1 (along the rows), the 64 interior rows
    for (int rp = warp; rp < PF_T / 2; rp += 8) pf_scan_line6(s + (PF_WARM + 2 * rp + half) * PF_P + 6 * l16, 1, pl, 3.0f);
    __syncthreads();
    const int X0 = tx0 + lane, X1 = tx0 + lane + 32;
    #pragma unroll 4
    for (int ly = warp; ly < PF_T; ly += 8) {
        const int Y = ty0 + ly;
        if (Y >= j.ny) break;
        const float* r = s + (PF_WARM + ly) * PF_P + PF_WARM;
        float* d = j.dst + (size_t)Y * j.nx;
        if (X0 < j.nx) d[X0] = r[lane];
        if (X1 < j.nx) d[X1] = r[lane + 32];
    }
}

int launch_prefilter2d(const PrefilterJob* h_jobs, int njobs, cudaStream_t st) {
    static PerDeviceOnce configured;
    size_t smem = sizeof(float) * PF_S * PF_P;
    if (configured.need()) {
        SR_CUDA(cudaFuncSetAttribute(k_prefilter2d, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    for (int first = 0; first < njobs; first += 8) {
        const int cnt = std::min(8, njobs - first);
        PrefilterJobs jobs{};
        int maxy = 0, maxx = 0;
        for (int i = 0; i < cnt; ++i) {
            jobs.job[i] = h_jobs[first + i];
            maxy = std::max(maxy, jobs.job[i].ny); maxx = std::max(maxx, jobs.job[i].nx);
        }
        dim3 grid((maxx + PF_T - 1) / PF_T, (maxy + PF_T - 1) / PF_T, cnt);
        k_prefilter2d<<<grid, 256, smem, st>>>(jobs);
        SR_LAUNCH_CHECK("k_prefilter2d");
    }
    return SR459_OK;
}

// ------------------------------------------------------------------------------------
// resampling kernels
// ------------------------------------------------------------------------------------
__device__ __forceinline__ float gather16(const float* __restrict__ c, int pitch, const AxisTap& a, const AxisTap& b) {
    float v = 0.0f;
    #pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float* row = c + (size_t)a.idx[i] * pitch;
        float r = 0.0f;
        #pragma unroll
        for (int k = 0; k < 4; ++k) r = fmaf(b.w[k], __ldg(row + b.idx[k]), r);
        v = fmaf(a.w[i], r, v);
    }
    return v;
}

__device__ __forceinline__ AxisTap load_tap(const AxisTap* p) {
    AxisTap t;
    const int4   i4 = __ldg(reinterpret_cast<const int4*>(p));
    const float4 w4 = __ldg(reinterpret_cast<const float4*>(p) + 1);
    t.idx[0] = i4.x; t.idx[1] = i4.y; t.idx[2] = i4.z; t.idx[3] = i4.w;
    t.w[0] = w4.x; t.w[1] = w4.y; t.w[2] = w4.z; t.w[3] = w4.w;
    return t;
}

__global__ void __launch_bounds__(256) k_resample_sum(ResampleSumArgs a) {
    int x = blockIdx.x * 32 + threadIdx.x;
    int y = blockIdx.y * 8 + threadIdx.y;
    if (x >= a.Wout || y >= a.Hout) return;
    float acc = 0.0f;
    for (int k = 0; k < a.n; ++k) {
        AxisTap ty = load_tap(a.ty + (size_t)k * a.Hout + y);
        AxisTap tx = load_tap(a.tx + (size_t)k * a.Wout + x);
        acc += gather16(a.coef[k], a.cpitch[k], ty, tx);
    }
    a.out[(size_t)y * a.Wout + x] = acc * a.scale;
}

int launch_resample_sum(const ResampleSumArgs& a, cudaStream_t st) {
    dim3 grid((a.Wout + 31) / 32, (a.Hout + 7) / 8), block(32, 8);
    k_resample_sum<<<grid, block, 0, st>>>(a);
    SR_LAUNCH_CHECK("k_resample_sum");
    return SR459_OK;
}

// ------------------------------------------------------------------------------------
// tiled separable resampling for UPSAMPLING tables (zoom, and zoom re-indexed by integer shifts)
// ------------------------------------------------------------------------------------
// out[y][x] = scale * sum_k sum_i wy_k[y][i] * ( sum_j wx_k[x][j] * coef_k[iy_k[y][i]][ix_k[x][j]] )
// One CTA owns a 64 x 128 output tile.  Per frame: the coefficient rows / columns the tile's table
// entries touch (at most 38 x 70 for a factor >= 2) are staged in shared memory, an x pass evaluates
// every staged row at the tile's 128 columns, and a y pass accumulates 8 rows x 4 columns per thread
// in registers.  8 FMA and ~6 shared-memory loads per output and frame instead of 16 + 16 global loads.
constexpr int ZT_H = 64, ZT_W = 128;
constexpr int ZR_R = ZT_H / 2 + 6, ZR_C = ZT_W / 2 + 6;     // staged coefficient region (rows x cols)
constexpr int ZR_P = ZR_C + 1;

// Per-frame table bounds of the tile: warp 0 the rows (also copies the row taps to shared memory), warp 1 the columns.
__device__ __forceinline__ void zt_bounds(const AxisTap* ty, const AxisTap* tx, int th, int tw, AxisTap* sty, int* bnd,
                                          int warp, int lane) {
    if (warp == 0) {
        int lo = 0x7fffffff, hi = -1;
        for (int y = lane; y < th; y += 32) {
            const AxisTap t = load_tap(ty + y);
            sty[y] = t;
            lo = min(lo, min(min(t.idx[0], t.idx[1]), min(t.idx[2], t.idx[3])));
            hi = max(hi, max(max(t.idx[0], t.idx[1]), max(t.idx[2], t.idx[3])));
        }
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) { lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o)); }
        if (lane == 0) { bnd[0] = lo; bnd[1] = hi; }
    } else if (warp == 1) {
        int lo = 0x7fffffff, hi = -1;
        for (int x = lane; x < tw; x += 32) {
            const int4 i4 = __ldg(reinterpret_cast<const int4*>(tx + x));
            lo = min(lo, min(min(i4.x, i4.y), min(i4.z, i4.w)));
            hi = max(hi, max(max(i4.x, i4.y), max(i4.z, i4.w)));
        }
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) { lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o)); }
        if (lane == 0) { bnd[2] = lo; bnd[3] = hi; }
    }
}

constexpr int ZT_NSR = (ZR_R + 7) / 8;          // staged rows per warp (warp w: rows w, w+8, ...)
constexpr int ZT_NSC = (ZR_C + 31) / 32;        // column chunks per lane (lane, lane+32, ...)

// all of a thread's loads of one frame's coefficient region (at most 38 x 70), issued back to back into registers
__device__ __forceinline__ void zt_fetch(float (&v)[ZT_NSR][ZT_NSC], const float* coef, int pitch, int rmin, int cmin, int nr, int nc,
                                         int warp, int lane) {
    const float* src = coef + (size_t)(rmin + warp) * pitch + cmin + lane;
    #pragma unroll
    for (int i = 0; i < ZT_NSR; ++i) {
        const bool rok = warp + 8 * i < nr;
        #pragma unroll
        for (int c = 0; c < ZT_NSC; ++c)
            v[i][c] = (rok && lane + 32 * c < nc) ? __ldg(src + 32 * c) : 0.0f;
        src += (size_t)8 * pitch;
    }
}
__device__ __forceinline__ void zt_stage(float* R, const float (&v)[ZT_NSR][ZT_NSC], int warp, int lane) {
    float* d = R + warp * ZR_P + lane;
    #pragma unroll
    for (int i = 0; i < ZT_NSR; ++i) {
        if (warp + 8 * i < ZR_R) {
            #pragma unroll
            for (int c = 0; c < ZT_NSC; ++c)
                if (lane + 32 * c < ZR_C) d[32 * c] = v[i][c];
        }
        d += 8 * ZR_P;
    }
}

// Software pipeline over the frames: while frame k runs its x and y passes out of shared memory, the table bounds of
// frame k+1 are reduced (warps 0/1, double-buffered) and its coefficient region is already in flight into registers.
__global__ void __launch_bounds__(256, 3) k_resample_tile(const __grid_constant__ ResampleSumArgs a) {
    __shared__ float R[ZR_R * ZR_P];            // staged coefficients
    __shared__ __align__(16) float T[ZR_R * ZT_W];   // x-passed rows
    __shared__ __align__(16) AxisTap sty[2][ZT_H];
    __shared__ int bnd[2][4];                   // rmin, rmax, cmin, cmax
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int y0 = blockIdx.y * ZT_H, x0 = blockIdx.x * ZT_W;
    const int th = min(ZT_H, a.Hout - y0), tw = min(ZT_W, a.Wout - x0);
    const int xq = tid & 31, yq = tid >> 5;     // y pass: columns 4xq..4xq+3, rows 8yq..8yq+7
    float acc[8][4];
    #pragma unroll
    for (int i = 0; i < 8; ++i) { acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.0f; }

    float v[ZT_NSR][ZT_NSC];
    zt_bounds(a.ty + y0, a.tx + x0, th, tw, sty[0], bnd[0], warp, lane);
    __syncthreads();
    int rmin = bnd[0][0], cmin = bnd[0][2];
    int nr = min(bnd[0][1] - rmin + 1, ZR_R), nc = min(bnd[0][3] - cmin + 1, ZR_C);
    zt_fetch(v, a.coef[0], a.cpitch[0], rmin, cmin, nr, nc, warp, lane);

    for (int k = 0; k < a.n; ++k) {
        const int cur = k & 1, nxt = cur ^ 1;
        const AxisTap* tx = a.tx + (size_t)k * a.Wout + x0;
        // ---- stage (the x pass of frame k-1 finished before its second barrier) ----------------------------
        zt_stage(R, v, warp, lane);
        __syncthreads();                          // also: frame k-1's y pass is done with T, sty[nxt] and bnd[nxt]
        if (k + 1 < a.n)      // tables of the next frame; read after the next barrier
            zt_bounds(a.ty + (size_t)(k + 1) * a.Hout + y0, a.tx + (size_t)(k + 1) * a.Wout + x0, th, tw, sty[nxt], bnd[nxt], warp, lane);
        // ---- x pass: T[rr][x] -----------------------------------------------------------------------------
        {
            const int x = tid & (ZT_W - 1), half = tid >> 7;
            if (x < tw) {
                const AxisTap t = load_tap(tx + x);
                // fully unrolled over the (at most 19) rows of this half: every address is a register plus an immediate
                const float* r0 = R + half * ZR_P + (t.idx[0] - cmin), * r1 = R + half * ZR_P + (t.idx[1] - cmin);
                const float* r2 = R + half * ZR_P + (t.idx[2] - cmin), * r3 = R + half * ZR_P + (t.idx[3] - cmin);
                float* tp = T + half * ZT_W + x;
                #pragma unroll
                for (int i = 0; i < ZR_R / 2; ++i) {
                    // rows >= nr of R hold staged zeros and the matching rows of T are never read: no guard
                    tp[2 * i * ZT_W] = fmaf(t.w[3], r3[2 * i * ZR_P], fmaf(t.w[2], r2[2 * i * ZR_P], fmaf(t.w[1], r1[2 * i * ZR_P], t.w[0] * r0[2 * i * ZR_P])));
                }
            } else {
                for (int rr = half; rr < ZR_R; rr += 2) T[rr * ZT_W + x] = 0.0f;
            }
        }
        __syncthreads();
        // ---- next frame's coefficients into flight, then the y pass of this one ---------------------------
        const int rmin_cur = rmin;
        if (k + 1 < a.n) {
            rmin = bnd[nxt][0]; cmin = bnd[nxt][2];
            nr = min(bnd[nxt][1] - rmin + 1, ZR_R); nc = min(bnd[nxt][3] - cmin + 1, ZR_C);
            zt_fetch(v, a.coef[k + 1], a.cpitch[k + 1], rmin, cmin, nr, nc, warp, lane);
        }
        #pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int y = 8 * yq + i;
            if (y < th) {
                const AxisTap t = sty[cur][y];
                const float4* tp = reinterpret_cast<const float4*>(T) + xq;
                #pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float4 vv = tp[(t.idx[q] - rmin_cur) * (ZT_W / 4)];
                    acc[i][0] = fmaf(t.w[q], vv.x, acc[i][0]); acc[i][1] = fmaf(t.w[q], vv.y, acc[i][1]);
                    acc[i][2] = fmaf(t.w[q], vv.z, acc[i][2]); acc[i][3] = fmaf(t.w[q], vv.w, acc[i][3]);
                }
            }
        }
    }
    // ---- store ------------------------------------------------------------------------------------------------
    const int gx = x0 + 4 * xq;
    #pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int gy = y0 + 8 * yq + i;
        if (gy >= a.Hout || gx >= a.Wout) continue;
        float* d = a.out + (size_t)gy * a.Wout + gx;
        if (gx + 3 < a.Wout && (a.Wout & 3) == 0) {
            *reinterpret_cast<float4*>(d) = make_float4(acc[i][0] * a.scale, acc[i][1] * a.scale, acc[i][2] * a.scale, acc[i][3] * a.scale);
        } else {
            for (int q = 0; q < 4 && gx + q < a.Wout; ++q) d[q] = acc[i][q] * a.scale;
        }
    }
}

int launch_resample_tile(const ResampleSumArgs& a, cudaStream_t st) {
    dim3 grid((a.Wout + ZT_W - 1) / ZT_W, (a.Hout + ZT_H - 1) / ZT_H);
    k_resample_tile<<<grid, 256, 0, st>>>(a);
    SR_LAUNCH_CHECK("k_resample_tile");
    return SR459_OK;
}

__global__ void __launch_bounds__(256) k_resample_err(ResampleErrArgs a) {
    const int k = blockIdx.z;
    int x = blockIdx.x * 32 + threadIdx.x;
    int y = blockIdx.y * 8 + threadIdx.y;
    float e2 = 0.0f;
    if (x < a.w && y < a.h) {
        AxisTap ty = load_tap(a.ty + (size_t)k * a.h + y);
        AxisTap tx = load_tap(a.tx + (size_t)k * a.w + x);
        float sim = gather16(a.coef, a.cpitch, ty, tx);
        size_t o = (size_t)y * a.w + x;
        if (a.lr[k]) {
            float e = __ldg(a.lr[k] + o) - sim;
            a.out[k][o] = e;
            e2 = (y >= a.mse_i0 && y < a.mse_i1) ? e * e : 0.0f;
        } else {
            a.out[k][o] = sim;
        }
    }
    if (a.partial) {
        __shared__ double red[8];
        double d = (double)e2;
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
        if (threadIdx.x == 0) red[threadIdx.y] = d;
        __syncthreads();
        if (threadIdx.x == 0 && threadIdx.y == 0) {
            double t = 0.0;
            #pragma unroll
            for (int i = 0; i < 8; ++i) t += red[i];
            a.partial[((size_t)k * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = t;
        }
    }
}

int launch_resample_err(const ResampleErrArgs& a, cudaStream_t st, int* nblocks_out) {
    dim3 grid((a.w + 31) / 32, (a.h + 7) / 8, a.n), block(32, 8);
    if (nblocks_out) *nblocks_out = (int)(grid.x * grid.y * grid.z);
    k_resample_err<<<grid, block, 0, st>>>(a);
    SR_LAUNCH_CHECK("k_resample_err");
    return SR459_OK;
}

// ---- the same, one CTA per 16 x 64 LR tile and ALL frames: the coefficient region the tile's taps touch (about 38 x 134 samples:
// the frames' sub-pixel shifts move it by a pixel or two) is staged ONCE in shared memory, de-interleaved by column parity so that
// the stride-2 reads of neighbouring lanes fall into consecutive banks; the row tables sit in shared memory, the column table of
// the thread's column in registers.  16 shared-memory loads per output instead of 16 cached global gathers + two 32-byte table
// loads (k_resample_err: 38 us at config 2; this one: see DESIGN.md).  Tables are used as they are (every boundary rule stays
// inside them); a tile whose taps span more than the staged region (never for |shift| < 4) falls back to global gathers.
constexpr int RE_SR = 2 * kRsTileH + 8, RE_SC = 2 * kRsTileW + 8, RE_HP = RE_SC / 2 + 1, RE_THREADS = 4 * kRsTileW;

__global__ void __launch_bounds__(RE_THREADS, 768 / RE_THREADS) k_resample_err_tile(const __grid_constant__ ResampleErrArgs a) {
    __shared__ float coef[RE_SR][2][RE_HP];
    __shared__ AxisTap sty[kMaxFrames][kRsTileH];
    __shared__ double red[RE_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int y0 = blockIdx.y * kRsTileH, x0 = blockIdx.x * kRsTileW;
    const int th = min(kRsTileH, a.h - y0), tw = min(kRsTileW, a.w - x0);
    // bounds of the coefficient region: the tables are monotone in the output index (coordinates grow with it, clamps keep the
    // order), so the first and last entries of every frame bound it -- uniform loads, all in flight at once
    int rlo = 0x7fffffff, rhi = -1, clo = 0x7fffffff, chi = -1;
    for (int k = 0; k < a.n; ++k) {
        const int4 r0 = __ldg(reinterpret_cast<const int4*>(a.ty + (size_t)k * a.h + y0));
        const int4 r1 = __ldg(reinterpret_cast<const int4*>(a.ty + (size_t)k * a.h + y0 + th - 1));
        const int4 c0 = __ldg(reinterpret_cast<const int4*>(a.tx + (size_t)k * a.w + x0));
        const int4 c1 = __ldg(reinterpret_cast<const int4*>(a.tx + (size_t)k * a.w + x0 + tw - 1));
        rlo = min(rlo, min(min(min(r0.x, r0.y), min(r0.z, r0.w)), min(min(r1.x, r1.y), min(r1.z, r1.w))));
        rhi = max(rhi, max(max(max(r0.x, r0.y), max(r0.z, r0.w)), max(max(r1.x, r1.y), max(r1.z, r1.w))));
        clo = min(clo, min(min(min(c0.x, c0.y), min(c0.z, c0.w)), min(min(c1.x, c1.y), min(c1.z, c1.w))));
        chi = max(chi, max(max(max(c0.x, c0.y), max(c0.z, c0.w)), max(max(c1.x, c1.y), max(c1.z, c1.w))));
    }
    const int nr = rhi - rlo + 1, nc = chi - clo + 1;
    const bool staged = nr <= RE_SR && nc <= RE_SC;
    const int cx = tid % kRsTileW, rg = tid / kRsTileW;             // column of the tile, group of four rows
    const int x = x0 + cx;
    // every global load the thread needs next is issued here, before the first dependent use: its share of the coefficient
    // region, its share of the row tables, the column taps of its column for the first frames
    constexpr int PER = (RE_SR * RE_SC + RE_THREADS - 1) / RE_THREADS, NPF = 4;
    AxisTap tyv;
    if (tid < a.n * th) { const int k = tid / th, r = tid - k * th; tyv = load_tap(a.ty + (size_t)k * a.h + y0 + r); }
    if (staged) {
        float cv[PER];
        #pragma unroll
        for (int i = 0; i < PER; ++i) {
            const int e = tid + RE_THREADS * i, r = e / RE_SC, c = e - r * RE_SC;
            cv[i] = (r < nr && c < nc) ? __ldg(a.coef + (size_t)(rlo + r) * a.cpitch + clo + c) : 0.0f;
        }
        #pragma unroll
        for (int i = 0; i < PER; ++i) {
            const int e = tid + RE_THREADS * i, r = e / RE_SC, c = e - r * RE_SC;
            if (r < RE_SR) coef[r][c & 1][c >> 1] = cv[i];
        }
    }
    if (tid < a.n * th) { const int k = tid / th, r = tid - k * th; sty[k][r] = tyv; }
    // the column taps and the LR samples of the first NPF frames travel while the CTA meets at the barrier
    AxisTap txv[NPF];
    float lrv[NPF][4];
    const bool col_ok = cx < tw;
    #pragma unroll
    for (int k = 0; k < NPF; ++k) {
        if (k < a.n && col_ok) {
            txv[k] = load_tap(a.tx + (size_t)k * a.w + x);
            #pragma unroll
            for (int r4 = 0; r4 < 4; ++r4) {
                const int r = rg * 4 + r4;
                lrv[k][r4] = (r < th && a.lr[k]) ? __ldg(a.lr[k] + (size_t)(y0 + r) * a.w + x) : 0.0f;
            }
        }
    }
    __syncthreads();
    float e2 = 0.0f;
    auto one_frame = [&](const int k, const AxisTap& tx, const float (&lrk)[4]) {
        int co[4];
        #pragma unroll
        for (int j = 0; j < 4; ++j) { const int c = tx.idx[j] - clo; co[j] = (c & 1) * RE_HP + (c >> 1); }
        #pragma unroll
        for (int r4 = 0; r4 < 4; ++r4) {
            const int r = rg * 4 + r4, y = y0 + r;
            if (r >= th) break;
            const AxisTap& ty = sty[k][r];
            float sim = 0.0f;
            if (staged) {
                #pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float* row = &coef[ty.idx[i] - rlo][0][0];
                    float v = 0.0f;
                    #pragma unroll
                    for (int j = 0; j < 4; ++j) v = fmaf(tx.w[j], row[co[j]], v);
                    sim = fmaf(ty.w[i], v, sim);
                }
            } else {
                sim = gather16(a.coef, a.cpitch, ty, tx);
            }
            const size_t o = (size_t)y * a.w + x;
            if (a.lr[k]) {
                const float e = lrk[r4] - sim;
                a.out[k][o] = e;
                if (y >= a.mse_i0 && y < a.mse_i1) e2 = fmaf(e, e, e2);
            } else {
                a.out[k][o] = sim;
            }
        }
    };
    if (col_ok) {
        #pragma unroll
        for (int k = 0; k < NPF; ++k) if (k < a.n) one_frame(k, txv[k], lrv[k]);
        for (int k = NPF; k < a.n; ++k) {                  // more than NPF frames: loaded in place
            const AxisTap tx = load_tap(a.tx + (size_t)k * a.w + x);
            float lrk[4];
            #pragma unroll
            for (int r4 = 0; r4 < 4; ++r4) {
                const int r = rg * 4 + r4;
                lrk[r4] = (r < th && a.lr[k]) ? __ldg(a.lr[k] + (size_t)(y0 + r) * a.w + x) : 0.0f;
            }
            one_frame(k, tx, lrk);
        }
    }
    if (a.partial) {
        double d = (double)e2;
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
        if (lane == 0) red[warp] = d;
        __syncthreads();
        if (tid == 0) {
            double t = 0.0;
            #pragma unroll
            for (int i = 0; i < RE_THREADS / 32; ++i) t += red[i];
            a.partial[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = t;
        }
    }
}

int launch_resample_err_tile(const ResampleErrArgs& a, cudaStream_t st) {
    dim3 grid((a.w + kRsTileW - 1) / kRsTileW, (a.h + kRsTileH - 1) / kRsTileH);
    k_resample_err_tile<<<grid, RE_THREADS, 0, st>>>(a);
    SR_LAUNCH_CHECK("k_resample_err_tile");
    return SR459_OK;
}

}  // namespace sr459
