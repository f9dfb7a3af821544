// ibp_lat.cu -- fused IBP iteration for LARGE or non-separable PSFs (reach <= 15, e.g. the 31x31
// measured pinhole PSF of BASELINE config 4) with integer HR shifts |2d| <= 1, lattice form.
//
// Same mathematics as ibp_dense.cu (mono_cal_target/run_sr.py:190-209 restated for integer shifts):
//     B   = conv0(hr, psf)                                  blur, zero padded       (:157-158)
//     err_k[i,j] = lr_k[i,j] - B[clamp(2i - sy_k), clamp(2j - sx_k)]                 (:161-165, :201)
//     acc[y,x]   = sum_k U(err_k)[clamp(y + sy_k), clamp(x + sx_k)]                  (:170-176)
//     hr'        = clip(hr + step/n * corr(acc, psf), lo, hi)                        (:178, :204-205)
// but the work is cut to what the parity lattices need (restructuring 1):
//   * frames whose HR shift has parity (P,Q) sample B only on the sites (2a+P, 2b+Q): the forward blur
//     is evaluated on one h x w site plane per lattice ("slot") -- 1/4 of the pixels;
//   * acc is non-zero only on those sites (plus clamp duplicates on HR row 0 / column 0), so the
//     transposed blur of an output pixel of parity (py,px) sees only the taps of matching parity.
// Both blurs then are sums of four stride-1 2-D correlations with 16x16 (or 16x15, 15x15) taps between
// de-interleaved planes: 961 MAC per site forward, 961 MAC per 4 pixels backward, ~480 MAC per HR pixel
// and iteration instead of the dense kernel's ~5000.  The path is FP32-FMA bound (SURVEY.md section 8d,
// "Exception C4"), so the inner loop is written for the FMA pipe: lanes own consecutive site ROWS
// (conflict-free 128-bit shared-memory loads with an odd row pitch), every lane keeps 16 outputs of its
// row in registers, taps are paired so that each FFMA2 (fma.rn.f32x2) does two MACs, and the weights come
// from the kernel-parameter constant bank through uniform registers (no LSU traffic at all).
//
// Per iteration: k_lat_fwd (site planes BL_s) -> k_lat_edge_b (B on HR row 0 / column 0, needed by the
// clamp terms) -> k_lat_resid (residual planes A_s = S1 - m*BL, clamp duplicates, exact f64 MSE partials)
// -> k_lat_bwd (transposed blur + update + clip).  Row bands (sr459_band_*) use the same kernels on the
// band's rows.
#include "internal.h"
#include <algorithm>
#include <string.h>
#include <stdlib.h>

namespace sr459 {

typedef unsigned long long u64;

__device__ __forceinline__ u64 lat_fma2(u64 a, u64 b, u64 c) {
    u64 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ float lat_sum2(u64 v) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
    return lo + hi;
}

constexpr int LT_ROWS = 32;             // site rows per tile: lane <-> row
constexpr int LT_WARPS = 4;
// NB = outputs per lane (consecutive site columns), a template parameter: 16 for whole images (tile = 32 x 64 sites; the best
// FFMA2 : LDS ratio), 8 for thin row bands (32 x 32 sites: twice the CTAs, so that a band of 1/8 of the image still gives every
// SM scheduler three warps -- one warp alone runs at about a third of the FFMA2 issue rate)

// Weight tables of one direction.  Tap sets per axis: parity 0 = the NT even taps, parity 1 = the NT-1 odd
// taps (+ one zero).  e[][][tu][m] = (w[2m], w[2m+1]); o[][][tu][m] = (w[2m-1], w[2m]) with w[-1] = w[NT] = 0:
// the two packings a lane needs for outputs at even / odd offsets from its 8-byte aligned window.
template <int NT>
struct alignas(16) LatTab {
    u64 w[2][2][NT][NT + 2];            // [0, NT/2): e pairs; [NT/2, NT+1): o pairs; one pad (16-byte rows)
};

struct LatGeom {
    int H, W, h, w;
    int hr_lo, hr_hi;                   // HR rows present in the local hr buffers
    int sa_lo, sa_hi;                   // site rows held by the planes (plane row 0 = site row sa_lo)
    int pitch;                          // plane pitch in floats (multiple of 4)
    int nslot;
    int slot_pq[kMaxSlots];             // P*2+Q
};

template <int NT, int NB>
__device__ __forceinline__ void lat_load_window(const float* p, u64 (&xp)[(NB + NT) / 2]) {
    const ulonglong2* q = reinterpret_cast<const ulonglong2*>(p);
    #pragma unroll
    for (int i = 0; i < (NB + NT) / 4; ++i) { const ulonglong2 t = q[i]; xp[2 * i] = t.x; xp[2 * i + 1] = t.y; }
}

// out[b] += sum_t w[t] * x[b + DELTA + t], b = 0..15; x = 32 floats as 16 aligned pairs; acc = (lo, hi) halves.
// Split by the parity of b + DELTA so that one weight packing is in use while the other one is being fetched.
template <int NT, int NB, int DELTA>
__device__ __forceinline__ void lat_conv_even(const u64 (&xp)[(NB + NT) / 2], const u64 (&we)[NT / 2], u64 (&acc)[NB]) {
    #pragma unroll
    for (int m = 0; m < NT / 2; ++m) {
        #pragma unroll
        for (int b = DELTA; b < NB; b += 2) acc[b] = lat_fma2(xp[(b + DELTA) / 2 + m], we[m], acc[b]);
    }
}
template <int NT, int NB, int DELTA>
__device__ __forceinline__ void lat_conv_odd(const u64 (&xp)[(NB + NT) / 2], const u64 (&wo)[NT / 2 + 2], u64 (&acc)[NB]) {
    #pragma unroll
    for (int m = 0; m <= NT / 2; ++m) {
        #pragma unroll
        for (int b = 1 - DELTA; b < NB; b += 2) acc[b] = lat_fma2(xp[(b + DELTA - 1) / 2 + m], wo[m], acc[b]);
    }
}
template <int NT>
__device__ __forceinline__ void lat_load_we(const u64* row, u64 (&we)[NT / 2]) {
    const ulonglong2* q = reinterpret_cast<const ulonglong2*>(row);
    #pragma unroll
    for (int i = 0; i < NT / 4; ++i) { const ulonglong2 t = q[i]; we[2 * i] = t.x; we[2 * i + 1] = t.y; }
}
template <int NT>
__device__ __forceinline__ void lat_load_wo(const u64* row, u64 (&wo)[NT / 2 + 2]) {
    const ulonglong2* q = reinterpret_cast<const ulonglong2*>(row + NT / 2);
    #pragma unroll
    for (int i = 0; i < NT / 4 + 1; ++i) { const ulonglong2 t = q[i]; wo[2 * i] = t.x; wo[2 * i + 1] = t.y; }
}

// one parity phase of a 2-D correlation: acc[b] += sum_{tu < ntu} sum_tv w[tu][tv] * x[row tu][b + DELTA + tv];
// software-pipelined: next window and next weight packing are fetched under the FFMA2 of the current ones
template <int NT, int NB, int DELTA>
__device__ __forceinline__ void lat_conv_phase(const float* base, int pitch, const u64 (*wt)[NT + 2], int ntu, u64 (&acc)[NB]) {
    u64 xa[(NB + NT) / 2], xb[(NB + NT) / 2], we[NT / 2], wo[NT / 2 + 2];
    lat_load_window<NT, NB>(base, xa);
    lat_load_we<NT>(wt[0], we);
    #pragma unroll 1
    for (int tu = 0; tu < ntu; tu += 2) {                    // two tap rows per trip: windows alternate xa / xb
        const int t1 = min(tu + 1, ntu - 1), t2 = min(tu + 2, ntu - 1);
        lat_load_wo<NT>(wt[tu], wo);
        lat_load_window<NT, NB>(base + t1 * pitch, xb);
        lat_conv_even<NT, NB, DELTA>(xa, we, acc);
        lat_load_we<NT>(wt[t1], we);
        lat_conv_odd<NT, NB, DELTA>(xa, wo, acc);
        if (tu + 1 < ntu) {                                  // warp-uniform
            lat_load_wo<NT>(wt[t1], wo);
            lat_load_window<NT, NB>(base + t2 * pitch, xa);
            lat_conv_even<NT, NB, DELTA>(xb, we, acc);
            lat_load_we<NT>(wt[t2], we);
            lat_conv_odd<NT, NB, DELTA>(xb, wo, acc);
        } else {
            lat_load_we<NT>(wt[t2], we);
        }
    }
}


template <int NT, int NB> struct LatDims {
    static constexpr int RT = NT - 1;               // padded PSF reach (odd)
    static constexpr int HB = (RT + 1) / 2;         // largest window offset magnitude
    static constexpr int SR = LT_ROWS + NT;         // staged rows per plane
    static constexpr int COLS = LT_WARPS * NB;      // site columns per tile
    static constexpr int SC = COLS + NT;            // staged columns per plane
    static constexpr int SP = (((SC + 4) / 4) & 1) ? SC + 4 : SC + 8;   // pitch: multiple of 4 floats, odd in float4 units
};

// =====================================================================================================
// forward: BL_s[a][b] = B[2a+P, 2b+Q] = sum_{u,v} psf[u][v] * hr0[2a+P+R-u, 2b+Q+R-v]   (hr0: zero outside)
// =====================================================================================================
template <int NT>
struct alignas(16) LatFwdArgs {
    const ulonglong2* tab;              // LatTab<NT>, device: w(tu,tv) = psf[2RT - us - 2tu][2RT - vs - 2tv]
    LatGeom g;
    const float* hr;                    // local buffer, row 0 = HR row hr_lo
    float* bl[kMaxSlots];
};

template <int NT, int NB>
__global__ void __launch_bounds__(LT_WARPS * 32, 3) k_lat_fwd(const __grid_constant__ LatFwdArgs<NT> a) {
    typedef LatDims<NT, NB> D;
    extern __shared__ __align__(16) float sm[];             // [4 phases][SR][SP] + weight table
    const LatGeom& g = a.g;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int s = blockIdx.z;
    u64 (*wsm)[2][NT][NT + 2] = reinterpret_cast<u64 (*)[2][NT][NT + 2]>(sm + 4 * D::SR * D::SP);
    for (int i = tid; i < 2 * NT * (NT + 2); i += LT_WARPS * 32) reinterpret_cast<ulonglong2*>(wsm)[i] = __ldg(a.tab + i);
    const int P = g.slot_pq[s] >> 1, Q = g.slot_pq[s] & 1;
    const int A0 = g.sa_lo + blockIdx.y * LT_ROWS, B0 = blockIdx.x * D::COLS;
    int du0[2], dv0[2];
    #pragma unroll
    for (int p = 0; p < 2; ++p) { du0[p] = -((D::RT + p - P) >> 1); dv0[p] = -((D::RT + p - Q) >> 1); }

    // ---- stage the four parity planes of hr around the tile (zero outside the image / the local rows) ----
    {
        // one warp per hr row, four rows (12 independent 8-byte loads per lane) in flight before the scatter
        const int rbase = 2 * (A0 - D::HB), cbase = B0 - D::HB;
        constexpr int NCP = D::SC + 1, NJ = (NCP + 31) / 32, RB = 4;
        const int o0 = D::HB + dv0[0], o1 = D::HB + dv0[1];             // plane column = cc2 - o
        for (int rq = warp * RB; rq < 2 * D::SR; rq += LT_WARPS * RB) {
            float2 v[RB][NJ];
            #pragma unroll
            for (int i = 0; i < RB; ++i) {
                const int r = rbase + rq + i;
                const bool rowok = r >= 0 && r < g.H && r >= g.hr_lo && r < g.hr_hi;
                const float2* src = reinterpret_cast<const float2*>(a.hr + (size_t)(r - g.hr_lo) * g.W) + cbase;
                #pragma unroll
                for (int j = 0; j < NJ; ++j) {
                    const int cc2 = lane + 32 * j, c = cbase + cc2;
                    v[i][j] = make_float2(0.f, 0.f);
                    if (rowok && cc2 < NCP && c >= 0 && c < g.w) v[i][j] = __ldg(src + cc2);
                }
            }
            #pragma unroll
            for (int i = 0; i < RB; ++i) {
                const int rr2 = rq + i, pr = rr2 & 1;                   // rbase is even
                const int rr = ((rr2 - pr) >> 1) - D::HB - du0[pr];
                if (rr < 0 || rr >= D::SR) continue;                    // warp-uniform
                float* d0 = sm + ((pr * 2 + 0) * D::SR + rr) * D::SP - o0;
                float* d1 = sm + ((pr * 2 + 1) * D::SR + rr) * D::SP - o1;
                #pragma unroll
                for (int j = 0; j < NJ; ++j) {
                    const int cc2 = lane + 32 * j;
                    if (cc2 >= o0 && cc2 < D::SC + o0) d0[cc2] = v[i][j].x;
                    if (cc2 >= o1 && cc2 < D::SC + o1) d1[cc2] = v[i][j].y;
                }
            }
        }
    }
    __syncthreads();

    u64 acc[NB];
    #pragma unroll
    for (int b = 0; b < NB; ++b) acc[b] = 0ull;
    #pragma unroll 1
    for (int ph = 0; ph < 4; ++ph) {
        const int pr = ph >> 1, pc = ph & 1;
        const int us = (P + D::RT - pr) & 1, vs = (Q + D::RT - pc) & 1;
        const float* base = sm + (ph * D::SR + lane) * D::SP + NB * warp;
        const int ntu = NT - us;                             // the odd tap set has NT-1 rows
        lat_conv_phase<NT, NB, 0>(base, D::SP, wsm[us][vs], ntu, acc);
    }
    const int ar = A0 + lane, bc = B0 + NB * warp;
    if (ar < g.sa_hi && bc < g.w) {
        float* d = a.bl[s] + (size_t)(ar - g.sa_lo) * g.pitch + bc;
        #pragma unroll
        for (int q = 0; q < NB / 4; ++q) {
            if (bc + 4 * q < g.pitch)                        // pitch is a multiple of 4 >= w: whole float4s
                *reinterpret_cast<float4*>(d + 4 * q) = make_float4(lat_sum2(acc[4 * q]), lat_sum2(acc[4 * q + 1]),
                                                                    lat_sum2(acc[4 * q + 2]), lat_sum2(acc[4 * q + 3]));
        }
    }
}

// =====================================================================================================
// B on HR row 0 and HR column 0 (the clamp targets of shift(mode='nearest')): one warp per point
// =====================================================================================================
struct LatEdgeBArgs {
    LatGeom g;
    const float* hr;
    const float* psf;                   // (2RT+1)^2 zero-padded PSF, device
    int RT;
    float* btop;                        // [W]  B[0, x]        (written when the local rows include row 0)
    float* bleft;                       // [H]  B[y, 0]        (written for the local rows)
};

__global__ void __launch_bounds__(256) k_lat_edge_b(const LatEdgeBArgs a) {
    const LatGeom& g = a.g;
    const int lane = threadIdx.x & 31;
    const int wid = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int ntop = (g.hr_lo == 0) ? g.W : 0;
    const int nleft = g.hr_hi - g.hr_lo;
    if (wid >= ntop + nleft) return;
    const int y = wid < ntop ? 0 : g.hr_lo + (wid - ntop);
    const int x = wid < ntop ? wid : 0;
    const int K = 2 * a.RT + 1;
    float v = 0.f;
    #pragma unroll 8
    for (int t = lane; t < K * K; t += 32) {
        const int u = t / K, vv = t - u * K;
        const int r = y + a.RT - u, c = x + a.RT - vv;
        if (r >= 0 && r < g.H && r >= g.hr_lo && r < g.hr_hi && c >= 0 && c < g.W)
            v = fmaf(__ldg(a.psf + t), __ldg(a.hr + (size_t)(r - g.hr_lo) * g.W + c), v);
    }
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) { if (wid < ntop) a.btop[x] = v; else a.bleft[y] = v; }
}

// =====================================================================================================
// residuals: A_s = sum_{k in s} (lr_k - BL_s) on the sites, clamp duplicates on HR row 0 / column 0,
// squared-error sums (f64) over the owned LR rows
// =====================================================================================================
struct LatResidArgs {
    LatGeom g;
    int n;
    int sy[kMaxFrames], sx[kMaxFrames], slot_of[kMaxFrames];
    const float* lr[kMaxFrames];        // local buffers, row 0 = LR row lr_lo
    int lr_lo, lr_hi, own_i0, own_i1;
    const float* bl[kMaxSlots];
    float* A[kMaxSlots];
    const float* btop; const float* bleft;
    float* arow;                        // [W] clamp duplicates on HR row 0
    float* acol;                        // [H] clamp duplicates on HR column 0 ((0,0): only frames with sy != -1)
    int rows_per_cta, nsite_cta;        // part A: CTAs per slot = ceil(rows / rows_per_cta)
    int nrow_cta, ncol_cta;             // parts B/C: CTAs of 256 columns along row 0 / 256 rows along column 0
    int first_cta;                      // blockIdx offset: nA when part A ran in k_lat_resid_s1 (only parts B/C are launched)
    double* partial;                    // one slot per CTA
};

__device__ __forceinline__ float lat_bval(const LatResidArgs& a, int k, int y, int x) {
    const LatGeom& g = a.g;
    if (y == 0 && (a.sy[k] & 1)) return __ldg(a.btop + x);       // off-lattice row 0
    if (x == 0 && (a.sx[k] & 1)) return __ldg(a.bleft + y);      // off-lattice column 0
    const int s = a.slot_of[k];
    const int P = g.slot_pq[s] >> 1, Q = g.slot_pq[s] & 1;
    const int ar = min(max((y - P) >> 1, g.sa_lo), g.sa_hi - 1);   // rows outside the planes only feed unused outputs
    return __ldg(a.bl[s] + (size_t)(ar - g.sa_lo) * g.pitch + ((x - Q) >> 1));
}

__global__ void __launch_bounds__(256) k_lat_resid(const LatResidArgs a) {
    const LatGeom& g = a.g;
    const int tid = threadIdx.x;
    double e2 = 0.0;
    const int nA = a.nsite_cta * g.nslot;
    const int bid = (int)blockIdx.x + a.first_cta;
    if (bid < nA) {
        // ---- part A: regular sites ------------------------------------------------------------------
        const int s = bid / a.nsite_cta, cb = bid - s * a.nsite_cta;
        const int P = g.slot_pq[s] >> 1, Q = g.slot_pq[s] & 1;
        const int r0 = g.sa_lo + cb * a.rows_per_cta, r1 = min(r0 + a.rows_per_cta, g.sa_hi);
        for (int ar = r0; ar < r1; ++ar) {
            const float* blrow = a.bl[s] + (size_t)(ar - g.sa_lo) * g.pitch;
            float* arow_out = a.A[s] + (size_t)(ar - g.sa_lo) * g.pitch;
            // per-frame row pointers (null: the frame has no sample on this site row); unrolled -> registers
            const float* rowp[kMaxFrames]; int jmax[kMaxFrames]; bool own[kMaxFrames];
            #pragma unroll
            for (int k = 0; k < kMaxFrames; ++k) {
                rowp[k] = nullptr; jmax[k] = 0; own[k] = false;
                if (k < a.n && a.slot_of[k] == s) {
                    const int i = ar + ((P + a.sy[k]) >> 1), ox = (Q + a.sx[k]) >> 1;
                    if (i < g.h && i >= a.lr_lo && i < a.lr_hi) {
                        rowp[k] = a.lr[k] + (size_t)(i - a.lr_lo) * g.w + ox;
                        jmax[k] = g.w - ox;                              // site columns b < jmax have a sample
                        own[k] = i >= a.own_i0 && i < a.own_i1;
                    }
                }
            }
            #pragma unroll 2
            for (int b = tid; b < g.w; b += 256) {
                const float bl = __ldg(blrow + b);
                float v[kMaxFrames];
                #pragma unroll
                for (int k = 0; k < kMaxFrames; ++k) v[k] = (rowp[k] && b < jmax[k]) ? __ldg(rowp[k] + b) : bl;   // all loads first
                float acc = 0.f, q = 0.f;
                #pragma unroll
                for (int k = 0; k < kMaxFrames; ++k) {
                    const float e = v[k] - bl;                           // exactly 0 where there is no sample
                    acc += e;
                    if (own[k]) q = fmaf(e, e, q);
                }
                arow_out[b] = acc;
                e2 += (double)q;
            }
        }
    } else if (bid < nA + a.nrow_cta) {
        const int c0 = (bid - nA) * 256;                       // this CTA's 256 columns
        // ---- parts B/C along HR row 0 / LR row 0 (only where the local rows include them) -----------------
        if (g.hr_lo == 0) {
            for (int x = c0 + tid; x < min(g.W, c0 + 256); x += 256) {
                float acc = 0.f;
                for (int k = 0; k < a.n; ++k) {
                    if (a.sy[k] != -1) continue;
                    const int X = min(max(x + a.sx[k], 0), g.W - 1);
                    if (X & 1) continue;
                    const int j = X >> 1;
                    const int bx = min(max(2 * j - a.sx[k], 0), g.W - 1);
                    acc += __ldg(a.lr[k] + j) - lat_bval(a, k, 1, bx);        // err_k[0][j]; LR row 0 is local row 0
                }
                a.arow[x] = acc;
            }
            if (a.own_i0 == 0) {
                for (int j = c0 + tid; j < min(g.w, c0 + 256); j += 256)
                    for (int k = 0; k < a.n; ++k) {
                        if (a.sy[k] != +1) continue;                          // i = 0 samples the clamped row 0
                        const int bx = min(max(2 * j - a.sx[k], 0), g.W - 1);
                        const float e = __ldg(a.lr[k] + j) - lat_bval(a, k, 0, bx);
                        e2 += (double)(e * e);
                    }
            }
        }
    } else {
        // ---- parts B/C along HR column 0 / LR column 0 ----------------------------------------------------------
        const int r0 = (bid - nA - a.nrow_cta) * 256;           // this CTA's 256 rows
        for (int y = g.hr_lo + r0 + tid; y < min(g.hr_hi, g.hr_lo + r0 + 256); y += 256) {
            float acc = 0.f;
            for (int k = 0; k < a.n; ++k) {
                if (a.sx[k] != -1) continue;
                if (y == 0 && a.sy[k] == -1) continue;                        // that corner term lives in arow[0]
                const int Y = y + a.sy[k];
                if (Y < 0 || Y > g.H - 1 || (Y & 1)) continue;
                const int i = Y >> 1;
                if (i < a.lr_lo || i >= a.lr_hi) continue;
                const int by = min(max(2 * i - a.sy[k], 0), g.H - 1);
                acc += __ldg(a.lr[k] + (size_t)(i - a.lr_lo) * g.w) - lat_bval(a, k, by, 1);
            }
            a.acol[y] = acc;
        }
        for (int i = a.lr_lo + r0 + tid; i < min(a.lr_hi, a.lr_lo + r0 + 256); i += 256) {
            if (i < a.own_i0 || i >= a.own_i1) continue;
            for (int k = 0; k < a.n; ++k) {
                if (a.sx[k] != +1) continue;                                  // j = 0 samples the clamped column 0
                if (i == 0 && a.sy[k] == +1) continue;                        // counted with row 0 above
                const int by = min(max(2 * i - a.sy[k], 0), g.H - 1);
                const float e = __ldg(a.lr[k] + (size_t)(i - a.lr_lo) * g.w) - lat_bval(a, k, by, 0);
                e2 += (double)(e * e);
            }
        }
    }
    __shared__ double red[8];
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) e2 += __shfl_xor_sync(0xffffffffu, e2, o);
    if ((tid & 31) == 0) red[tid >> 5] = e2;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        #pragma unroll
        for (int i = 0; i < 8; ++i) t += red[i];
        a.partial[bid] = t;
    }
}

// part A from pre-summed planes: A_s = S1_s - m B_s, squared error (S1 - m B)^2 / m over the site rows this launch owns (the
// within-site variance V of the frames is constant over the iterations: k_lattice_sums, added once by the caller).
// One CTA per site row and slot, as part A of k_lat_resid (same partial-sum slots).
struct LatResidS1Args {
    LatGeom g;
    int n;
    int sy[kMaxFrames], sx[kMaxFrames], slot_of[kMaxFrames];
    int lr_lo, lr_hi, own_i0, own_i1;
    const float* bl[kMaxSlots];
    float* A[kMaxSlots];
    const float* s1[kMaxSlots];         // by lattice slot
    int s1_row0, s1_rows;
    int nsite_cta;
    double* partial;
};

__global__ void __launch_bounds__(256) k_lat_resid_s1(const __grid_constant__ LatResidS1Args a) {
    const LatGeom& g = a.g;
    const int tid = threadIdx.x;
    const int s = blockIdx.x / a.nsite_cta, ar = g.sa_lo + (blockIdx.x - s * a.nsite_cta);
    const int P = g.slot_pq[s] >> 1, Q = g.slot_pq[s] & 1;
    int m_all = 0, m_last = 0;                          // frames with a sample at a site of this row: columns < w-1 / column w-1
    #pragma unroll
    for (int k = 0; k < kMaxFrames; ++k) {
        if (k < a.n && a.slot_of[k] == s) {
            const int i = ar + ((P + a.sy[k]) >> 1), ox = (Q + a.sx[k]) >> 1;
            if (i < g.h && i >= a.lr_lo && i < a.lr_hi) { ++m_all; if (ox == 0) ++m_last; }
        }
    }
    const bool own = ar >= a.own_i0 && ar < a.own_i1;
    const bool have = ar >= a.s1_row0 && ar < a.s1_row0 + a.s1_rows;
    const float* blrow = a.bl[s] + (size_t)(ar - g.sa_lo) * g.pitch;
    const float* srow = a.s1[s] + (size_t)(have ? ar - a.s1_row0 : 0) * g.pitch;
    float* out = a.A[s] + (size_t)(ar - g.sa_lo) * g.pitch;
    const float mf = have ? (float)m_all : 0.0f, ml = have ? (float)m_last : 0.0f;
    const float wf = (own && mf > 0.0f) ? 1.0f / mf : 0.0f, wl = (own && ml > 0.0f) ? 1.0f / ml : 0.0f;
    double e2 = 0.0;
    for (int b4 = tid; 4 * b4 < g.w; b4 += 256) {                          // pitch is a multiple of 4: whole float4s
        const float4 bl = __ldg(reinterpret_cast<const float4*>(blrow) + b4);
        const float4 sv = have ? __ldg(reinterpret_cast<const float4*>(srow) + b4) : make_float4(0.f, 0.f, 0.f, 0.f);
        const int b = 4 * b4;
        float4 r; float q = 0.0f;
        {
            const bool l0 = b == g.w - 1, l1 = b + 1 == g.w - 1, l2 = b + 2 == g.w - 1, l3 = b + 3 == g.w - 1;
            r.x = fmaf(-(l0 ? ml : mf), bl.x, sv.x); q = fmaf(r.x * (l0 ? wl : wf), r.x, q);
            r.y = (b + 1 < g.w) ? fmaf(-(l1 ? ml : mf), bl.y, sv.y) : 0.0f; q = fmaf(r.y * (l1 ? wl : wf), r.y, q);
            r.z = (b + 2 < g.w) ? fmaf(-(l2 ? ml : mf), bl.z, sv.z) : 0.0f; q = fmaf(r.z * (l2 ? wl : wf), r.z, q);
            r.w = (b + 3 < g.w) ? fmaf(-(l3 ? ml : mf), bl.w, sv.w) : 0.0f; q = fmaf(r.w * (l3 ? wl : wf), r.w, q);
        }
        *(reinterpret_cast<float4*>(out) + b4) = r;
        e2 += (double)q;
    }
    __shared__ double red[8];
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) e2 += __shfl_xor_sync(0xffffffffu, e2, o);
    if ((tid & 31) == 0) red[tid >> 5] = e2;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        #pragma unroll
        for (int i = 0; i < 8; ++i) t += red[i];
        a.partial[blockIdx.x] = t;
    }
}

// =====================================================================================================
// backward: hr'[Y,X] = clip(hr[Y,X] + alpha * sum_{u,v} psf[u][v] * acc[Y-R+u, X-R+v])
// acc = A_s on the sites of every slot + arow on HR row 0 + acol on HR column 0
// =====================================================================================================
template <int NT>
struct alignas(16) LatBwdArgs {
    const ulonglong2* tab;              // LatTab<NT>, device: w(tu,tv) = psf[2tu + us][2tv + vs]
    LatGeom g;
    const float* A[kMaxSlots];
    const float* hr_in; float* hr_out;  // local buffers, row 0 = HR row hr_lo
    int m_lo, m_hi;                     // HR row pairs to update: rows [2 m_lo, 2 m_hi)
    float alpha, lo, hi;
    const float* arow; const float* acol; const float* psf;   // clamp terms (generic path near row 0 / column 0)
    int has_row0, has_col0;
};

struct LatEdgeRefs { const float* arow; const float* acol; const float* psf; int W, y_lo, y_hi, has_row0, has_col0; };

template <int NT>
__device__ __noinline__ float lat_edge_terms(LatEdgeRefs e, int Y, int X) {
    constexpr int RT = NT - 1, K = 2 * RT + 1;
    float v = 0.f;
    if (e.has_row0 && Y <= RT) {
        const float* pr = e.psf + (RT - Y) * K;
        for (int t = 0; t < K; ++t) {
            const int x = X - RT + t;
            if (x >= 0 && x < e.W) v = fmaf(__ldg(pr + t), __ldg(e.arow + x), v);
        }
    }
    if (e.has_col0 && X <= RT) {
        for (int t = 0; t < K; ++t) {
            const int y = Y - RT + t;
            if (y >= e.y_lo && y < e.y_hi) v = fmaf(__ldg(e.psf + t * K + (RT - X)), __ldg(e.acol + y), v);
        }
    }
    return v;
}

template <int NT, int NB>
__global__ void __launch_bounds__(LT_WARPS * 32, 4) k_lat_bwd(const __grid_constant__ LatBwdArgs<NT> a) {
    typedef LatDims<NT, NB> D;
    extern __shared__ __align__(16) float sm[];             // [nslot][SR][SP] + weight table
    const LatGeom& g = a.g;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    u64 (*wsm)[2][NT][NT + 2] = reinterpret_cast<u64 (*)[2][NT][NT + 2]>(sm + g.nslot * D::SR * D::SP);
    for (int i = tid; i < 2 * NT * (NT + 2); i += LT_WARPS * 32) reinterpret_cast<ulonglong2*>(wsm)[i] = __ldg(a.tab + i);
    const int M0 = a.m_lo + blockIdx.y * LT_ROWS, N0 = blockIdx.x * D::COLS;

    // ---- stage the residual planes around the tile: rows [M0-HB, M0-HB+SR), cols [N0-HB, N0-HB+SC) -------
    for (int s = 0; s < g.nslot; ++s) {
        const float* src = a.A[s];
        float* dst = sm + s * D::SR * D::SP;
        constexpr int NC4 = D::SP / 4, NEL = D::SR * NC4, NTH = LT_WARPS * 32, BATCH = (NEL + NTH - 1) / NTH;
        float4 v[BATCH];
        #pragma unroll
        for (int k = 0; k < BATCH; ++k) {                            // all loads first (independent), then the stores
            const int e = tid + k * NTH;
            const int rr = e / NC4, c4 = e - rr * NC4;
            const int ar = M0 - D::HB + rr, b = N0 - D::HB + 4 * c4;     // b is a multiple of 4, pitch too
            v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (e < NEL && ar >= g.sa_lo && ar < g.sa_hi && b >= 0 && b < g.w) {
                v[k] = __ldg(reinterpret_cast<const float4*>(src + (size_t)(ar - g.sa_lo) * g.pitch + b));
                if (b + 1 >= g.w) v[k].y = 0.f;
                if (b + 2 >= g.w) v[k].z = 0.f;
                if (b + 3 >= g.w) v[k].w = 0.f;
            }
        }
        #pragma unroll
        for (int k = 0; k < BATCH; ++k) {
            const int e = tid + k * NTH;
            if (e < NEL) *reinterpret_cast<float4*>(dst + 4 * e) = v[k];   // rr * SP + 4 * c4 == 4 * e
        }
    }
    __syncthreads();

    const int m = M0 + lane;
    const int n0 = N0 + NB * warp;
    #pragma unroll 1
    for (int py = 0; py < 2; ++py) {
        float v0[NB];
        #pragma unroll 1
        for (int px = 0; px < 2; ++px) {
            u64 acc[NB];
            #pragma unroll
            for (int b = 0; b < NB; ++b) acc[b] = 0ull;
            for (int s = 0; s < g.nslot; ++s) {
                const int P = g.slot_pq[s] >> 1, Q = g.slot_pq[s] & 1;
                const int qy = D::RT + P - py, qx = D::RT + Q - px;
                const int us = qy & 1, vs = qx & 1;
                const int ro = D::HB - (qy >> 1), delta = D::HB - (qx >> 1);     // 0 or 1
                const float* base = sm + (s * D::SR + lane + ro) * D::SP + NB * warp;
                const int ntu = NT - us;
                if (delta == 0) lat_conv_phase<NT, NB, 0>(base, D::SP, wsm[us][vs], ntu, acc);
                else            lat_conv_phase<NT, NB, 1>(base, D::SP, wsm[us][vs], ntu, acc);
            }
            if (px == 0) {
                #pragma unroll
                for (int b = 0; b < NB; ++b) v0[b] = lat_sum2(acc[b]);
            } else if (m < a.m_hi && n0 < g.w) {
                // ---- epilogue of HR row Y = 2m + py: 32 consecutive pixels per lane ---------------------------
                const int Y = 2 * m + py, X0 = 2 * n0;
                const float* hin = a.hr_in + (size_t)(Y - g.hr_lo) * g.W + X0;
                float* hout = a.hr_out + (size_t)(Y - g.hr_lo) * g.W + X0;
                const bool edge = (a.has_row0 && Y <= D::RT) || (a.has_col0 && X0 <= D::RT);
                #pragma unroll
                for (int q = 0; q < NB / 2; ++q) {
                    if (X0 + 4 * q >= g.W) break;
                    const float4 hv = __ldg(reinterpret_cast<const float4*>(hin) + q);
                    float c0 = v0[2 * q], c1 = lat_sum2(acc[2 * q]), c2 = v0[2 * q + 1], c3 = lat_sum2(acc[2 * q + 1]);
                    if (edge) {
                        const LatEdgeRefs er{a.arow, a.acol, a.psf, g.W, max(g.hr_lo, 0), min(g.hr_hi, g.H), a.has_row0, a.has_col0};
                        c0 += lat_edge_terms<NT>(er, Y, X0 + 4 * q);     c1 += lat_edge_terms<NT>(er, Y, X0 + 4 * q + 1);
                        c2 += lat_edge_terms<NT>(er, Y, X0 + 4 * q + 2); c3 += lat_edge_terms<NT>(er, Y, X0 + 4 * q + 3);
                    }
                    float4 r;
                    r.x = fminf(fmaxf(fmaf(a.alpha, c0, hv.x), a.lo), a.hi);
                    r.y = fminf(fmaxf(fmaf(a.alpha, c1, hv.y), a.lo), a.hi);
                    r.z = fminf(fmaxf(fmaf(a.alpha, c2, hv.z), a.lo), a.hi);
                    r.w = fminf(fmaxf(fmaf(a.alpha, c3, hv.w), a.lo), a.hi);
                    *(reinterpret_cast<float4*>(hout) + q) = r;
                }
            }
        }
    }
}

// =====================================================================================================
// host side
// =====================================================================================================
static inline u64 pack_pair(float lo, float hi) {
    uint32_t a, b;
    memcpy(&a, &lo, 4); memcpy(&b, &hi, 4);
    return (u64)a | ((u64)b << 32);
}

template <int NT>
static void build_tab(const std::vector<float>& psfP, bool forward, LatTab<NT>& t) {
    constexpr int RT = NT - 1, K = 2 * RT + 1;
    for (int us = 0; us < 2; ++us)
        for (int vs = 0; vs < 2; ++vs)
            for (int tu = 0; tu < NT; ++tu) {
                const int u = forward ? 2 * RT - us - 2 * tu : 2 * tu + us;
                float w[NT + 2];                                   // w[-1] .. w[NT]
                w[0] = 0.f; w[NT + 1] = 0.f;
                for (int tv = 0; tv < NT; ++tv) {
                    const int v = forward ? 2 * RT - vs - 2 * tv : 2 * tv + vs;
                    w[tv + 1] = (u >= 0 && u < K && v >= 0 && v < K) ? psfP[(size_t)u * K + v] : 0.f;
                }
                for (int m = 0; m < NT / 2; ++m)  t.w[us][vs][tu][m] = pack_pair(w[2 * m + 1], w[2 * m + 2]);
                for (int m = 0; m <= NT / 2; ++m) t.w[us][vs][tu][NT / 2 + m] = pack_pair(w[2 * m], w[2 * m + 1]);
                t.w[us][vs][tu][NT + 1] = 0ull;
            }
}

struct LatPlan {
    int NT = 16, RT = 15;
    std::vector<float> psfP;            // (2RT+1)^2, zero padded, centred
    LatTab<16> f16, b16;
    LatTab<8>  f8, b8;
    float* d_psf = nullptr;
    void *d_tabf = nullptr, *d_tabb = nullptr;   // device copies of the forward / backward tables
    // planes (allocated for one geometry)
    int rows = 0, pitch = 0, nslot = 0, H = 0, W = 0;
    float* d_bl[kMaxSlots] = {nullptr, nullptr, nullptr, nullptr};
    float* d_A[kMaxSlots] = {nullptr, nullptr, nullptr, nullptr};
    float *d_btop = nullptr, *d_bleft = nullptr, *d_arow = nullptr, *d_acol = nullptr;
    int64_t bytes = 0;
    bool attr_done[2] = {false, false};   // per tile width (NB = 8 / 16)
};

bool lat_supported(int kh, int kw, int smax, int W) {
    return (kh & 1) && (kw & 1) && kh / 2 <= 15 && kw / 2 <= 15 && smax <= 1 && (W % 4) == 0;
}

void lat_destroy(LatPlan* L);

int lat_create(LatPlan** out, const double* psf, int kh, int kw) {
    LatPlan* L = new (std::nothrow) LatPlan();
    if (!L) { set_error("out of host memory"); return SR459_ENOMEM; }
    const int rh = kh / 2, rw = kw / 2, reach = std::max(rh, rw);
    L->NT = reach <= 7 ? 8 : 16; L->RT = L->NT - 1;
    const int K = 2 * L->RT + 1;
    L->psfP.assign((size_t)K * K, 0.f);
    for (int u = 0; u < kh; ++u)
        for (int v = 0; v < kw; ++v) L->psfP[(size_t)(u + L->RT - rh) * K + (v + L->RT - rw)] = (float)psf[(size_t)u * kw + v];
    if (L->NT == 16) { build_tab<16>(L->psfP, true, L->f16); build_tab<16>(L->psfP, false, L->b16); }
    else             { build_tab<8>(L->psfP, true, L->f8);   build_tab<8>(L->psfP, false, L->b8); }
    if (gmalloc((void**)&L->d_psf, sizeof(float) * K * K) != cudaSuccess ||
        cudaMemcpy(L->d_psf, L->psfP.data(), sizeof(float) * K * K, cudaMemcpyHostToDevice) != cudaSuccess) {
        set_error("lat_create: device allocation failed"); delete L; return SR459_ECUDA;
    }
    L->bytes += (int64_t)sizeof(float) * K * K;
    const void* hf = L->NT == 16 ? (const void*)&L->f16 : (const void*)&L->f8;
    const void* hb = L->NT == 16 ? (const void*)&L->b16 : (const void*)&L->b8;
    const size_t tb = L->NT == 16 ? sizeof(LatTab<16>) : sizeof(LatTab<8>);
    if (gmalloc(&L->d_tabf, tb) != cudaSuccess || gmalloc(&L->d_tabb, tb) != cudaSuccess ||
        cudaMemcpy(L->d_tabf, hf, tb, cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMemcpy(L->d_tabb, hb, tb, cudaMemcpyHostToDevice) != cudaSuccess) {
        set_error("lat_create: device allocation failed"); lat_destroy(L); return SR459_ECUDA;
    }
    L->bytes += 2 * (int64_t)tb;
    *out = L;
    return SR459_OK;
}

static void lat_free_planes(LatPlan* L) {
    for (int s = 0; s < kMaxSlots; ++s) {
        if (L->d_bl[s]) gfree(L->d_bl[s]);
        if (L->d_A[s]) gfree(L->d_A[s]);
        L->d_bl[s] = L->d_A[s] = nullptr;
    }
    for (float** p : {&L->d_btop, &L->d_bleft, &L->d_arow, &L->d_acol}) { if (*p) gfree(*p); *p = nullptr; }
    L->rows = 0;
}

void lat_destroy(LatPlan* L) {
    if (!L) return;
    lat_free_planes(L);
    if (L->d_psf) gfree(L->d_psf);
    if (L->d_tabf) gfree(L->d_tabf);
    if (L->d_tabb) gfree(L->d_tabb);
    delete L;
}

int64_t lat_scratch_bytes(const LatPlan* L) { return L ? L->bytes : 0; }

struct LatSetup { LatGeom g; int slot_of[kMaxFrames]; int has_row0, has_col0; int nsite_cta, rows_per_cta, nrow_cta, ncol_cta; };

static void lat_setup(const LatPlan* L, const LatCall& c, LatSetup& S) {
    LatGeom& g = S.g;
    g.H = c.H; g.W = c.W; g.h = c.h; g.w = c.w; g.hr_lo = c.hr_lo; g.hr_hi = c.hr_hi;
    g.sa_lo = std::max(0, (c.upd_y0 - L->RT - 1) / 2 - 1);
    g.sa_hi = std::min(c.h, (c.upd_y1 + L->RT) / 2 + 1);
    g.pitch = (c.w + 3) & ~3;
    g.nslot = 0;
    for (int k = 0; k < c.n; ++k) {
        const int pq = ((c.sy[k] & 1) << 1) | (c.sx[k] & 1);
        int s = -1;
        for (int t = 0; t < g.nslot; ++t) if (g.slot_pq[t] == pq) s = t;
        if (s < 0) { s = g.nslot++; g.slot_pq[s] = pq; }
        S.slot_of[k] = s;
    }
    S.has_row0 = 0; S.has_col0 = 0;
    for (int k = 0; k < c.n; ++k) { if (c.sy[k] == -1 && c.hr_lo == 0) S.has_row0 = 1; if (c.sx[k] == -1) S.has_col0 = 1; }
    S.rows_per_cta = 1;
    S.nsite_cta = (g.sa_hi - g.sa_lo + S.rows_per_cta - 1) / S.rows_per_cta;
    S.nrow_cta = (g.W + 255) / 256;
    S.ncol_cta = (g.hr_hi - g.hr_lo + 255) / 256;
}

int lat_per_iter(const LatPlan* L, const LatCall& c) {
    LatSetup S; lat_setup(L, c, S);
    return S.nsite_cta * S.g.nslot + S.nrow_cta + S.ncol_cta;
}

static int lat_ensure_planes(LatPlan* L, const LatGeom& g) {
    const int rows = g.sa_hi - g.sa_lo;
    if (L->rows >= rows && L->pitch == g.pitch && L->nslot >= g.nslot && L->H == g.H && L->W == g.W) return SR459_OK;
    lat_free_planes(L);
    const size_t cnt = (size_t)rows * g.pitch;
    for (int s = 0; s < g.nslot; ++s) {
        SR_CUDA(gmalloc((void**)&L->d_bl[s], cnt * sizeof(float)));
        SR_CUDA(gmalloc((void**)&L->d_A[s], cnt * sizeof(float)));
        L->bytes += 2 * (int64_t)cnt * 4;
    }
    SR_CUDA(gmalloc((void**)&L->d_btop, sizeof(float) * g.W));
    SR_CUDA(gmalloc((void**)&L->d_arow, sizeof(float) * g.W));
    SR_CUDA(gmalloc((void**)&L->d_bleft, sizeof(float) * g.H));
    SR_CUDA(gmalloc((void**)&L->d_acol, sizeof(float) * g.H));
    SR_CUDA(cudaMemset(L->d_btop, 0, sizeof(float) * g.W));
    SR_CUDA(cudaMemset(L->d_arow, 0, sizeof(float) * g.W));
    SR_CUDA(cudaMemset(L->d_bleft, 0, sizeof(float) * g.H));
    SR_CUDA(cudaMemset(L->d_acol, 0, sizeof(float) * g.H));
    L->bytes += 8 * (int64_t)(g.W + g.H);
    L->rows = rows; L->pitch = g.pitch; L->nslot = g.nslot; L->H = g.H; L->W = g.W;
    return SR459_OK;
}

template <int NT, int NB>
static int lat_iterate_t(LatPlan* L, const LatCall& c, const LatSetup& S, const LatTab<NT>& ft, const LatTab<NT>& bt, cudaStream_t st) {
    typedef LatDims<NT, NB> D;
    const LatGeom& g = S.g;
    const size_t smem_f = sizeof(float) * 4 * D::SR * D::SP + sizeof(LatTab<NT>), smem_b = sizeof(float) * g.nslot * D::SR * D::SP + sizeof(LatTab<NT>);
    if (!L->attr_done[NB == 16]) {
        SR_CUDA(cudaFuncSetAttribute(k_lat_fwd<NT, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_f));
        SR_CUDA(cudaFuncSetAttribute(k_lat_bwd<NT, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(float) * kMaxSlots * D::SR * D::SP + sizeof(LatTab<NT>))));
        L->attr_done[NB == 16] = true;
    }
    const int tiles_x = (g.w + D::COLS - 1) / D::COLS;
    if (c.side != nullptr && c.ev_fork != nullptr && c.ev_join != nullptr) {
        SR_CUDA(cudaEventRecord(c.ev_fork, st)); SR_CUDA(cudaStreamWaitEvent(c.side, c.ev_fork, 0));
    }
    {   // forward
        LatFwdArgs<NT> a;
        a.tab = reinterpret_cast<const ulonglong2*>(L->d_tabf); a.g = g; a.hr = c.hr_in;
        for (int s = 0; s < kMaxSlots; ++s) a.bl[s] = L->d_bl[s];
        dim3 grid(tiles_x, (g.sa_hi - g.sa_lo + LT_ROWS - 1) / LT_ROWS, g.nslot);
        k_lat_fwd<NT, NB><<<grid, LT_WARPS * 32, smem_f, st>>>(a);
        SR_LAUNCH_CHECK("k_lat_fwd");
    }
    const bool forked = c.side != nullptr && c.ev_fork != nullptr && c.ev_join != nullptr;
    {   // B on row 0 / column 0: independent of k_lat_fwd (both only read hr) -> on the side stream beside it when there is one
        LatEdgeBArgs a{g, c.hr_in, L->d_psf, L->RT, L->d_btop, L->d_bleft};
        const int npts = (g.hr_lo == 0 ? g.W : 0) + (g.hr_hi - g.hr_lo);
        k_lat_edge_b<<<(npts + 7) / 8, 256, 0, forked ? c.side : st>>>(a);
        SR_LAUNCH_CHECK("k_lat_edge_b");
    }
    bool joined = !forked;
    {   // residuals
        LatResidArgs a{};
        a.g = g; a.n = c.n;
        for (int k = 0; k < c.n; ++k) { a.sy[k] = c.sy[k]; a.sx[k] = c.sx[k]; a.slot_of[k] = S.slot_of[k]; a.lr[k] = c.lr[k]; }
        a.lr_lo = c.lr_lo; a.lr_hi = c.lr_hi; a.own_i0 = c.own_i0; a.own_i1 = c.own_i1;
        for (int s = 0; s < kMaxSlots; ++s) { a.bl[s] = L->d_bl[s]; a.A[s] = L->d_A[s]; }
        a.btop = L->d_btop; a.bleft = L->d_bleft; a.arow = L->d_arow; a.acol = L->d_acol;
        a.rows_per_cta = S.rows_per_cta; a.nsite_cta = S.nsite_cta; a.nrow_cta = S.nrow_cta; a.ncol_cta = S.ncol_cta;
        a.partial = c.partial;
        const int nA = S.nsite_cta * g.nslot;
        // pre-summed planes for every lattice of this call: part A is one elementwise pass (k_lat_resid_s1)
        const float* s1_of[kMaxSlots] = {nullptr, nullptr, nullptr, nullptr};
        bool have_s1 = c.s1_n > 0 && S.rows_per_cta == 1;
        for (int t = 0; t < g.nslot && have_s1; ++t) {
            for (int q = 0; q < c.s1_n; ++q) if (c.s1_pq[q] == g.slot_pq[t]) s1_of[t] = c.s1[q];
            if (!s1_of[t]) have_s1 = false;
        }
        if (have_s1) {
            LatResidS1Args r{};
            r.g = g; r.n = c.n;
            for (int k = 0; k < c.n; ++k) { r.sy[k] = c.sy[k]; r.sx[k] = c.sx[k]; r.slot_of[k] = S.slot_of[k]; }
            r.lr_lo = c.lr_lo; r.lr_hi = c.lr_hi; r.own_i0 = c.own_i0; r.own_i1 = c.own_i1;
            for (int s = 0; s < kMaxSlots; ++s) { r.bl[s] = L->d_bl[s]; r.A[s] = L->d_A[s]; r.s1[s] = s1_of[s]; }
            r.s1_row0 = c.s1_row0; r.s1_rows = c.s1_rows; r.nsite_cta = S.nsite_cta; r.partial = c.partial;
            // the clamp / MSE-only terms of row 0 and column 0 (parts B / C) need k_lat_fwd's planes and k_lat_edge_b's vectors: on the
            // side stream, behind k_lat_edge_b and a second fork after k_lat_fwd, beside the elementwise pass
            cudaStream_t sbc = st;
            if (forked) { SR_CUDA(cudaEventRecord(c.ev_fork, st)); SR_CUDA(cudaStreamWaitEvent(c.side, c.ev_fork, 0)); sbc = c.side; }
            k_lat_resid_s1<<<nA, 256, 0, st>>>(r);
            SR_LAUNCH_CHECK("k_lat_resid_s1");
            a.first_cta = nA;
            k_lat_resid<<<S.nrow_cta + S.ncol_cta, 256, 0, sbc>>>(a);
            SR_LAUNCH_CHECK("k_lat_resid");
            if (forked) { SR_CUDA(cudaEventRecord(c.ev_join, c.side)); SR_CUDA(cudaStreamWaitEvent(st, c.ev_join, 0)); joined = true; }
        } else {
            if (!joined) { SR_CUDA(cudaEventRecord(c.ev_join, c.side)); SR_CUDA(cudaStreamWaitEvent(st, c.ev_join, 0)); joined = true; }
            a.first_cta = 0;
            k_lat_resid<<<nA + S.nrow_cta + S.ncol_cta, 256, 0, st>>>(a);
            SR_LAUNCH_CHECK("k_lat_resid");
        }
    }
    {   // backward + update
        LatBwdArgs<NT> a;
        a.tab = reinterpret_cast<const ulonglong2*>(L->d_tabb); a.g = g;
        for (int s = 0; s < kMaxSlots; ++s) a.A[s] = L->d_A[s];
        a.hr_in = c.hr_in; a.hr_out = c.hr_out;
        a.m_lo = c.upd_y0 / 2; a.m_hi = c.upd_y1 / 2;
        a.alpha = c.alpha; a.lo = c.lo; a.hi = c.hi;
        a.arow = L->d_arow; a.acol = L->d_acol; a.psf = L->d_psf;
        a.has_row0 = S.has_row0; a.has_col0 = S.has_col0;
        dim3 grid(tiles_x, (a.m_hi - a.m_lo + LT_ROWS - 1) / LT_ROWS);
        k_lat_bwd<NT, NB><<<grid, LT_WARPS * 32, smem_b, st>>>(a);
        SR_LAUNCH_CHECK("k_lat_bwd");
    }
    return SR459_OK;
}

int lat_iterate(LatPlan* L, const LatCall& c, cudaStream_t st) {
    LatSetup S; lat_setup(L, c, S);
    SR_TRY(lat_ensure_planes(L, S.g));
    // thin updates (row bands): the narrow tile doubles the CTA count as long as the wide one cannot fill the machine
    int n_sm = 148; { int dev = 0; if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev); }
    const long long wide_tiles = (long long)((S.g.w + 4 * 16 - 1) / (4 * 16)) * (((c.upd_y1 - c.upd_y0) / 2 + LT_ROWS - 1) / LT_ROWS);
    static const int force_nb = getenv("SR459_LAT_NB") ? atoi(getenv("SR459_LAT_NB")) : 0;
    const bool narrow = force_nb ? force_nb == 8 : wide_tiles < 3LL * n_sm;
    if (L->NT == 16) return narrow ? lat_iterate_t<16, 8>(L, c, S, L->f16, L->b16, st) : lat_iterate_t<16, 16>(L, c, S, L->f16, L->b16, st);
    return narrow ? lat_iterate_t<8, 8>(L, c, S, L->f8, L->b8, st) : lat_iterate_t<8, 16>(L, c, S, L->f8, L->b8, st);
}

}  // namespace sr459
