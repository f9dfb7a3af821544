// ibp_sep.cu -- fused IBP iteration for integer HR shifts and a SEPARABLE PSF: the throughput path.
//
// The mathematics is the body of ibp() (mono_cal_target/run_sr.py:190-209) in the closed form of
// ibp_dense.cu.  Three exact restructurings make one iteration cost ~9 bytes and ~11 FMA per HR pixel:
//
//  (1) polyphase lattices.  With factor 2 and |HR shift| <= 1, frame k samples the blurred image
//      B = conv0(hr, psf) only at the sites (y, x) = (2i - sy_k, 2j - sx_k), one of the four
//      parity lattices (P, Q) = (sy_k & 1, sx_k & 1), and its zero-inserted, back-shifted residual
//      (:170-176) lands on the same sites.  The nominal +-0.5 px corner shifts (quirk Q3) put all
//      four frames on lattice (1, 1).
//  (2) co-sited frames are pre-summed.  All frames of a lattice see the same B(site), so
//          sum_k (lr_k - B)   = S1 - m B                    S1 = sum_k lr_k(site), m = #frames
//          sum_k (lr_k - B)^2 = V + (S1 - m B)^2 / m        V  = sum_k (lr_k - S1/m)^2
//      S1 (one h x w plane per active lattice) and the scalar sum of V are constant over the
//      iterations: k_lattice_sums computes them once per ibp() call, and an iteration then reads one
//      S1 plane per lattice instead of n LR frames (1 instead of n bytes per HR pixel).
//  (3) the Gaussian PSF of make_gaussian_psf (:104-111) is exactly rank 1, psf = acol (x) brow, so
//      both blurs run as a column pass and a row pass that only touch lattice rows / columns.
//
// One CTA (512 threads) owns a 64 x 128 HR tile:
//   stage   TMA: hr tile + 8 px halo (80 x 144, zero fill outside the image = fftconvolve's zero
//           padding) and the 36 x 72 S1 box of every active lattice -> shared memory, one mbarrier.
//   forward column pass  hs -> C  (lattice rows only, float4 down the columns)
//           row pass     C  -> B at the sites; residual S1 - m B in place over the S1 box,
//                        residual energy of the sites the tile owns
//   backward row pass    residual plane -> W (lattice rows, all tile columns)
//           column pass  W -> correction, hr' = clip(hr + step/n * corr), float4 stores.
// The ndi_shift(mode='nearest') clamps only matter on the first tile row / column ("fix-ups",
// evaluated literally from the raw LR frames as in ibp_fast.cu); tiles that touch an image or band
// border run the bounds-checked instantiation of the two passes that need it.
#include "internal.h"
#include "tma.h"

#include <algorithm>
#include <mutex>

namespace sr459 {

constexpr int ST_H = 64, ST_W = 128;              // HR tile
constexpr int SH_TOP  = 7;                        // staged hr rows y0-7 .. y0+70 (the 78 rows a 7x7 PSF reaches),
constexpr int SH_ROWS = ST_H + 14, SH_COLS = ST_W + 16;   //   cols x0-8 .. x0+135 (TMA wants a 16 B aligned first column)
constexpr int SA_R = ST_H / 2 + 4;                // 36 lattice rows   ys = y0-4+2a+P
constexpr int SA_C = ST_W / 2 + 4;                // 68 lattice cols   xs = x0-4+2b+Q
constexpr int SA_P = SA_C + 4;                    // pitch of a lattice plane: TMA needs a 16 B aligned first column, so
constexpr int SA_O = 2;                           //   the S1 box starts 2 sites early and site b sits at column b + SA_O
constexpr int SC_P = 148;                         // pitch of the column-pass plane (== 20 mod 32: conflict-free LDS.128)
constexpr int SW_P = ST_W;                        // pitch of the backward row-pass plane
#ifndef SR459_SEP_THREADS
#define SR459_SEP_THREADS 256
#define SR459_SEP_CTAS 3
#endif
constexpr int S_THREADS = SR459_SEP_THREADS;
constexpr int S_CTAS    = SR459_SEP_CTAS;          // resident CTAs per SM the shared-memory budget is cut for

constexpr int SM_HS   = SH_ROWS * SH_COLS;                        // floats
constexpr int SM_CPL  = ((SA_R * SC_P * 4 + 127) / 128) * 32;     // floats, 128 B multiple
constexpr int SM_ACC  = ((SA_R * SA_P * 4 + 127) / 128) * 32;
constexpr int SM_WD   = SA_R * SW_P;
constexpr int SM_FIX  = SM_WD;                    // edge fix-up vectors live in the tail of the column-pass plane
constexpr int FX_R0 = SM_FIX, FX_C0 = FX_R0 + ST_W + 8, FX_WR = FX_C0 + ST_H + 8, FX_WC = FX_WR + ST_W, FX_END = FX_WC + ST_H;
static_assert(SM_CPL >= FX_END, "the first backward plane and the fix-up vectors alias the column-pass plane");
static_assert((SM_HS * 4) % 128 == 0, "TMA destinations must be 128 B aligned");

void ibp_sep_tile(int* th, int* tw) { *th = ST_H; *tw = ST_W; }

bool ibp_sep_supported(int kh, int kw, int smax, bool separable, int W) {
    return separable && kh == kw && (kh == 1 || kh == 3 || kh == 5 || kh == 7) && smax <= 1 && (W % 4) == 0;
}

// ---- tensor maps ---------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int encode_tile_map_f32(CUtensorMap* out, const void* base, uint64_t cols, uint64_t rows, uint64_t pitch_elems,
                        uint32_t box_c, uint32_t box_r) {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    if (!fn) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return SR459_ECUDA; }
    if ((reinterpret_cast<uintptr_t>(base) & 15) || ((pitch_elems * 4) & 15)) {
        set_error("TMA needs a 16-byte aligned base and row pitch (base %p, pitch %llu floats)", base, (unsigned long long)pitch_elems);
        return SR459_EINVAL;
    }
    cuuint64_t gdim[2] = {cols, rows};
    cuuint64_t gstr[1] = {pitch_elems * 4};
    cuuint32_t box[2] = {box_c, box_r};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return SR459_ECUDA; }
    return SR459_OK;
}

int sep_make_hr_map(CUtensorMap* out, const float* base, int W, int rows) {
    return encode_tile_map_f32(out, base, (uint64_t)W, (uint64_t)rows, (uint64_t)W, SH_COLS, SH_ROWS);
}
int sep_make_s1_map(CUtensorMap* out, const float* base, int w, int rows, int pitch) {
    return encode_tile_map_f32(out, base, (uint64_t)w, (uint64_t)rows, (uint64_t)pitch, SA_P, SA_R);
}

// ---- lattice sums (once per ibp() call) ----------------------------------------------------------------
constexpr int LS_ROWS = 8;                             // plane rows per CTA: 8 x n independent loads per thread, one atomic per CTA
// NF = compile-time bound on the frames of a lattice (>= every slot_n): the per-frame code is fully unrolled and predicated,
// so a 4-frame set must not pay for kMaxFrames slots
template <int NF>
__global__ void __launch_bounds__(256) k_lattice_sums(const __grid_constant__ LatticeArgs a) {
    const int slot = blockIdx.z;
    const int b = blockIdx.x * 256 + threadIdx.x;
    double vd = 0.0;
    if (b < a.w) {
        #pragma unroll (NF <= 4 ? 4 : 2)
        for (int rr = 0; rr < LS_ROWS; ++rr) {
            const int ar = blockIdx.y * LS_ROWS + rr;      // plane row; site row index a = a.row0 + ar
            if (ar >= a.rows) break;
            const int ag = a.row0 + ar;
            float x[NF];
            int m = 0;
            float s = 0.0f;
            #pragma unroll
            for (int f = 0; f < NF; ++f) {
                x[f] = 0.0f;
                if (f < a.slot_n[slot]) {
                    const int k = a.slot_k[slot][f];
                    const int i = ag + a.slot_oy[slot][f], j = b + a.slot_ox[slot][f];
                    if (i >= a.lr_lo && i < a.lr_hi && i < a.h && j < a.w) {
                        x[f] = __ldg(a.lr[k] + (size_t)(i - a.lr_lo) * a.w + j);
                        m |= 1 << f;
                    }
                }
            }
            #pragma unroll
            for (int f = 0; f < NF; ++f) s += x[f];      // frame order, as before (absent frames add 0)
            a.s1[slot][(size_t)ar * a.pitch + b] = s;
            const int cnt = __popc(m);
            if (cnt > 0 && ag >= a.own_a0 && ag < a.own_a1) {
                const float mu = s / (float)cnt;
                float v = 0.0f;
                #pragma unroll
                for (int f = 0; f < NF; ++f)
                    if (m & (1 << f)) { const float d = x[f] - mu; v = fmaf(d, d, v); }
                vd += (double)v;                               // per-site f32, across sites f64 (as before)
            }
        }
    }
    __shared__ double red[8];
    double d = vd;
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = d;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        #pragma unroll
        for (int i = 0; i < 8; ++i) t += red[i];
        if (t != 0.0) atomicAdd(a.vsum, t);
    }
}

int launch_lattice_sums(const LatticeArgs& a, cudaStream_t st) {
    if (a.nslot <= 0 || a.rows <= 0) return SR459_OK;
    SR_CUDA(cudaMemsetAsync(a.vsum, 0, sizeof(double), st));
    dim3 grid((a.w + 255) / 256, (a.rows + LS_ROWS - 1) / LS_ROWS, a.nslot);
    int nmax = 1;
    for (int sl = 0; sl < a.nslot; ++sl) nmax = std::max(nmax, a.slot_n[sl]);
    if (nmax <= 1)      k_lattice_sums<1><<<grid, 256, 0, st>>>(a);
    else if (nmax <= 2) k_lattice_sums<2><<<grid, 256, 0, st>>>(a);
    else if (nmax <= 4) k_lattice_sums<4><<<grid, 256, 0, st>>>(a);
    else                k_lattice_sums<kMaxFrames><<<grid, 256, 0, st>>>(a);
    SR_LAUNCH_CHECK("k_lattice_sums");
    return SR459_OK;
}

// ---- the iteration kernels ------------------------------------------------------------------------------
struct SepTile {
    int y0, x0;                 // tile origin (global HR)
    int oy0, oy1, ox0, ox1;     // pixels / sites this tile owns (global HR, half open)
    int ilo, ihi;               // LR rows that exist and are local
};

// the threads that work on one tile synchronise through this: the whole CTA (k_ibp_sep) or one
// consumer group of the pipelined kernel (a named barrier)
template <int NT, int BAR>
struct GroupSync {
    __device__ __forceinline__ static void sync() {
        if (BAR == 0) __syncthreads();
        else asm volatile("bar.sync %0, %1;\n" :: "r"(BAR), "n"(NT) : "memory");
    }
};

// packed FP32 pair FMA (Blackwell FFMA2): (d0, d1) += w * (x0, x1), one issue slot instead of two
__device__ __forceinline__ void ffma2(float& d0, float& d1, float w, float x0, float x1) {
    asm("{\n.reg .b64 a, b, c;\nmov.b64 a, {%2, %2};\nmov.b64 b, {%3, %4};\nmov.b64 c, {%0, %1};\n"
        "fma.rn.f32x2 c, a, b, c;\nmov.b64 {%0, %1}, c;\n}\n" : "+f"(d0), "+f"(d1) : "f"(w), "f"(x0), "f"(x1));
}

// forward column pass: C[a][lx] = sum_u acol[u] * hs[ly_s+3-u][lx], ly_s = 3+2a+P, 36 lattice rows x 144 staged columns
template <int P, int NT>
__device__ __forceinline__ void fwd_colpass(const SepArgs& a, int tid, const float* __restrict__ hs, float* __restrict__ cpl) {
    #pragma unroll 1
    for (int u = tid; u < (SA_R / 3) * (SH_COLS / 4); u += NT) {
        const int rg = u / (SH_COLS / 4), g = u - rg * (SH_COLS / 4);
        const int a0 = 3 * rg;
        const float4* src = reinterpret_cast<const float4*>(hs + (2 * a0 + P) * SH_COLS) + g;
        float4 v[11];
        #pragma unroll
        for (int r = 0; r < 11; ++r) v[r] = src[r * (SH_COLS / 4)];
        #pragma unroll
        for (int j = 0; j < 3; ++j) {
            float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
            #pragma unroll
            for (int t = 0; t < 7; ++t) {
                const float wgt = a.acol[t];
                const float4 x = v[6 + 2 * j - t];
                ffma2(o.x, o.y, wgt, x.x, x.y);
                ffma2(o.z, o.w, wgt, x.z, x.w);
            }
            *reinterpret_cast<float4*>(cpl + (a0 + j) * SC_P + 4 * g) = o;
        }
    }
}

// forward row pass + residual: B(a,b) = sum_v brow[v] * C[a][4+2b+Q+3-v];  acc(a,b) = S1 - m B (in place)
template <int Q, bool EDGE, int NT>
__device__ __forceinline__ float fwd_rowpass(const SepArgs& a, int tid, const SepTile& c, int slot, int P,
                                             const float* __restrict__ cpl, float* __restrict__ acc) {
    const IbpGeom& g = a.g;
    float e2 = 0.0f;
    const float m_int = a.mtab[slot][0];
    #pragma unroll 1
    for (int u = tid; u < SA_R * (SA_C / 4); u += NT) {
        int r, cg;
        if (u < 576) {                       // quarter-warp = 2 rows x 4 column groups: conflict-free LDS.128
            const int su = u >> 3, l = u & 7;
            r = 2 * (su >> 2) + (l >> 2);
            cg = 4 * (su & 3) + (l & 3);
        } else {
            r = u - 576; cg = 16;
        }
        const float4* p = reinterpret_cast<const float4*>(cpl + r * SC_P + 8 * cg);
        float f[16];
        #pragma unroll
        for (int q = 0; q < 4; ++q) { const float4 t = p[q]; f[4 * q] = t.x; f[4 * q + 1] = t.y; f[4 * q + 2] = t.z; f[4 * q + 3] = t.w; }
        float B[4];
        #pragma unroll
        for (int j = 0; j < 4; ++j) {
            float s = 0.0f;
            #pragma unroll
            for (int v = 0; v < 7; ++v) s = fmaf(a.brow[v], f[7 + 2 * j + Q - v], s);
            B[j] = s;
        }
        float2* sp = reinterpret_cast<float2*>(acc + r * SA_P + SA_O + 4 * cg);
        const float2 s2a = sp[0], s2b = sp[1];
        const float s1[4] = {s2a.x, s2a.y, s2b.x, s2b.y};
        float res[4];
        if (!EDGE) {
            const bool rown = (r >= 2) && (r < SA_R - 2);
            #pragma unroll
            for (int j = 0; j < 4; ++j) {
                res[j] = fmaf(-m_int, B[j], s1[j]);
                const int b = 4 * cg + j;
                if (rown && b >= 2 && b < SA_C - 2) e2 = fmaf(res[j], res[j], e2);
            }
        } else {
            const int ys = c.y0 - 4 + 2 * r + P;
            const int ag = (ys - P) >> 1;
            const bool rv = (ys >= 0) && (ys < g.H) && (ag >= c.ilo) && (ag < c.ihi);
            const bool rown = rv && (ys >= c.oy0) && (ys < c.oy1);
            const int ra = (ag == g.h - 1) ? 2 : 0;
            #pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int xs = c.x0 - 4 + 2 * (4 * cg + j) + Q;
                const bool valid = rv && (xs >= 0) && (xs < g.W);
                const int bg = (xs - Q) >> 1;
                const float m = a.mtab[slot][ra + ((bg == g.w - 1) ? 1 : 0)];
                res[j] = valid ? fmaf(-m, B[j], s1[j]) : 0.0f;
                if (valid && rown && xs >= c.ox0 && xs < c.ox1 && m > 0.0f) e2 += res[j] * res[j] / m;
            }
        }
        sp[0] = make_float2(res[0], res[1]);
        sp[1] = make_float2(res[2], res[3]);
    }
    return EDGE ? e2 : e2 / m_int;
}

// backward row pass: W[a][x] (+)= sum_b brow[3 + xs_b - x] * acc[a][b], lattice rows x 128 tile columns
template <int Q, int NT>
__device__ __forceinline__ void bwd_rowpass(const SepArgs& a, int tid, const float* __restrict__ acc, float* __restrict__ wd, bool add) {
    #pragma unroll 1
    for (int u = tid; u < SA_R * (ST_W / 4); u += NT) {
        const int ar = u >> 5, g = u & 31;
        const float2* p = reinterpret_cast<const float2*>(acc + ar * SA_P + SA_O + 2 * g);
        float t[6];
        #pragma unroll
        for (int q = 0; q < 3; ++q) { const float2 x = p[q]; t[2 * q] = x.x; t[2 * q + 1] = x.y; }
        float o[4] = {0.f, 0.f, 0.f, 0.f};
        #pragma unroll
        for (int tt = 0; tt < 6; ++tt) {
            #pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int diff = 2 * tt + Q - 4 - i;          // xs_b - x for b = 2g + tt, x = x0 + 4g + i
                if (diff >= -3 && diff <= 3) o[i] = fmaf(a.brow[3 + diff], t[tt], o[i]);
            }
        }
        float4* d = reinterpret_cast<float4*>(wd + ar * SW_P + 4 * g);
        if (add) { const float4 e = *d; o[0] += e.x; o[1] += e.y; o[2] += e.z; o[3] += e.w; }
        *d = make_float4(o[0], o[1], o[2], o[3]);
    }
}

// backward column pass contribution of the lattice rows of parity P to a 4 x 4 pixel block
template <int P>
__device__ __forceinline__ void bwd_colpass(const SepArgs& a, const float* __restrict__ wd, int rg, int g, float (&o)[4][4]) {
    const float4* p = reinterpret_cast<const float4*>(wd + (2 * rg) * SW_P) + g;
    #pragma unroll
    for (int tt = 0; tt < 6; ++tt) {
        bool used = false;
        #pragma unroll
        for (int jj = 0; jj < 4; ++jj) { const int diff = 2 * tt + P - 4 - jj; used = used || (diff >= -3 && diff <= 3); }
        if (!used) continue;
        const float4 x = p[tt * (SW_P / 4)];
        #pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
            const int diff = 2 * tt + P - 4 - jj;             // ys_a - y for a = 2rg + tt, y = y0 + 4rg + jj
            if (diff >= -3 && diff <= 3) {
                const float wgt = a.acol[3 + diff];
                ffma2(o[jj][0], o[jj][1], wgt, x.x, x.y);
                ffma2(o[jj][2], o[jj][3], wgt, x.z, x.w);
            }
        }
    }
}

// literal B = conv0(hr, psf) at one global site from the staged tile (edge fix-ups only)
__device__ __noinline__ float sep_b_direct(const SepArgs& a, const float* __restrict__ hs, int y0, int x0, int gy, int gx) {
    const float* c = hs + (gy - (y0 - SH_TOP)) * SH_COLS + (gx - (x0 - 8));
    float v = 0.0f;
    for (int u = 0; u < 7; ++u)
        for (int w = 0; w < 7; ++w) v = fmaf(a.psf[u * 7 + w], c[(3 - u) * SH_COLS + (3 - w)], v);
    return v;
}

__device__ __forceinline__ int sclamp(int v, int lo, int hi) { return min(max(v, lo), hi); }

// the terms ndi_shift's clamp creates at the top / left image edges (first tile row / column only)
template <int NT>
__device__ __noinline__ float edge_fixups(const SepArgs& a, int tid, const SepTile& c, const float* __restrict__ hs,
                                          float* __restrict__ R0, float* __restrict__ C0, bool top, bool left) {
    const IbpGeom& g = a.g;
    float e2 = 0.0f;
    if (top) {
        // residuals deposited a second time on HR row 0 by frames with sy = -1
        for (int t = tid; t < ST_W + 8; t += NT) {
            const int x = c.x0 - 4 + t;
            float v = 0.0f;
            if (x >= 0 && x < g.W) {
                for (int k = 0; k < g.n; ++k) {
                    if (g.sy[k] != -1) continue;
                    const int X = sclamp(x + g.sx[k], 0, g.W - 1);
                    if (X & 1) continue;
                    const int j = X >> 1;
                    if (j >= g.w || c.ilo > 0) continue;
                    v += __ldg(a.lr[k] + (size_t)(0 - g.lr_lo) * g.w + j) -
                         sep_b_direct(a, hs, c.y0, c.x0, 1, sclamp(2 * j - g.sx[k], 0, g.W - 1));
                }
            }
            R0[t] = v;
        }
        // LR row 0 of frames with sy = +1 reads B[clamp(-1)] = B[0]: MSE only
        if (g.upd_y0 == 0 && c.ilo == 0) {
            for (int k = 0; k < g.n; ++k) {
                if (g.sy[k] != 1) continue;
                for (int j = tid; j < g.w; j += NT) {
                    const int cx = sclamp(2 * j - g.sx[k], 0, g.W - 1);
                    if (cx < c.ox0 || cx >= c.ox1) continue;
                    const float e = __ldg(a.lr[k] + (size_t)(0 - g.lr_lo) * g.w + j) - sep_b_direct(a, hs, c.y0, c.x0, 0, cx);
                    e2 = fmaf(e, e, e2);
                }
            }
        }
    }
    if (left) {
        for (int t = tid; t < ST_H + 8; t += NT) {
            const int y = c.y0 - 4 + t;
            float v = 0.0f;
            if (y >= 0 && y < g.H) {
                for (int k = 0; k < g.n; ++k) {
                    if (g.sx[k] != -1) continue;
                    if (y + g.sy[k] < 0) continue;               // corner: already counted in R0
                    const int Y = sclamp(y + g.sy[k], 0, g.H - 1);
                    if (Y & 1) continue;
                    const int i = Y >> 1;
                    if (i >= c.ihi || i < c.ilo) continue;
                    v += __ldg(a.lr[k] + (size_t)(i - g.lr_lo) * g.w + 0) -
                         sep_b_direct(a, hs, c.y0, c.x0, sclamp(2 * i - g.sy[k], 0, g.H - 1), 1);
                }
            }
            C0[t] = v;
        }
        // LR col 0 of frames with sx = +1 reads B[., clamp(-1)] = B[., 0]: MSE only
        for (int k = 0; k < g.n; ++k) {
            if (g.sx[k] != 1) continue;
            for (int t = tid; t < ST_H / 2 + 2; t += NT) {
                const int i = (c.y0 >> 1) - 1 + t;
                if (i < c.ilo || i >= c.ihi) continue;
                const int cy = 2 * i - g.sy[k];
                if (cy < 0 || cy < c.oy0 || cy >= c.oy1) continue;
                const float e = __ldg(a.lr[k] + (size_t)(i - g.lr_lo) * g.w + 0) - sep_b_direct(a, hs, c.y0, c.x0, cy, 0);
                e2 = fmaf(e, e, e2);
            }
        }
    }
    return e2;
}

__device__ __forceinline__ SepTile make_tile(const SepArgs& a, int by, int bx) {
    const IbpGeom& g = a.g;
    SepTile c;
    c.y0 = a.ty_origin + by * ST_H;
    c.x0 = bx * ST_W;
    c.oy0 = max(c.y0, g.upd_y0); c.oy1 = min(min(c.y0 + ST_H, g.upd_y1), g.H);
    c.ox0 = c.x0;                c.ox1 = min(c.x0 + ST_W, g.W);
    c.ilo = max(0, g.lr_lo);     c.ihi = min(g.h, g.lr_hi);
    return c;
}

// everything between "the TMA boxes have landed" and "the tile is stored"; returns this thread's share of
// the tile's residual energy.  hs: staged hr, cpl: column-pass plane (later W plane + fix-up vectors),
// accs: S1 boxes (in place: residual planes), wd2: second W plane (both row parities active)
template <int NT, int BAR>
__device__ __forceinline__ float process_tile(const SepArgs& a, int tid, const SepTile& c, const float* __restrict__ hs,
                                              float* __restrict__ cpl, float* __restrict__ accs, float* __restrict__ wd2) {
    using G = GroupSync<NT, BAR>;
    const IbpGeom& g = a.g;
    float* R0 = cpl + FX_R0;             // duplicates deposited on HR row 0, x = x0-4 .. x0+131
    float* C0 = cpl + FX_C0;             // duplicates deposited on HR col 0, y = y0-4 .. y0+67
    float* WR = cpl + FX_WR;             // their row-passed / column-passed forms
    float* WC = cpl + FX_WC;
    const bool top = (c.y0 == 0), left = (c.x0 == 0);
    const bool edge = (c.y0 < 8) || (c.x0 < 8) || (c.y0 + ST_H + 5 > g.H) || (c.x0 + ST_W + 5 > g.W) ||
                      (c.y0 + ST_H > g.upd_y1) || (c.y0 < g.upd_y0) ||
                      (((c.y0 - 6) >> 1) < c.ilo) || (((c.y0 + ST_H + 4) >> 1) + 1 > c.ihi);

    // ---- forward --------------------------------------------------------------------------------------------
    float e2 = 0.0f;
    int lastP = -1;
    for (int s = 0; s < a.nph; ++s) {
        const int P = a.ph_pq[s] >> 1, Q = a.ph_pq[s] & 1;
        if (P != lastP) {
            if (s) G::sync();                          // the previous lattice has finished reading C
            if (P) fwd_colpass<1, NT>(a, tid, hs, cpl); else fwd_colpass<0, NT>(a, tid, hs, cpl);
            lastP = P;
            G::sync();
        }
        float* acc = accs + s * SM_ACC;
        if (edge) e2 += Q ? fwd_rowpass<1, true, NT>(a, tid, c, s, P, cpl, acc) : fwd_rowpass<0, true, NT>(a, tid, c, s, P, cpl, acc);
        else      e2 += Q ? fwd_rowpass<1, false, NT>(a, tid, c, s, P, cpl, acc) : fwd_rowpass<0, false, NT>(a, tid, c, s, P, cpl, acc);
    }
    if (top || left) {
        G::sync();                                     // R0 / C0 alias the column-pass plane
        e2 += edge_fixups<NT>(a, tid, c, hs, R0, C0, top, left);
    }
    G::sync();                                         // residual planes complete, C free

    // ---- backward row pass: one W plane per active row parity ----------------------------------------------
    int wsel0 = -1, wsel1 = -1;                        // which plane (0: cpl, 1: wd2) holds row parity 0 / 1
    {
        int nplanes = 0;
        for (int s = 0; s < a.nph; ++s) {
            const int P = a.ph_pq[s] >> 1, Q = a.ph_pq[s] & 1;
            int sel = P ? wsel1 : wsel0;
            const bool add = sel >= 0;
            if (!add) {
                sel = nplanes++;
                if (P) wsel1 = sel; else wsel0 = sel;
            }
            float* w = sel ? wd2 : cpl;
            if (Q) bwd_rowpass<1, NT>(a, tid, accs + s * SM_ACC, w, add); else bwd_rowpass<0, NT>(a, tid, accs + s * SM_ACC, w, add);
        }
    }
    // separable form of the duplicates: corr[y][x] += acol[3-y] * WR[x] (y <= 3) + brow[3-x] * WC[y] (x <= 3)
    if (top) {
        for (int t = tid; t < ST_W; t += NT) {
            float v = 0.0f;
            #pragma unroll
            for (int tb = 0; tb < 7; ++tb) v = fmaf(a.brow[tb], R0[t + tb + 1], v);       // R0 index of x + tb - 3
            WR[t] = v;
        }
    }
    if (left) {
        for (int t = tid; t < ST_H; t += NT) {
            float v = 0.0f;
            #pragma unroll
            for (int ta = 0; ta < 7; ++ta) v = fmaf(a.acol[ta], C0[t + ta + 1], v);       // C0 index of y + ta - 3
            WC[t] = v;
        }
    }
    G::sync();

    // ---- backward column pass, update, clip, store ----------------------------------------------------------
    #pragma unroll 1
    for (int u = tid; u < (ST_H / 4) * (ST_W / 4); u += NT) {
        const int rg = u >> 5, gq = u & 31;
        float o[4][4];
        #pragma unroll
        for (int jj = 0; jj < 4; ++jj) { o[jj][0] = o[jj][1] = o[jj][2] = o[jj][3] = 0.f; }
        if (wsel0 >= 0) bwd_colpass<0>(a, wsel0 ? wd2 : cpl, rg, gq, o);
        if (wsel1 >= 0) bwd_colpass<1>(a, wsel1 ? wd2 : cpl, rg, gq, o);
        const int gx = c.x0 + 4 * gq;
        float* outp = a.hr_out + (size_t)(c.y0 + 4 * rg - g.hr_lo) * g.W + gx;
        const float* hp = hs + (4 * rg + SH_TOP) * SH_COLS + 4 * gq + 8;
        #pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
            const int gy = c.y0 + 4 * rg + jj;
            if (edge) {
                if (gy < c.oy0 || gy >= c.oy1 || gx >= c.ox1) continue;
                if (top && gy <= 3) {
                    const float wgt = a.acol[3 - gy];
                    const float4 wr = *reinterpret_cast<const float4*>(WR + 4 * gq);
                    o[jj][0] = fmaf(wgt, wr.x, o[jj][0]); o[jj][1] = fmaf(wgt, wr.y, o[jj][1]);
                    o[jj][2] = fmaf(wgt, wr.z, o[jj][2]); o[jj][3] = fmaf(wgt, wr.w, o[jj][3]);
                }
                if (left && gq == 0) {
                    const float cv = WC[4 * rg + jj];
                    #pragma unroll
                    for (int i = 0; i < 4; ++i) o[jj][i] = fmaf(a.brow[3 - i], cv, o[jj][i]);
                }
            }
            float4 r = *reinterpret_cast<const float4*>(hp + jj * SH_COLS);
            ffma2(r.x, r.y, g.alpha, o[jj][0], o[jj][1]);
            ffma2(r.z, r.w, g.alpha, o[jj][2], o[jj][3]);
            r.x = fminf(fmaxf(r.x, g.lo), g.hi);
            r.y = fminf(fmaxf(r.y, g.lo), g.hi);
            r.z = fminf(fmaxf(r.z, g.lo), g.hi);
            r.w = fminf(fmaxf(r.w, g.lo), g.hi);
            *reinterpret_cast<float4*>(outp + (size_t)jj * g.W) = r;
        }
    }
    return e2;
}

// ---- one tile per CTA (any number of active lattices) ----------------------------------------------------
__global__ void __launch_bounds__(S_THREADS, S_CTAS) k_ibp_sep(const __grid_constant__ SepArgs a) {
    extern __shared__ __align__(128) float sm[];
    __shared__ __align__(8) uint64_t mbar;
    __shared__ float red[S_THREADS / 32];
    float* hs   = sm;
    float* cpl  = sm + SM_HS;
    float* accs = cpl + SM_CPL;
    const int tid = threadIdx.x;
    const SepTile c = make_tile(a, blockIdx.y, blockIdx.x);

    if (tid == 0) {                                    // one thread arms the mbarrier and issues the TMA boxes
        mbar_init(&mbar, 1);
        mbar_fence_init();
        mbar_expect_tx(&mbar, (uint32_t)(4 * (SM_HS + a.nph * SA_R * SA_P)));
        tma_load_2d(hs, &a.tm_hr, c.x0 - 8, c.y0 - SH_TOP - a.g.hr_lo, &mbar);
        for (int s = 0; s < a.nph; ++s)
            tma_load_2d(accs + s * SM_ACC, &a.tm_s1[s], (c.x0 >> 1) - 2 - SA_O, (c.y0 >> 1) - 2 - a.s1_row0, &mbar);
    }
    __syncthreads();                                   // the barrier is initialised before anyone polls it
    mbar_wait(&mbar, 0);

    float e2 = process_tile<S_THREADS, 0>(a, tid, c, hs, cpl, accs, accs + a.nph * SM_ACC);

    #pragma unroll
    for (int off = 16; off > 0; off >>= 1) e2 += __shfl_xor_sync(0xffffffffu, e2, off);
    if ((tid & 31) == 0) red[tid >> 5] = e2;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        #pragma unroll
        for (int i = 0; i < S_THREADS / 32; ++i) t += (double)red[i];
        a.partial[blockIdx.y * gridDim.x + blockIdx.x] = t;
    }
}

size_t ibp_sep_smem(int nph, int nrowpar) {
    return sizeof(float) * ((size_t)SM_HS + SM_CPL + (size_t)nph * SM_ACC + (nrowpar > 1 ? SM_WD : 0));
}

// ---- persistent, pipelined variant (one active lattice): the headline configuration -----------------------
// One CTA per SM: a producer warp streams tiles through a 3-stage ring of {hr tile, S1 box} buffers with
// TMA; two consumer groups of PG_T threads each take every other tile.  While both groups compute, the
// third stage is already in flight, so no warp ever waits for HBM; tiles are handed out dynamically
// (one global counter per iteration), which also removes the wave-quantisation tail of a plain grid.
constexpr int PG_T = 384;                         // threads per consumer group
constexpr int PG_N = 2;                           // consumer groups
constexpr int PG_STAGES = 3;
constexpr int PG_THREADS = PG_N * PG_T + 32;      // + producer warp
constexpr int SM_STAGE = SM_HS + SM_ACC;          // floats per stage
static_assert((SM_STAGE * 4) % 128 == 0, "stage size must keep TMA destinations 128 B aligned");

__global__ void __launch_bounds__(PG_THREADS, 1) k_ibp_sep_pipe(const __grid_constant__ SepArgs a) {
    extern __shared__ __align__(128) float sm[];
    __shared__ __align__(8) uint64_t full[PG_STAGES];
    __shared__ __align__(8) uint64_t empty[PG_STAGES];
    __shared__ int tile_of[PG_STAGES];
    __shared__ float red[PG_N][PG_T / 32];
    const int tid = threadIdx.x;
    const int ntiles = a.tiles_x * a.tiles_y;

    if (tid == 0) {
        for (int s = 0; s < PG_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        mbar_fence_init();
    }
    __syncthreads();

    if (tid >= PG_N * PG_T) {
        // ===== producer warp (one elected lane) =====
        if (tid == PG_N * PG_T) {
            int ended = 0;
            for (int k = 0; ended < PG_N; ++k) {
                const int s = k % PG_STAGES;
                if (k >= PG_STAGES) mbar_wait(&empty[s], ((k / PG_STAGES) - 1) & 1);
                int t = -1;
                if (!ended) {
                    t = atomicAdd(a.tile_counter, 1);
                    if (t >= ntiles) t = -1;
                }
                tile_of[s] = t;
                if (t >= 0) {
                    const int by = t / a.tiles_x, bx = t - by * a.tiles_x;
                    const int y0 = a.ty_origin + by * ST_H, x0 = bx * ST_W;
                    float* st = sm + s * SM_STAGE;
                    mbar_expect_tx(&full[s], (uint32_t)(4 * (SM_HS + SA_R * SA_P)));
                    tma_load_2d(st, &a.tm_hr, x0 - 8, y0 - SH_TOP - a.g.hr_lo, &full[s]);
                    tma_load_2d(st + SM_HS, &a.tm_s1[0], (x0 >> 1) - 2 - SA_O, (y0 >> 1) - 2 - a.s1_row0, &full[s]);
                } else {
                    ++ended;
                    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" :: "r"(smem_u32(&full[s])) : "memory");
                }
            }
        }
        return;
    }

    // ===== consumer groups =====
    const int grp = tid / PG_T, gt = tid - grp * PG_T;
    float* cpl = sm + PG_STAGES * SM_STAGE + grp * SM_CPL;
    for (int k = grp; ; k += PG_N) {
        const int s = k % PG_STAGES;
        mbar_wait(&full[s], (k / PG_STAGES) & 1);
        const int t = tile_of[s];
        if (t < 0) break;
        const int by = t / a.tiles_x, bx = t - by * a.tiles_x;
        const SepTile c = make_tile(a, by, bx);
        float* st = sm + s * SM_STAGE;
        float e2 = grp == 0 ? process_tile<PG_T, 1>(a, gt, c, st, cpl, st + SM_HS, nullptr)
                            : process_tile<PG_T, 2>(a, gt, c, st, cpl, st + SM_HS, nullptr);
        #pragma unroll
        for (int off = 16; off > 0; off >>= 1) e2 += __shfl_xor_sync(0xffffffffu, e2, off);
        if ((gt & 31) == 0) red[grp][gt >> 5] = e2;
        if (grp == 0) GroupSync<PG_T, 1>::sync(); else GroupSync<PG_T, 2>::sync();     // stage and red[] are done with
        if (gt == 0) {
            double sum = 0.0;
            #pragma unroll
            for (int i = 0; i < PG_T / 32; ++i) sum += (double)red[grp][i];
            a.partial[t] = sum;
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" :: "r"(smem_u32(&empty[s])) : "memory");
        }
    }
}

size_t ibp_sep_pipe_smem() { return sizeof(float) * ((size_t)PG_STAGES * SM_STAGE + (size_t)PG_N * SM_CPL); }

int launch_ibp_sep(const SepArgs& a, cudaStream_t st) {
    if (a.tiles_x <= 0 || a.tiles_y <= 0) return SR459_OK;
    if (a.nph == 1 && a.tile_counter && a.n_sm > 0) {
        const size_t smem = ibp_sep_pipe_smem();
        static PerDeviceOnce configured;
        if (configured.need()) {
            SR_CUDA(cudaFuncSetAttribute(k_ibp_sep_pipe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        }
        const int ntiles = a.tiles_x * a.tiles_y;
        const int grid = std::min(a.n_sm, (ntiles + PG_N - 1) / PG_N);
        k_ibp_sep_pipe<<<grid, PG_THREADS, smem, st>>>(a);
        SR_LAUNCH_CHECK("k_ibp_sep_pipe");
        return SR459_OK;
    }
    int rowpar = 0;
    for (int s = 0; s < a.nph; ++s) rowpar |= 1 << (a.ph_pq[s] >> 1);
    const size_t smem = ibp_sep_smem(a.nph, rowpar == 3 ? 2 : 1);
    static PerDeviceMax configured;
    if (configured.raise(smem)) {
        SR_CUDA(cudaFuncSetAttribute(k_ibp_sep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    dim3 grid(a.tiles_x, a.tiles_y);
    k_ibp_sep<<<grid, S_THREADS, smem, st>>>(a);
    SR_LAUNCH_CHECK("k_ibp_sep");
    return SR459_OK;
}

}  // namespace sr459
