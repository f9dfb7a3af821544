// ibp_fast.cu -- fused IBP iteration, integer HR shifts, polyphase kernel (the throughput path).
//
// Same mathematics as ibp_dense.cu (the body of ibp(), mono_cal_target/run_sr.py:190-209,
// for integer factor*shift), specialised for what the reference actually runs:
// factor 2, |HR shift| <= 1, odd square PSF up to 7x7.  Where no clamp of
// ndi_shift(mode='nearest') is active,
//   * frame k samples the blurred image B only on ONE polyphase lattice, the sites whose
//     (y,x) parity equals the parity of its shift (sy_k, sx_k)                    (:163-165)
//   * its zero-inserted, back-shifted residual lives on that same lattice          (:170-176)
// so B is evaluated only on the active lattices (1 of 4 for the nominal +-0.5 px corner
// shifts) and the transposed blur (:178) only visits taps that can meet a non-zero site:
// ~12 MAC per HR pixel each way instead of 49.
//
// One CTA owns a 64x128 HR tile:
//   stage 1  hr tile + 8 px halo -> shared memory (float4 loads, zero outside the image =
//            fftconvolve's zero padding).
//   stage 2  per active lattice: a thread walks a 4-site column strip -- 13 staged rows x 4
//            LDS.64 (unit-stride across the warp: conflict free), PSF taps as constant-bank
//            FFMA operands -- fetches the matching LR samples coalesced from global,
//            accumulates the residual energy of the sites the tile owns, and stores the
//            summed residuals to a 36x68 shared-memory lattice plane.
//   stage 3  transposed blur from the lattice plane(s), update, clip, float2 stores.
// The ndi_shift clamps only bite at the top and left image edges when |shift| <= 1:
//   * LR row 0 of a frame with sy=+1 reads B[clamp(-1)] = B[0]: a residual that enters the
//     MSE (:202) but is never back-projected;
//   * HR row 0 receives a duplicate of the residuals frames with sy=-1 deposit on row 1;
// (same for columns).  CTAs on the first tile row / column evaluate those few terms
// literally ("fix-ups") and all edge CTAs run a bounds-checked instantiation; the other
// ~90 % of the CTAs run with no checks at all.
#include "internal.h"

namespace sr459 {

constexpr int FT_H    = 64;             // tile rows (HR)
constexpr int FT_W    = 128;            // tile cols (HR)
constexpr int FH_ROWS = FT_H + 16;      // staged hr rows  y0-8 .. y0+71
constexpr int FH_COLS = FT_W + 16;      // staged hr cols  x0-8 .. x0+135
constexpr int FA_R    = FT_H / 2 + 4;   // 36: lattice-plane rows   y' = y0-4+2a+P
constexpr int FA_C    = FT_W / 2 + 4;   // 68: lattice-plane cols   x' = x0-4+2b+Q
constexpr int F_THREADS = 256;
constexpr int F_PRE   = 4;              // frames per lattice whose LR samples are prefetched

void ibp_fast_tile(int* th, int* tw) { *th = FT_H; *tw = FT_W; }

bool ibp_fast_supported(int kh, int kw, int smax) {
    return kh == kw && (kh == 1 || kh == 3 || kh == 5 || kh == 7) && smax <= 1;
}

struct TileCtx {
    int y0, x0;                 // tile origin (global HR)
    int oy0, oy1, ox0, ox1;     // sites / pixels this tile owns (global HR, half open)
    int ilo, ihi;               // LR rows that exist and are local
};

// ---- stage 2: B on lattice (P,Q), residuals, accumulation ------------------------------------------
// M > 0: exactly M (<= 4) frames live on this lattice and the tile needs no bounds checks: everything is
//        compile-time unrolled.  M == 0: general instantiation (edge tiles, more than 4 frames per lattice).
// RR: the PSF is zero outside |tap offset| <= RR (the measured 7 x 7 PSF of the reference has an empty outer ring: RR = 2)
template <int P, int Q, int M, int RR>
__device__ __forceinline__ void stage_forward(const FastArgs& a, const float* __restrict__ hs,
                                              float* __restrict__ plane, const TileCtx& c, float& e2,
                                              bool wait_stage1) {
    constexpr bool CHK = (M == 0);
    constexpr int  NF  = CHK ? F_PRE : M;               // prefetched frames
    const IbpGeom& g = a.g;
    constexpr int ph = P * 2 + Q;
    const int m = a.ph_n[ph];
    const int mpre = m < F_PRE ? m : F_PRE;
    const int ibase = (c.y0 - 4) >> 1, jbase = (c.x0 - 4) >> 1;
    const int w = g.w;
    constexpr int NT = (FA_R / 4) * FA_C;

    // per-frame LR base pointers for this tile: sample of site (a,b) is pk[f][a*w + b]
    const float* pk[NF];
    int oyf[NF], oxf[NF];
    #pragma unroll
    for (int f = 0; f < NF; ++f) {
        const int k = a.ph_k[ph][f < mpre ? f : 0];
        oyf[f] = (P + g.sy[k]) >> 1; oxf[f] = (Q + g.sx[k]) >> 1;
        pk[f] = a.lr[k] + (ptrdiff_t)(ibase + oyf[f] - g.lr_lo) * w + (jbase + oxf[f]);
    }

    // LR samples of task t: fetched one task ahead so their global latency hides behind the FMA block
    auto fetch_lr = [&](int t, float (&dst)[4][NF]) {
        const int grp = t / FA_C, b = t - grp * FA_C;
        const int a0 = grp * 4;
        const int off = a0 * w + b;
        #pragma unroll
        for (int f = 0; f < NF; ++f) {
            #pragma unroll
            for (int s = 0; s < 4; ++s) {
                if (CHK) {
                    const int i = ibase + oyf[f] + a0 + s, j = jbase + oxf[f] + b;
                    const bool ok = (f < mpre) && (i >= c.ilo) && (i < c.ihi) && (j >= 0) && (j < w) &&
                                    (c.y0 - 4 + 2 * (a0 + s) + P >= 0) && (c.x0 - 4 + 2 * b + Q >= 0);
                    dst[s][f] = ok ? __ldg(pk[f] + (off + s * w)) : 0.0f;
                } else {
                    dst[s][f] = __ldg(pk[f] + (off + s * w));
                }
            }
        }
    };

    float lrn[4][NF];
    if (threadIdx.x < NT) fetch_lr(threadIdx.x, lrn);
    if (wait_stage1) {                 // first lattice: the staged hr tile is needed from here on
        asm volatile("cp.async.wait_all;\n" ::: "memory");
        __syncthreads();
    }

    #pragma unroll 1
    for (int t = threadIdx.x; t < NT; t += F_THREADS) {
        const int grp = t / FA_C, b = t - grp * FA_C;
        const int a0 = grp * 4;
        float lrv[4][NF];
        #pragma unroll
        for (int f = 0; f < NF; ++f) {
            #pragma unroll
            for (int s = 0; s < 4; ++s) lrv[s][f] = lrn[s][f];
        }
        if (t + F_THREADS < NT) fetch_lr(t + F_THREADS, lrn);

        // B at the 4 sites: staged rows 8*grp+P+1 .. +12, staged cols 2b+Q+1 .. 2b+Q+7
        float B[4] = {0.f, 0.f, 0.f, 0.f};
        const float* base = hs + (2 * a0 + P + 1) * FH_COLS + 2 * b + (Q ? 2 : 0);
        #pragma unroll
        for (int rr = 0; rr < 13; ++rr) {
            if (rr < 3 - RR || rr > 9 + RR) continue;       // no site of the four needs this staged row
            float v[8];
            #pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float2 t2 = *reinterpret_cast<const float2*>(base + rr * FH_COLS + 2 * q);
                v[2 * q] = t2.x; v[2 * q + 1] = t2.y;
            }
            #pragma unroll
            for (int s = 0; s < 4; ++s) {
                const int dy = rr - 3 - 2 * s;           // staged row - site centre row
                if (dy >= -RR && dy <= RR) {
                    #pragma unroll
                    for (int dx = -RR; dx <= RR; ++dx)
                        B[s] = fmaf(a.psf[(3 - dy) * 7 + (3 - dx)], v[dx + (Q ? 3 : 4)], B[s]);
                }
            }
        }

        if (!CHK) {
            // owned sites: plane rows 2..33 and cols 2..65 (the rest is halo recomputed by neighbours)
            const float cm = (b >= 2 && b < FA_C - 2) ? 1.0f : 0.0f;
            const float mlo = (grp == 0) ? 0.0f : cm, mhi = (grp == FA_R / 4 - 1) ? 0.0f : cm;
            #pragma unroll
            for (int s = 0; s < 4; ++s) {
                float acc = 0.0f, sq = 0.0f;
                #pragma unroll
                for (int f = 0; f < NF; ++f) {
                    const float e = lrv[s][f] - B[s];
                    acc += e;
                    sq = fmaf(e, e, sq);
                }
                e2 = fmaf(s < 2 ? mlo : mhi, sq, e2);
                plane[(a0 + s) * FA_C + b] = acc;
            }
        } else {
            #pragma unroll
            for (int s = 0; s < 4; ++s) {
                const int yg = c.y0 - 4 + 2 * (a0 + s) + P, xg = c.x0 - 4 + 2 * b + Q;
                const bool own = (yg >= c.oy0) && (yg < c.oy1) && (xg >= c.ox0) && (xg < c.ox1);
                float acc = 0.0f, sq = 0.0f;
                #pragma unroll
                for (int f = 0; f < NF; ++f) {
                    const int i = ibase + oyf[f] + a0 + s, j = jbase + oxf[f] + b;
                    const bool ok = (f < mpre) && (i >= c.ilo) && (i < c.ihi) && (j >= 0) && (j < w) && (yg >= 0) && (xg >= 0);
                    if (ok) {
                        const float e = lrv[s][f] - B[s];
                        acc += e;
                        sq = fmaf(e, e, sq);
                    }
                }
                for (int f = F_PRE; f < m; ++f) {            // more than 4 frames on one lattice: rare
                    const int k = a.ph_k[ph][f];
                    const int i = ibase + ((P + g.sy[k]) >> 1) + a0 + s, j = jbase + ((Q + g.sx[k]) >> 1) + b;
                    if ((i >= c.ilo) && (i < c.ihi) && (j >= 0) && (j < w) && (yg >= 0) && (xg >= 0)) {
                        const float e = __ldg(a.lr[k] + (ptrdiff_t)(i - g.lr_lo) * w + j) - B[s];
                        acc += e;
                        sq = fmaf(e, e, sq);
                    }
                }
                if (own) e2 += sq;
                plane[(a0 + s) * FA_C + b] = acc;
            }
        }
    }
}

template <int P, int Q, int RR>
__device__ __forceinline__ void forward_dispatch(const FastArgs& a, const float* __restrict__ hs, float* __restrict__ plane,
                                                 const TileCtx& c, float& e2, bool wait_stage1, bool edge) {
    const int m = a.ph_n[P * 2 + Q];
    if (edge || m > F_PRE) { stage_forward<P, Q, 0, RR>(a, hs, plane, c, e2, wait_stage1); return; }
    switch (m) {
        case 1: stage_forward<P, Q, 1, RR>(a, hs, plane, c, e2, wait_stage1); break;
        case 2: stage_forward<P, Q, 2, RR>(a, hs, plane, c, e2, wait_stage1); break;
        case 3: stage_forward<P, Q, 3, RR>(a, hs, plane, c, e2, wait_stage1); break;
        default: stage_forward<P, Q, 4, RR>(a, hs, plane, c, e2, wait_stage1); break;
    }
}

// ---- stage 3: transposed-blur contribution of lattice (P,Q) to a 4-cell column strip --------------
template <int P, int Q, int RR>
__device__ __forceinline__ void stage_backward(const FastArgs& a, const float* __restrict__ plane,
                                               int i0, int j, float (&o)[4][2][2]) {
    #pragma unroll
    for (int ar = 0; ar < 8; ++ar) {
        float v[5];
        #pragma unroll
        for (int cc = 0; cc < 5; ++cc) v[cc] = plane[(i0 + ar) * FA_C + j + cc];
        #pragma unroll
        for (int cell = 0; cell < 4; ++cell) {
            #pragma unroll
            for (int oy = 0; oy < 2; ++oy) {
                #pragma unroll
                for (int ta = 3 - RR; ta <= 3 + RR; ++ta) {
                    if (((ta + P + 3 + oy) & 1) != 0) continue;
                    const int da = (oy + ta + 1 - P) / 2;      // exact: numerator is even
                    if (cell + da != ar) continue;
                    #pragma unroll
                    for (int ox = 0; ox < 2; ++ox) {
                        #pragma unroll
                        for (int tb = 3 - RR; tb <= 3 + RR; ++tb) {
                            if (((tb + Q + 3 + ox) & 1) != 0) continue;
                            const int db = (ox + tb + 1 - Q) / 2;
                            o[cell][oy][ox] = fmaf(a.psf[ta * 7 + tb], v[db], o[cell][oy][ox]);
                        }
                    }
                }
            }
        }
    }
}

// literal B = conv0(hr, psf) at one global site, from the staged tile (edge fix-ups only)
__device__ __noinline__ float b_direct(const FastArgs& a, const float* __restrict__ hs, int y0, int x0, int gy, int gx) {
    const float* c = hs + (gy - (y0 - 8)) * FH_COLS + (gx - (x0 - 8));
    float v = 0.0f;
    for (int u = 0; u < 7; ++u)
        for (int w = 0; w < 7; ++w) v = fmaf(a.psf[u * 7 + w], c[(3 - u) * FH_COLS + (3 - w)], v);
    return v;
}

__device__ __forceinline__ int clampi2(int v, int lo, int hi) { return min(max(v, lo), hi); }

template <int RR>
__global__ void __launch_bounds__(F_THREADS, 3) k_ibp_fast(const __grid_constant__ FastArgs a) {
    extern __shared__ __align__(16) float sm[];
    __shared__ float R0[FT_W + 8];       // duplicates deposited on HR row 0, x = x0-4 .. x0+131
    __shared__ float C0[FT_H + 8];       // duplicates deposited on HR col 0, y = y0-4 .. y0+67
    __shared__ float red[F_THREADS / 32];
    const IbpGeom& g = a.g;
    float* hs = sm;                                        // FH_ROWS x FH_COLS
    float* planes = sm + FH_ROWS * FH_COLS;                // one FA_R x FA_C plane per active lattice
    const int tid = threadIdx.x;
    TileCtx c;
    c.y0 = a.ty_origin + blockIdx.y * FT_H;
    c.x0 = blockIdx.x * FT_W;
    c.oy0 = max(c.y0, g.upd_y0); c.oy1 = min(min(c.y0 + FT_H, g.upd_y1), g.H);
    c.ox0 = c.x0;                c.ox1 = min(c.x0 + FT_W, g.W);
    c.ilo = max(0, g.lr_lo);     c.ihi = min(g.h, g.lr_hi);

    // ---- stage 1: stage hr (zero outside image / local band) with 16-byte cp.async -------------------
    // (LDGSTS: no register staging, every copy of a thread in flight at once; src-size 0 zero-fills).
    // One warp per staged row: 32 lanes x 16 B + 4 lanes x 16 B = the 144-float row.
    {
        const int lane = tid & 31, warp = tid >> 5;
        const int ylo = max(0, g.hr_lo), yhi = min(g.H, g.hr_hi);
        const float* src0 = a.hr_in + (ptrdiff_t)(c.y0 - 8 - g.hr_lo) * g.W + (c.x0 - 8);
        const unsigned dst0 = (unsigned)__cvta_generic_to_shared(hs);
        const int gxa = c.x0 - 8 + 4 * lane, gxb = c.x0 - 8 + 128 + 4 * lane;
        const bool ca = (gxa >= 0) && (gxa < g.W), cb = (lane < 4) && (gxb < g.W);
        #pragma unroll
        for (int r = 0; r < FH_ROWS / 8; ++r) {
            const int ly = warp + 8 * r;
            const int gy = c.y0 - 8 + ly;
            const bool rok = (gy >= ylo) && (gy < yhi);
            const float* rs = src0 + (ptrdiff_t)ly * g.W;
            const unsigned d = dst0 + (unsigned)(ly * FH_COLS + 4 * lane) * 4u;
            {
                const bool in = rok && ca;
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" :: "r"(d), "l"(in ? rs + 4 * lane : a.hr_in), "r"(in ? 16 : 0) : "memory");
            }
            if (lane < 4) {
                const bool in = rok && cb;
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" :: "r"(d + 512u), "l"(in ? rs + 128 + 4 * lane : a.hr_in), "r"(in ? 16 : 0) : "memory");
            }
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    }

    // a tile is "edge" when any of its lattice sites could pair with a missing LR sample, sit outside
    // the image, or fall outside the rows this launch owns
    const bool top = (c.y0 == 0), left = (c.x0 == 0);
    const bool edge = (c.y0 < 8) || (c.x0 < 8) || (c.y0 + FT_H + 5 > g.H) || (c.x0 + FT_W + 5 > g.W) ||
                      (c.y0 + FT_H > g.upd_y1) || (((c.y0 - 6) >> 1) < c.ilo) || (((c.y0 + FT_H + 4) >> 1) + 1 > c.ihi);

    float e2 = 0.0f;
    {
        int slot = 0;
        if (a.ph_n[0]) { forward_dispatch<0, 0, RR>(a, hs, planes + slot * FA_R * FA_C, c, e2, slot == 0, edge); ++slot; }
        if (a.ph_n[1]) { forward_dispatch<0, 1, RR>(a, hs, planes + slot * FA_R * FA_C, c, e2, slot == 0, edge); ++slot; }
        if (a.ph_n[2]) { forward_dispatch<1, 0, RR>(a, hs, planes + slot * FA_R * FA_C, c, e2, slot == 0, edge); ++slot; }
        if (a.ph_n[3]) { forward_dispatch<1, 1, RR>(a, hs, planes + slot * FA_R * FA_C, c, e2, slot == 0, edge); ++slot; }
        if (slot == 0) {                 // no frames at all: still finish stage 1
            asm volatile("cp.async.wait_all;\n" ::: "memory");
            __syncthreads();
        }
    }

    // ---- edge fix-ups: the terms ndi_shift's clamp creates at the top / left image edges ----------------
    if (top) {
        // residuals deposited a second time on HR row 0 by frames with sy = -1
        for (int t = tid; t < FT_W + 8; t += F_THREADS) {
            const int x = c.x0 - 4 + t;
            float v = 0.0f;
            if (x >= 0 && x < g.W) {
                for (int k = 0; k < g.n; ++k) {
                    if (g.sy[k] != -1) continue;
                    const int X = clampi2(x + g.sx[k], 0, g.W - 1);
                    if (X & 1) continue;
                    const int j = X >> 1;
                    if (j >= g.w || c.ilo > 0) continue;
                    v += __ldg(a.lr[k] + (size_t)(0 - g.lr_lo) * g.w + j) -
                         b_direct(a, hs, c.y0, c.x0, 1, clampi2(2 * j - g.sx[k], 0, g.W - 1));
                }
            }
            R0[t] = v;
        }
        // LR row 0 of frames with sy = +1 reads B[clamp(-1)] = B[0]: MSE only
        if (g.upd_y0 == 0 && c.ilo == 0) {
            for (int k = 0; k < g.n; ++k) {
                if (g.sy[k] != 1) continue;
                for (int j = tid; j < g.w; j += F_THREADS) {
                    const int cx = clampi2(2 * j - g.sx[k], 0, g.W - 1);
                    if (cx < c.ox0 || cx >= c.ox1) continue;
                    const float e = __ldg(a.lr[k] + (size_t)(0 - g.lr_lo) * g.w + j) - b_direct(a, hs, c.y0, c.x0, 0, cx);
                    e2 = fmaf(e, e, e2);
                }
            }
        }
    }
    if (left) {
        for (int t = tid; t < FT_H + 8; t += F_THREADS) {
            const int y = c.y0 - 4 + t;
            float v = 0.0f;
            if (y >= 0 && y < g.H) {
                for (int k = 0; k < g.n; ++k) {
                    if (g.sx[k] != -1) continue;
                    if (y + g.sy[k] < 0) continue;               // corner: already counted in R0
                    const int Y = clampi2(y + g.sy[k], 0, g.H - 1);
                    if (Y & 1) continue;
                    const int i = Y >> 1;
                    if (i >= c.ihi || i < c.ilo) continue;
                    v += __ldg(a.lr[k] + (size_t)(i - g.lr_lo) * g.w + 0) -
                         b_direct(a, hs, c.y0, c.x0, clampi2(2 * i - g.sy[k], 0, g.H - 1), 1);
                }
            }
            C0[t] = v;
        }
        // LR col 0 of frames with sx = +1 reads B[., clamp(-1)] = B[., 0]: MSE only (row-virtual ones were
        // counted by the top fix-up)
        for (int k = 0; k < g.n; ++k) {
            if (g.sx[k] != 1) continue;
            for (int t = tid; t < FT_H / 2 + 2; t += F_THREADS) {
                const int i = (c.y0 >> 1) - 1 + t;
                if (i < c.ilo || i >= c.ihi) continue;
                const int cy = 2 * i - g.sy[k];
                if (cy < 0 || cy < c.oy0 || cy >= c.oy1) continue;
                const float e = __ldg(a.lr[k] + (size_t)(i - g.lr_lo) * g.w + 0) - b_direct(a, hs, c.y0, c.x0, cy, 0);
                e2 = fmaf(e, e, e2);
            }
        }
    }
    __syncthreads();

    // ---- stage 3 -----------------------------------------------------------------------------------------
    for (int t = tid; t < (FT_H / 8) * (FT_W / 2); t += F_THREADS) {
        const int grp = t / (FT_W / 2), j = t - grp * (FT_W / 2);
        const int i0 = grp * 4;
        float o[4][2][2];
        #pragma unroll
        for (int cc = 0; cc < 4; ++cc) { o[cc][0][0] = o[cc][0][1] = o[cc][1][0] = o[cc][1][1] = 0.f; }
        int slot = 0;
        if (a.ph_n[0]) { stage_backward<0, 0, RR>(a, planes + slot * FA_R * FA_C, i0, j, o); ++slot; }
        if (a.ph_n[1]) { stage_backward<0, 1, RR>(a, planes + slot * FA_R * FA_C, i0, j, o); ++slot; }
        if (a.ph_n[2]) { stage_backward<1, 0, RR>(a, planes + slot * FA_R * FA_C, i0, j, o); ++slot; }
        if (a.ph_n[3]) { stage_backward<1, 1, RR>(a, planes + slot * FA_R * FA_C, i0, j, o); ++slot; }

        const int gx = c.x0 + 2 * j;
        #pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
            #pragma unroll
            for (int oy = 0; oy < 2; ++oy) {
                const int ly = 2 * (i0 + cc) + oy + 8;
                const int gy = c.y0 + 2 * (i0 + cc) + oy;
                float o0 = o[cc][oy][0], o1 = o[cc][oy][1];
                if (edge) {
                    if (gy >= c.oy1 || gx >= c.ox1) continue;
                    if (top && gy <= 3) {
                        for (int tb = 0; tb < 7; ++tb) {
                            const float w = a.psf[(3 - gy) * 7 + tb];
                            o0 = fmaf(w, R0[gx + tb - 3 - (c.x0 - 4)], o0);
                            o1 = fmaf(w, R0[gx + 1 + tb - 3 - (c.x0 - 4)], o1);
                        }
                    }
                    if (left && gx <= 3) {
                        for (int ta = 0; ta < 7; ++ta) {
                            const float cv = C0[gy + ta - 3 - (c.y0 - 4)];
                            if (3 - gx >= 0) o0 = fmaf(a.psf[ta * 7 + (3 - gx)], cv, o0);
                            if (3 - (gx + 1) >= 0) o1 = fmaf(a.psf[ta * 7 + (3 - (gx + 1))], cv, o1);
                        }
                    }
                }
                const float2 hv = *reinterpret_cast<const float2*>(hs + ly * FH_COLS + 2 * j + 8);
                float2 r;
                r.x = fminf(fmaxf(fmaf(g.alpha, o0, hv.x), g.lo), g.hi);
                r.y = fminf(fmaxf(fmaf(g.alpha, o1, hv.y), g.lo), g.hi);
                *reinterpret_cast<float2*>(a.hr_out + (size_t)(gy - g.hr_lo) * g.W + gx) = r;
            }
        }
    }

    // ---- residual energy of this tile --------------------------------------------------------------------
    #pragma unroll
    for (int off = 16; off > 0; off >>= 1) e2 += __shfl_xor_sync(0xffffffffu, e2, off);
    if ((tid & 31) == 0) red[tid >> 5] = e2;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        #pragma unroll
        for (int i = 0; i < F_THREADS / 32; ++i) t += (double)red[i];
        a.partial[blockIdx.y * gridDim.x + blockIdx.x] = t;
    }
}

int launch_ibp_fast(const FastArgs& a, cudaStream_t st) {
    if (a.tiles_x <= 0 || a.tiles_y <= 0) return SR459_OK;
    int nph = 0;
    for (int p = 0; p < 4; ++p) nph += a.ph_n[p] > 0;
    size_t smem = sizeof(float) * (FH_ROWS * FH_COLS + (size_t)nph * FA_R * FA_C);
    static PerDeviceMax configured;
    if (configured.raise(smem)) {
        SR_CUDA(cudaFuncSetAttribute(k_ibp_fast<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        SR_CUDA(cudaFuncSetAttribute(k_ibp_fast<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    // effective reach of the PSF: an empty outer ring halves the multiply-adds (25 instead of 49 taps)
    bool ring_empty = true;
    for (int i = 0; i < 7; ++i) ring_empty = ring_empty && a.psf[i] == 0.f && a.psf[42 + i] == 0.f && a.psf[7 * i] == 0.f && a.psf[7 * i + 6] == 0.f;
    dim3 grid(a.tiles_x, a.tiles_y);
    if (ring_empty) k_ibp_fast<2><<<grid, F_THREADS, smem, st>>>(a);
    else            k_ibp_fast<3><<<grid, F_THREADS, smem, st>>>(a);
    SR_LAUNCH_CHECK("k_ibp_fast");
    return SR459_OK;
}

}  // namespace sr459
