// ibp_ws.cu -- fused IBP iteration, integer HR shifts, separable PSF: warp-synchronous form.
//
// Same mathematics and the same three restructurings as ibp_sep.cu (polyphase lattices, co-sited
// frames pre-summed into S1 planes, rank-1 PSF); what changes is how the work is laid on the SM.
// ibp_sep.cu runs four CTA-wide phases separated by barriers and exchanges every intermediate
// through shared memory; at 24 resident warps per SM that is latency bound (each warp issues once
// in ~12 cycles).  Here a warp carries a lattice row from the staged hr tile all the way to the
// row-passed residual without leaving its registers:
//
//   lane l owns staged columns 4l..4l+3 (x = x0-8+4l..), i.e. two sites of every lattice row
//   column pass   7 staged rows (sliding register window, LDS.128)  -> C[4] per lane   (FFMA2)
//   row pass      C of the two neighbour lanes by warp shuffle      -> B at the lane's 2 sites
//   residual      S1 (LDS.64 from the TMA box) - m B                -> 2 values per lane
//   back row pass residuals of the neighbour lanes by shuffle       -> W[4] per lane -> shared
// One CTA barrier later the transposed column pass, update, clip and float4 stores run with
// lanes on 4-column groups and warps on 8-row bands.  Per tile: 1 mbarrier wait + 1 barrier.
//
// Tile 64 x 112 HR (the 128 staged columns are exactly one float4 per lane), 256 threads,
// ~70 KB of shared memory -> 3 CTAs per SM.  Reference: ibp(), mono_cal_target/run_sr.py:190-209.
#include "internal.h"
#include "tma.h"

#include <algorithm>

namespace sr459 {

constexpr int WT_H = 64, WT_W = 112;              // HR tile
constexpr int WH_TOP = 7;                         // staged hr rows y0-7 .. y0+70
constexpr int WH_ROWS = WT_H + 14, WH_COLS = 128; //   cols x0-8 .. x0+119
constexpr int WA_R = WT_H / 2 + 4;                // 36 lattice rows  ys = y0-4+2a+P
constexpr int WA_C = 64;                          // S1 box columns: lane l reads sites 2l, 2l+1 (staged col 4l+Q, 4l+2+Q)
constexpr int WW_P = 128;                         // pitch of the W planes
constexpr int W_THREADS = 256;
constexpr int WSM_HS  = WH_ROWS * WH_COLS;        // floats
constexpr int WSM_ACC = WA_R * WA_C;
constexpr int WSM_WD  = WA_R * WW_P;
static_assert((WSM_HS * 4) % 128 == 0 && (WSM_ACC * 4) % 128 == 0, "TMA destinations must be 128 B aligned");

void ibp_ws_tile(int* th, int* tw) { *th = WT_H; *tw = WT_W; }
int ws_make_hr_map(CUtensorMap* out, const float* base, int W, int rows) {
    return encode_tile_map_f32(out, base, (uint64_t)W, (uint64_t)rows, (uint64_t)W, WH_COLS, WH_ROWS);
}
int ws_make_s1_map(CUtensorMap* out, const float* base, int w, int rows, int pitch) {
    return encode_tile_map_f32(out, base, (uint64_t)w, (uint64_t)rows, (uint64_t)pitch, WA_C, WA_R);
}

struct WsTile {
    int y0, x0;                 // tile origin (global HR)
    int oy0, oy1, ox0, ox1;     // pixels / sites this tile owns (global HR, half open)
    int ilo, ihi;               // site rows whose S1 sums exist
};

__device__ __forceinline__ void wffma2(float& d0, float& d1, float w, float x0, float x1) {
    asm("{\n.reg .b64 a, b, c;\nmov.b64 a, {%2, %2};\nmov.b64 b, {%3, %4};\nmov.b64 c, {%0, %1};\n"
        "fma.rn.f32x2 c, a, b, c;\nmov.b64 {%0, %1}, c;\n}\n" : "+f"(d0), "+f"(d1) : "f"(w), "f"(x0), "f"(x1));
}

// forward + backward row pass of R consecutive lattice rows a0 .. a0+R-1 of lattice (P, Q); one warp
template <int P, int Q, int R, bool EDGE>
__device__ __forceinline__ float ws_rows(const SepArgs& a, const WsTile& c, int slot, int lane, int a0,
                                         const float* __restrict__ hs, const float* __restrict__ s1box,
                                         float* __restrict__ wd, bool add) {
    const IbpGeom& g = a.g;
    const unsigned FULL = 0xffffffffu;
    const float m_int = a.mtab[slot][0];
    const float4* hp = reinterpret_cast<const float4*>(hs + (2 * a0 + P) * WH_COLS) + lane;
    float4 v[2 * R + 5];                              // sliding window: rows 2j .. 2j+6 are live in step j
    #pragma unroll
    for (int r = 0; r < 7; ++r) v[r] = hp[r * (WH_COLS / 4)];
    float e2 = 0.0f;
    #pragma unroll
    for (int j = 0; j < R; ++j) {
        if (j + 1 < R) {                                // prefetch the two rows the next step adds
            v[2 * j + 7] = hp[(2 * j + 7) * (WH_COLS / 4)];
            v[2 * j + 8] = hp[(2 * j + 8) * (WH_COLS / 4)];
        }
        // ---- column pass: C[x] = sum_u acol[u] * hs[ly_s + 3 - u][x],  ly_s - (2 a0 + P) = 3 + 2j ----------
        float4 C = make_float4(0.f, 0.f, 0.f, 0.f);
        #pragma unroll
        for (int t = 0; t < 7; ++t) {
            const float4 x = v[6 + 2 * j - t];
            wffma2(C.x, C.y, a.acol[t], x.x, x.y);
            wffma2(C.z, C.w, a.acol[t], x.z, x.w);
        }
        // ---- row pass: B at the lane's sites (staged cols 4l+Q, 4l+2+Q) --------------------------------------
        float f[9];                                     // staged cols 4l+Q-3 .. 4l+Q+5
        if (Q == 0) {
            f[0] = __shfl_up_sync(FULL, C.y, 1); f[1] = __shfl_up_sync(FULL, C.z, 1); f[2] = __shfl_up_sync(FULL, C.w, 1);
            f[3] = C.x; f[4] = C.y; f[5] = C.z; f[6] = C.w;
            f[7] = __shfl_down_sync(FULL, C.x, 1); f[8] = __shfl_down_sync(FULL, C.y, 1);
        } else {
            f[0] = __shfl_up_sync(FULL, C.z, 1); f[1] = __shfl_up_sync(FULL, C.w, 1);
            f[2] = C.x; f[3] = C.y; f[4] = C.z; f[5] = C.w;
            f[6] = __shfl_down_sync(FULL, C.x, 1); f[7] = __shfl_down_sync(FULL, C.y, 1); f[8] = __shfl_down_sync(FULL, C.z, 1);
        }
        float B1 = 0.0f, B2 = 0.0f;
        #pragma unroll
        for (int t = 0; t < 7; ++t) {                   // B(site) = sum_v brow[v] * C[site + 3 - v]
            B1 = fmaf(a.brow[t], f[6 - t], B1);
            B2 = fmaf(a.brow[t], f[8 - t], B2);
        }
        // ---- residual ---------------------------------------------------------------------------------------------
        const int ar = a0 + j;
        const float2 s1 = *reinterpret_cast<const float2*>(s1box + ar * WA_C + 2 * lane);
        float r1, r2;
        if (!EDGE) {
            r1 = fmaf(-m_int, B1, s1.x);
            r2 = fmaf(-m_int, B2, s1.y);
            if (ar >= 2 && ar < WA_R - 2 && lane >= 2 && lane < 30) e2 = fmaf(r1, r1, fmaf(r2, r2, e2));
        } else {
            const int ys = c.y0 - 4 + 2 * ar + P;
            const int ag = (ys - P) >> 1;
            const bool rv = (ys >= 0) && (ys < g.H) && (ag >= c.ilo) && (ag < c.ihi);
            const bool rown = rv && (ys >= c.oy0) && (ys < c.oy1);
            const int ra = (ag == g.h - 1) ? 2 : 0;
            const int xs1 = c.x0 - 8 + 4 * lane + Q, xs2 = xs1 + 2;
            const bool v1 = rv && (xs1 >= 0) && (xs1 < g.W), v2 = rv && (xs2 >= 0) && (xs2 < g.W);
            const float m1 = a.mtab[slot][ra + ((((xs1 - Q) >> 1) == g.w - 1) ? 1 : 0)];
            const float m2 = a.mtab[slot][ra + ((((xs2 - Q) >> 1) == g.w - 1) ? 1 : 0)];
            r1 = v1 ? fmaf(-m1, B1, s1.x) : 0.0f;
            r2 = v2 ? fmaf(-m2, B2, s1.y) : 0.0f;
            if (v1 && rown && xs1 >= c.ox0 && xs1 < c.ox1 && m1 > 0.0f) e2 += r1 * r1 / m1;
            if (v2 && rown && xs2 >= c.ox0 && xs2 < c.ox1 && m2 > 0.0f) e2 += r2 * r2 / m2;
        }
        // ---- backward row pass: W[x] = sum_sites brow[3 + xs - x] * res(xs), |xs - x| <= 3, x = 4l..4l+3 ---------
        float4 Wv;
        if (Q == 0) {                                   // sites at staged cols 4l-2 | 4l, 4l+2 | 4l+4, 4l+6
            const float L2 = __shfl_up_sync(FULL, r2, 1), R1 = __shfl_down_sync(FULL, r1, 1), R2 = __shfl_down_sync(FULL, r2, 1);
            Wv.x = a.brow[1] * L2 + a.brow[3] * r1 + a.brow[5] * r2;
            Wv.y = a.brow[0] * L2 + a.brow[2] * r1 + a.brow[4] * r2 + a.brow[6] * R1;
            Wv.z = a.brow[1] * r1 + a.brow[3] * r2 + a.brow[5] * R1;
            Wv.w = a.brow[0] * r1 + a.brow[2] * r2 + a.brow[4] * R1 + a.brow[6] * R2;
        } else {                                        // sites at staged cols 4l-3, 4l-1 | 4l+1, 4l+3 | 4l+5
            const float L1 = __shfl_up_sync(FULL, r1, 1), L2 = __shfl_up_sync(FULL, r2, 1), R1 = __shfl_down_sync(FULL, r1, 1);
            Wv.x = a.brow[0] * L1 + a.brow[2] * L2 + a.brow[4] * r1 + a.brow[6] * r2;
            Wv.y = a.brow[1] * L2 + a.brow[3] * r1 + a.brow[5] * r2;
            Wv.z = a.brow[0] * L2 + a.brow[2] * r1 + a.brow[4] * r2 + a.brow[6] * R1;
            Wv.w = a.brow[1] * r1 + a.brow[3] * r2 + a.brow[5] * R1;
        }
        float4* d = reinterpret_cast<float4*>(wd + ar * WW_P) + lane;
        if (add) { const float4 e = *d; Wv.x += e.x; Wv.y += e.y; Wv.z += e.z; Wv.w += e.w; }
        *d = Wv;
    }
    return EDGE ? e2 : e2 / m_int;
}

template <int P, int Q, bool EDGE>
__device__ __forceinline__ float ws_forward(const SepArgs& a, const WsTile& c, int slot, int warp, int lane,
                                            const float* hs, const float* s1box, float* wd, bool add) {
    // 36 lattice rows over 8 warps: 5, 5, 5, 5, 4, 4, 4, 4
    if (warp < 4) return ws_rows<P, Q, 5, EDGE>(a, c, slot, lane, 5 * warp, hs, s1box, wd, add);
    return ws_rows<P, Q, 4, EDGE>(a, c, slot, lane, 20 + 4 * (warp - 4), hs, s1box, wd, add);
}

// transposed column pass of the lattice rows of parity P into an 8-row x 4-column block
template <int P>
__device__ __forceinline__ void ws_colpass(const SepArgs& a, const float* __restrict__ wd, int warp, int lane, float (&o)[8][4]) {
    const float4* p = reinterpret_cast<const float4*>(wd + (4 * warp) * WW_P) + lane;
    #pragma unroll
    for (int tt = 0; tt < 8; ++tt) {
        bool used = false;
        #pragma unroll
        for (int jj = 0; jj < 8; ++jj) { const int diff = 2 * tt + P - 4 - jj; used = used || (diff >= -3 && diff <= 3); }
        if (!used) continue;
        const float4 x = p[tt * (WW_P / 4)];
        #pragma unroll
        for (int jj = 0; jj < 8; ++jj) {
            const int diff = 2 * tt + P - 4 - jj;             // ys_a - y for a = 4 warp + tt, y = y0 + 8 warp + jj
            if (diff >= -3 && diff <= 3) {
                wffma2(o[jj][0], o[jj][1], a.acol[3 + diff], x.x, x.y);
                wffma2(o[jj][2], o[jj][3], a.acol[3 + diff], x.z, x.w);
            }
        }
    }
}

__device__ __noinline__ float ws_b_direct(const SepArgs& a, const float* __restrict__ hs, int y0, int x0, int gy, int gx) {
    const float* c = hs + (gy - (y0 - WH_TOP)) * WH_COLS + (gx - (x0 - 8));
    float v = 0.0f;
    for (int u = 0; u < 7; ++u)
        for (int w = 0; w < 7; ++w) v = fmaf(a.psf[u * 7 + w], c[(3 - u) * WH_COLS + (3 - w)], v);
    return v;
}

__device__ __forceinline__ int wclamp(int v, int lo, int hi) { return min(max(v, lo), hi); }

// the terms ndi_shift's clamp creates at the top / left image edges (first tile row / column only);
// see ibp_fast.cu for the derivation.  R0[t]: x = x0-4+t, C0[t]: y = y0-4+t.
__device__ __noinline__ float ws_edge_fixups(const SepArgs& a, int tid, int y0, int x0, int oy0, int oy1, int ox0, int ox1,
                                             int ilo, int ihi, const float* __restrict__ hs,
                                             float* __restrict__ R0, float* __restrict__ C0, bool top, bool left) {
    const IbpGeom& g = a.g;
    float e2 = 0.0f;
    if (top) {
        for (int t = tid; t < WT_W + 8; t += W_THREADS) {
            const int x = x0 - 4 + t;
            float v = 0.0f;
            if (x >= 0 && x < g.W) {
                for (int k = 0; k < g.n; ++k) {
                    if (g.sy[k] != -1) continue;
                    const int X = wclamp(x + g.sx[k], 0, g.W - 1);
                    if (X & 1) continue;
                    const int j = X >> 1;
                    if (j >= g.w || ilo > 0) continue;
                    v += __ldg(a.lr[k] + (size_t)(0 - g.lr_lo) * g.w + j) -
                         ws_b_direct(a, hs, y0, x0, 1, wclamp(2 * j - g.sx[k], 0, g.W - 1));
                }
            }
            R0[t] = v;
        }
        if (g.upd_y0 == 0 && ilo == 0) {               // LR row 0 of frames with sy = +1 reads B[0]: MSE only
            for (int k = 0; k < g.n; ++k) {
                if (g.sy[k] != 1) continue;
                for (int j = tid; j < g.w; j += W_THREADS) {
                    const int cx = wclamp(2 * j - g.sx[k], 0, g.W - 1);
                    if (cx < ox0 || cx >= ox1) continue;
                    const float e = __ldg(a.lr[k] + (size_t)(0 - g.lr_lo) * g.w + j) - ws_b_direct(a, hs, y0, x0, 0, cx);
                    e2 = fmaf(e, e, e2);
                }
            }
        }
    }
    if (left) {
        for (int t = tid; t < WT_H + 8; t += W_THREADS) {
            const int y = y0 - 4 + t;
            float v = 0.0f;
            if (y >= 0 && y < g.H) {
                for (int k = 0; k < g.n; ++k) {
                    if (g.sx[k] != -1) continue;
                    if (y + g.sy[k] < 0) continue;               // corner: already counted in R0
                    const int Y = wclamp(y + g.sy[k], 0, g.H - 1);
                    if (Y & 1) continue;
                    const int i = Y >> 1;
                    if (i >= ihi || i < ilo) continue;
                    v += __ldg(a.lr[k] + (size_t)(i - g.lr_lo) * g.w + 0) -
                         ws_b_direct(a, hs, y0, x0, wclamp(2 * i - g.sy[k], 0, g.H - 1), 1);
                }
            }
            C0[t] = v;
        }
        for (int k = 0; k < g.n; ++k) {                // LR col 0 of frames with sx = +1 reads B[., 0]: MSE only
            if (g.sx[k] != 1) continue;
            for (int t = tid; t < WT_H / 2 + 2; t += W_THREADS) {
                const int i = (y0 >> 1) - 1 + t;
                if (i < ilo || i >= ihi) continue;
                const int cy = 2 * i - g.sy[k];
                if (cy < 0 || cy < oy0 || cy >= oy1) continue;
                const float e = __ldg(a.lr[k] + (size_t)(i - g.lr_lo) * g.w + 0) - ws_b_direct(a, hs, y0, x0, cy, 0);
                e2 = fmaf(e, e, e2);
            }
        }
    }
    return e2;
}

__global__ void __launch_bounds__(W_THREADS, 3) k_ibp_ws(const __grid_constant__ SepArgs a) {
    extern __shared__ __align__(128) float sm[];
    __shared__ __align__(8) uint64_t mbar;
    __shared__ float R0[WT_W + 8], C0[WT_H + 8];
    __shared__ __align__(16) float WR[WT_W];
    __shared__ float WC[WT_H];
    __shared__ float red[W_THREADS / 32];
    const IbpGeom& g = a.g;
    float* hs   = sm;
    float* accs = sm + WSM_HS;                         // S1 boxes, one per active lattice
    float* wds  = accs + a.nph * WSM_ACC;              // W planes, one per active row parity
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    WsTile c;
    c.y0 = a.ty_origin + blockIdx.y * WT_H;
    c.x0 = blockIdx.x * WT_W;
    c.oy0 = max(c.y0, g.upd_y0); c.oy1 = min(min(c.y0 + WT_H, g.upd_y1), g.H);
    c.ox0 = c.x0;                c.ox1 = min(c.x0 + WT_W, g.W);
    c.ilo = max(0, g.lr_lo);     c.ihi = min(g.h, g.lr_hi);

    if (tid == 0 && a.dbg != 2) {
        mbar_init(&mbar, 1);
        mbar_fence_init();
        mbar_expect_tx(&mbar, (uint32_t)(4 * (WSM_HS + a.nph * WSM_ACC)));
        tma_load_2d(hs, &a.tm_hr, c.x0 - 8, c.y0 - WH_TOP - g.hr_lo, &mbar);
        for (int s = 0; s < a.nph; ++s)
            tma_load_2d(accs + s * WSM_ACC, &a.tm_s1[s], (c.x0 >> 1) - 4, (c.y0 >> 1) - 2 - a.s1_row0, &mbar);
    }
    const bool top = (c.y0 == 0), left = (c.x0 == 0);
    const bool edge = (c.y0 < 8) || (c.x0 < 8) || (c.y0 + WT_H + 5 > g.H) || (c.x0 + WT_W + 9 > g.W) ||
                      (c.y0 + WT_H > g.upd_y1) || (c.y0 < g.upd_y0) ||
                      (((c.y0 - 6) >> 1) < c.ilo) || (((c.y0 + WT_H + 4) >> 1) + 1 > c.ihi);
    __syncthreads();
    if (a.dbg != 2) mbar_wait(&mbar, 0);
    if (a.dbg == 1) {                                  // probe: memory path only
        if (lane >= 2 && lane < 30) {
            const int gx = c.x0 - 8 + 4 * lane;
            for (int jj = 0; jj < 8; ++jj) {
                const int gy = c.y0 + 8 * warp + jj;
                if (gy < c.oy1 && gx < c.ox1)
                    *reinterpret_cast<float4*>(a.hr_out + (size_t)(gy - g.hr_lo) * g.W + gx) =
                        *reinterpret_cast<const float4*>(hs + (8 * warp + jj + WH_TOP) * WH_COLS + 4 * lane);
            }
        }
        return;
    }

    // ---- forward + backward row pass, lattice by lattice (W planes: parity-0 rows first) -------------------
    float e2 = 0.0f;
    int wsel0 = -1, wsel1 = -1, nplanes = 0;
    for (int s = 0; s < a.nph; ++s) {
        const int P = a.ph_pq[s] >> 1, Q = a.ph_pq[s] & 1;
        int sel = P ? wsel1 : wsel0;
        const bool add = sel >= 0;
        if (!add) { sel = nplanes++; if (P) wsel1 = sel; else wsel0 = sel; }
        float* wd = wds + sel * WSM_WD;
        const float* s1box = accs + s * WSM_ACC;
        const int pq = a.ph_pq[s];
        if (edge) {
            switch (pq) {
                case 0: e2 += ws_forward<0, 0, true>(a, c, s, warp, lane, hs, s1box, wd, add); break;
                case 1: e2 += ws_forward<0, 1, true>(a, c, s, warp, lane, hs, s1box, wd, add); break;
                case 2: e2 += ws_forward<1, 0, true>(a, c, s, warp, lane, hs, s1box, wd, add); break;
                default: e2 += ws_forward<1, 1, true>(a, c, s, warp, lane, hs, s1box, wd, add); break;
            }
        } else {
            switch (pq) {
                case 0: e2 += ws_forward<0, 0, false>(a, c, s, warp, lane, hs, s1box, wd, add); break;
                case 1: e2 += ws_forward<0, 1, false>(a, c, s, warp, lane, hs, s1box, wd, add); break;
                case 2: e2 += ws_forward<1, 0, false>(a, c, s, warp, lane, hs, s1box, wd, add); break;
                default: e2 += ws_forward<1, 1, false>(a, c, s, warp, lane, hs, s1box, wd, add); break;
            }
        }
        (void)Q;
    }
    if (top || left) {
        e2 += ws_edge_fixups(a, tid, c.y0, c.x0, c.oy0, c.oy1, c.ox0, c.ox1, c.ilo, c.ihi, hs, R0, C0, top, left);
        __syncthreads();
        // separable form of the duplicates: corr[y][x] += acol[3-y] * WR[x] (y <= 3) + brow[3-x] * WC[y] (x <= 3)
        if (top) {
            for (int t = tid; t < WT_W; t += W_THREADS) {
                float v = 0.0f;
                #pragma unroll
                for (int tb = 0; tb < 7; ++tb) v = fmaf(a.brow[tb], R0[t + tb + 1], v);
                WR[t] = v;
            }
        }
        if (left) {
            for (int t = tid; t < WT_H; t += W_THREADS) {
                float v = 0.0f;
                #pragma unroll
                for (int ta = 0; ta < 7; ++ta) v = fmaf(a.acol[ta], C0[t + ta + 1], v);
                WC[t] = v;
            }
        }
    }
    __syncthreads();

    // ---- transposed column pass, update, clip, store: warp = 8 output rows, lane = 4 columns -----------------
    if (lane >= 2 && lane < 30) {
        float o[8][4];
        #pragma unroll
        for (int jj = 0; jj < 8; ++jj) { o[jj][0] = o[jj][1] = o[jj][2] = o[jj][3] = 0.f; }
        if (wsel0 >= 0) ws_colpass<0>(a, wds + wsel0 * WSM_WD, warp, lane, o);
        if (wsel1 >= 0) ws_colpass<1>(a, wds + wsel1 * WSM_WD, warp, lane, o);
        const int gx = c.x0 - 8 + 4 * lane;
        float* outp = a.hr_out + (size_t)(c.y0 + 8 * warp - g.hr_lo) * g.W + gx;
        const float* hp = hs + (8 * warp + WH_TOP) * WH_COLS + 4 * lane;
        #pragma unroll
        for (int jj = 0; jj < 8; ++jj) {
            const int gy = c.y0 + 8 * warp + jj;
            if (edge) {
                if (gy < c.oy0 || gy >= c.oy1 || gx >= c.ox1) continue;
                if (top && gy <= 3) {
                    const float wgt = a.acol[3 - gy];
                    const float4 wr = *reinterpret_cast<const float4*>(WR + 4 * (lane - 2));
                    o[jj][0] = fmaf(wgt, wr.x, o[jj][0]); o[jj][1] = fmaf(wgt, wr.y, o[jj][1]);
                    o[jj][2] = fmaf(wgt, wr.z, o[jj][2]); o[jj][3] = fmaf(wgt, wr.w, o[jj][3]);
                }
                if (left && lane == 2) {
                    const float cv = WC[8 * warp + jj];
                    #pragma unroll
                    for (int i = 0; i < 4; ++i) o[jj][i] = fmaf(a.brow[3 - i], cv, o[jj][i]);
                }
            }
            float4 r = *reinterpret_cast<const float4*>(hp + jj * WH_COLS);
            wffma2(r.x, r.y, g.alpha, o[jj][0], o[jj][1]);
            wffma2(r.z, r.w, g.alpha, o[jj][2], o[jj][3]);
            r.x = fminf(fmaxf(r.x, g.lo), g.hi);
            r.y = fminf(fmaxf(r.y, g.lo), g.hi);
            r.z = fminf(fmaxf(r.z, g.lo), g.hi);
            r.w = fminf(fmaxf(r.w, g.lo), g.hi);
            *reinterpret_cast<float4*>(outp + (size_t)jj * g.W) = r;
        }
    }

    // ---- residual energy of this tile ----------------------------------------------------------------------
    #pragma unroll
    for (int off = 16; off > 0; off >>= 1) e2 += __shfl_xor_sync(0xffffffffu, e2, off);
    if (lane == 0) red[warp] = e2;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        #pragma unroll
        for (int i = 0; i < W_THREADS / 32; ++i) t += (double)red[i];
        a.partial[blockIdx.y * gridDim.x + blockIdx.x] = t;
    }
}

int launch_ibp_ws(const SepArgs& a, cudaStream_t st) {
    if (a.tiles_x <= 0 || a.tiles_y <= 0) return SR459_OK;
    int rowpar = 0;
    for (int s = 0; s < a.nph; ++s) rowpar |= 1 << (a.ph_pq[s] >> 1);
    const size_t smem = sizeof(float) * ((size_t)WSM_HS + (size_t)a.nph * WSM_ACC + (size_t)(rowpar == 3 ? 2 : 1) * WSM_WD);
    static PerDeviceMax configured;
    if (configured.raise(smem)) {
        SR_CUDA(cudaFuncSetAttribute(k_ibp_ws, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    dim3 grid(a.tiles_x, a.tiles_y);
    k_ibp_ws<<<grid, W_THREADS, smem, st>>>(a);
    SR_LAUNCH_CHECK("k_ibp_ws");
    return SR459_OK;
}

}  // namespace sr459
