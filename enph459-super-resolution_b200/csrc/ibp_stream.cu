// ibp_stream.cu -- fused IBP iteration, integer HR shifts, rank-1 PSF, one parity lattice:
// register-streaming form.  Reference: ibp(), mono_cal_target/run_sr.py:190-209 (closed form in
// SURVEY.md section 8a; the three exact restructurings are described in ibp_sep.cu / DESIGN.md 3.1).
//
// Every WARP is an independent pipeline that walks DOWN a 128-column strip of the HR image:
//
//   TMA      hr rows (8 per stage, zero fill outside the image = fftconvolve's zero padding) and the
//            S1 rows of the lattice (4 per stage) land in a 3-stage shared-memory ring private to
//            the warp; one mbarrier per stage, issued by lane 0 one whole stage ahead of use.
//   window   lane l owns staged columns 4l..4l+3 and keeps a sliding window of 7 hr rows in
//            registers (two LDS.128 per step).
//   step     one lattice row ys per step:
//              column pass  7 window rows -> C[4]                       (packed FFMA2)
//              row pass     C of the neighbour lanes by shuffle -> B at the lane's 2 sites
//              residual     S1 - m B, residual energy
//              transposed row pass  residuals of the neighbour lanes by shuffle -> W[4]
//              transposed column pass over a 4-row register ring of W -> HR rows ys-3, ys-2
//              update, clip, float4 stores
// There is no CTA barrier and no intermediate ever touches shared memory; a 128-thread CTA is just
// four such warps, three CTAs per SM.  The ndi_shift clamp terms (duplicates on HR row 0 / column 0,
// MSE-only terms on LR row 0 / column 0) enter through four pre-summed edge vectors (k_edge_sums,
// once per ibp() call) in the first step of the top segments and in the left strip.
#include "internal.h"
#include "tma.h"

#include <stdlib.h>
#include <algorithm>

namespace sr459 {

typedef unsigned long long u64;

constexpr int SS_COLS   = 128;                 // staged columns per strip
constexpr int SS_VALID  = 112;                 // columns a strip updates (lanes 2..29)
constexpr int SS_ROWS   = 8;                   // hr rows per stage
constexpr int SS_SROWS  = 4;                   // S1 rows per stage
constexpr int SS_STAGES = 3;
constexpr int SS_HR_FLOATS = SS_ROWS * SS_COLS;        // 1024 floats = 4 KB
constexpr int SS_S1_FLOATS = SS_SROWS * 64;            // 256 floats  = 1 KB
constexpr int SS_STAGE_FLOATS = SS_HR_FLOATS + SS_S1_FLOATS;
constexpr int SS_WARPS   = 4;
constexpr int SS_THREADS = 32 * SS_WARPS;
constexpr int SS_CTAS    = 3;

int stream_make_hr_map(CUtensorMap* out, const float* base, int W, int rows) {
    return encode_tile_map_f32(out, base, (uint64_t)W, (uint64_t)rows, (uint64_t)W, SS_COLS, SS_ROWS);
}
int stream_make_s1_map(CUtensorMap* out, const float* base, int w, int rows, int pitch) {
    return encode_tile_map_f32(out, base, (uint64_t)w, (uint64_t)rows, (uint64_t)pitch, 64, SS_SROWS);
}
bool ibp_stream_supported(int kh, int kw, int smax, bool separable, int W, int nslot) {
    return separable && kh == kw && (kh == 1 || kh == 3 || kh == 5 || kh == 7) && smax <= 1 && (W % 4) == 0 && nslot == 1;
}

// ---- packed f32x2 helpers ------------------------------------------------------------------------------
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(u64 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

struct Row { u64 lo, hi; };                     // four consecutive floats: (x, y) (z, w)
__device__ __forceinline__ Row lds_row(const float* p) {
    const ulonglong2 t = *reinterpret_cast<const ulonglong2*>(p);
    Row r; r.lo = t.x; r.hi = t.y; return r;
}

// ---- edge sums: the frames' samples that ndi_shift's clamp sends to row 0 / column 0 ------------------------
// (see the header comment of ibp_fast.cu for the derivation; all frames sit on ONE lattice here)
//   colS1[a]  = sum over frames with sx = -1 of lr[a + oy][0]   duplicate residuals on HR column 0
//   leftS1[a] = sum over frames with sx = +1 of lr[a + oy][0]   LR column 0 reads B[., 0]: MSE only
//   rowS1[b]  = sum over frames with sy = -1 of lr[0][b + ox]   duplicate residuals on HR row 0
//   topS1[b]  = sum over frames with sy = +1 of lr[0][b + ox]   LR row 0 reads B[0, .]: MSE only
//   scal[0]   = sum over frames (sy, sx) = (-1, -1) of lr[0][0] duplicate at HR (0, 0)
//   scal[1]   = sum over frames (sy, sx) = (+1, +1) of lr[0][0] reads B[0, 0]: MSE only
// and the within-group variances of the MSE-only groups are added to *vsum (sum of V of k_lattice_sums).
struct EdgeSumArgs {
    const float* lr[kMaxFrames];
    int n, h, w;
    int sy[kMaxFrames], sx[kMaxFrames];
    int oy[kMaxFrames], ox[kMaxFrames];
    float* colS1; float* leftS1; float* rowS1; float* topS1; float* scal;
    double* vsum;
    int own_a0, own_a1;       // local site rows whose MSE-only column terms this launch counts (row bands)
    int top_owned;            // the band holds LR row 0: the row vectors' and the corner's variances count
};

__global__ void __launch_bounds__(256) k_edge_sums(const __grid_constant__ EdgeSumArgs a) {
    const int t = blockIdx.x * 256 + threadIdx.x;
    double var = 0.0;
    if (t < a.h) {                                    // column vectors, site row index a = t
        float dup = 0.f, s = 0.f, x[kMaxFrames]; int m = 0;
        #pragma unroll
        for (int k = 0; k < kMaxFrames; ++k) {
            x[k] = 0.f;
            if (k >= a.n) continue;
            const int i = t + a.oy[k];
            if (i >= a.h) continue;
            const float v = __ldg(a.lr[k] + (size_t)i * a.w);
            if (a.sx[k] == -1) dup += v;
            if (a.sx[k] == +1) { x[k] = v; s += v; m |= 1 << k; }
        }
        a.colS1[t] = dup; a.leftS1[t] = s;
        if (m && t >= a.own_a0 && t < a.own_a1) { const float mu = s / (float)__popc(m);
            #pragma unroll
            for (int k = 0; k < kMaxFrames; ++k) if (m & (1 << k)) { const float d = x[k] - mu; var += (double)d * d; } }
    }
    if (t < a.w) {                                    // row vectors, site column index b = t
        float dup = 0.f, s = 0.f, x[kMaxFrames]; int m = 0;
        #pragma unroll
        for (int k = 0; k < kMaxFrames; ++k) {
            x[k] = 0.f;
            if (k >= a.n) continue;
            const int j = t + a.ox[k];
            if (j >= a.w) continue;
            const float v = __ldg(a.lr[k] + j);
            if (a.sy[k] == -1) dup += v;
            if (a.sy[k] == +1) { x[k] = v; s += v; m |= 1 << k; }
        }
        a.rowS1[t] = dup; a.topS1[t] = s;
        if (m && a.top_owned) { const float mu = s / (float)__popc(m);
            #pragma unroll
            for (int k = 0; k < kMaxFrames; ++k) if (m & (1 << k)) { const float d = x[k] - mu; var += (double)d * d; } }
    }
    if (t == 0) {
        float dup = 0.f, s = 0.f, x[kMaxFrames]; int m = 0;
        for (int k = 0; k < a.n; ++k) {
            const float v = __ldg(a.lr[k]);
            x[k] = v;
            if (a.sy[k] == -1 && a.sx[k] == -1) dup += v;
            if (a.sy[k] == +1 && a.sx[k] == +1) { s += v; m |= 1 << k; }
        }
        a.scal[0] = dup; a.scal[1] = s;
        if (m && a.top_owned) { const float mu = s / (float)__popc(m);
            for (int k = 0; k < a.n; ++k) if (m & (1 << k)) { const float d = x[k] - mu; var += (double)d * d; } }
    }
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) var += __shfl_xor_sync(0xffffffffu, var, o);
    if ((threadIdx.x & 31) == 0 && var != 0.0) atomicAdd(a.vsum, var);
}

int launch_edge_sums(const EdgeSumArgs& a, cudaStream_t st) {
    const int n = std::max(a.h, a.w);
    k_edge_sums<<<(n + 255) / 256, 256, 0, st>>>(a);
    SR_LAUNCH_CHECK("k_edge_sums");
    return SR459_OK;
}

// ---- the streaming kernel ----------------------------------------------------------------------------------
struct LaneConst {
    u64 A[7];                 // (acol[u], acol[u])
    float b[7];               // brow
    u64 alpha2;
    float lo, hi;
    float mI1, mI2, mL1, mL2; // frames with a sample at the lane's two sites: interior rows / last site row
    float wI1, wI2, wL1, wL2; // 1/m if the site is owned (MSE), else 0
    float wlI, wlL;           // LEFT strip, lane 2: 1/ml (interior rows / last site row), else 0
    float lb[4];              // LEFT strip, lane 2: brow[3], brow[2], brow[1], brow[0]; else 0
    bool  st_ok;              // this lane stores (owned columns inside the image)
};

struct State {
    Row v[8];                 // hr window: slot (row - sb) mod 8
    Row wr[4];                // W ring: slot step mod 4
    Row w0;                   // W of the duplicate site row ys = 0 (top segments only)
    float e2;
};

// B at the lane's two sites (staged cols 4l+Q, 4l+2+Q) from the column-passed C of this and the neighbour lanes
template <int Q>
__device__ __forceinline__ void row_pass(const LaneConst& k, float Cx, float Cy, float Cz, float Cw, float& B1, float& B2) {
    const unsigned FULL = 0xffffffffu;
    float f[9];                                         // staged cols 4l+Q-3 .. 4l+Q+5
    if (Q == 0) {
        f[0] = __shfl_up_sync(FULL, Cy, 1); f[1] = __shfl_up_sync(FULL, Cz, 1); f[2] = __shfl_up_sync(FULL, Cw, 1);
        f[3] = Cx; f[4] = Cy; f[5] = Cz; f[6] = Cw;
        f[7] = __shfl_down_sync(FULL, Cx, 1); f[8] = __shfl_down_sync(FULL, Cy, 1);
    } else {
        f[0] = __shfl_up_sync(FULL, Cz, 1); f[1] = __shfl_up_sync(FULL, Cw, 1);
        f[2] = Cx; f[3] = Cy; f[4] = Cz; f[5] = Cw;
        f[6] = __shfl_down_sync(FULL, Cx, 1); f[7] = __shfl_down_sync(FULL, Cy, 1); f[8] = __shfl_down_sync(FULL, Cz, 1);
    }
    // B(site) = sum_v brow[v] * C[site + 3 - v]; two partial sums per site keep the chains short
    float a1 = k.b[0] * f[6], a2 = k.b[0] * f[8], c1 = k.b[1] * f[5], c2 = k.b[1] * f[7];
    a1 = fmaf(k.b[2], f[4], a1); a2 = fmaf(k.b[2], f[6], a2); c1 = fmaf(k.b[3], f[3], c1); c2 = fmaf(k.b[3], f[5], c2);
    a1 = fmaf(k.b[4], f[2], a1); a2 = fmaf(k.b[4], f[4], a2); c1 = fmaf(k.b[5], f[1], c1); c2 = fmaf(k.b[5], f[3], c2);
    a1 = fmaf(k.b[6], f[0], a1); a2 = fmaf(k.b[6], f[2], a2);
    B1 = a1 + c1; B2 = a2 + c2;
}

// W[x] = sum_sites brow[3 + xs - x] * res(xs), |xs - x| <= 3, at the lane's columns x = 4l..4l+3
template <int Q>
__device__ __forceinline__ void back_row(const LaneConst& k, float r1, float r2, float& Wx, float& Wy, float& Wz, float& Ww) {
    const unsigned FULL = 0xffffffffu;
    if (Q == 0) {                                       // sites at staged cols 4l-2 | 4l, 4l+2 | 4l+4, 4l+6
        const float L2 = __shfl_up_sync(FULL, r2, 1), R1 = __shfl_down_sync(FULL, r1, 1), R2 = __shfl_down_sync(FULL, r2, 1);
        Wx = fmaf(k.b[1], L2, fmaf(k.b[3], r1, k.b[5] * r2));
        Wy = fmaf(k.b[0], L2, k.b[2] * r1) + fmaf(k.b[4], r2, k.b[6] * R1);
        Wz = fmaf(k.b[1], r1, fmaf(k.b[3], r2, k.b[5] * R1));
        Ww = fmaf(k.b[0], r1, k.b[2] * r2) + fmaf(k.b[4], R1, k.b[6] * R2);
    } else {                                            // sites at staged cols 4l-3, 4l-1 | 4l+1, 4l+3 | 4l+5
        const float L1 = __shfl_up_sync(FULL, r1, 1), L2 = __shfl_up_sync(FULL, r2, 1), R1 = __shfl_down_sync(FULL, r1, 1);
        Wx = fmaf(k.b[0], L1, k.b[2] * L2) + fmaf(k.b[4], r1, k.b[6] * r2);
        Wy = fmaf(k.b[1], L2, fmaf(k.b[3], r1, k.b[5] * r2));
        Wz = fmaf(k.b[0], L2, k.b[2] * r1) + fmaf(k.b[4], r2, k.b[6] * R1);
        Ww = fmaf(k.b[1], r1, fmaf(k.b[3], r2, k.b[5] * R1));
    }
}

__device__ __forceinline__ float clipf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

// ---- row bands: system-scope flag accesses (the arrival counters are written by the neighbouring GPUs) ----------------------
__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys_u32(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long gtimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// what a warp knows about itself (uniform across the warp)
struct WarpCtx {
    int ya, yb;               // HR rows this warp updates (and whose sites it owns for the MSE)
    int a0;                   // site-row index of step 0 (may be negative)
    int h;                    // LR rows (site rows are 0..h-1)
    bool top;                 // ya == 0 and the lattice has row parity 1: the clamp terms of HR row 0 apply
    int x0;                   // first owned column
    bool push[2];             // row bands: some of this warp's rows also go to the band above / below
};

// One lattice row.  T = step mod 4 fixes every register slot at compile time:
//   hr window   row (ys + 3 - u) sits in slot (2T + 1 - u) mod 8; this step's new rows are slots 2T, 2T+1
//   W ring      this step writes slot T; W of step j-d sits in slot (T - d) mod 4
// MODE: M_GEN = every row / ownership / store decision at run time (segment ends, image borders);
// otherwise the site row is a valid interior one and MODE says whether the warp owns it for the MSE
// (M_OWN) and whether this step's two HR rows are stored (M_ST).
constexpr int M_OWN = 1, M_ST = 2, M_GEN = 4;
template <int T, int P, int Q, int MODE, bool LEFT, bool BAND>
__device__ __forceinline__ void stream_step(State& s, const LaneConst& k, const WarpCtx& c, const StreamArgs& a,
                                            const int aj, const float* __restrict__ s1row, const float* __restrict__ pf,
                                            float*& outp, const float cS, const float lS, const int lane) {
    #define WIN(u) s.v[(2 * T + 1 - (u) + 8) & 7]
    // ---- column pass ---------------------------------------------------------------------------------------
    u64 Clo = mul2(k.A[0], WIN(0).lo), Chi = mul2(k.A[0], WIN(0).hi);
    u64 Dlo = mul2(k.A[1], WIN(1).lo), Dhi = mul2(k.A[1], WIN(1).hi);
    Clo = fma2(k.A[2], WIN(2).lo, Clo); Chi = fma2(k.A[2], WIN(2).hi, Chi);
    Dlo = fma2(k.A[3], WIN(3).lo, Dlo); Dhi = fma2(k.A[3], WIN(3).hi, Dhi);
    Clo = fma2(k.A[4], WIN(4).lo, Clo); Chi = fma2(k.A[4], WIN(4).hi, Chi);
    Dlo = fma2(k.A[5], WIN(5).lo, Dlo); Dhi = fma2(k.A[5], WIN(5).hi, Dhi);
    Clo = fma2(k.A[6], WIN(6).lo, Clo); Chi = fma2(k.A[6], WIN(6).hi, Chi);
    Clo = add2(Clo, Dlo); Chi = add2(Chi, Dhi);
    constexpr bool GEN = (MODE & M_GEN) != 0;
    constexpr bool OWN = GEN || (MODE & M_OWN) != 0;     // the residual energy is accumulated
    constexpr bool ST  = GEN || (MODE & M_ST) != 0;      // the transposed column pass / update / stores run
    const Row h1 = WIN(6), h2 = WIN(5);                 // hr rows ys-3, ys-2: updated at the end of this step

    // top segments, site row ys = 1: B on HR row 0 (LR row 0 of the frames with sy = +1 reads it)
    u64 C0lo = 0, C0hi = 0;
    const bool top_step = GEN && (P == 1) && (T == 2) && c.top && (aj == 0);   // a0 = -2 in a top segment: always step 2
    if (top_step) {
        C0lo = mul2(k.A[0], WIN(1).lo); C0hi = mul2(k.A[0], WIN(1).hi);
        C0lo = fma2(k.A[1], WIN(2).lo, C0lo); C0hi = fma2(k.A[1], WIN(2).hi, C0hi);
        C0lo = fma2(k.A[2], WIN(3).lo, C0lo); C0hi = fma2(k.A[2], WIN(3).hi, C0hi);
        C0lo = fma2(k.A[3], WIN(4).lo, C0lo); C0hi = fma2(k.A[3], WIN(4).hi, C0hi);
    }
    // ---- the two rows the next step adds ----------------------------------------------------------------------
    s.v[(2 * T + 2) & 7] = lds_row(pf);
    s.v[(2 * T + 3) & 7] = lds_row(pf + SS_COLS);
    #undef WIN

    // ---- row pass, residual ------------------------------------------------------------------------------------
    float Cx, Cy, Cz, Cw, B1, B2;
    upk(Clo, Cx, Cy); upk(Chi, Cz, Cw);
    row_pass<Q>(k, Cx, Cy, Cz, Cw, B1, B2);
    const float2 s1 = *reinterpret_cast<const float2*>(s1row);
    float m1 = k.mI1, m2 = k.mI2, w1 = k.wI1, w2 = k.wI2;
    bool valid = true, last = false, own = true;
    if (GEN) {
        valid = (aj >= 0) && (aj < c.h);
        last = (aj == c.h - 1);
        const int ys = 2 * aj + P;
        own = valid && (ys >= c.ya) && (ys < c.yb);
        m1 = valid ? (last ? k.mL1 : k.mI1) : 0.0f;
        m2 = valid ? (last ? k.mL2 : k.mI2) : 0.0f;
        w1 = own ? (last ? k.wL1 : k.wI1) : 0.0f;
        w2 = own ? (last ? k.wL2 : k.wI2) : 0.0f;
    }
    const float r1 = fmaf(-m1, B1, s1.x), r2 = fmaf(-m2, B2, s1.y);
    if (OWN) {
        s.e2 = fmaf(r1 * w1, r1, s.e2);
        s.e2 = fmaf(r2 * w2, r2, s.e2);
    }

    // ---- transposed row pass ------------------------------------------------------------------------------------
    float Wx, Wy, Wz, Ww;
    back_row<Q>(k, r1, r2, Wx, Wy, Wz, Ww);
    if (LEFT) {                                         // Q == 1, x0 == 0: lane 2 holds columns 0..3, its first site is xs = 1
        const float mc = GEN ? (valid ? (last ? a.mc[1] : a.mc[0]) : 0.0f) : a.mc[0];
        const float ml = GEN ? (valid ? (last ? a.ml[1] : a.ml[0]) : 0.0f) : a.ml[0];
        const float dcol = fmaf(-mc, B1, cS);           // duplicate residual at (ys, 0)
        Wx = fmaf(k.lb[0], dcol, Wx); Wy = fmaf(k.lb[1], dcol, Wy);
        Wz = fmaf(k.lb[2], dcol, Wz); Ww = fmaf(k.lb[3], dcol, Ww);
        // LR column 0 of the frames with sx = +1 reads B[ys, 0] (columns < 0 are zero): MSE only
        const float Bc0 = fmaf(k.b[0], Cw, fmaf(k.b[1], Cz, fmaf(k.b[2], Cy, k.b[3] * Cx)));
        const float rl = fmaf(-ml, Bc0, lS);
        const float wl = GEN ? (own ? (last ? k.wlL : k.wlI) : 0.0f) : k.wlI;
        if (OWN) s.e2 = fmaf(rl * wl, rl, s.e2);
    }
    s.wr[T].lo = pk(Wx, Wy); s.wr[T].hi = pk(Wz, Ww);

    if (top_step) {
        float C0x, C0y, C0z, C0w, B01, B02;
        upk(C0lo, C0x, C0y); upk(C0hi, C0z, C0w);
        row_pass<Q>(k, C0x, C0y, C0z, C0w, B01, B02);
        const int b1 = ((c.x0 - 8) >> 1) + 2 * lane, b2 = b1 + 1;
        const bool v1 = (b1 >= 0) && (b1 < a.w), v2 = (b2 >= 0) && (b2 < a.w);
        const int xs1 = 2 * b1 + Q, xs2 = 2 * b2 + Q;
        const bool o1 = v1 && xs1 >= c.x0 && xs1 < c.x0 + SS_VALID, o2 = v2 && xs2 >= c.x0 && xs2 < c.x0 + SS_VALID;
        const float tS1 = v1 ? __ldg(a.topS1 + b1) : 0.0f, tS2 = v2 ? __ldg(a.topS1 + b2) : 0.0f;
        const float rS1 = v1 ? __ldg(a.rowS1 + b1) : 0.0f, rS2 = v2 ? __ldg(a.rowS1 + b2) : 0.0f;
        const float mt1 = a.mt[b1 == a.w - 1], mt2 = a.mt[b2 == a.w - 1];
        const float mr1 = v1 ? a.mr[b1 == a.w - 1] : 0.0f, mr2 = v2 ? a.mr[b2 == a.w - 1] : 0.0f;
        if (o1 && mt1 > 0.0f) { const float e = fmaf(-mt1, B01, tS1); s.e2 += e * e / mt1; }
        if (o2 && mt2 > 0.0f) { const float e = fmaf(-mt2, B02, tS2); s.e2 += e * e / mt2; }
        if (LEFT && lane == 2 && a.m_msec > 0.0f) {     // LR (0, 0) of the frames (+1, +1) reads B[0, 0]
            const float B00 = fmaf(k.b[0], C0w, fmaf(k.b[1], C0z, fmaf(k.b[2], C0y, k.b[3] * C0x)));
            const float e = fmaf(-a.m_msec, B00, __ldg(a.scal + 1));
            s.e2 += e * e / a.m_msec;
        }
        const float d1 = fmaf(-mr1, B1, rS1), d2 = fmaf(-mr2, B2, rS2);   // duplicates on HR row 0
        float Vx, Vy, Vz, Vw;
        back_row<Q>(k, d1, d2, Vx, Vy, Vz, Vw);
        if (LEFT) {                                     // the duplicate at (0, 0): frames (-1, -1), reads B[1, 1]
            const float d00 = fmaf(-a.m_dupc, B1, __ldg(a.scal + 0));
            Vx = fmaf(k.lb[0], d00, Vx); Vy = fmaf(k.lb[1], d00, Vy);
            Vz = fmaf(k.lb[2], d00, Vz); Vw = fmaf(k.lb[3], d00, Vw);
        }
        s.w0.lo = pk(Vx, Vy); s.w0.hi = pk(Vz, Vw);
    }

    if (ST) {
        // ---- transposed column pass: HR rows y1 = ys - 3 (sites ys-6, ys-4, ys-2, ys) and y2 = ys - 2 ----------------
        const Row& wa = s.wr[(T + 1) & 3]; const Row& wb = s.wr[(T + 2) & 3]; const Row& wc = s.wr[(T + 3) & 3]; const Row& wd = s.wr[T];
        u64 o1lo = add2(fma2(k.A[2], wb.lo, mul2(k.A[0], wa.lo)), fma2(k.A[6], wd.lo, mul2(k.A[4], wc.lo)));
        u64 o1hi = add2(fma2(k.A[2], wb.hi, mul2(k.A[0], wa.hi)), fma2(k.A[6], wd.hi, mul2(k.A[4], wc.hi)));
        u64 o2lo = fma2(k.A[5], wd.lo, fma2(k.A[3], wc.lo, mul2(k.A[1], wb.lo)));
        u64 o2hi = fma2(k.A[5], wd.hi, fma2(k.A[3], wc.hi, mul2(k.A[1], wb.hi)));
        const int y1 = 2 * aj + P - 3;
        if (GEN && c.top) {                                 // HR rows 0..3 also see the duplicate site row ys = 0
            if (y1 >= 0 && y1 < 4) { const float g = a.acol[3 - y1]; const u64 g2 = pk(g, g); o1lo = fma2(g2, s.w0.lo, o1lo); o1hi = fma2(g2, s.w0.hi, o1hi); }
            if (y1 + 1 >= 0 && y1 + 1 < 4) { const float g = a.acol[2 - y1]; const u64 g2 = pk(g, g); o2lo = fma2(g2, s.w0.lo, o2lo); o2hi = fma2(g2, s.w0.hi, o2hi); }
        }
        // ---- update, clip, store ----------------------------------------------------------------------------------------
        float4 q1, q2;
        upk(fma2(k.alpha2, o1lo, h1.lo), q1.x, q1.y); upk(fma2(k.alpha2, o1hi, h1.hi), q1.z, q1.w);
        upk(fma2(k.alpha2, o2lo, h2.lo), q2.x, q2.y); upk(fma2(k.alpha2, o2hi, h2.hi), q2.z, q2.w);
        q1.x = clipf(q1.x, k.lo, k.hi); q1.y = clipf(q1.y, k.lo, k.hi); q1.z = clipf(q1.z, k.lo, k.hi); q1.w = clipf(q1.w, k.lo, k.hi);
        q2.x = clipf(q2.x, k.lo, k.hi); q2.y = clipf(q2.y, k.lo, k.hi); q2.z = clipf(q2.z, k.lo, k.hi); q2.w = clipf(q2.w, k.lo, k.hi);
        bool st1 = k.st_ok, st2 = k.st_ok;
        if (GEN) { st1 = st1 && (y1 >= c.ya) && (y1 < c.yb); st2 = st2 && (y1 + 1 >= c.ya) && (y1 + 1 < c.yb); }
        if (st1) *reinterpret_cast<float4*>(outp) = q1;
        if (st2) *reinterpret_cast<float4*>(outp + a.W) = q2;
        if (GEN && BAND) {                                  // row bands: boundary rows go straight into the neighbours' halo rows
            #pragma unroll
            for (int d = 0; d < 2; ++d) {
                if (!c.push[d] || (a.bp.probe & 2)) continue;
                float* pp = a.bp.peer_out[d] + (outp - a.hr_out);
                if (st1 && y1 >= a.bp.push_y0[d] && y1 < a.bp.push_y1[d]) *reinterpret_cast<float4*>(pp) = q1;
                if (st2 && y1 + 1 >= a.bp.push_y0[d] && y1 + 1 < a.bp.push_y1[d]) *reinterpret_cast<float4*>(pp + a.W) = q2;
            }
        }
    }
    outp += 2 * a.W;
}

// Four steps; M0..M3 are the steps' MODEs (see stream_step)
template <int P, int Q, int M0, int M1, int M2, int M3, bool LEFT, bool BAND>
__device__ __forceinline__ void stream_iteration(State& s, const LaneConst& k, const WarpCtx& c, const StreamArgs& a,
                                                 const int it, const float* __restrict__ ring, uint64_t* bars,
                                                 float*& outp, float& curC, float& curL, const int lane) {
    const int slot = (it + 1) % SS_STAGES, nslot = (it + 2) % SS_STAGES;
    const float* hrs = ring + slot * SS_STAGE_FLOATS + 4 * lane;             // this iteration's stage, the lane's columns
    const float* s1s = ring + slot * SS_STAGE_FLOATS + SS_HR_FLOATS + 2 * lane;
    const float* nxt = ring + nslot * SS_STAGE_FLOATS + 4 * lane;
    const int aj = c.a0 + 4 * it;
    float nC = 0.0f, nL = 0.0f;
    if (LEFT) {                                          // edge vectors of the NEXT iteration's four steps (lanes 0..3)
        const int an = aj + 4 + lane;
        if (lane < 4 && an >= 0 && an < c.h) { nC = __ldg(a.colS1 + an); nL = __ldg(a.leftS1 + an); }
    }
    const unsigned FULL = 0xffffffffu;
    #define EDGE(t) (LEFT ? __shfl_sync(FULL, curC, t) : 0.0f), (LEFT ? __shfl_sync(FULL, curL, t) : 0.0f)
    stream_step<0, P, Q, M0, LEFT, BAND>(s, k, c, a, aj + 0, s1s + 0 * 64, hrs + 2 * SS_COLS, outp, EDGE(0), lane);
    stream_step<1, P, Q, M1, LEFT, BAND>(s, k, c, a, aj + 1, s1s + 1 * 64, hrs + 4 * SS_COLS, outp, EDGE(1), lane);
    stream_step<2, P, Q, M2, LEFT, BAND>(s, k, c, a, aj + 2, s1s + 2 * 64, hrs + 6 * SS_COLS, outp, EDGE(2), lane);
    mbar_wait(&bars[nslot], (uint32_t)(((it + 2) / SS_STAGES) & 1));          // the next stage: its first two rows are prefetched now
    stream_step<3, P, Q, M3, LEFT, BAND>(s, k, c, a, aj + 3, s1s + 3 * 64, nxt, outp, EDGE(3), lane);
    #undef EDGE
    curC = nC; curL = nL;
}

__device__ __forceinline__ void stream_issue(const StreamArgs& a, float* ring, uint64_t* bars, int stage, int xs0, int sb, int a0) {
    const int slot = stage % SS_STAGES;
    float* dst = ring + slot * SS_STAGE_FLOATS;
    mbar_expect_tx(&bars[slot], 4u * SS_STAGE_FLOATS);
    tma_load_2d(dst, &a.tm_hr, xs0, sb + SS_ROWS * (stage - 1), &bars[slot]);
    tma_load_2d(dst + SS_HR_FLOATS, &a.tm_s1, xs0 >> 1, a0 + SS_SROWS * (stage - 1), &bars[slot]);
}

// BAND: the launch works on a row band whose halo rows it exchanges itself; the whole-image instantiation carries none of that code
template <int P, int Q, bool BAND>
__global__ void __launch_bounds__(SS_THREADS, SS_CTAS) k_ibp_stream(const __grid_constant__ StreamArgs a) {
    extern __shared__ __align__(128) float sm[];
    __shared__ __align__(8) uint64_t bars_all[SS_WARPS][SS_STAGES];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int gw = blockIdx.x * SS_WARPS + warp;
    const int nbw = BAND ? (a.b_top + a.b_bot) * a.nstrip : 0;      // border warps come first (row bands)
    if (gw >= nbw + a.nseg * a.nstrip) return;           // whole warps only; nothing below synchronises the CTA
    float* ring = sm + warp * (SS_STAGES * SS_STAGE_FLOATS);
    uint64_t* bars = bars_all[warp];

    WarpCtx c;
    int strip, nit;
    const int brows = 8 * a.b_nit - 6;
    if (BAND && gw < nbw) {
        strip = gw % a.nstrip; nit = a.b_nit;
        const bool is_top = a.b_top && gw < a.nstrip;
        c.ya = is_top ? a.y_base : a.y_end - brows;
        c.yb = c.ya + brows;
    } else {
        const int g2 = gw - nbw;
        strip = g2 % a.nstrip; nit = a.nit;
        c.ya = a.y_base + (BAND ? a.b_top * brows : 0) + (g2 / a.nstrip) * a.seg_rows;   // segments partition the rows between the border zones
        c.yb = min(c.ya + a.seg_rows, a.y_end - (BAND ? a.b_bot * brows : 0));
    }
    c.x0 = strip * SS_VALID;
    c.h = a.h;
    c.top = (P == 1) && (c.ya == 0);
    const int ys0 = c.ya - 3 + (1 - P);                  // first site row the segment needs (parity P)
    c.a0 = (ys0 - P) >> 1;                               // exact: ys0 - P is even
    const int sb = ys0 + 2;                              // first NEW hr row of step 0
    const int xs0 = c.x0 - 8;

    // programmatic dependent launch: the next iteration's grid may start (and run everything above and the
    // barrier set-up below) while this one drains; nothing written by the previous grid is touched before the wait
    asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory");
    if (lane == 0) {
        #pragma unroll
        for (int i = 0; i < SS_STAGES; ++i) mbar_init(&bars[i], 1);
        mbar_fence_init();
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    }
    asm volatile("griddepcontrol.wait;\n" ::: "memory");
    // row bands: the warps whose windows reach into a halo wait until the neighbour's rows of the previous iteration have
    // arrived (its counter reaches BASE + it); everybody else starts at once
    c.push[0] = c.push[1] = false;
    if (BAND && a.bp.flags != nullptr) {
        c.push[0] = a.bp.has[0] && c.ya < a.bp.push_y1[0];
        c.push[1] = a.bp.has[1] && c.yb > a.bp.push_y0[1];
        const bool need_up = a.bp.has[0] && (c.ya - 8 < a.bp.push_y0[0]);
        const bool need_dn = a.bp.has[1] && (c.yb + 8 > a.bp.push_y1[1]);
        if ((need_up || need_dn) && lane == 0 && !(a.bp.probe & 1)) {
            const unsigned want = (a.bp.flags[BF_BASE] + (unsigned)a.bp.it) * (unsigned)a.bp.unit;
            const unsigned long long t0 = gtimer_ns();
            #pragma unroll
            for (int d = 0; d < 2; ++d) {
                if (!(d == 0 ? need_up : need_dn)) continue;
                const unsigned* f = a.bp.flags + (d == 0 ? BF_ARR_UP : BF_ARR_DOWN);
                while ((int)(ld_acquire_sys_u32(f) - want) < 0) {
                    if (gtimer_ns() - t0 > a.bp.timeout_ns) { a.bp.flags[BF_ERROR] = 1u; break; }
                    __nanosleep(64);
                }
            }
        }
        __syncwarp();
    }
    if (lane == 0) {
        #pragma unroll
        for (int st = 0; st < SS_STAGES; ++st) stream_issue(a, ring, bars, st, xs0, sb, c.a0);
    }
    __syncwarp();

    // ---- per-lane constants -------------------------------------------------------------------------------------------
    LaneConst k;
    #pragma unroll
    for (int u = 0; u < 7; ++u) { k.A[u] = pk(a.acol[u], a.acol[u]); k.b[u] = a.brow[u]; }
    k.alpha2 = pk(a.alpha, a.alpha); k.lo = a.lo; k.hi = a.hi;
    {
        const int b1 = (xs0 >> 1) + 2 * lane, b2 = b1 + 1;
        const bool v1 = (b1 >= 0) && (b1 < a.w), v2 = (b2 >= 0) && (b2 < a.w);
        const int xs1 = 2 * b1 + Q, xs2 = 2 * b2 + Q;
        const bool o1 = v1 && xs1 >= c.x0 && xs1 < c.x0 + SS_VALID, o2 = v2 && xs2 >= c.x0 && xs2 < c.x0 + SS_VALID;
        const int l1 = (b1 == a.w - 1), l2 = (b2 == a.w - 1);
        k.mI1 = v1 ? a.mtab[l1] : 0.0f;     k.mI2 = v2 ? a.mtab[l2] : 0.0f;
        k.mL1 = v1 ? a.mtab[2 + l1] : 0.0f; k.mL2 = v2 ? a.mtab[2 + l2] : 0.0f;
        k.wI1 = (o1 && k.mI1 > 0.0f) ? 1.0f / k.mI1 : 0.0f; k.wI2 = (o2 && k.mI2 > 0.0f) ? 1.0f / k.mI2 : 0.0f;
        k.wL1 = (o1 && k.mL1 > 0.0f) ? 1.0f / k.mL1 : 0.0f; k.wL2 = (o2 && k.mL2 > 0.0f) ? 1.0f / k.mL2 : 0.0f;
    }
    const bool left = (Q == 1) && (c.x0 == 0) && a.has_left;
    #pragma unroll
    for (int i = 0; i < 4; ++i) k.lb[i] = (left && lane == 2) ? a.brow[3 - i] : 0.0f;
    k.wlI = (left && lane == 2 && a.ml[0] > 0.0f) ? 1.0f / a.ml[0] : 0.0f;
    k.wlL = (left && lane == 2 && a.ml[1] > 0.0f) ? 1.0f / a.ml[1] : 0.0f;
    const int gx = xs0 + 4 * lane;
    k.st_ok = (lane >= 2) && (lane < 30) && (gx < a.W);

    State s;
    #pragma unroll
    for (int i = 0; i < 4; ++i) { s.wr[i].lo = 0; s.wr[i].hi = 0; }
    s.w0.lo = 0; s.w0.hi = 0; s.e2 = 0.0f;
    float* outp = a.hr_out + (ptrdiff_t)(ys0 - 3) * a.W + gx;   // row of step 0's first output (never dereferenced outside the image)
    float curC = 0.0f, curL = 0.0f;
    if (left && lane < 4) {
        const int an = c.a0 + lane;
        if (an >= 0 && an < c.h) { curC = __ldg(a.colS1 + an); curL = __ldg(a.leftS1 + an); }
    }

    // ---- prologue: the five window rows above step 0's new rows, and those two rows ---------------------------------------
    mbar_wait(&bars[0], 0);
    mbar_wait(&bars[1], 0);
    {
        const float* p0 = ring + 4 * lane;                                  // stage 0: rows sb-8 .. sb-1
        const float* p1 = ring + SS_STAGE_FLOATS + 4 * lane;                // stage 1: rows sb .. sb+7
        s.v[2].lo = 0; s.v[2].hi = 0;
        #pragma unroll
        for (int r = 3; r < 8; ++r) s.v[r] = lds_row(p0 + r * SS_COLS);
        s.v[0] = lds_row(p1);
        s.v[1] = lds_row(p1 + SS_COLS);
    }
    __syncwarp();                                        // stage 0 is consumed: its slot takes stage 3
    if (lane == 0 && SS_STAGES <= nit + 1) stream_issue(a, ring, bars, SS_STAGES, xs0, sb, c.a0);

    // A REGULAR segment (full length, every site row interior, no clamp rows, no peer stores) has a fixed schedule of step kinds:
    // the first iteration is the warm-up (kind 1), the last one the tail (kind 2) for P = 1 -- both general (kind 3) for P = 0 --
    // and everything between is the fast form (verified against the per-iteration classification below for every geometry).
    // That classification cost 13 % of the kernel's instructions when it ran in every iteration of every warp.
    const bool regular = !c.top && (c.yb - c.ya == 8 * nit - 6) && (c.a0 >= 0) && (c.a0 + 4 * nit - 1 < c.h - 1) &&
                         !(BAND && (c.push[0] || c.push[1]));
    for (int it = 0; it < nit; ++it) {
        // classify the four steps: interior site row? owned (MSE)? both HR rows stored, or none?
        int kind = 0;                                    // 0 all fast, 1 first (warm-up), 2 last, 3 general
        if (regular) {
            kind = (it == 0) ? (P == 1 ? 1 : 3) : (it == nit - 1 ? (P == 1 ? 2 : 3) : 0);
        } else {
            int nf = 0, pat = 0;
            #pragma unroll
            for (int t = 0; t < 4; ++t) {
                const int aj = c.a0 + 4 * it + t, ys = 2 * aj + P, y1 = ys - 3;
                const bool interior = (aj >= 0) && (aj < c.h - 1);
                const bool own = (ys >= c.ya) && (ys < c.yb);
                const bool both = (y1 >= c.ya) && (y1 + 1 < c.yb) && !(c.top && y1 < 4);
                const bool none = (y1 + 1 < c.ya);
                int m = -1;
                if (interior && both) m = own ? (M_OWN | M_ST) : M_ST;
                else if (interior && none) m = own ? M_OWN : 0;
                if (m < 0) nf = 1;
                pat |= (m & 3) << (2 * t);
            }
            const int FAST = M_OWN | M_ST;
            if (nf || (BAND && (c.push[0] || c.push[1]))) kind = 3;     // pushing warps take the general form (it has the peer stores)
            else if (pat == (FAST | FAST << 2 | FAST << 4 | FAST << 6)) kind = 0;
            else if (pat == (0 | 0 << 2 | M_OWN << 4 | FAST << 6)) kind = 1;
            else if (pat == (FAST | FAST << 2 | FAST << 4 | M_ST << 6)) kind = 2;
            else kind = 3;
        }
        constexpr int F = M_OWN | M_ST;
        if (left) {
            if (kind == 0)      stream_iteration<P, Q, F, F, F, F, true, BAND>(s, k, c, a, it, ring, bars, outp, curC, curL, lane);
            else if (kind == 1) stream_iteration<P, Q, 0, 0, M_OWN, F, true, BAND>(s, k, c, a, it, ring, bars, outp, curC, curL, lane);
            else if (kind == 2) stream_iteration<P, Q, F, F, F, M_ST, true, BAND>(s, k, c, a, it, ring, bars, outp, curC, curL, lane);
            else                stream_iteration<P, Q, M_GEN, M_GEN, M_GEN, M_GEN, true, BAND>(s, k, c, a, it, ring, bars, outp, curC, curL, lane);
        } else {
            if (kind == 0)      stream_iteration<P, Q, F, F, F, F, false, BAND>(s, k, c, a, it, ring, bars, outp, curC, curL, lane);
            else if (kind == 1) stream_iteration<P, Q, 0, 0, M_OWN, F, false, BAND>(s, k, c, a, it, ring, bars, outp, curC, curL, lane);
            else if (kind == 2) stream_iteration<P, Q, F, F, F, M_ST, false, BAND>(s, k, c, a, it, ring, bars, outp, curC, curL, lane);
            else                stream_iteration<P, Q, M_GEN, M_GEN, M_GEN, M_GEN, false, BAND>(s, k, c, a, it, ring, bars, outp, curC, curL, lane);
        }
        // stage it+1 is consumed: refill its slot with stage it+4
        __syncwarp();
        if (lane == 0 && it + 4 <= nit + 1) stream_issue(a, ring, bars, it + 4, xs0, sb, c.a0);
    }

    float e2 = s.e2;
    #pragma unroll
    for (int off = 16; off > 0; off >>= 1) e2 += __shfl_xor_sync(0xffffffffu, e2, off);
    if (lane == 0) a.partial[gw] = (double)e2;

    // row bands: once my peer stores are visible system-wide, count my arrival in the neighbour's arena (one remote
    // atomic per pushing warp; the neighbour waits for `unit` of them per iteration)
    if (BAND && (c.push[0] || c.push[1]) && !(a.bp.probe & 4)) {
        __threadfence_system();
        __syncwarp();
        if (lane == 0) {
            if (c.push[0]) atomicAdd_system(a.bp.peer_flag[0], 1u);
            if (c.push[1]) atomicAdd_system(a.bp.peer_flag[1], 1u);
        }
    }
}

void stream_geometry(int H, int W, int n_sm, int* nit_out, int* seg_rows, int* nseg_out, int* nstrip_out) {
    const int nstrip = (W + SS_VALID - 1) / SS_VALID;
    const int slots = std::max(1, n_sm) * SS_CTAS;
    double best = 1e30; int best_nit = 2;
    for (int nit = 2; nit <= 128; ++nit) {
        const int L = 8 * nit - 6;
        const int nseg = (H + L - 1) / L;
        const int ctas = (nseg * nstrip + SS_WARPS - 1) / SS_WARPS;
        const int waves = (ctas + slots - 1) / slots;
        const double cost = (double)waves * (nit + 1.5);
        if (cost < best - 1e-9) { best = cost; best_nit = nit; }
        if (L >= H) break;
    }
    *nit_out = best_nit; *seg_rows = 8 * best_nit - 6;
    *nseg_out = (H + *seg_rows - 1) / *seg_rows; *nstrip_out = nstrip;
}

int launch_stream_edge_sums(const float* const* lr, int n, int h, int w, const int* sy, const int* sx, int P, int Q,
                            float* buf, double* vsum, int own_a0, int own_a1, int top_owned, cudaStream_t st) {
    EdgeSumArgs e{};
    e.n = n; e.h = h; e.w = w;
    e.own_a0 = own_a0; e.own_a1 = own_a1; e.top_owned = top_owned;
    for (int k = 0; k < n; ++k) {
        e.lr[k] = lr[k]; e.sy[k] = sy[k]; e.sx[k] = sx[k];
        e.oy[k] = (P + sy[k]) >> 1; e.ox[k] = (Q + sx[k]) >> 1;
    }
    e.colS1 = buf; e.leftS1 = buf + h; e.rowS1 = buf + 2 * h; e.topS1 = buf + 2 * h + w; e.scal = buf + 2 * h + 2 * w;
    e.vsum = vsum;
    return launch_edge_sums(e, st);
}

int launch_ibp_stream(const StreamArgs& a, int P, int Q, cudaStream_t st) {
    const int warps = (a.nseg + a.b_top + a.b_bot) * a.nstrip;
    if (warps <= 0) return SR459_OK;
    const size_t smem = sizeof(float) * (size_t)SS_WARPS * SS_STAGES * SS_STAGE_FLOATS;
    static PerDeviceOnce configured;
    if (configured.need()) {
        SR_CUDA(cudaFuncSetAttribute(k_ibp_stream<0, 0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        SR_CUDA(cudaFuncSetAttribute(k_ibp_stream<0, 1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        SR_CUDA(cudaFuncSetAttribute(k_ibp_stream<1, 0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        SR_CUDA(cudaFuncSetAttribute(k_ibp_stream<1, 1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        SR_CUDA(cudaFuncSetAttribute(k_ibp_stream<0, 0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        SR_CUDA(cudaFuncSetAttribute(k_ibp_stream<0, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        SR_CUDA(cudaFuncSetAttribute(k_ibp_stream<1, 0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        SR_CUDA(cudaFuncSetAttribute(k_ibp_stream<1, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    const int grid = (warps + SS_WARPS - 1) / SS_WARPS;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(SS_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    { static const bool no_pdl = getenv("SR459_NO_PDL") != nullptr; cfg.attrs = attr; cfg.numAttrs = no_pdl ? 0 : 1; }
    const bool band = a.bp.flags != nullptr || a.b_top || a.b_bot;
    cudaError_t le;
    switch (P * 2 + Q + (band ? 4 : 0)) {
        case 0: le = cudaLaunchKernelEx(&cfg, k_ibp_stream<0, 0, false>, a); break;
        case 1: le = cudaLaunchKernelEx(&cfg, k_ibp_stream<0, 1, false>, a); break;
        case 2: le = cudaLaunchKernelEx(&cfg, k_ibp_stream<1, 0, false>, a); break;
        case 3: le = cudaLaunchKernelEx(&cfg, k_ibp_stream<1, 1, false>, a); break;
        case 4: le = cudaLaunchKernelEx(&cfg, k_ibp_stream<0, 0, true>, a); break;
        case 5: le = cudaLaunchKernelEx(&cfg, k_ibp_stream<0, 1, true>, a); break;
        case 6: le = cudaLaunchKernelEx(&cfg, k_ibp_stream<1, 0, true>, a); break;
        default: le = cudaLaunchKernelEx(&cfg, k_ibp_stream<1, 1, true>, a); break;
    }
    if (le != cudaSuccess) { set_error("launch of k_ibp_stream failed: %s", cudaGetErrorString(le)); return SR459_ECUDA; }
    SR_LAUNCH_CHECK("k_ibp_stream");
    return SR459_OK;
}

}  // namespace sr459
