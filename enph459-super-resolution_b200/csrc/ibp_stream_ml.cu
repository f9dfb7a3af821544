// ibp_stream_ml.cu -- fused IBP iteration, integer HR shifts, rank-1 PSF, TWO TO FOUR parity lattices:
// the register-streaming form of ibp_stream.cu for frame sets whose shifts fall on several lattices, e.g.
// mono_cal_target's own geometry (centre frame on the even/even lattice + four +-0.5 px frames on the odd/odd one,
// mono_cal_target/run_sr.py:60-66).  Reference: ibp(), run_sr.py:190-209; closed form SURVEY.md section 8a.
//
// Same pipeline per warp as k_ibp_stream (TMA ring of hr rows -> register window -> column pass -> row pass by shuffle
// -> residual -> transposed row pass -> W ring -> transposed column pass -> update / clip / float4 stores; no CTA
// barrier, no intermediate in shared memory), with one step = one PAIR of HR rows (2a, 2a+1):
//   window     8 hr rows 2a-3 .. 2a+4 in registers (slot = row mod 8 relative to the segment)
//   column pass once per row parity that carries a lattice (even row: rows 2a-3..2a+3, odd row: 2a-2..2a+4)
//   per lattice (P, Q)  row pass at the lane's two sites of column parity Q, residual against the lattice's S1 plane,
//              transposed row pass; the contributions of the lattices of one row parity add up in ONE W row
//   W ring     8 rows 2a-6 .. 2a+1; transposed column pass with all 7 taps (rows of an inactive parity skipped at
//              compile time) -> HR rows 2a-3, 2a-2
// The ndi_shift clamp terms exist per lattice exactly as in the one-lattice kernel: duplicates on HR row 0 and the
// MSE-only terms of LR row 0 for lattices with P = 1 (first step of the top segments), duplicates on HR column 0 and
// the MSE-only terms of LR column 0 for lattices with Q = 1 (lane 2 of the left strip); their pre-summed edge vectors
// come from k_edge_sums, once per lattice and ibp() call.  MASK (bit 2P+Q) is a template parameter.
#include "internal.h"
#include "tma.h"

#include <stdlib.h>
#include <algorithm>

namespace sr459 {

namespace {

typedef unsigned long long u64;

constexpr int ML_COLS   = 128;                 // staged columns per strip
constexpr int ML_VALID  = 112;                 // columns a strip updates (lanes 2..29)
constexpr int ML_ROWS   = 8;                   // hr rows per stage
constexpr int ML_SROWS  = 4;                   // S1 rows per stage and lattice
constexpr int ML_STAGES = 3;
constexpr int ML_HR_FLOATS = ML_ROWS * ML_COLS;
constexpr int ML_S1_FLOATS = ML_SROWS * 64;
constexpr int ML_WARPS   = 4;
constexpr int ML_THREADS = 32 * ML_WARPS;
constexpr int ML_CTAS    = 2;

__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(u64 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

struct Row { u64 lo, hi; };
__device__ __forceinline__ Row lds_row(const float* p) {
    const ulonglong2 t = *reinterpret_cast<const ulonglong2*>(p);
    Row r; r.lo = t.x; r.hi = t.y; return r;
}
__device__ __forceinline__ float clipf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

__host__ __device__ constexpr int popc4(int m) { return (m & 1) + ((m >> 1) & 1) + ((m >> 2) & 1) + ((m >> 3) & 1); }
__host__ __device__ constexpr int slot_of(int mask, int pq) { return popc4(mask & ((1 << pq) - 1)); }

// B at the lane's two sites (staged cols 4l+Q, 4l+2+Q) from the column-passed C of this and the neighbour lanes
template <int Q>
__device__ __forceinline__ void row_pass(const float (&b)[7], float Cx, float Cy, float Cz, float Cw, float& B1, float& B2) {
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
    float a1 = b[0] * f[6], a2 = b[0] * f[8], c1 = b[1] * f[5], c2 = b[1] * f[7];
    a1 = fmaf(b[2], f[4], a1); a2 = fmaf(b[2], f[6], a2); c1 = fmaf(b[3], f[3], c1); c2 = fmaf(b[3], f[5], c2);
    a1 = fmaf(b[4], f[2], a1); a2 = fmaf(b[4], f[4], a2); c1 = fmaf(b[5], f[1], c1); c2 = fmaf(b[5], f[3], c2);
    a1 = fmaf(b[6], f[0], a1); a2 = fmaf(b[6], f[2], a2);
    B1 = a1 + c1; B2 = a2 + c2;
}

// W[x] += sum_sites brow[3 + xs - x] * res(xs), |xs - x| <= 3, at the lane's columns x = 4l..4l+3
template <int Q>
__device__ __forceinline__ void back_row_acc(const float (&b)[7], float r1, float r2, float& Wx, float& Wy, float& Wz, float& Ww) {
    const unsigned FULL = 0xffffffffu;
    if (Q == 0) {
        const float L2 = __shfl_up_sync(FULL, r2, 1), R1 = __shfl_down_sync(FULL, r1, 1), R2 = __shfl_down_sync(FULL, r2, 1);
        Wx += fmaf(b[1], L2, fmaf(b[3], r1, b[5] * r2));
        Wy += fmaf(b[0], L2, b[2] * r1) + fmaf(b[4], r2, b[6] * R1);
        Wz += fmaf(b[1], r1, fmaf(b[3], r2, b[5] * R1));
        Ww += fmaf(b[0], r1, b[2] * r2) + fmaf(b[4], R1, b[6] * R2);
    } else {
        const float L1 = __shfl_up_sync(FULL, r1, 1), L2 = __shfl_up_sync(FULL, r2, 1), R1 = __shfl_down_sync(FULL, r1, 1);
        Wx += fmaf(b[0], L1, b[2] * L2) + fmaf(b[4], r1, b[6] * r2);
        Wy += fmaf(b[1], L2, fmaf(b[3], r1, b[5] * r2));
        Wz += fmaf(b[0], L2, b[2] * r1) + fmaf(b[4], r2, b[6] * R1);
        Ww += fmaf(b[1], r1, fmaf(b[3], r2, b[5] * R1));
    }
}

struct LaneConst {
    u64 A[7];                 // (acol[u], acol[u])
    float b[7];               // brow
    u64 alpha2;
    float lo, hi;
    float mI1[kMaxSlots], mI2[kMaxSlots], mL1[kMaxSlots], mL2[kMaxSlots];   // frames with a sample at the lane's two sites: interior / last site row
    float wI1[kMaxSlots], wI2[kMaxSlots], wL1[kMaxSlots], wL2[kMaxSlots];   // 1/m if the site is owned (MSE), else 0
    float lb[4];              // LEFT strip, lane 2: brow[3], brow[2], brow[1], brow[0]; else 0
    bool  lane2;              // LEFT strip, lane 2
    bool  st_ok;              // this lane stores (owned columns inside the image)
};

struct State {
    Row v[8];                 // hr window: row 2a-3+u of step T in slot (2T + u) mod 8
    Row wr[8];                // W ring: row 2a+p of step T in slot (2T + p) mod 8
    Row w0;                   // W of the duplicate site row ys = 0 (top segments; lattices with P = 1, summed)
    float e2;
};

struct WarpCtx {
    int ya, yb;               // HR rows this warp updates (and whose sites it owns for the MSE)
    int a0;                   // site-row index of step 0 (may be negative)
    int h;
    bool top;                 // ya == 0: the clamp terms of HR row 0 apply (lattices with P = 1)
    int x0;
};

// One lattice of the row whose column pass is (Cx..Cw): residuals, energy, W contribution, clamp terms.
template <int MASK, int PQ, bool GEN, bool LEFT>
__device__ __forceinline__ void lattice_part(State& s, const LaneConst& k, const WarpCtx& c, const StreamMLArgs& a, const int aj,
                                             const float* __restrict__ s1stage, const float Cx, const float Cy, const float Cz, const float Cw,
                                             const float C0x, const float C0y, const float C0z, const float C0w, const bool top_step,
                                             const bool own_fast, float& Wx, float& Wy, float& Wz, float& Ww, const int lane) {
    constexpr int P = PQ >> 1, Q = PQ & 1, S = slot_of(MASK, PQ);
    float B1, B2;
    row_pass<Q>(k.b, Cx, Cy, Cz, Cw, B1, B2);
    const float2 s1 = *reinterpret_cast<const float2*>(s1stage + S * ML_S1_FLOATS);
    float m1 = k.mI1[S], m2 = k.mI2[S], w1 = own_fast ? k.wI1[S] : 0.0f, w2 = own_fast ? k.wI2[S] : 0.0f;
    bool valid = true, last = false, own = own_fast;
    if (GEN) {
        valid = (aj >= 0) && (aj < c.h);
        last = (aj == c.h - 1);
        const int ys = 2 * aj + P;
        own = valid && (ys >= c.ya) && (ys < c.yb);
        m1 = valid ? (last ? k.mL1[S] : k.mI1[S]) : 0.0f;
        m2 = valid ? (last ? k.mL2[S] : k.mI2[S]) : 0.0f;
        w1 = own ? (last ? k.wL1[S] : k.wI1[S]) : 0.0f;
        w2 = own ? (last ? k.wL2[S] : k.wI2[S]) : 0.0f;
    }
    const float r1 = fmaf(-m1, B1, s1.x), r2 = fmaf(-m2, B2, s1.y);
    s.e2 = fmaf(r1 * w1, r1, s.e2);
    s.e2 = fmaf(r2 * w2, r2, s.e2);
    back_row_acc<Q>(k.b, r1, r2, Wx, Wy, Wz, Ww);

    const float* edge = a.edge[S];                      // colS1[h], leftS1[h], rowS1[w], topS1[w], scal[4]
    if (LEFT && Q == 1) {                               // x0 == 0: lane 2 holds columns 0..3, its first site is xs = 1
        float cS = 0.0f, lS = 0.0f;
        if (k.lane2 && valid) { cS = __ldg(edge + aj); lS = __ldg(edge + a.h + aj); }
        const float mc = valid ? a.mc[S][last] : 0.0f, ml = valid ? a.ml[S][last] : 0.0f;
        const float dcol = fmaf(-mc, B1, cS);           // duplicate residual at (ys, 0)
        Wx = fmaf(k.lb[0], dcol, Wx); Wy = fmaf(k.lb[1], dcol, Wy);
        Wz = fmaf(k.lb[2], dcol, Wz); Ww = fmaf(k.lb[3], dcol, Ww);
        // LR column 0 of the frames with sx = +1 reads B[ys, 0] (columns < 0 are zero): MSE only
        const float Bc0 = fmaf(k.b[0], Cw, fmaf(k.b[1], Cz, fmaf(k.b[2], Cy, k.b[3] * Cx)));
        const float rl = fmaf(-ml, Bc0, lS);
        if (k.lane2 && own && ml > 0.0f) s.e2 += rl * rl / ml;
    }
    if (GEN && P == 1 && top_step) {                    // aj == 0 in a top segment
        float B01, B02;
        row_pass<Q>(k.b, C0x, C0y, C0z, C0w, B01, B02);
        const int b1 = ((c.x0 - 8) >> 1) + 2 * lane, b2 = b1 + 1;
        const bool v1 = (b1 >= 0) && (b1 < a.w), v2 = (b2 >= 0) && (b2 < a.w);
        const int xs1 = 2 * b1 + Q, xs2 = 2 * b2 + Q;
        const bool o1 = v1 && xs1 >= c.x0 && xs1 < c.x0 + ML_VALID, o2 = v2 && xs2 >= c.x0 && xs2 < c.x0 + ML_VALID;
        const float* rowS1 = edge + 2 * a.h; const float* topS1 = rowS1 + a.w; const float* scal = topS1 + a.w;
        const float tS1 = v1 ? __ldg(topS1 + b1) : 0.0f, tS2 = v2 ? __ldg(topS1 + b2) : 0.0f;
        const float rS1 = v1 ? __ldg(rowS1 + b1) : 0.0f, rS2 = v2 ? __ldg(rowS1 + b2) : 0.0f;
        const float mt1 = a.mt[S][b1 == a.w - 1], mt2 = a.mt[S][b2 == a.w - 1];
        const float mr1 = v1 ? a.mr[S][b1 == a.w - 1] : 0.0f, mr2 = v2 ? a.mr[S][b2 == a.w - 1] : 0.0f;
        if (o1 && mt1 > 0.0f) { const float e = fmaf(-mt1, B01, tS1); s.e2 += e * e / mt1; }
        if (o2 && mt2 > 0.0f) { const float e = fmaf(-mt2, B02, tS2); s.e2 += e * e / mt2; }
        if (LEFT && Q == 1 && k.lane2 && a.m_msec[S] > 0.0f) {       // LR (0, 0) of the frames (+1, +1) reads B[0, 0]
            const float B00 = fmaf(k.b[0], C0w, fmaf(k.b[1], C0z, fmaf(k.b[2], C0y, k.b[3] * C0x)));
            const float e = fmaf(-a.m_msec[S], B00, __ldg(scal + 1));
            s.e2 += e * e / a.m_msec[S];
        }
        const float d1 = fmaf(-mr1, B1, rS1), d2 = fmaf(-mr2, B2, rS2);   // duplicates on HR row 0
        float Vx = 0.f, Vy = 0.f, Vz = 0.f, Vw = 0.f;
        back_row_acc<Q>(k.b, d1, d2, Vx, Vy, Vz, Vw);
        if (LEFT && Q == 1) {                           // the duplicate at (0, 0): frames (-1, -1), reads B[1, 1]
            const float d00 = fmaf(-a.m_dupc[S], B1, k.lane2 ? __ldg(scal + 0) : 0.0f);
            Vx = fmaf(k.lb[0], d00, Vx); Vy = fmaf(k.lb[1], d00, Vy);
            Vz = fmaf(k.lb[2], d00, Vz); Vw = fmaf(k.lb[3], d00, Vw);
        }
        s.w0.lo = add2(s.w0.lo, pk(Vx, Vy)); s.w0.hi = add2(s.w0.hi, pk(Vz, Vw));
    }
}

// One pair of HR rows.  T = step mod 4 fixes every register slot at compile time.
template <int MASK, int T, bool GEN, bool LEFT>
__device__ __forceinline__ void ml_step(State& s, const LaneConst& k, const WarpCtx& c, const StreamMLArgs& a, const int aj,
                                        const float* __restrict__ s1row, const float* __restrict__ pf, float*& outp, const int lane) {
    constexpr bool ROW0 = (MASK & 3) != 0, ROW1 = (MASK & 12) != 0;
    #define WIN(u) s.v[(2 * T + (u)) & 7]
    const bool top_step = GEN && ROW1 && c.top && (aj == 0);
    // fast form (interior site rows away from the image's top rows): ownership and stores still depend on where the pair sits in the
    // warp's segment (warm-up and tail iterations) -- warp-uniform flags instead of a second code path
    const bool ownE = (2 * aj >= c.ya) && (2 * aj < c.yb), ownO = (2 * aj + 1 >= c.ya) && (2 * aj + 1 < c.yb);
    // ---- column passes: even row 2a over window rows u = 0..6 (weight acol[6-u]), odd row 2a+1 over u = 1..7 (acol[7-u])
    u64 Elo = 0, Ehi = 0, Olo = 0, Ohi = 0;
    if (ROW0 || top_step) {                             // (for a = 0 the even row's column pass IS the one of HR row 0)
        u64 Dlo, Dhi;
        Elo = mul2(k.A[6], WIN(0).lo); Ehi = mul2(k.A[6], WIN(0).hi);
        Dlo = mul2(k.A[5], WIN(1).lo); Dhi = mul2(k.A[5], WIN(1).hi);
        Elo = fma2(k.A[4], WIN(2).lo, Elo); Ehi = fma2(k.A[4], WIN(2).hi, Ehi);
        Dlo = fma2(k.A[3], WIN(3).lo, Dlo); Dhi = fma2(k.A[3], WIN(3).hi, Dhi);
        Elo = fma2(k.A[2], WIN(4).lo, Elo); Ehi = fma2(k.A[2], WIN(4).hi, Ehi);
        Dlo = fma2(k.A[1], WIN(5).lo, Dlo); Dhi = fma2(k.A[1], WIN(5).hi, Dhi);
        Elo = fma2(k.A[0], WIN(6).lo, Elo); Ehi = fma2(k.A[0], WIN(6).hi, Ehi);
        Elo = add2(Elo, Dlo); Ehi = add2(Ehi, Dhi);
    }
    if (ROW1) {
        u64 Dlo, Dhi;
        Olo = mul2(k.A[6], WIN(1).lo); Ohi = mul2(k.A[6], WIN(1).hi);
        Dlo = mul2(k.A[5], WIN(2).lo); Dhi = mul2(k.A[5], WIN(2).hi);
        Olo = fma2(k.A[4], WIN(3).lo, Olo); Ohi = fma2(k.A[4], WIN(3).hi, Ohi);
        Dlo = fma2(k.A[3], WIN(4).lo, Dlo); Dhi = fma2(k.A[3], WIN(4).hi, Dhi);
        Olo = fma2(k.A[2], WIN(5).lo, Olo); Ohi = fma2(k.A[2], WIN(5).hi, Ohi);
        Dlo = fma2(k.A[1], WIN(6).lo, Dlo); Dhi = fma2(k.A[1], WIN(6).hi, Dhi);
        Olo = fma2(k.A[0], WIN(7).lo, Olo); Ohi = fma2(k.A[0], WIN(7).hi, Ohi);
        Olo = add2(Olo, Dlo); Ohi = add2(Ohi, Dhi);
    }
    const Row h1 = WIN(0), h2 = WIN(1);                 // hr rows 2a-3, 2a-2: updated at the end of this step
    // ---- the two rows the next step adds take the slots of those two -------------------------------------------------------
    s.v[(2 * T + 8) & 7] = lds_row(pf);
    s.v[(2 * T + 9) & 7] = lds_row(pf + ML_COLS);
    #undef WIN

    float Ex, Ey, Ez, Ew, Ox, Oy, Oz, Ow;
    upk(Elo, Ex, Ey); upk(Ehi, Ez, Ew); upk(Olo, Ox, Oy); upk(Ohi, Oz, Ow);
    // ---- lattices of the even row ---------------------------------------------------------------------------------------------
    if (ROW0) {
        float Wx = 0.f, Wy = 0.f, Wz = 0.f, Ww = 0.f;
        if (MASK & 1) lattice_part<MASK, 0, GEN, LEFT>(s, k, c, a, aj, s1row, Ex, Ey, Ez, Ew, 0.f, 0.f, 0.f, 0.f, false, ownE, Wx, Wy, Wz, Ww, lane);
        if (MASK & 2) lattice_part<MASK, 1, GEN, LEFT>(s, k, c, a, aj, s1row, Ex, Ey, Ez, Ew, 0.f, 0.f, 0.f, 0.f, false, ownE, Wx, Wy, Wz, Ww, lane);
        s.wr[(2 * T) & 7].lo = pk(Wx, Wy); s.wr[(2 * T) & 7].hi = pk(Wz, Ww);
    }
    if (ROW1) {
        float Wx = 0.f, Wy = 0.f, Wz = 0.f, Ww = 0.f;
        if (MASK & 4) lattice_part<MASK, 2, GEN, LEFT>(s, k, c, a, aj, s1row, Ox, Oy, Oz, Ow, Ex, Ey, Ez, Ew, top_step, ownO, Wx, Wy, Wz, Ww, lane);
        if (MASK & 8) lattice_part<MASK, 3, GEN, LEFT>(s, k, c, a, aj, s1row, Ox, Oy, Oz, Ow, Ex, Ey, Ez, Ew, top_step, ownO, Wx, Wy, Wz, Ww, lane);
        s.wr[(2 * T + 1) & 7].lo = pk(Wx, Wy); s.wr[(2 * T + 1) & 7].hi = pk(Wz, Ww);
    }

    // ---- transposed column pass: HR row y1 = 2a-3 from W rows 2a-6+d, y2 = 2a-2 from W rows 2a-5+d (weight acol[d]) ----------------
    #define WR(r) s.wr[(2 * T + 8 + (r)) & 7]           // W row 2a + r, r = -6 .. 1
    u64 o1lo = 0, o1hi = 0, o2lo = 0, o2hi = 0;
    {
        u64 plo = 0, phi = 0, qlo = 0, qhi = 0;         // two partial sums per row keep the chains short
        #pragma unroll
        for (int d = 0; d < 7; ++d) {
            const bool act1 = (d & 1) ? ROW1 : ROW0;    // parity of row 2a-6+d
            if (act1) { if (d & 2) { plo = fma2(k.A[d], WR(d - 6).lo, plo); phi = fma2(k.A[d], WR(d - 6).hi, phi); }
                        else       { o1lo = fma2(k.A[d], WR(d - 6).lo, o1lo); o1hi = fma2(k.A[d], WR(d - 6).hi, o1hi); } }
            const bool act2 = (d & 1) ? ROW0 : ROW1;    // parity of row 2a-5+d
            if (act2) { if (d & 2) { qlo = fma2(k.A[d], WR(d - 5).lo, qlo); qhi = fma2(k.A[d], WR(d - 5).hi, qhi); }
                        else       { o2lo = fma2(k.A[d], WR(d - 5).lo, o2lo); o2hi = fma2(k.A[d], WR(d - 5).hi, o2hi); } }
        }
        o1lo = add2(o1lo, plo); o1hi = add2(o1hi, phi); o2lo = add2(o2lo, qlo); o2hi = add2(o2hi, qhi);
    }
    #undef WR
    const int y1 = 2 * aj - 3;
    if (GEN && ROW1 && c.top) {                         // HR rows 0..3 also see the duplicate site row ys = 0
        if (y1 >= 0 && y1 < 4) { const float g = a.acol[3 - y1]; const u64 g2 = pk(g, g); o1lo = fma2(g2, s.w0.lo, o1lo); o1hi = fma2(g2, s.w0.hi, o1hi); }
        if (y1 + 1 >= 0 && y1 + 1 < 4) { const float g = a.acol[2 - y1]; const u64 g2 = pk(g, g); o2lo = fma2(g2, s.w0.lo, o2lo); o2hi = fma2(g2, s.w0.hi, o2hi); }
    }
    float4 q1, q2;
    upk(fma2(k.alpha2, o1lo, h1.lo), q1.x, q1.y); upk(fma2(k.alpha2, o1hi, h1.hi), q1.z, q1.w);
    upk(fma2(k.alpha2, o2lo, h2.lo), q2.x, q2.y); upk(fma2(k.alpha2, o2hi, h2.hi), q2.z, q2.w);
    q1.x = clipf(q1.x, k.lo, k.hi); q1.y = clipf(q1.y, k.lo, k.hi); q1.z = clipf(q1.z, k.lo, k.hi); q1.w = clipf(q1.w, k.lo, k.hi);
    q2.x = clipf(q2.x, k.lo, k.hi); q2.y = clipf(q2.y, k.lo, k.hi); q2.z = clipf(q2.z, k.lo, k.hi); q2.w = clipf(q2.w, k.lo, k.hi);
    const bool st1 = k.st_ok && (y1 >= c.ya) && (y1 < c.yb), st2 = k.st_ok && (y1 + 1 >= c.ya) && (y1 + 1 < c.yb);
    if (st1) *reinterpret_cast<float4*>(outp) = q1;
    if (st2) *reinterpret_cast<float4*>(outp + a.W) = q2;
    outp += 2 * a.W;
}

template <int MASK, bool GEN, bool LEFT>
__device__ __forceinline__ void ml_iteration(State& s, const LaneConst& k, const WarpCtx& c, const StreamMLArgs& a, const int it,
                                             const float* __restrict__ ring, uint64_t* bars, float*& outp, const int lane) {
    constexpr int NL = popc4(MASK), STAGE_FLOATS = ML_HR_FLOATS + NL * ML_S1_FLOATS;
    const int slot = (it + 1) % ML_STAGES, nslot = (it + 2) % ML_STAGES;
    const float* hrs = ring + slot * STAGE_FLOATS + 4 * lane;
    const float* s1s = ring + slot * STAGE_FLOATS + ML_HR_FLOATS + 2 * lane;
    const float* nxt = ring + nslot * STAGE_FLOATS + 4 * lane;
    const int aj = c.a0 + 4 * it;
    ml_step<MASK, 0, GEN, LEFT>(s, k, c, a, aj + 0, s1s + 0 * 64, hrs + 2 * ML_COLS, outp, lane);
    ml_step<MASK, 1, GEN, LEFT>(s, k, c, a, aj + 1, s1s + 1 * 64, hrs + 4 * ML_COLS, outp, lane);
    ml_step<MASK, 2, GEN, LEFT>(s, k, c, a, aj + 2, s1s + 2 * 64, hrs + 6 * ML_COLS, outp, lane);
    mbar_wait(&bars[nslot], (uint32_t)(((it + 2) / ML_STAGES) & 1));
    ml_step<MASK, 3, GEN, LEFT>(s, k, c, a, aj + 3, s1s + 3 * 64, nxt, outp, lane);
}

template <int MASK>
__device__ __forceinline__ void ml_issue(const StreamMLArgs& a, float* ring, uint64_t* bars, int stage, int xs0, int sb, int a0) {
    constexpr int NL = popc4(MASK), STAGE_FLOATS = ML_HR_FLOATS + NL * ML_S1_FLOATS;
    const int slot = stage % ML_STAGES;
    float* dst = ring + slot * STAGE_FLOATS;
    mbar_expect_tx(&bars[slot], 4u * STAGE_FLOATS);
    tma_load_2d(dst, &a.tm_hr, xs0, sb + ML_ROWS * (stage - 1), &bars[slot]);
    #pragma unroll
    for (int s = 0; s < NL; ++s)
        tma_load_2d(dst + ML_HR_FLOATS + s * ML_S1_FLOATS, &a.tm_s1[s], xs0 >> 1, a0 + ML_SROWS * (stage - 1), &bars[slot]);
}

template <int MASK>
__global__ void __launch_bounds__(ML_THREADS, ML_CTAS) k_ibp_stream_ml(const __grid_constant__ StreamMLArgs a) {
    constexpr int NL = popc4(MASK), STAGE_FLOATS = ML_HR_FLOATS + NL * ML_S1_FLOATS;
    constexpr bool HASQ1 = (MASK & 10) != 0;
    extern __shared__ __align__(128) float sm[];
    __shared__ __align__(8) uint64_t bars_all[ML_WARPS][ML_STAGES];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int gw = blockIdx.x * ML_WARPS + warp;
    if (gw >= a.nseg * a.nstrip) return;                 // whole warps only; nothing below synchronises the CTA
    float* ring = sm + warp * (ML_STAGES * STAGE_FLOATS);
    uint64_t* bars = bars_all[warp];

    WarpCtx c;
    const int strip = gw % a.nstrip, nit = a.nit;
    c.ya = (gw / a.nstrip) * a.seg_rows;
    c.yb = min(c.ya + a.seg_rows, a.H);
    c.x0 = strip * ML_VALID;
    c.h = a.h;
    c.top = (c.ya == 0);
    c.a0 = (c.ya >> 1) - 2;                              // step t writes HR rows ya-7+2t, ya-6+2t: row ya is the second row of step 3
    const int sb = 2 * c.a0 + 3;                         // first NEW hr row of step 0 (its window is rows 2a0-3 .. 2a0+4)
    const int xs0 = c.x0 - 8;

    asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory");
    if (lane == 0) {
        #pragma unroll
        for (int i = 0; i < ML_STAGES; ++i) mbar_init(&bars[i], 1);
        mbar_fence_init();
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    }
    asm volatile("griddepcontrol.wait;\n" ::: "memory");
    if (lane == 0) {
        #pragma unroll
        for (int st = 0; st < ML_STAGES; ++st) ml_issue<MASK>(a, ring, bars, st, xs0, sb, c.a0);
    }
    __syncwarp();

    LaneConst k;
    #pragma unroll
    for (int u = 0; u < 7; ++u) { k.A[u] = pk(a.acol[u], a.acol[u]); k.b[u] = a.brow[u]; }
    k.alpha2 = pk(a.alpha, a.alpha); k.lo = a.lo; k.hi = a.hi;
    {
        const int b1 = (xs0 >> 1) + 2 * lane, b2 = b1 + 1;
        const bool v1 = (b1 >= 0) && (b1 < a.w), v2 = (b2 >= 0) && (b2 < a.w);
        const int l1 = (b1 == a.w - 1), l2 = (b2 == a.w - 1);
        #pragma unroll
        for (int pq = 0; pq < 4; ++pq) {
            if (!(MASK & (1 << pq))) continue;
            const int S = slot_of(MASK, pq), Q = pq & 1;
            const int xs1 = 2 * b1 + Q, xs2 = 2 * b2 + Q;
            const bool o1 = v1 && xs1 >= c.x0 && xs1 < c.x0 + ML_VALID, o2 = v2 && xs2 >= c.x0 && xs2 < c.x0 + ML_VALID;
            k.mI1[S] = v1 ? a.mtab[S][l1] : 0.0f;     k.mI2[S] = v2 ? a.mtab[S][l2] : 0.0f;
            k.mL1[S] = v1 ? a.mtab[S][2 + l1] : 0.0f; k.mL2[S] = v2 ? a.mtab[S][2 + l2] : 0.0f;
            k.wI1[S] = (o1 && k.mI1[S] > 0.0f) ? 1.0f / k.mI1[S] : 0.0f; k.wI2[S] = (o2 && k.mI2[S] > 0.0f) ? 1.0f / k.mI2[S] : 0.0f;
            k.wL1[S] = (o1 && k.mL1[S] > 0.0f) ? 1.0f / k.mL1[S] : 0.0f; k.wL2[S] = (o2 && k.mL2[S] > 0.0f) ? 1.0f / k.mL2[S] : 0.0f;
        }
    }
    const bool left = HASQ1 && (c.x0 == 0);
    k.lane2 = left && (lane == 2);
    #pragma unroll
    for (int i = 0; i < 4; ++i) k.lb[i] = k.lane2 ? a.brow[3 - i] : 0.0f;
    const int gx = xs0 + 4 * lane;
    k.st_ok = (lane >= 2) && (lane < 30) && (gx < a.W);

    State s;
    #pragma unroll
    for (int i = 0; i < 8; ++i) { s.wr[i].lo = 0; s.wr[i].hi = 0; }
    s.w0.lo = 0; s.w0.hi = 0; s.e2 = 0.0f;
    float* outp = a.hr_out + (ptrdiff_t)(2 * c.a0 - 3) * a.W + gx;    // row of step 0's first output (never dereferenced outside the image)

    // ---- prologue: window rows 2a0-3 .. 2a0+2 (stage 0 rows 2..7) and step 0's two new rows (stage 1 rows 0, 1) ---------------------
    mbar_wait(&bars[0], 0);
    mbar_wait(&bars[1], 0);
    {
        const float* p0 = ring + 4 * lane;
        const float* p1 = ring + STAGE_FLOATS + 4 * lane;
        #pragma unroll
        for (int u = 0; u < 6; ++u) s.v[u] = lds_row(p0 + (u + 2) * ML_COLS);
        s.v[6] = lds_row(p1);
        s.v[7] = lds_row(p1 + ML_COLS);
    }
    __syncwarp();                                        // stage 0 is consumed: its slot takes stage 3
    if (lane == 0 && ML_STAGES <= nit + 1) ml_issue<MASK>(a, ring, bars, ML_STAGES, xs0, sb, c.a0);

    for (int it = 0; it < nit; ++it) {
        // fast form: the four pairs are interior site rows (all frames present) and none of the rows written sees the top clamp terms
        const int aj0 = c.a0 + 4 * it, aj3 = aj0 + 3;
        const bool fast = (aj0 >= 0) && (aj3 < c.h - 1) && !(c.top && 2 * aj0 - 3 < 4);
        if (left) {
            if (fast) ml_iteration<MASK, false, true>(s, k, c, a, it, ring, bars, outp, lane);
            else      ml_iteration<MASK, true, true>(s, k, c, a, it, ring, bars, outp, lane);
        } else {
            if (fast) ml_iteration<MASK, false, false>(s, k, c, a, it, ring, bars, outp, lane);
            else      ml_iteration<MASK, true, false>(s, k, c, a, it, ring, bars, outp, lane);
        }
        __syncwarp();                                    // stage it+1 is consumed: refill its slot with stage it+4
        if (lane == 0 && it + 4 <= nit + 1) ml_issue<MASK>(a, ring, bars, it + 4, xs0, sb, c.a0);
    }

    float e2 = s.e2;
    #pragma unroll
    for (int off = 16; off > 0; off >>= 1) e2 += __shfl_xor_sync(0xffffffffu, e2, off);
    if (lane == 0) a.partial[gw] = (double)e2;
}

template <int MASK>
int launch_mask(const StreamMLArgs& a, cudaStream_t st) {
    constexpr int NL = popc4(MASK);
    const int warps = a.nseg * a.nstrip;
    const size_t smem = sizeof(float) * (size_t)ML_WARPS * ML_STAGES * (ML_HR_FLOATS + NL * ML_S1_FLOATS);
    static PerDeviceOnce configured;
    if (configured.need()) SR_CUDA(cudaFuncSetAttribute(k_ibp_stream_ml<MASK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((warps + ML_WARPS - 1) / ML_WARPS); cfg.blockDim = dim3(ML_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    { static const bool no_pdl = getenv("SR459_NO_PDL") != nullptr; cfg.attrs = attr; cfg.numAttrs = no_pdl ? 0 : 1; }
    const cudaError_t le = cudaLaunchKernelEx(&cfg, k_ibp_stream_ml<MASK>, a);
    if (le != cudaSuccess) { set_error("launch of k_ibp_stream_ml failed: %s", cudaGetErrorString(le)); return SR459_ECUDA; }
    SR_LAUNCH_CHECK("k_ibp_stream_ml");
    return SR459_OK;
}

}  // namespace

bool ibp_stream_ml_supported(int kh, int kw, int smax, bool separable, int W, int nslot) {
    return separable && kh == kw && (kh == 1 || kh == 3 || kh == 5 || kh == 7) && smax <= 1 && (W % 4) == 0 && nslot >= 2 && nslot <= kMaxSlots;
}

// rows per warp: seg_rows = 8 nit - 8 (the first 7 and the last output row of a warp's 8 nit rows fall outside its segment)
void stream_ml_geometry(int H, int W, int n_sm, int* nit_out, int* seg_rows, int* nseg_out, int* nstrip_out) {
    const int nstrip = (W + ML_VALID - 1) / ML_VALID;
    const int slots = std::max(1, n_sm) * ML_CTAS;
    double best = 1e30; int best_nit = 2;
    for (int nit = 2; nit <= 128; ++nit) {
        const int L = 8 * nit - 8;
        const int nseg = (H + L - 1) / L;
        const int ctas = (nseg * nstrip + ML_WARPS - 1) / ML_WARPS;
        const int waves = (ctas + slots - 1) / slots;
        const double cost = (double)waves * (nit + 1.5);
        if (cost < best - 1e-9) { best = cost; best_nit = nit; }
        if (L >= H) break;
    }
    *nit_out = best_nit; *seg_rows = 8 * best_nit - 8;
    *nseg_out = (H + *seg_rows - 1) / *seg_rows; *nstrip_out = nstrip;
}

int stream_ml_make_hr_map(CUtensorMap* out, const float* base, int W, int rows) {
    return encode_tile_map_f32(out, base, (uint64_t)W, (uint64_t)rows, (uint64_t)W, ML_COLS, ML_ROWS);
}

int launch_ibp_stream_ml(const StreamMLArgs& a, int mask, cudaStream_t st) {
    if (a.nseg * a.nstrip <= 0) return SR459_OK;
    switch (mask) {
        case 3:  return launch_mask<3>(a, st);
        case 5:  return launch_mask<5>(a, st);
        case 6:  return launch_mask<6>(a, st);
        case 7:  return launch_mask<7>(a, st);
        case 9:  return launch_mask<9>(a, st);
        case 10: return launch_mask<10>(a, st);
        case 11: return launch_mask<11>(a, st);
        case 12: return launch_mask<12>(a, st);
        case 13: return launch_mask<13>(a, st);
        case 14: return launch_mask<14>(a, st);
        case 15: return launch_mask<15>(a, st);
        default: break;
    }
    set_error("k_ibp_stream_ml: lattice mask %d has fewer than two lattices", mask);
    return SR459_EUNSUPPORTED;
}

}  // namespace sr459
