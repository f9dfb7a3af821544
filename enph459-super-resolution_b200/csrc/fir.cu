// fir.cu -- back-projection with FRACTIONAL shifts as one table-driven FIR kernel.
//
// The reference's back_project (mono_cal_target/run_sr.py:168-178) zero-inserts the LR residual onto the HR grid and
// calls scipy.ndimage.shift(up, (-2dy, -2dx), order=3, mode='nearest'): edge-pad by 12, cubic-B-spline prefilter
// (an all-pole IIR filter), 4-tap spline evaluation at o + s + 12.  Every step is linear and, per axis, the chain
// is  out[o] = sum_j w_j(o) * sum_d h[d] * up[clamp(p_j(o) - d - 12)]  with h[d] = sqrt(3) z^|d| (the prefilter's impulse
// response, z = sqrt(3) - 2; |d| <= 16 keeps everything above 7e-10) and up[Y] = err[Y/f] on multiples of the factor f,
// zero elsewhere.  Folding the zero pattern, the clamps of the edge padding and the spline weights together gives,
// for every HR output index o, at most 20 taps on consecutive LR samples:  out[o] = sum_a M[o][a] * err[i0[o] + a].
// The tables M (float, built in double on the host) carry ALL boundary rules; the kernel is a plain separable
// 20 x 20-tap gather per frame, accumulated over the frames -- no full-resolution prefilter (4 frames x 2 passes over
// (H+24) x (W+24)) and no 64-tap coefficient gather (k_prefilter2d + k_resample_sum of the generic path).
#include "internal.h"
#include <math.h>
#include <algorithm>

namespace sr459 {

static inline void fir_cubic_w(double t, double w[4]) {     // scipy's cubic B-spline weights at fraction t
    const double u = 1.0 - t;
    w[0] = u * u * u / 6.0;
    w[1] = (t * t * (t - 2.0) * 3.0 + 4.0) / 6.0;
    w[2] = (u * u * (u - 2.0) * 3.0 + 4.0) / 6.0;
    w[3] = 1.0 - w[0] - w[1] - w[2];
}

// One axis: HR length N = f * n_lr, shift s (HR pixels, the argument scipy.ndimage.shift receives).
// i0[o] (first LR tap) and w[o * kFirTaps + a].  Returns false if the axis is too short for the fixed tap count.
bool build_fir_bp_axis(int n_lr, int f, double s, std::vector<int>& i0, std::vector<float>& w) {
    const int N = n_lr * f, NP = N + 2 * kPad, R = 16;
    if (n_lr < kFirTaps) return false;
    const double z = sqrt(3.0) - 2.0;
    double h[2 * 16 + 1];
    for (int d = -R; d <= R; ++d) h[d + R] = sqrt(3.0) * pow(z, abs(d));
    i0.assign(N, 0);
    w.assign((size_t)N * kFirTaps, 0.0f);
    std::vector<double> acc(kFirTaps);
    for (int o = 0; o < N; ++o) {
        double x = (double)o - s + (double)kPad;
        x = std::min(std::max(x, 0.0), (double)(NP - 1));           // mode='nearest' coordinate clip
        const double fl = floor(x);
        double cw[4];
        fir_cubic_w(x - fl, cw);
        const long long base = (long long)fl - 1;
        // first LR sample any tap can touch (before clamping), then clamp the window into the array
        long long ylo = base - R - kPad;
        long long first = ylo <= 0 ? 0 : (ylo + f - 1) / f;
        first = std::min<long long>(std::max<long long>(first, 0), n_lr - kFirTaps);
        std::fill(acc.begin(), acc.end(), 0.0);
        for (int j = 0; j < 4; ++j) {
            const long long p = std::min<long long>(std::max<long long>(base + j, 0), NP - 1);   // tap index clamp
            for (int d = -R; d <= R; ++d) {
                long long Y = p - d - kPad;                           // index into the unpadded HR array
                Y = std::min<long long>(std::max<long long>(Y, 0), N - 1);     // edge padding + 'nearest' extension
                if (Y % f) continue;                                  // zero-inserted sample
                const long long i = Y / f;
                const long long a = i - first;
                if (a < 0 || a >= kFirTaps) return false;             // cannot happen for |s| << 12; refuse rather than drop
                acc[(size_t)a] += cw[j] * h[d + R];
            }
        }
        i0[o] = (int)first;
        for (int a = 0; a < kFirTaps; ++a) w[(size_t)o * kFirTaps + a] = (float)acc[(size_t)a];
    }
    return true;
}

// ---------------------------------------------------------------------------------------------------------------
// kernel: out[Y][X] = sum_k sum_a My_k[Y][a] * sum_b Mx_k[X][b] * err_k[iy0_k[Y] + a][ix0_k[X] + b]
// One CTA (256 threads) owns a 32 x 64 output tile (narrow: with k_fir_poly taking the interior these are the image
// borders, one latency-bound wave); per frame: stage the LR region, x pass into T, y pass into register accumulators.
// ---------------------------------------------------------------------------------------------------------------
constexpr int FB_TH = 32, FB_TW = 64;
constexpr int FB_LPR = FB_TW / 4;                  // y pass: lanes per output row (4 columns per lane)
constexpr int FB_RPI = 8 * (32 / FB_LPR);           // y pass: output rows per CTA iteration
constexpr int FB_NI  = FB_TH / FB_RPI;
constexpr int FB_RR = FB_TH / 2 + kFirTaps + 2;     // 38 staged LR rows   (factor >= 2)
constexpr int FB_RC = FB_TW / 2 + kFirTaps + 4;     // 56 staged LR columns
constexpr int FB_RP = FB_RC + 1;

__global__ void __launch_bounds__(256, 3) k_fir_bp(const __grid_constant__ FirBpArgs a) {
    __shared__ float R[FB_RR * FB_RP];
    __shared__ __align__(16) float T[FB_RR * FB_TW];
    __shared__ float wy[FB_TH * kFirTaps];
    __shared__ int iy[FB_TH];
    __shared__ int bnd[4];
    const int tid = threadIdx.x, lane = tid & 31;
    const int y0 = blockIdx.y * FB_TH, x0 = blockIdx.x * FB_TW;
    if (y0 >= a.skip_y0 && y0 < a.skip_y1 && x0 >= a.skip_x0 && x0 < a.skip_x1) return;    // k_fir_poly's tiles
    const int th = min(FB_TH, a.H - y0), tw = min(FB_TW, a.W - x0);
    float acc[FB_NI][4];
    #pragma unroll
    for (int i = 0; i < FB_NI; ++i) { acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f; }
    const int yrow = (tid / FB_LPR), ycol = tid % FB_LPR;      // y pass: row within an iteration, 4-column group

    for (int k = 0; k < a.n; ++k) {
        const int* iy0 = a.iy0 + (size_t)k * a.H + y0;
        const int* ix0 = a.ix0 + (size_t)k * a.W + x0;
        const float* wyk = a.wy + ((size_t)k * a.H + y0) * kFirTaps;
        const float* wxk = a.wx + ((size_t)k * a.W + x0) * kFirTaps;
        __syncthreads();                                   // previous frame's passes are done with R / T / wy
        if (tid < 64) {                                    // bounds of the LR region (i0 is monotone, but stay general)
            const int* src = tid < 32 ? iy0 : ix0;
            const int cnt = tid < 32 ? th : tw;
            int lo = 0x7fffffff, hi = -1;
            for (int q = lane; q < cnt; q += 32) { const int v = __ldg(src + q); lo = min(lo, v); hi = max(hi, v); }
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1) { lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o)); }
            if (lane == 0) { bnd[(tid >> 5) * 2] = lo; bnd[(tid >> 5) * 2 + 1] = hi; }
        }
        for (int q = tid; q < th * kFirTaps; q += 256) wy[q] = __ldg(wyk + q);
        if (tid < th) iy[tid] = __ldg(iy0 + tid);
        __syncthreads();
        const int rmin = bnd[0], cmin = bnd[2];
        const int nr = min(bnd[1] - rmin + kFirTaps, FB_RR), nc = min(bnd[3] - cmin + kFirTaps, FB_RC);
        {   // stage: all loads of a thread first
            constexpr int NST = (FB_RR * FB_RC + 255) / 256;
            float v[NST];
            const float* src = a.err[k] + (size_t)rmin * a.w + cmin;
            #pragma unroll
            for (int q = 0; q < NST; ++q) {
                const int e = tid + 256 * q, rr = e / FB_RC, cc = e - rr * FB_RC;
                v[q] = (rr < nr && cc < nc && rmin + rr < a.h && cmin + cc < a.w) ? __ldg(src + (size_t)rr * a.w + cc) : 0.f;
            }
            #pragma unroll
            for (int q = 0; q < NST; ++q) {
                const int e = tid + 256 * q, rr = e / FB_RC, cc = e - rr * FB_RC;
                if (rr < FB_RR) R[rr * FB_RP + cc] = v[q];
            }
        }
        __syncthreads();
        {   // x pass: T[rr][x], one column per thread, two row phases
            constexpr int NPH = 256 / FB_TW;
            const int x = tid & (FB_TW - 1), half = tid / FB_TW;
            if (x < tw) {
                float wreg[kFirTaps];
                #pragma unroll
                for (int b = 0; b < kFirTaps; ++b) wreg[b] = __ldg(wxk + (size_t)x * kFirTaps + b);
                const int off = __ldg(ix0 + x) - cmin;
                for (int rr = half; rr < nr; rr += NPH) {
                    const float* r = R + rr * FB_RP + off;
                    float s = 0.f;
                    #pragma unroll
                    for (int b = 0; b < kFirTaps; ++b) s = fmaf(wreg[b], r[b], s);
                    T[rr * FB_TW + x] = s;
                }
            } else {
                for (int rr = half; rr < nr; rr += NPH) T[rr * FB_TW + x] = 0.f;
            }
        }
        __syncthreads();
        {   // y pass: rows yrow, yrow + FB_RPI, ...; columns 4*ycol .. 4*ycol+3
            const float4* tp = reinterpret_cast<const float4*>(T) + ycol;
            #pragma unroll
            for (int i = 0; i < FB_NI; ++i) {
                const int y = yrow + FB_RPI * i;
                if (y < th) {
                    const float* wrow = wy + y * kFirTaps;
                    const float4* t0 = tp + (iy[y] - rmin) * (FB_TW / 4);
                    #pragma unroll
                    for (int q = 0; q < kFirTaps; ++q) {
                        const float4 v = t0[q * (FB_TW / 4)];
                        const float wq = wrow[q];
                        acc[i][0] = fmaf(wq, v.x, acc[i][0]); acc[i][1] = fmaf(wq, v.y, acc[i][1]);
                        acc[i][2] = fmaf(wq, v.z, acc[i][2]); acc[i][3] = fmaf(wq, v.w, acc[i][3]);
                    }
                }
            }
        }
    }
    const int gx = x0 + 4 * ycol;
    #pragma unroll
    for (int i = 0; i < FB_NI; ++i) {
        const int gy = y0 + yrow + FB_RPI * i;
        if (gy >= a.H || gx >= a.W) continue;
        float* d = a.out + (size_t)gy * a.W + gx;
        if (gx + 3 < a.W && (a.W & 3) == 0) *reinterpret_cast<float4*>(d) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        else for (int q = 0; q < 4 && gx + q < a.W; ++q) d[q] = acc[i][q];
    }
}

int launch_fir_bp(const FirBpArgs& a, cudaStream_t st) {
    dim3 grid((a.W + FB_TW - 1) / FB_TW, (a.H + FB_TH - 1) / FB_TH);
    k_fir_bp<<<grid, 256, 0, st>>>(a);
    SR_LAUNCH_CHECK("k_fir_bp");
    return SR459_OK;
}


// ---------------------------------------------------------------------------------------------------------------
// polyphase interior kernel
// ---------------------------------------------------------------------------------------------------------------
// One CTA (256 threads) owns an 80 x 128 output tile.  Per frame: stage 64 x 88 LR samples; x pass: lane <-> staged row
// (odd pitches: conflict-free), one task = 16 consecutive outputs of a row from a 28-sample register window, the 20
// weights of the parity in registers (FMA : LDS = 11 : 1 instead of 1 : 1 in k_fir_bp); y pass: thread <-> column,
// 40 output rows per thread in two chunks, a 30-sample register window down the column.
constexpr int FP_TH = 80, FP_TW = 128;
constexpr int FP_RR = 64, FP_RC = 88, FP_RP = 89;      // staged rows (FP_TH/2 + 21 = 61 used), columns (64 + 21 = 85 used), pitch
constexpr int FP_TP = 129;                             // pitch of the x-passed rows

template <int D0, int D1>
__device__ __forceinline__ void fp_xpass(const float* __restrict__ r, float* __restrict__ t, const float* w0, const float* w1) {
    float v[28], wr[kFirTaps];
    #pragma unroll
    for (int j = 0; j < 28; ++j) v[j] = r[j];
    #pragma unroll
    for (int b = 0; b < kFirTaps; ++b) wr[b] = w0[b];
    #pragma unroll
    for (int m = 0; m < 8; ++m) {
        float s = 0.f;
        #pragma unroll
        for (int b = 0; b < kFirTaps; ++b) s = fmaf(wr[b], v[m + D0 + b], s);
        t[2 * m] = s;
    }
    #pragma unroll
    for (int b = 0; b < kFirTaps; ++b) wr[b] = w1[b];
    #pragma unroll
    for (int m = 0; m < 8; ++m) {
        float s = 0.f;
        #pragma unroll
        for (int b = 0; b < kFirTaps; ++b) s = fmaf(wr[b], v[m + D1 + b], s);
        t[2 * m + 1] = s;
    }
}

template <int D0, int D1>
__device__ __forceinline__ void fp_ypass(const float* __restrict__ tc, float (&acc)[2][2][10], const float* w0, const float* w1) {
    #pragma unroll
    for (int c = 0; c < 2; ++c) {
        float v[30], wr[kFirTaps];
        #pragma unroll
        for (int j = 0; j < 30; ++j) v[j] = tc[(10 * c + j) * FP_TP];
        #pragma unroll
        for (int b = 0; b < kFirTaps; ++b) wr[b] = w0[b];
        #pragma unroll
        for (int m = 0; m < 10; ++m) {
            float s = acc[c][0][m];
            #pragma unroll
            for (int b = 0; b < kFirTaps; ++b) s = fmaf(wr[b], v[m + D0 + b], s);
            acc[c][0][m] = s;
        }
        #pragma unroll
        for (int b = 0; b < kFirTaps; ++b) wr[b] = w1[b];
        #pragma unroll
        for (int m = 0; m < 10; ++m) {
            float s = acc[c][1][m];
            #pragma unroll
            for (int b = 0; b < kFirTaps; ++b) s = fmaf(wr[b], v[m + D1 + b], s);
            acc[c][1][m] = s;
        }
    }
}

__global__ void __launch_bounds__(256, 2) k_fir_poly(const __grid_constant__ FirPolyArgs a) {
    extern __shared__ float fp_smem[];
    float* R = fp_smem;                      // FP_RR x FP_RP
    float* T = fp_smem + FP_RR * FP_RP;      // FP_RR x FP_TP
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ty0 = a.y0 + blockIdx.y * FP_TH, tx0 = a.x0 + blockIdx.x * FP_TW;
    const int th = min(FP_TH, a.y1 - ty0), tw = min(FP_TW, a.x1 - tx0);
    const int x = tid & (FP_TW - 1), hh = tid >> 7;
    float acc[2][2][10];
    #pragma unroll
    for (int c = 0; c < 2; ++c)
        #pragma unroll
        for (int m = 0; m < 10; ++m) acc[c][0][m] = acc[c][1][m] = 0.f;

    for (int k = 0; k < a.n; ++k) {
        const FirPolyFrame fr = a.fr[k];
        const int rbase = (ty0 >> 1) + fr.cy, cbase = (tx0 >> 1) + fr.cx;
        __syncthreads();                                 // the previous frame's passes are done with R and T
        {   // stage: warp w takes rows w, w+8, ...; two batches of 4 rows x 3 column chunks, loads first
            const float* src = a.err[k];
            #pragma unroll
            for (int half = 0; half < 2; ++half) {
                float v[4][3];
                #pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int rr = warp + 8 * (4 * half + i), gr = rbase + rr;
                    #pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        const int cc = lane + 32 * c, gc = cbase + cc;
                        v[i][c] = (cc < FP_RC && gr >= 0 && gr < a.h && gc >= 0 && gc < a.w) ? __ldg(src + (size_t)gr * a.w + gc) : 0.f;
                    }
                }
                #pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int rr = warp + 8 * (4 * half + i);
                    #pragma unroll
                    for (int c = 0; c < 3; ++c)
                        if (lane + 32 * c < FP_RC) R[rr * FP_RP + lane + 32 * c] = v[i][c];
                }
            }
        }
        __syncthreads();
        {   // x pass: lane <-> staged row, warp <-> 16-output segment, two row halves
            const float* w0 = a.wx[k][0]; const float* w1 = a.wx[k][1];
            const int dsel = fr.dx[0] * 2 + fr.dx[1];
            #pragma unroll 1
            for (int half = 0; half < 2; ++half) {
                const int rr = lane + 32 * half;
                const float* r = R + rr * FP_RP + 8 * warp;
                float* t = T + rr * FP_TP + 16 * warp;
                if (dsel == 0)      fp_xpass<0, 0>(r, t, w0, w1);
                else if (dsel == 1) fp_xpass<0, 1>(r, t, w0, w1);
                else if (dsel == 2) fp_xpass<1, 0>(r, t, w0, w1);
                else                fp_xpass<1, 1>(r, t, w0, w1);
            }
        }
        __syncthreads();
        {   // y pass: output rows 40 hh + 2m + p read x-passed rows 20 hh + m + d[p] + tap
            const float* w0 = a.wy[k][0]; const float* w1 = a.wy[k][1];
            const float* tc = T + (20 * hh) * FP_TP + x;
            const int dsel = fr.dy[0] * 2 + fr.dy[1];
            if (dsel == 0)      fp_ypass<0, 0>(tc, acc, w0, w1);
            else if (dsel == 1) fp_ypass<0, 1>(tc, acc, w0, w1);
            else if (dsel == 2) fp_ypass<1, 0>(tc, acc, w0, w1);
            else                fp_ypass<1, 1>(tc, acc, w0, w1);
        }
    }
    if (x >= tw) return;                                // ragged last tile column of the region
    float* o = a.out + (size_t)ty0 * a.W + tx0 + x;
    #pragma unroll
    for (int c = 0; c < 2; ++c)
        #pragma unroll
        for (int m = 0; m < 10; ++m) {
            const int yy = 40 * hh + 20 * c + 2 * m;
            if (yy < th)     o[(size_t)yy * a.W]       = acc[c][0][m];
            if (yy + 1 < th) o[(size_t)(yy + 1) * a.W] = acc[c][1][m];
        }
}

// Interior (2-periodic) range of one axis table: pattern taken at the centre, extended outwards while it holds.
static bool fp_axis_range(int N, const int* i0, const float* w, int c[2], float W[2][kFirTaps], int* lo, int* hi) {
    const int oc = (N / 2) & ~1;
    if (oc + 1 >= N) return false;
    for (int p = 0; p < 2; ++p) {
        c[p] = i0[oc + p] - oc / 2;
        for (int a = 0; a < kFirTaps; ++a) W[p][a] = w[(size_t)(oc + p) * kFirTaps + a];
    }
    auto ok = [&](int o) {
        const int p = o & 1;
        if (i0[o] != (o - p) / 2 + c[p]) return false;
        for (int a = 0; a < kFirTaps; ++a)
            if (fabsf(w[(size_t)o * kFirTaps + a] - W[p][a]) > 1e-7f) return false;
        return true;
    };
    int l = oc, h = oc;
    while (l > 0 && ok(l - 1)) --l;
    while (h < N && ok(h)) ++h;
    *lo = l; *hi = h;
    return h > l;
}

bool fir_poly_analyse(int n, int H, int W, const std::vector<int>& iy0, const std::vector<int>& ix0,
                      const std::vector<float>& wy, const std::vector<float>& wx, FirPolyArgs& out) {
    if (n < 1 || n > 8) return false;
    int ylo = 0, yhi = H, xlo = 0, xhi = W;
    for (int k = 0; k < n; ++k) {
        int c[2], lo, hi;
        if (!fp_axis_range(H, iy0.data() + (size_t)k * H, wy.data() + (size_t)k * H * kFirTaps, c, out.wy[k], &lo, &hi)) return false;
        out.fr[k].cy = std::min(c[0], c[1]); out.fr[k].dy[0] = c[0] - out.fr[k].cy; out.fr[k].dy[1] = c[1] - out.fr[k].cy;
        ylo = std::max(ylo, lo); yhi = std::min(yhi, hi);
        if (!fp_axis_range(W, ix0.data() + (size_t)k * W, wx.data() + (size_t)k * W * kFirTaps, c, out.wx[k], &lo, &hi)) return false;
        out.fr[k].cx = std::min(c[0], c[1]); out.fr[k].dx[0] = c[0] - out.fr[k].cx; out.fr[k].dx[1] = c[1] - out.fr[k].cx;
        xlo = std::max(xlo, lo); xhi = std::min(xhi, hi);
        if (out.fr[k].dy[0] > 1 || out.fr[k].dy[1] > 1 || out.fr[k].dx[0] > 1 || out.fr[k].dx[1] > 1) return false;
    }
    // the region is cut to k_fir_bp's tile grid (32 x 128): that kernel skips exactly these tiles
    out.y0 = (ylo + FB_TH - 1) / FB_TH * FB_TH; out.y1 = yhi / FB_TH * FB_TH;
    out.x0 = (xlo + FB_TW - 1) / FB_TW * FB_TW; out.x1 = xhi / FB_TW * FB_TW;
    out.n = n; out.H = H; out.W = W;
    return out.y1 > out.y0 && out.x1 > out.x0;
}

int launch_fir_poly(const FirPolyArgs& a, cudaStream_t st) {
    static PerDeviceOnce configured;
    const size_t smem = sizeof(float) * (size_t)FP_RR * (FP_RP + FP_TP);
    if (configured.need()) {
        SR_CUDA(cudaFuncSetAttribute(k_fir_poly, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    dim3 grid((a.x1 - a.x0 + FP_TW - 1) / FP_TW, (a.y1 - a.y0 + FP_TH - 1) / FP_TH);
    k_fir_poly<<<grid, 256, smem, st>>>(a);
    SR_LAUNCH_CHECK("k_fir_poly");
    return SR459_OK;
}

}  // namespace sr459
