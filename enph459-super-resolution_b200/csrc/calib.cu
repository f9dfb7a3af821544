// calib.cu -- shift calibration and PSF characterisation on the device.
//
// The reference measures where a pinhole image sits (calibration_beam_shift/calibrate_shift_grid.py:66-96,
// find_pinhole_center) with the helpers of data_collection/psf_mtf_utils.py: find_peak :48-51 (argmax of the
// sigma = 2 Gaussian-smoothed frame), extract_psf :54-71 (ROI minus a percentile of its outer ring, clamped at 0),
// fit_gaussian_psf :115-131 (bounded 7-parameter rotated-Gaussian least squares; scipy curve_fit there, a projected
// Levenberg-Marquardt iteration with the analytic Jacobian here) and compute_mtf :134-170 (zero-padded 2-D FFT
// magnitude of the PSF, azimuthal mean).  A calibration run is 30 pinhole frames per XPR corner: the batch is the
// natural grid dimension.
//
//   k_smooth_argmax   one 32 x 32 tile per CTA: both axes of scipy's gaussian_filter (boundary 'reflect', axis 0 first,
//                     the taps added in the restatement's order without FMA contraction, float64) and the tile's
//                     first-maximum candidate; k_argmax_final picks the frame's first maximum
//   k_pinhole_fit     one CTA per frame: ROI, bitonic-sorted outer ring -> percentile, background removal, centre of
//                     mass start point, the whole LM iteration (normal equations by block reduction, the 7 x 7 solve by
//                     one thread), fall-back centre of mass
//   k_psf_mtf_rows / k_psf_mtf_cols / k_radial   direct (zero-skipping) 2-D DFT of the padded PSF and its ring averages
// Everything is float64: these run once per calibration on 31 x 31 ROIs; what matters is that every integer decision
// (peak coordinates, ring membership, percentile order statistics) equals the reference's.
#include "internal.h"
#include "plan.h"

#include <math.h>
#include <vector>

namespace sr459 {

constexpr int CAL_T = 32;                 // smoothing tile
constexpr int CAL_MAXR = 12;              // smoothing radius limit (sigma <= 3)
constexpr int CAL_ROI = 31;               // ROI side limit (crop radius <= 15)
constexpr int CAL_FIT_THREADS = 256;

__host__ __device__ inline int reflect_idx(int i, int n) {      // scipy 'reflect' == numpy 'symmetric'
    while (i < 0 || i >= n) { if (i < 0) i = -i - 1; if (i >= n) i = 2 * n - 1 - i; }
    return i;
}

struct SmoothArgs {
    const void* imgs; int is_f32; int n, H, W, r;
    double w[2 * CAL_MAXR + 1];
    double* bval; int* bidx;              // per tile candidates [n][tiles]
    double* smooth_out;                   // optional full smoothed stack (tests), else null
};

template <typename T>
__global__ void __launch_bounds__(256) k_smooth_argmax(const __grid_constant__ SmoothArgs a) {
    __shared__ double sin_[(CAL_T + 2 * CAL_MAXR)][CAL_T + 2 * CAL_MAXR + 1];
    __shared__ double smid[CAL_T][CAL_T + 2 * CAL_MAXR + 1];
    __shared__ double rv[8]; __shared__ int ri[8];
    const int r = a.r, side = CAL_T + 2 * r;
    const int x0 = blockIdx.x * CAL_T, y0 = blockIdx.y * CAL_T, img = blockIdx.z;
    const T* src = (const T*)a.imgs + (size_t)img * a.H * a.W;
    for (int t = threadIdx.x; t < side * side; t += 256) {
        const int yy = t / side, xx = t - yy * side;
        const int gy = reflect_idx(y0 - r + yy, a.H), gx = reflect_idx(x0 - r + xx, a.W);
        sin_[yy][xx] = (double)src[(size_t)gy * a.W + gx];
    }
    __syncthreads();
    // axis 0 (rows of the tile, all staged columns)
    for (int t = threadIdx.x; t < CAL_T * side; t += 256) {
        const int yy = t / side, xx = t - yy * side;
        double acc = 0.0;
        for (int k = 0; k <= 2 * r; ++k) acc = __dadd_rn(acc, __dmul_rn(a.w[k], sin_[yy + k][xx]));
        smid[yy][xx] = acc;
    }
    __syncthreads();
    double best = -INFINITY; int bi = 0x7fffffff;
    for (int t = threadIdx.x; t < CAL_T * CAL_T; t += 256) {
        const int yy = t / CAL_T, xx = t - yy * CAL_T;
        const int gy = y0 + yy, gx = x0 + xx;
        if (gy >= a.H || gx >= a.W) continue;
        double acc = 0.0;
        for (int k = 0; k <= 2 * r; ++k) acc = __dadd_rn(acc, __dmul_rn(a.w[k], smid[yy][xx + k]));
        if (a.smooth_out) a.smooth_out[(size_t)img * a.H * a.W + (size_t)gy * a.W + gx] = acc;
        const int gi = gy * a.W + gx;
        if (acc > best || (acc == best && gi < bi)) { best = acc; bi = gi; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, best, o); const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if ((threadIdx.x & 31) == 0) { rv[threadIdx.x >> 5] = best; ri[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < 8; ++i) if (rv[i] > best || (rv[i] == best && ri[i] < bi)) { best = rv[i]; bi = ri[i]; }
        const int tiles = gridDim.x * gridDim.y, tile = blockIdx.y * gridDim.x + blockIdx.x;
        a.bval[(size_t)img * tiles + tile] = best; a.bidx[(size_t)img * tiles + tile] = bi;
    }
}

__global__ void __launch_bounds__(256) k_argmax_final(const double* bval, const int* bidx, int tiles, int* out) {
    __shared__ double rv[8]; __shared__ int ri[8];
    const int img = blockIdx.x;
    double best = -INFINITY; int bi = 0x7fffffff;
    for (int t = threadIdx.x; t < tiles; t += 256) {
        const double v = bval[(size_t)img * tiles + t]; const int i = bidx[(size_t)img * tiles + t];
        if (v > best || (v == best && i < bi)) { best = v; bi = i; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, best, o); const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if ((threadIdx.x & 31) == 0) { rv[threadIdx.x >> 5] = best; ri[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < 8; ++i) if (rv[i] > best || (rv[i] == best && ri[i] < bi)) { best = rv[i]; bi = ri[i]; }
        out[img] = bi;
    }
}

// ---- block helpers for the fit kernel -----------------------------------------------------------------------------------
__device__ __forceinline__ double block_sum(double v, double* red) {     // all threads get the sum; red: 8 doubles
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double s = 0.0;
    for (int i = 0; i < CAL_FIT_THREADS / 32; ++i) s += red[i];
    return s;
}
__device__ __forceinline__ double block_max(double v, double* red) {
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double s = red[0];
    for (int i = 1; i < CAL_FIT_THREADS / 32; ++i) s = fmax(s, red[i]);
    return s;
}
// ascending bitonic sort of 1024 doubles in shared memory (pad with +inf)
__device__ void bitonic_1024(double* v) {
    for (int k = 2; k <= 1024; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            __syncthreads();
            for (int i = threadIdx.x; i < 1024; i += CAL_FIT_THREADS) {
                const int l = i ^ j;
                if (l > i) {
                    const double x = v[i], y = v[l];
                    const bool up = (i & k) == 0;
                    if ((x > y) == up) { v[i] = y; v[l] = x; }
                }
            }
        }
    __syncthreads();
}
// numpy percentile, method 'linear', of the sorted v[0..n-1]
__device__ double percentile_sorted(const double* v, int n, double q) {
    const double pos = q / 100.0 * (double)(n - 1);
    int lo = (int)floor(pos); if (lo > n - 1) lo = n - 1; if (lo < 0) lo = 0;
    const int hi = min(lo + 1, n - 1);
    const double t = pos - (double)lo, a = v[lo], b = v[hi], d = b - a;
    return (t >= 0.5) ? b - d * (1.0 - t) : a + d * t;
}

struct FitArgs {
    const void* imgs; int is_f32; int n, H, W;
    const int* peaks;                 // [n] flat index of find_peak (device)
    int radius; double bg_percentile; int max_iter;   // bg_percentile < 0: the ROI is taken as it is (no background removal)
    double* out;                      // [n][12]: cx, cy, amp, x0, y0, sx, sy, theta, offset, status, pr, pc
    double* roi_out;                  // optional [n][31*31] background-subtracted ROIs (row pitch = ROI width)
    int* roi_dims;                    // optional [n][4]: rh, rw, r0, c0
};

// residuals / normal equations of the rotated Gaussian at parameter vector p over this thread's ROI pixels
__device__ void fit_accumulate(const double* roi, int rh, int rw, const double* p, double* acc /*36: A upper 28, g 7, cost*/) {
    for (int i = 0; i < 36; ++i) acc[i] = 0.0;
    const double amp = p[0], x0 = p[1], y0 = p[2], sx = p[3], sy = p[4], th = p[5], off = p[6];
    const double cth = cos(th), sth = sin(th), ct = cth * cth, st = sth * sth, s2 = sin(2.0 * th), c2 = cos(2.0 * th);
    const double qa = ct / (2 * sx * sx) + st / (2 * sy * sy), qb = -s2 / (4 * sx * sx) + s2 / (4 * sy * sy),
                 qc = st / (2 * sx * sx) + ct / (2 * sy * sy);
    const double sx3 = sx * sx * sx, sy3 = sy * sy * sy;
    const double da_sx = -ct / sx3, db_sx = s2 / (2 * sx3), dc_sx = -st / sx3;
    const double da_sy = -st / sy3, db_sy = -s2 / (2 * sy3), dc_sy = -ct / sy3;
    const double kk = 1 / (2 * sy * sy) - 1 / (2 * sx * sx);
    const double da_th = s2 * kk, db_th = c2 * kk, dc_th = -s2 * kk;
    for (int t = threadIdx.x; t < rh * rw; t += CAL_FIT_THREADS) {
        const int yy = t / rw, xx = t - yy * rw;
        const double dx = (double)xx - x0, dy = (double)yy - y0;
        const double e = exp(-(qa * dx * dx + 2 * qb * dx * dy + qc * dy * dy)), g = amp * e;
        double J[7];
        J[0] = e; J[1] = g * (2 * qa * dx + 2 * qb * dy); J[2] = g * (2 * qb * dx + 2 * qc * dy);
        J[3] = -g * (da_sx * dx * dx + 2 * db_sx * dx * dy + dc_sx * dy * dy);
        J[4] = -g * (da_sy * dx * dx + 2 * db_sy * dx * dy + dc_sy * dy * dy);
        J[5] = -g * (da_th * dx * dx + 2 * db_th * dx * dy + dc_th * dy * dy);
        J[6] = 1.0;
        const double res = off + g - roi[t];
        int q = 0;
        #pragma unroll
        for (int i = 0; i < 7; ++i)
            #pragma unroll
            for (int j = i; j < 7; ++j) acc[q++] += J[i] * J[j];
        #pragma unroll
        for (int i = 0; i < 7; ++i) acc[28 + i] += J[i] * res;
        acc[35] += res * res;
    }
}

// (A + lam diag(max(diag A, 1e-12))) dp = -g by Gaussian elimination with partial pivoting; false if singular
__device__ bool solve7(const double* A /*7x7 full*/, const double* g, double lam, double* dp) {
    double M[7][8];
    for (int i = 0; i < 7; ++i) {
        for (int j = 0; j < 7; ++j) M[i][j] = A[i * 7 + j];
        M[i][i] += lam * fmax(A[i * 7 + i], 1e-12);
        M[i][7] = -g[i];
    }
    for (int c = 0; c < 7; ++c) {
        int piv = c; double best = fabs(M[c][c]);
        for (int r = c + 1; r < 7; ++r) if (fabs(M[r][c]) > best) { best = fabs(M[r][c]); piv = r; }
        if (!(best > 0.0) || !isfinite(best)) return false;
        if (piv != c) for (int j = 0; j < 8; ++j) { const double t = M[c][j]; M[c][j] = M[piv][j]; M[piv][j] = t; }
        for (int r = c + 1; r < 7; ++r) {
            const double f = M[r][c] / M[c][c];
            for (int j = c; j < 8; ++j) M[r][j] -= f * M[c][j];
        }
    }
    for (int i = 6; i >= 0; --i) {
        double s = M[i][7];
        for (int j = i + 1; j < 7; ++j) s -= M[i][j] * dp[j];
        dp[i] = s / M[i][i];
        if (!isfinite(dp[i])) return false;
    }
    return true;
}

template <typename T>
__global__ void __launch_bounds__(CAL_FIT_THREADS) k_pinhole_fit(const __grid_constant__ FitArgs a) {
    __shared__ double roi[CAL_ROI * CAL_ROI];
    __shared__ double srt[1024];
    __shared__ double red[8];
    __shared__ double sums[36 * (CAL_FIT_THREADS / 32)];
    __shared__ double tot[36];
    __shared__ double P[7], PN[7], LO[7], HI[7], Afull[49], Gv[7];
    __shared__ double s_cost, s_lam; __shared__ int s_flag, s_cnt;
    const int img = blockIdx.x, tid = threadIdx.x, R = a.radius;
    const T* src = (const T*)a.imgs + (size_t)img * a.H * a.W;
    const int pk = a.peaks[img], pr = pk / a.W, pc = pk - pr * a.W;
    const int r0 = max(pr - R, 0), r1 = min(pr + R + 1, a.H), c0 = max(pc - R, 0), c1 = min(pc + R + 1, a.W);
    const int rh = r1 - r0, rw = c1 - c0, npx = rh * rw;
    for (int t = tid; t < npx; t += CAL_FIT_THREADS) { const int yy = t / rw, xx = t - yy * rw; roi[t] = (double)src[(size_t)(r0 + yy) * a.W + c0 + xx]; }
    if (tid == 0) s_cnt = 0;
    for (int t = tid; t < 1024; t += CAL_FIT_THREADS) srt[t] = INFINITY;
    __syncthreads();
    // outer ring (>= 0.6 radius from the ROI centre) -> percentile
    {
        const int cy = rh / 2, cx = rw / 2; const double lim = (R * 0.6) * (R * 0.6);
        for (int t = tid; t < npx; t += CAL_FIT_THREADS) {
            const int yy = t / rw, xx = t - yy * rw;
            if ((double)((yy - cy) * (yy - cy) + (xx - cx) * (xx - cx)) >= lim) srt[atomicAdd(&s_cnt, 1)] = roi[t];
        }
    }
    if (a.bg_percentile >= 0.0) {
        bitonic_1024(srt);
        const int nring = s_cnt;
        const double bg = nring > 0 ? percentile_sorted(srt, nring, a.bg_percentile) : 0.0;
        __syncthreads();
        for (int t = tid; t < npx; t += CAL_FIT_THREADS) roi[t] = fmax(roi[t] - bg, 0.0);
    }
    __syncthreads();
    if (a.roi_out) for (int t = tid; t < npx; t += CAL_FIT_THREADS) a.roi_out[(size_t)img * CAL_ROI * CAL_ROI + t] = roi[t];
    if (a.roi_dims && tid == 0) { int* d = a.roi_dims + 4 * img; d[0] = rh; d[1] = rw; d[2] = r0; d[3] = c0; }

    // start point: peak value, centre of mass above 10 % of the peak (subpixel_centre, :74-78)
    double mx = -INFINITY;
    for (int t = tid; t < npx; t += CAL_FIT_THREADS) mx = fmax(mx, roi[t]);
    mx = block_max(mx, red);
    double m0 = 0.0, my = 0.0, mxx = 0.0;
    for (int t = tid; t < npx; t += CAL_FIT_THREADS) {
        const int yy = t / rw, xx = t - yy * rw;
        const double v = roi[t] > mx * 0.1 ? roi[t] : 0.0;
        m0 += v; my += v * yy; mxx += v * xx;
    }
    m0 = block_sum(m0, red); my = block_sum(my, red); mxx = block_sum(mxx, red);
    double* out = a.out + 12 * (size_t)img;
    bool ok = (mx > 0.0) && isfinite(m0) && m0 > 0.0;
    if (tid == 0) {
        P[0] = mx; P[1] = mxx / m0; P[2] = my / m0; P[3] = 2.0; P[4] = 2.0; P[5] = 0.0; P[6] = 0.0;
        const double pi = 3.14159265358979323846;
        LO[0] = 0; LO[1] = 0; LO[2] = 0; LO[3] = 0.3; LO[4] = 0.3; LO[5] = -pi; LO[6] = -INFINITY;
        HI[0] = mx * 2; HI[1] = rw; HI[2] = rh; HI[3] = rw / 2.0; HI[4] = rh / 2.0; HI[5] = pi; HI[6] = mx * 0.5;
        for (int i = 0; i < 7; ++i) P[i] = fmin(fmax(P[i], LO[i]), HI[i]);
        s_lam = 1e-3;
    }
    __syncthreads();
    double acc[36];
    auto evaluate = [&](const double* p) {           // -> tot[36] visible to all threads
        fit_accumulate(roi, rh, rw, p, acc);
        for (int i = 0; i < 36; ++i) { double v = acc[i]; for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o); if ((tid & 31) == 0) sums[i * 8 + (tid >> 5)] = v; }
        __syncthreads();
        if (tid < 36) { double s = 0.0; for (int w = 0; w < 8; ++w) s += sums[tid * 8 + w]; tot[tid] = s; }
        __syncthreads();
    };
    int status = 1;                                    // 0 = Gaussian fit, 1 = centre-of-mass fall-back, 2 = peak
    if (ok) {
        evaluate(P);
        if (tid == 0) {
            int q = 0;
            for (int i = 0; i < 7; ++i) for (int j = i; j < 7; ++j) { Afull[i * 7 + j] = Afull[j * 7 + i] = tot[q++]; }
            for (int i = 0; i < 7; ++i) Gv[i] = tot[28 + i];
            s_cost = tot[35];
        }
        __syncthreads();
        bool finished = false;
        for (int it = 0; it < a.max_iter && !finished; ++it) {
            bool step_ok = false;
            for (int tr = 0; tr < 30; ++tr) {
                if (tid == 0) {
                    double dp[7];
                    s_flag = solve7(Afull, Gv, s_lam, dp) ? 1 : 0;
                    if (s_flag) for (int i = 0; i < 7; ++i) PN[i] = fmin(fmax(P[i] + dp[i], LO[i]), HI[i]);
                    else s_lam *= 10.0;
                }
                __syncthreads();
                if (!s_flag) continue;
                evaluate(PN);
                if (tot[35] < s_cost) { step_ok = true; break; }
                __syncthreads();
                if (tid == 0) s_lam *= 10.0;
                __syncthreads();
            }
            if (!step_ok) { finished = true; status = 0; break; }
            __syncthreads();
            if (tid == 0) {
                const double cn = tot[35];
                double dmax = 0.0;
                for (int i = 0; i < 7; ++i) dmax = fmax(dmax, fabs(PN[i] - P[i]));
                s_flag = ((s_cost - cn) <= 1e-15 * fmax(s_cost, 1e-300)) && (dmax <= 1e-10);
                for (int i = 0; i < 7; ++i) P[i] = PN[i];
                int q = 0;
                for (int i = 0; i < 7; ++i) for (int j = i; j < 7; ++j) { Afull[i * 7 + j] = Afull[j * 7 + i] = tot[q++]; }
                for (int i = 0; i < 7; ++i) Gv[i] = tot[28 + i];
                s_cost = cn; s_lam = fmax(s_lam / 10.0, 1e-12);
            }
            __syncthreads();
            if (s_flag) { finished = true; status = 0; }
        }
        // max_iter exhausted without convergence: the reference's curve_fit raises -> fall back
    }
    if (status == 0) {
        if (tid == 0) {
            out[0] = P[1] + c0; out[1] = P[2] + r0;
            for (int i = 0; i < 7; ++i) out[2 + i] = P[i];
            out[9] = 0.0; out[10] = pr; out[11] = pc;
        }
        return;
    }
    // calibrate_shift_grid.py:88-96: centre of mass of the ROI above its median, floor at 10 % of the maximum
    for (int t = tid; t < 1024; t += CAL_FIT_THREADS) srt[t] = t < npx ? roi[t] : INFINITY;
    bitonic_1024(srt);
    const double med = (npx & 1) ? srt[npx / 2] : 0.5 * (srt[npx / 2 - 1] + srt[npx / 2]);
    double tm = 0.0;
    for (int t = tid; t < npx; t += CAL_FIT_THREADS) tm = fmax(tm, fmax(roi[t] - med, 0.0));
    tm = block_max(tm, red);
    double f0 = 0.0, fy = 0.0, fx = 0.0;
    for (int t = tid; t < npx; t += CAL_FIT_THREADS) {
        const int yy = t / rw, xx = t - yy * rw;
        double v = fmax(roi[t] - med, 0.0); if (v < tm * 0.1) v = 0.0;
        f0 += v; fy += v * yy; fx += v * xx;
    }
    f0 = block_sum(f0, red); fy = block_sum(fy, red); fx = block_sum(fx, red);
    if (tid == 0) {
        if (f0 == 0.0) { out[0] = pc; out[1] = pr; out[9] = 2.0; }
        else { out[0] = fx / f0 + c0; out[1] = fy / f0 + r0; out[9] = 1.0; }
        for (int i = 0; i < 7; ++i) out[2 + i] = NAN;
        out[10] = pr; out[11] = pc;
    }
}

// ---- compute_mtf: |DFT2| of the PSF zero-padded (centred) to pad x pad, origin moved to the centre (ifftshift / fftshift) ----
// big[r0 + y][c0 + x] = psf[y][x] / sum;  F[u][v] = sum big[m][n] exp(-2 pi i (u m' + v n') / pad) with m' = (m - pad/2) mod pad;
// output index (U, V) = fftshift -> frequency u = U - pad/2.
__global__ void k_psf_mtf_rows(const double* psf, int ph, int pw, int pad, int c0, double inv_tot, double2* rowF) {
    // rowF[y][V] = sum_x psf[y][x]/tot * exp(-2 pi i v n' / pad), v = V - pad/2, n' = c0 + x - pad/2
    const int V = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (V >= pad) return;
    const int v = V - pad / 2;
    double re = 0.0, im = 0.0;
    for (int x = 0; x < pw; ++x) {
        const int np = c0 + x - pad / 2;
        const long long k = ((long long)v * np) % pad;
        double s, c; sincospi(-2.0 * (double)k / (double)pad, &s, &c);
        const double val = psf[y * pw + x] * inv_tot;
        re += val * c; im += val * s;
    }
    rowF[y * pad + V] = make_double2(re, im);
}
__global__ void k_psf_mtf_cols(const double2* rowF, int ph, int pad, int r0, double* mtf) {
    const int V = blockIdx.x * blockDim.x + threadIdx.x, U = blockIdx.y;
    if (V >= pad) return;
    const int u = U - pad / 2;
    double re = 0.0, im = 0.0;
    for (int y = 0; y < ph; ++y) {
        const int mp = r0 + y - pad / 2;
        const long long k = ((long long)u * mp) % pad;
        double s, c; sincospi(-2.0 * (double)k / (double)pad, &s, &c);
        const double2 f = rowF[y * pad + V];
        re += f.x * c - f.y * s; im += f.x * s + f.y * c;
    }
    mtf[(size_t)U * pad + V] = sqrt(re * re + im * im);
}
// ring sums of data / peak over integer radii about (cy, cx); one CTA
__global__ void __launch_bounds__(256) k_radial(const double* data, int h, int w, double cy, double cx, int max_radius,
                                                double* prof, double* peak_out) {
    extern __shared__ double sh[];                 // tot[max_radius], cnt[max_radius]
    __shared__ double red[8];
    double* tot = sh; double* cnt = sh + max_radius;
    for (int i = threadIdx.x; i < 2 * max_radius; i += 256) sh[i] = 0.0;
    double mx = 0.0;
    for (int t = threadIdx.x; t < h * w; t += 256) mx = fmax(mx, data[t]);
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
    __syncthreads();
    mx = red[0]; for (int i = 1; i < 8; ++i) mx = fmax(mx, red[i]);
    for (int t = threadIdx.x; t < h * w; t += 256) {
        const int yy = t / w, xx = t - yy * w;
        const double dx = xx - cx, dy = yy - cy;
        const int ring = (int)sqrt(dx * dx + dy * dy);
        if (ring < max_radius) { atomicAdd(&tot[ring], mx > 0.0 ? data[t] / mx : data[t]); atomicAdd(&cnt[ring], 1.0); }   // mtf_2d /= mtf_2d.max()
    }
    __syncthreads();
    for (int i = threadIdx.x; i < max_radius; i += 256) prof[i] = cnt[i] > 0.0 ? tot[i] / cnt[i] : 0.0;
    if (threadIdx.x == 0 && peak_out) *peak_out = mx;
}
__global__ void k_scale(double* d, int n, const double* peak) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && *peak > 0.0) d[i] = d[i] / *peak;
}

}  // namespace sr459

using namespace sr459;

extern "C" {

int sr459_find_peaks(const void* d_imgs, int is_f32, int n_img, int H, int W, double sigma, int* h_peaks, double* d_smooth_out,
                     void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!d_imgs || n_img < 1 || H < 1 || W < 1 || !(sigma > 0.0) || !h_peaks) { set_error("find_peaks: bad arguments"); return SR459_EINVAL; }
    if ((int64_t)H * W >= (1LL << 31)) { set_error("find_peaks: image too large"); return SR459_EUNSUPPORTED; }
    SmoothArgs a{};
    a.r = (int)(4.0 * sigma + 0.5);                       // scipy: truncate = 4
    if (a.r > CAL_MAXR) { set_error("find_peaks: sigma %.3f needs radius %d > %d", sigma, a.r, CAL_MAXR); return SR459_EUNSUPPORTED; }
    if (a.r > H || a.r > W) { set_error("find_peaks: image smaller than the smoothing radius"); return SR459_EUNSUPPORTED; }
    {   // numpy: w = exp(-0.5 / sigma^2 * x^2); w /= w.sum()   (sequential sum is what np.sum does below 8 elements; pairwise above:
        // restated by np_pairwise_sum in api.cu -- 17 taps at sigma = 2 are summed in one unrolled-by-8 block + remainder)
        double s = 0.0, w[2 * CAL_MAXR + 1];
        for (int k = 0; k <= 2 * a.r; ++k) { const double x = (double)(k - a.r); w[k] = exp(-0.5 / (sigma * sigma) * (x * x)); }
        s = np_pairwise_sum(w, (size_t)(2 * a.r + 1));
        for (int k = 0; k <= 2 * a.r; ++k) a.w[k] = w[k] / s;
    }
    a.imgs = d_imgs; a.is_f32 = is_f32; a.n = n_img; a.H = H; a.W = W; a.smooth_out = d_smooth_out;
    dim3 grid((W + CAL_T - 1) / CAL_T, (H + CAL_T - 1) / CAL_T, n_img);
    const int tiles = grid.x * grid.y;
    double* bval = nullptr; int* bidx = nullptr; int* d_out = nullptr;
    SR_CUDA(gmalloc((void**)&bval, sizeof(double) * (size_t)tiles * n_img));
    SR_CUDA(gmalloc((void**)&bidx, sizeof(int) * (size_t)tiles * n_img));
    SR_CUDA(gmalloc((void**)&d_out, sizeof(int) * n_img));
    a.bval = bval; a.bidx = bidx;
    if (is_f32) k_smooth_argmax<float><<<grid, 256, 0, st>>>(a); else k_smooth_argmax<uint8_t><<<grid, 256, 0, st>>>(a);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) { count_launch(); k_argmax_final<<<n_img, 256, 0, st>>>(bval, bidx, tiles, d_out); e = cudaGetLastError(); }
    if (e == cudaSuccess) { count_launch(); e = cudaMemcpyAsync(h_peaks, d_out, sizeof(int) * n_img, cudaMemcpyDeviceToHost, st); }
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    gfree(bval); gfree(bidx); gfree(d_out);
    if (e != cudaSuccess) { set_error("find_peaks: %s", cudaGetErrorString(e)); return SR459_ECUDA; }
    return SR459_OK;
}

int sr459_pinhole_fit(const void* d_imgs, int is_f32, int n_img, int H, int W, const int* h_peaks, int radius, double bg_percentile,
                      int max_iter, double* h_out, double* h_roi, int* h_roi_dims, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!d_imgs || n_img < 1 || H < 1 || W < 1 || !h_peaks || !h_out || radius < 1 || max_iter < 1) { set_error("pinhole_fit: bad arguments"); return SR459_EINVAL; }
    if (2 * radius + 1 > CAL_ROI) { set_error("pinhole_fit: crop radius %d > %d", radius, (CAL_ROI - 1) / 2); return SR459_EUNSUPPORTED; }
    for (int k = 0; k < n_img; ++k) if (h_peaks[k] < 0 || (int64_t)h_peaks[k] >= (int64_t)H * W) { set_error("pinhole_fit: peak %d outside the image", k); return SR459_EINVAL; }
    FitArgs a{};
    a.imgs = d_imgs; a.is_f32 = is_f32; a.n = n_img; a.H = H; a.W = W; a.radius = radius; a.bg_percentile = bg_percentile; a.max_iter = max_iter;
    int* d_peaks = nullptr; double* d_out = nullptr; double* d_roi = nullptr; int* d_dims = nullptr;
    SR_CUDA(gmalloc((void**)&d_peaks, sizeof(int) * n_img));
    SR_CUDA(gmalloc((void**)&d_out, sizeof(double) * 12 * n_img));
    SR_CUDA(gmalloc((void**)&d_roi, sizeof(double) * CAL_ROI * CAL_ROI * (size_t)n_img));
    SR_CUDA(gmalloc((void**)&d_dims, sizeof(int) * 4 * n_img));
    a.peaks = d_peaks; a.out = d_out; a.roi_out = d_roi; a.roi_dims = d_dims;
    cudaError_t e = cudaMemcpyAsync(d_peaks, h_peaks, sizeof(int) * n_img, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
        if (is_f32 == 2)  k_pinhole_fit<double><<<n_img, CAL_FIT_THREADS, 0, st>>>(a);
        else if (is_f32)  k_pinhole_fit<float><<<n_img, CAL_FIT_THREADS, 0, st>>>(a);
        else              k_pinhole_fit<uint8_t><<<n_img, CAL_FIT_THREADS, 0, st>>>(a);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) { count_launch(); e = cudaMemcpyAsync(h_out, d_out, sizeof(double) * 12 * n_img, cudaMemcpyDeviceToHost, st); }
    if (e == cudaSuccess && h_roi) e = cudaMemcpyAsync(h_roi, d_roi, sizeof(double) * CAL_ROI * CAL_ROI * (size_t)n_img, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess && h_roi_dims) e = cudaMemcpyAsync(h_roi_dims, d_dims, sizeof(int) * 4 * n_img, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    gfree(d_peaks); gfree(d_out); gfree(d_roi); gfree(d_dims);
    if (e != cudaSuccess) { set_error("pinhole_fit: %s", cudaGetErrorString(e)); return SR459_ECUDA; }
    return SR459_OK;
}

int sr459_psf_mtf(const double* h_psf, int ph, int pw, int pad, double* h_mtf2d, double* h_radial, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!h_psf || ph < 1 || pw < 1 || pad < ph || pad < pw || (pad & 1) || !h_radial) { set_error("psf_mtf: bad arguments"); return SR459_EINVAL; }
    if (pad > 4096) { set_error("psf_mtf: pad %d too large", pad); return SR459_EUNSUPPORTED; }
    // compute_mtf :147-149: the zero-padded array is summed (pairwise over pad*pad elements); zeros do not change a
    // floating-point sum, but the pairwise blocks do depend on where the non-zeros sit -> build the padded array here
    const int r0 = (pad - ph) / 2, c0 = (pad - pw) / 2;
    double tot;
    {
        std::vector<double> big((size_t)pad * pad, 0.0);
        for (int y = 0; y < ph; ++y) for (int x = 0; x < pw; ++x) big[(size_t)(r0 + y) * pad + c0 + x] = h_psf[y * pw + x];
        tot = np_pairwise_sum(big.data(), big.size());
    }
    double* d_psf = nullptr; double2* d_row = nullptr; double* d_mtf = nullptr; double* d_prof = nullptr; double* d_peak = nullptr;
    SR_CUDA(gmalloc((void**)&d_psf, sizeof(double) * ph * pw));
    SR_CUDA(gmalloc((void**)&d_row, sizeof(double2) * (size_t)ph * pad));
    SR_CUDA(gmalloc((void**)&d_mtf, sizeof(double) * (size_t)pad * pad));
    SR_CUDA(gmalloc((void**)&d_prof, sizeof(double) * (pad / 2)));
    SR_CUDA(gmalloc((void**)&d_peak, sizeof(double)));
    cudaError_t e = cudaMemcpyAsync(d_psf, h_psf, sizeof(double) * ph * pw, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
        k_psf_mtf_rows<<<dim3((pad + 127) / 128, ph), 128, 0, st>>>(d_psf, ph, pw, pad, c0, tot > 0.0 ? 1.0 / tot : 1.0, d_row);
        k_psf_mtf_cols<<<dim3((pad + 127) / 128, pad), 128, 0, st>>>(d_row, ph, pad, r0, d_mtf);
        k_radial<<<1, 256, sizeof(double) * 2 * (pad / 2), st>>>(d_mtf, pad, pad, pad / 2.0, pad / 2.0, pad / 2, d_prof, d_peak);
        k_scale<<<(pad * pad + 255) / 256, 256, 0, st>>>(d_mtf, pad * pad, d_peak);
        e = cudaGetLastError();
        if (e == cudaSuccess) count_launch(4);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(h_radial, d_prof, sizeof(double) * (pad / 2), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess && h_mtf2d) e = cudaMemcpyAsync(h_mtf2d, d_mtf, sizeof(double) * (size_t)pad * pad, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    gfree(d_psf); gfree(d_row); gfree(d_mtf); gfree(d_prof); gfree(d_peak);
    if (e != cudaSuccess) { set_error("psf_mtf: %s", cudaGetErrorString(e)); return SR459_ECUDA; }
    return SR459_OK;
}

}  // extern "C"
