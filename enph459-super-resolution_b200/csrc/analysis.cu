// analysis.cu -- quality analytics on HR outputs, on the device.
//
// The reference judges its Native-2x / SAA / SAA+IBP images by the slanted-edge method (mono_cal_target/analysis.ipynb,
// cell 7: slanted_edge_esf + esf_to_mtf; cell 10: mtf_at_fraction -> MTF50 / MTF10) and by the local contrast of a bar
// profile (cell 4: local_contrast).  Here the whole chain runs on the image where it already lies (device memory, float32
// or uint8, any pitch), in float64 like the notebook:
//
//   k_roi_smooth     gaussian_filter(roi, 1.5): both axes, scipy 'reflect' boundary inside the ROI, taps added in order
//                    without FMA contraction (the smoothed field equals the numpy restatement bit for bit)
//   k_sobel_mag      scipy.ndimage.sobel along both axes + gradient magnitude
//   k_edge_esf       ONE CTA: 85th-percentile threshold by radix select (two order statistics + numpy's lerp), line fit of
//                    the edge pixels, side selection, second fit, signed distance of every ROI pixel, quarter-pixel
//                    half-open bins exactly as np.arange lays them out, bin means, gap interpolation, orientation, then
//                    LSF = np.gradient (both of numpy's spacing branches), Hann window, DFT magnitude, MTF50 / MTF10
//   k_local_contrast sliding-window (max - min) / (max + min + 1e-9)
// Decisions that are integers upstream (edge-pixel set, row/column orientation, bin membership, bin count) are made with
// the same comparisons on the same float64 values.
#include "internal.h"

#include <math.h>
#include <vector>

namespace sr459 {

constexpr int AN_T = 32, AN_MAXR = 8, AN_THREADS = 1024, AN_MAXBINS = 160;

__host__ __device__ inline int an_reflect(int i, int n) {
    while (i < 0 || i >= n) { if (i < 0) i = -i - 1; if (i >= n) i = 2 * n - 1 - i; }
    return i;
}

struct RoiSmoothArgs {
    const void* base; int is_f32; int pitch; int rh, rw, r;
    double w[2 * AN_MAXR + 1];
    double* out;                            // rh x rw
};

template <typename T>
__global__ void __launch_bounds__(256) k_roi_smooth(const __grid_constant__ RoiSmoothArgs a) {
    __shared__ double sin_[AN_T + 2 * AN_MAXR][AN_T + 2 * AN_MAXR + 1];
    __shared__ double smid[AN_T][AN_T + 2 * AN_MAXR + 1];
    const int r = a.r, side = AN_T + 2 * r, x0 = blockIdx.x * AN_T, y0 = blockIdx.y * AN_T;
    const T* src = (const T*)a.base;
    for (int t = threadIdx.x; t < side * side; t += 256) {
        const int yy = t / side, xx = t - yy * side;
        sin_[yy][xx] = (double)src[(size_t)an_reflect(y0 - r + yy, a.rh) * a.pitch + an_reflect(x0 - r + xx, a.rw)];
    }
    __syncthreads();
    for (int t = threadIdx.x; t < AN_T * side; t += 256) {
        const int yy = t / side, xx = t - yy * side;
        double acc = 0.0;
        for (int k = 0; k <= 2 * r; ++k) acc = __dadd_rn(acc, __dmul_rn(a.w[k], sin_[yy + k][xx]));
        smid[yy][xx] = acc;
    }
    __syncthreads();
    for (int t = threadIdx.x; t < AN_T * AN_T; t += 256) {
        const int yy = t / AN_T, xx = t - yy * AN_T, gy = y0 + yy, gx = x0 + xx;
        if (gy >= a.rh || gx >= a.rw) continue;
        double acc = 0.0;
        for (int k = 0; k <= 2 * r; ++k) acc = __dadd_rn(acc, __dmul_rn(a.w[k], smid[yy][xx + k]));
        a.out[(size_t)gy * a.rw + gx] = acc;
    }
}

// sobel(s, axis): derivative [-1 0 1] along the axis, then [1 2 1] along the other one; mag = sqrt(gx^2 + gy^2)
__global__ void __launch_bounds__(256) k_sobel_mag(const double* __restrict__ s, int rh, int rw, double* __restrict__ mag) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= rw || y >= rh) return;
    double v[3][3];
    #pragma unroll
    for (int j = 0; j < 3; ++j)
        #pragma unroll
        for (int i = 0; i < 3; ++i) v[j][i] = s[(size_t)an_reflect(y - 1 + j, rh) * rw + an_reflect(x - 1 + i, rw)];
    // axis 0 derivative per column, then smoothing along axis 1:  ((d0 + 2 d1) + d2)
    double d[3];
    #pragma unroll
    for (int i = 0; i < 3; ++i) d[i] = __dadd_rn(-v[0][i], v[2][i]);
    const double gy = __dadd_rn(__dadd_rn(d[0], __dmul_rn(2.0, d[1])), d[2]);
    #pragma unroll
    for (int j = 0; j < 3; ++j) d[j] = __dadd_rn(-v[j][0], v[j][2]);
    const double gx = __dadd_rn(__dadd_rn(d[0], __dmul_rn(2.0, d[1])), d[2]);
    mag[(size_t)y * rw + x] = sqrt(__dadd_rn(__dmul_rn(gx, gx), __dmul_rn(gy, gy)));
}

struct EdgeArgs {
    const void* base; int is_f32; int pitch; int rh, rw; int right_side;
    const double* mag; double* dist;        // rh x rw scratch
    double pitch_mm;
    double* out_x; double* out_y; double* out_lsf; double* out_freq; double* out_mtf;   // device, AN_MAXBINS each
    double* out_scal;                       // [8]: status, n_esf, angle, mtf50, mtf10, n_edge, n_side, thresh
};

__device__ __forceinline__ double an_bsum(double v, double* red) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double s = 0.0;
    for (int i = 0; i < AN_THREADS / 32; ++i) s += red[i];
    return s;
}
__device__ __forceinline__ double an_bmin(double v, double* red) {
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double s = red[0];
    for (int i = 1; i < AN_THREADS / 32; ++i) s = fmin(s, red[i]);
    return s;
}

// k-th smallest (0-based) of n non-negative doubles: 8-bit radix select on the bit patterns
__device__ unsigned long long an_select(const double* v, int n, int k, int* hist, unsigned long long* sh64, int* shk) {
    unsigned long long prefix = 0, mask = 0;
    if (threadIdx.x == 0) *shk = k;
    for (int shift = 56; shift >= 0; shift -= 8) {
        for (int i = threadIdx.x; i < 256; i += AN_THREADS) hist[i] = 0;
        __syncthreads();
        for (int t = threadIdx.x; t < n; t += AN_THREADS) {
            const unsigned long long key = (unsigned long long)__double_as_longlong(v[t]);
            if ((key & mask) == prefix) atomicAdd(&hist[(int)((key >> shift) & 255ull)], 1);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int kk = *shk, b = 0;
            for (; b < 255; ++b) { if (kk < hist[b]) break; kk -= hist[b]; }
            *shk = kk; *sh64 = prefix | ((unsigned long long)b << shift);
        }
        __syncthreads();
        prefix = *sh64; mask |= 0xffull << shift;
        __syncthreads();
    }
    return prefix;
}

template <typename T>
__global__ void __launch_bounds__(AN_THREADS) k_edge_esf(const __grid_constant__ EdgeArgs a) {
    __shared__ double red[32];
    __shared__ int hist[256];
    __shared__ unsigned long long sh64; __shared__ int shk;
    __shared__ double bsum[AN_MAXBINS], bcnt[AN_MAXBINS];
    __shared__ double ex[AN_MAXBINS], ey[AN_MAXBINS], lsf[AN_MAXBINS];
    __shared__ int s_n;
    const int tid = threadIdx.x, n = a.rh * a.rw, rw = a.rw;
    const T* src = (const T*)a.base;
    double* sc = a.out_scal;

    // ---- thresh = np.percentile(mag, 85) -----------------------------------------------------------------------------------
    const double pos = (85.0 / 100.0) * (double)(n - 1);
    const int klo = (int)floor(pos), khi = min(klo + 1, n - 1);
    const double vlo = __longlong_as_double((long long)an_select(a.mag, n, klo, hist, &sh64, &shk));
    double vhi = vlo;
    if (khi > klo) {
        double cle = 0.0, nxt = INFINITY;
        for (int t = tid; t < n; t += AN_THREADS) { const double m = a.mag[t]; if (m <= vlo) cle += 1.0; else nxt = fmin(nxt, m); }
        cle = an_bsum(cle, red); nxt = an_bmin(nxt, red);
        vhi = ((int)cle >= khi + 1) ? vlo : nxt;
    }
    const double gam = pos - (double)klo, dba = vhi - vlo;
    const double thresh = (gam >= 0.5) ? vhi - dba * (1.0 - gam) : vlo + dba * gam;

    // ---- edge pixels: mag > thresh; extent, first line fit -------------------------------------------------------------------
    double cnt = 0, sr = 0, scx = 0, rmin = INFINITY, rmax = -INFINITY, cmin = INFINITY, cmax = -INFINITY;
    for (int t = tid; t < n; t += AN_THREADS) if (a.mag[t] > thresh) {
        const int r = t / rw, c = t - r * rw;
        cnt += 1.0; sr += r; scx += c; rmin = fmin(rmin, (double)r); rmax = fmax(rmax, (double)r); cmin = fmin(cmin, (double)c); cmax = fmax(cmax, (double)c);
    }
    cnt = an_bsum(cnt, red); sr = an_bsum(sr, red); scx = an_bsum(scx, red);
    rmin = an_bmin(rmin, red); rmax = -an_bmin(-rmax, red); cmin = an_bmin(cmin, red); cmax = -an_bmin(-cmax, red);
    if (tid == 0) { sc[5] = cnt; sc[7] = thresh; }
    if (cnt < 20.0) { if (tid == 0) sc[0] = 1.0; return; }            // "Too few edge pixels detected"
    const bool row_x = (rmax - rmin) >= (cmax - cmin);
    const double mr = sr / cnt, mc = scx / cnt;
    double sxx = 0, sxy = 0;
    for (int t = tid; t < n; t += AN_THREADS) if (a.mag[t] > thresh) {
        const int r = t / rw, c = t - r * rw;
        const double x = (row_x ? r : c) - (row_x ? mr : mc), y = (row_x ? c : r) - (row_x ? mc : mr);
        sxx += x * x; sxy += x * y;
    }
    sxx = an_bsum(sxx, red); sxy = an_bsum(sxy, red);
    const double m1 = sxy / sxx, b1 = (row_x ? mc : mr) - m1 * (row_x ? mr : mc);
    const double nrm1 = sqrt(1.0 + m1 * m1);
    // ---- the chosen side of the line, second fit ------------------------------------------------------------------------------
    auto on_side = [&](int r, int c) {
        const double x = row_x ? r : c, y = row_x ? c : r;
        const double ed = (y - m1 * x - b1) / nrm1;
        return a.right_side ? (ed > 0.0) : (ed < 0.0);
    };
    double c2 = 0, sx2 = 0, sy2 = 0;
    for (int t = tid; t < n; t += AN_THREADS) if (a.mag[t] > thresh) {
        const int r = t / rw, c = t - r * rw;
        if (on_side(r, c)) { c2 += 1.0; sx2 += row_x ? r : c; sy2 += row_x ? c : r; }
    }
    c2 = an_bsum(c2, red); sx2 = an_bsum(sx2, red); sy2 = an_bsum(sy2, red);
    if (tid == 0) sc[6] = c2;
    if (c2 < 10.0) { if (tid == 0) sc[0] = 2.0; return; }             // "Too few edge pixels on <side> side"
    const double mx2 = sx2 / c2, my2 = sy2 / c2;
    sxx = 0; sxy = 0;
    for (int t = tid; t < n; t += AN_THREADS) if (a.mag[t] > thresh) {
        const int r = t / rw, c = t - r * rw;
        if (on_side(r, c)) { const double x = (row_x ? r : c) - mx2, y = (row_x ? c : r) - my2; sxx += x * x; sxy += x * y; }
    }
    sxx = an_bsum(sxx, red); sxy = an_bsum(sxy, red);
    const double m = sxy / sxx, b2 = my2 - m * mx2;
    const double angle = (row_x ? atan2(1.0, m) : atan2(m, 1.0)) * (180.0 / 3.14159265358979323846);
    const double nrm = sqrt(1.0 + m * m);

    // ---- signed distance of every pixel; window (-8, 10); bins = arange(dmin, dmax + 1/4, 1/4) ----------------------------------
    double dmin = INFINITY, dmax = -INFINITY;
    for (int t = tid; t < n; t += AN_THREADS) {
        const int r = t / rw, c = t - r * rw;
        const double x = row_x ? r : c, y = row_x ? c : r;
        const double d = __ddiv_rn(__dsub_rn(__dsub_rn(y, __dmul_rn(m, x)), b2), nrm);
        a.dist[t] = d;
        if (d > -8.0 && d < 10.0) { dmin = fmin(dmin, d); dmax = fmax(dmax, d); }
    }
    dmin = an_bmin(dmin, red); dmax = -an_bmin(-dmax, red);
    if (!(dmin <= dmax)) { if (tid == 0) sc[0] = 3.0; return; }
    const double bw = 0.25;
    const int nbins_edges = (int)ceil((dmax + bw - dmin) / bw);        // len(np.arange(start, stop, step))
    const int nb = nbins_edges - 1;                                      // number of bins
    if (nb < 1 || nb > AN_MAXBINS - 1) { if (tid == 0) sc[0] = 4.0; return; }
    const double e1 = dmin + bw, delta = e1 - dmin;                     // numpy fills arange as start + i * (buffer[1] - buffer[0])
    auto edge = [&](int i) { return i == 0 ? dmin : (i == 1 ? e1 : __dadd_rn(dmin, __dmul_rn((double)i, delta))); };
    for (int i = tid; i < AN_MAXBINS; i += AN_THREADS) { bsum[i] = 0.0; bcnt[i] = 0.0; }
    __syncthreads();
    for (int t = tid; t < n; t += AN_THREADS) {
        const double d = a.dist[t];
        if (!(d > -8.0 && d < 10.0)) continue;
        int i = (int)floor((d - dmin) / bw);
        i = max(0, min(i, nb));
        while (i > 0 && d < edge(i)) --i;
        while (i < nb && d >= edge(i + 1)) ++i;
        if (i >= nb || d < edge(i)) continue;                           // beyond the last edge: in no bin
        const int r = t / rw, c = t - r * rw;
        atomicAdd(&bsum[i], (double)src[(size_t)r * a.pitch + c]);
        atomicAdd(&bcnt[i], 1.0);
    }
    __syncthreads();
    // ---- bin centres / means, gap interpolation (np.interp), orientation ---------------------------------------------------------
    if (tid == 0) {
        int nvalid = 0;
        for (int i = 0; i < nb; ++i) {
            ex[i] = 0.5 * (edge(i) + edge(i + 1));
            ey[i] = bcnt[i] > 0.0 ? bsum[i] / bcnt[i] : NAN;
            nvalid += bcnt[i] > 0.0;
        }
        if (nvalid > 2) {
            int first = 0, last = nb - 1;
            while (bcnt[first] <= 0.0) ++first;
            while (bcnt[last] <= 0.0) --last;
            for (int i = 0; i < nb; ++i) {
                if (bcnt[i] > 0.0) continue;
                if (i < first) { ey[i] = ey[first]; continue; }
                if (i > last) { ey[i] = ey[last]; continue; }
                int j = i, k = i;
                while (bcnt[j] <= 0.0) --j;
                while (bcnt[k] <= 0.0) ++k;
                const double slope = (ey[k] - ey[j]) / (ex[k] - ex[j]);
                ey[i] = slope * (ex[i] - ex[j]) + ey[j];
            }
        }
        if (ey[nb - 1] < ey[0]) {                                       // rising edge wanted: esf_x = -esf_x[::-1], esf_y reversed
            for (int i = 0; i < nb / 2; ++i) {
                const double tx = ex[i], ty = ey[i];
                ex[i] = -ex[nb - 1 - i]; ey[i] = ey[nb - 1 - i]; ex[nb - 1 - i] = -tx; ey[nb - 1 - i] = ty;
            }
            if (nb & 1) ex[nb / 2] = -ex[nb / 2];
        }
        // LSF = np.gradient(esf_y, esf_x): uniform branch only if every spacing is bit-equal, else the second-order
        // non-uniform interior formula; first-order one-sided differences at both ends
        bool uniform = true;
        const double d0 = ex[1] - ex[0];
        for (int i = 1; i + 1 < nb; ++i) if ((ex[i + 1] - ex[i]) != d0) { uniform = false; break; }
        if (nb >= 2) {
            lsf[0] = (ey[1] - ey[0]) / (uniform ? d0 : (ex[1] - ex[0]));
            lsf[nb - 1] = (ey[nb - 1] - ey[nb - 2]) / (uniform ? d0 : (ex[nb - 1] - ex[nb - 2]));
            for (int i = 1; i + 1 < nb; ++i) {
                if (uniform) lsf[i] = (ey[i + 1] - ey[i - 1]) / (2.0 * d0);
                else {
                    const double dx1 = ex[i] - ex[i - 1], dx2 = ex[i + 1] - ex[i];
                    const double ca = -(dx2) / (dx1 * (dx1 + dx2)), cb = (dx2 - dx1) / (dx1 * dx2), cc = dx1 / (dx2 * (dx1 + dx2));
                    lsf[i] = ca * ey[i - 1] + cb * ey[i] + cc * ey[i + 1];
                }
            }
        }
        s_n = nb;
    }
    __syncthreads();
    const int ne = s_n;
    if (ne < 4) { if (tid == 0) sc[0] = 5.0; return; }
    for (int i = tid; i < ne; i += AN_THREADS) { a.out_x[i] = ex[i]; a.out_y[i] = ey[i]; a.out_lsf[i] = lsf[i]; }
    __syncthreads();
    // ---- MTF = |DFT(lsf * hanning)|[: n // 2] / MTF[0];  freq = fftfreq(n, mean(diff(esf_x)))[: n // 2] ----------------------------
    for (int i = tid; i < ne; i += AN_THREADS) {
        const double h = 0.5 + 0.5 * cos(3.14159265358979323846 * (double)(1 - ne + 2 * i) / (double)(ne - 1));   // np.hanning
        bsum[i] = lsf[i] * h;
    }
    __syncthreads();
    const int nh = ne / 2;
    for (int k = tid; k < nh; k += AN_THREADS) {
        double re = 0.0, im = 0.0;
        for (int j = 0; j < ne; ++j) {
            const int ph = (int)(((long long)k * j) % ne);
            double s, c; sincospi(-2.0 * (double)ph / (double)ne, &s, &c);
            re += bsum[j] * c; im += bsum[j] * s;
        }
        bcnt[k] = sqrt(re * re + im * im);
    }
    __syncthreads();
    if (tid == 0) {
        const double m0 = bcnt[0];
        double dsum = 0.0;
        for (int i = 0; i + 1 < ne; ++i) dsum += ex[i + 1] - ex[i];
        const double dx = dsum / (double)(ne - 1), val = 1.0 / ((double)ne * dx);
        for (int k = 0; k < nh; ++k) { const double mv = m0 > 0.0 ? bcnt[k] / m0 : bcnt[k]; bcnt[k] = mv; a.out_mtf[k] = mv; a.out_freq[k] = (double)k * val; }
        // MTF50 / MTF10 over the positive frequencies, in cycles / mm (frequency axis divided by the pixel pitch first)
        for (int q = 0; q < 2; ++q) {
            const double frac = q == 0 ? 0.5 : 0.1;
            double res = NAN;
            bool any = false, all = true;
            for (int k = 1; k < nh; ++k) { const bool ab = bcnt[k] >= frac; any |= ab; all &= ab; }
            if (any && !all)
                for (int k = 1; k + 1 < nh; ++k)
                    if (bcnt[k] >= frac && !(bcnt[k + 1] >= frac)) {
                        const double f0 = a.out_freq[k] / a.pitch_mm, f1 = a.out_freq[k + 1] / a.pitch_mm, v0 = bcnt[k], v1 = bcnt[k + 1];
                        res = fabs(v1 - v0) < 1e-12 ? f0 : f0 + (frac - v0) * (f1 - f0) / (v1 - v0);
                        break;
                    }
            sc[3 + q] = res;
        }
        sc[0] = 0.0; sc[1] = (double)ne; sc[2] = angle;
    }
}

__global__ void k_local_contrast(const double* __restrict__ p, int n, int hw, double* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double r = 0.0;
    if (hw > 0 && i >= hw && i < n - hw) {
        double mn = p[i - hw], mx = mn;
        for (int j = i - hw + 1; j < i + hw; ++j) { mn = fmin(mn, p[j]); mx = fmax(mx, p[j]); }
        r = (mx - mn) / (mx + mn + 1e-9);
    }
    out[i] = r;
}

static void gaussian_weights(double sigma, int r, double* w) {           // scipy _gaussian_kernel1d, order 0
    double t[2 * 32 + 1];
    for (int k = 0; k <= 2 * r; ++k) { const double x = (double)(k - r); t[k] = exp(-0.5 / (sigma * sigma) * (x * x)); }
    const double s = np_pairwise_sum(t, (size_t)(2 * r + 1));
    for (int k = 0; k <= 2 * r; ++k) w[k] = t[k] / s;
}

}  // namespace sr459

using namespace sr459;

extern "C" {

int sr459_edge_mtf(const void* d_img, int is_f32, int pitch, int ry0, int rx0, int rh, int rw, int right_side, double pitch_mm,
                   double* h_esf_x, double* h_esf_y, double* h_lsf, double* h_freq, double* h_mtf, double* h_scal, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!d_img || pitch < rw || ry0 < 0 || rx0 < 0 || rh < 1 || rw < 1 || !h_scal || !(pitch_mm > 0.0)) { set_error("edge_mtf: bad arguments"); return SR459_EINVAL; }
    if (rh < 8 || rw < 8) { set_error("Too few edge pixels detected"); return SR459_EINVAL; }
    if ((int64_t)rh * rw > (1 << 22)) { set_error("edge_mtf: ROI too large"); return SR459_EUNSUPPORTED; }
    const size_t esz = is_f32 ? 4 : 1;
    const char* base = (const char*)d_img + ((size_t)ry0 * pitch + rx0) * esz;
    const size_t n = (size_t)rh * rw;
    double* scratch = nullptr;
    SR_CUDA(gmalloc((void**)&scratch, sizeof(double) * (3 * n + 5 * AN_MAXBINS + 8)));
    double* smooth = scratch; double* mag = scratch + n; double* dist = scratch + 2 * n; double* outs = scratch + 3 * n;
    RoiSmoothArgs sa{};
    sa.base = base; sa.is_f32 = is_f32; sa.pitch = pitch; sa.rh = rh; sa.rw = rw; sa.r = (int)(4.0 * 1.5 + 0.5); sa.out = smooth;
    gaussian_weights(1.5, sa.r, sa.w);
    EdgeArgs ea{};
    ea.base = base; ea.is_f32 = is_f32; ea.pitch = pitch; ea.rh = rh; ea.rw = rw; ea.right_side = right_side; ea.mag = mag; ea.dist = dist;
    ea.pitch_mm = pitch_mm;
    ea.out_x = outs; ea.out_y = outs + AN_MAXBINS; ea.out_lsf = outs + 2 * AN_MAXBINS; ea.out_freq = outs + 3 * AN_MAXBINS;
    ea.out_mtf = outs + 4 * AN_MAXBINS; ea.out_scal = outs + 5 * AN_MAXBINS;
    cudaError_t e = cudaMemsetAsync(outs, 0, sizeof(double) * (5 * AN_MAXBINS + 8), st);
    if (e == cudaSuccess) {
        dim3 g((rw + AN_T - 1) / AN_T, (rh + AN_T - 1) / AN_T);
        if (is_f32) k_roi_smooth<float><<<g, 256, 0, st>>>(sa); else k_roi_smooth<uint8_t><<<g, 256, 0, st>>>(sa);
        k_sobel_mag<<<dim3((rw + 31) / 32, (rh + 7) / 8), 256, 0, st>>>(smooth, rh, rw, mag);
        if (is_f32) k_edge_esf<float><<<1, AN_THREADS, 0, st>>>(ea); else k_edge_esf<uint8_t><<<1, AN_THREADS, 0, st>>>(ea);
        e = cudaGetLastError();
        if (e == cudaSuccess) count_launch(3);
    }
    std::vector<double> host(5 * AN_MAXBINS + 8);
    if (e == cudaSuccess) e = cudaMemcpyAsync(host.data(), outs, sizeof(double) * host.size(), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    gfree(scratch);
    if (e != cudaSuccess) { set_error("edge_mtf: %s", cudaGetErrorString(e)); return SR459_ECUDA; }
    const double* sc = host.data() + 5 * AN_MAXBINS;
    for (int i = 0; i < 8; ++i) h_scal[i] = sc[i];
    const int status = (int)sc[0];
    if (status == 1) { set_error("Too few edge pixels detected"); return SR459_EINVAL; }
    if (status == 2) { set_error("Too few edge pixels on %s side", right_side ? "right" : "left"); return SR459_EINVAL; }
    if (status != 0) { set_error("edge_mtf: degenerate edge-spread function (status %d)", status); return SR459_EINVAL; }
    const int ne = (int)sc[1];
    if (h_esf_x) memcpy(h_esf_x, host.data(), sizeof(double) * ne);
    if (h_esf_y) memcpy(h_esf_y, host.data() + AN_MAXBINS, sizeof(double) * ne);
    if (h_lsf) memcpy(h_lsf, host.data() + 2 * AN_MAXBINS, sizeof(double) * ne);
    if (h_freq) memcpy(h_freq, host.data() + 3 * AN_MAXBINS, sizeof(double) * (ne / 2));
    if (h_mtf) memcpy(h_mtf, host.data() + 4 * AN_MAXBINS, sizeof(double) * (ne / 2));
    return SR459_OK;
}

int sr459_edge_mtf_capacity(void) { return AN_MAXBINS; }

int sr459_local_contrast(const double* h_profile, int n, int window, double* h_out, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!h_profile || !h_out || n < 1 || window < 0) { set_error("local_contrast: bad arguments"); return SR459_EINVAL; }
    double* d = nullptr;
    SR_CUDA(gmalloc((void**)&d, sizeof(double) * 2 * (size_t)n));
    cudaError_t e = cudaMemcpyAsync(d, h_profile, sizeof(double) * n, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
        const int hw = window / 2;
        k_local_contrast<<<(n + 255) / 256, 256, 0, st>>>(d, n, (n > 2 * hw) ? hw : 0, d + n);
        e = cudaGetLastError();
        if (e == cudaSuccess) count_launch();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(h_out, d + n, sizeof(double) * n, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    gfree(d);
    if (e != cudaSuccess) { set_error("local_contrast: %s", cudaGetErrorString(e)); return SR459_ECUDA; }
    return SR459_OK;
}

}  // extern "C"
