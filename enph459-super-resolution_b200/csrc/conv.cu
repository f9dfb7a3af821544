// conv.cu -- zero-padded 2-D convolution / correlation with an arbitrary PSF.
//
// Replaces blur() = scipy.signal.fftconvolve(img, kernel, mode='same')
// (mono_cal_target/run_sr.py:157-158) and its use with the flipped kernel in
// back_project (:178).  'same' with a kh x kw kernel is
//     out[y,x] = sum_{u,v} k[u,v] * img0[y + cy - u, x + cx - v],  cy=(kh-1)/2, cx=(kw-1)/2
// with img0 the zero extension of img.  The host turns that into a list of
// non-zero taps (dy, dx, w) -- the measured 7x7 PSF has only 19 non-zero taps of
// 49, the 31x31 one 830 of 961 -- and the kernel evaluates the sparse
// correlation from a shared-memory tile with halo.  The IBP update
// hr <- clip(hr + step/n * corr, 0, 255) (run_sr.py:204-205) is an optional
// fused epilogue.
#include "internal.h"
#include <algorithm>
#include <cstdlib>

namespace sr459 {

void build_taps_host(const double* k, int kh, int kw, bool flip, TapList& t) {
    t.w.clear(); t.dy.clear(); t.dx.clear();
    const int cy = (kh - 1) / 2, cx = (kw - 1) / 2;
    int reach = 0;
    for (int u = 0; u < kh; ++u)
        for (int v = 0; v < kw; ++v) {
            double val = flip ? k[(kh - 1 - u) * kw + (kw - 1 - v)] : k[u * kw + v];
            if (val == 0.0) continue;
            t.w.push_back((float)val);
            t.dy.push_back(cy - u);
            t.dx.push_back(cx - v);
            reach = std::max(reach, std::max(std::abs(cy - u), std::abs(cx - v)));
        }
    t.ntaps = (int)t.w.size();
    t.reach = reach;
}

int upload_taps(TapList& t) {
    free_taps(t);
    size_t n = std::max(1, t.ntaps);
    SR_CUDA(gmalloc((void**)&t.d_w, n * sizeof(float)));
    SR_CUDA(gmalloc((void**)&t.d_dy, n * sizeof(int)));
    SR_CUDA(gmalloc((void**)&t.d_dx, n * sizeof(int)));
    if (t.ntaps) {
        SR_CUDA(cudaMemcpy(t.d_w, t.w.data(), t.ntaps * sizeof(float), cudaMemcpyHostToDevice));
        SR_CUDA(cudaMemcpy(t.d_dy, t.dy.data(), t.ntaps * sizeof(int), cudaMemcpyHostToDevice));
        SR_CUDA(cudaMemcpy(t.d_dx, t.dx.data(), t.ntaps * sizeof(int), cudaMemcpyHostToDevice));
    }
    return SR459_OK;
}

void free_taps(TapList& t) {
    if (t.d_w) gfree(t.d_w);
    if (t.d_dy) gfree(t.d_dy);
    if (t.d_dx) gfree(t.d_dx);
    t.d_w = nullptr; t.d_dy = nullptr; t.d_dx = nullptr;
}

constexpr int CV_TW = 64;   // tile width  (outputs)
constexpr int CV_TH = 32;   // tile height (outputs)
constexpr int CV_RPT = 4;   // consecutive rows per thread

// 256 threads: 64 columns x 4 row-groups, each thread 8 rows in two chunks of 4.
__global__ void __launch_bounds__(256)
k_corr2d(const float* __restrict__ in, int H, int W, const float* __restrict__ tw,
         const int* __restrict__ tdy, const int* __restrict__ tdx, int ntaps, int R,
         float* __restrict__ out, int mode, const float* __restrict__ base, float alpha,
         float lo, float hi, int sy0, int sy1) {
    extern __shared__ float sm[];
    const int SW = CV_TW + 2 * R, SH = CV_TH + 2 * R;
    float* tile = sm;                                  // SH x SW
    float* sw   = tile + SH * SW;                      // ntaps weights
    int*   soff = reinterpret_cast<int*>(sw + ntaps);  // ntaps smem offsets
    const int tid = threadIdx.x;
    const int y0 = blockIdx.y * CV_TH, x0 = blockIdx.x * CV_TW;

    for (int t = tid; t < SH * SW; t += 256) {
        int ly = t / SW, lx = t - ly * SW;
        int gy = y0 - R + ly, gx = x0 - R + lx;
        tile[t] = (gy >= 0 && gy < H && gx >= 0 && gx < W) ? __ldg(in + (size_t)gy * W + gx) : 0.0f;
    }
    for (int t = tid; t < ntaps; t += 256) {
        sw[t]   = tw[t];
        soff[t] = tdy[t] * SW + tdx[t];
    }
    __syncthreads();

    const int lx = tid & 63, rg = tid >> 6;            // rg 0..3
    #pragma unroll
    for (int chunk = 0; chunk < 2; ++chunk) {
        const int ly = chunk * 16 + rg * CV_RPT;
        const float* c = tile + (ly + R) * SW + lx + R;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        for (int t = 0; t < ntaps; ++t) {
            const float w = sw[t];
            const float* p = c + soff[t];
            a0 = fmaf(w, p[0], a0);
            a1 = fmaf(w, p[SW], a1);
            a2 = fmaf(w, p[2 * SW], a2);
            a3 = fmaf(w, p[3 * SW], a3);
        }
        float v[4] = {a0, a1, a2, a3};
        const int gx = x0 + lx;
        if (gx < W) {
            #pragma unroll
            for (int r = 0; r < 4; ++r) {
                int gy = y0 + ly + r;
                if (gy < H && gy >= sy0 && gy < sy1) {
                    size_t o = (size_t)gy * W + gx;
                    float val = v[r];
                    if (mode == 1) val = fminf(fmaxf(fmaf(alpha, val, __ldg(base + o)), lo), hi);
                    out[o] = val;
                }
            }
        }
    }
}

int launch_corr2d(const CorrArgs& a, cudaStream_t st) {
    const TapList& t = *a.taps;
    const int R = t.reach;
    size_t smem = sizeof(float) * (size_t)(CV_TW + 2 * R) * (CV_TH + 2 * R) + (size_t)t.ntaps * 8;
    if (smem > 200 * 1024) { set_error("PSF too large for the shared-memory correlation tile (reach %d)", R); return SR459_EUNSUPPORTED; }
    if (smem > 48 * 1024) {
        SR_CUDA(cudaFuncSetAttribute(k_corr2d, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    dim3 grid((a.W + CV_TW - 1) / CV_TW, (a.H + CV_TH - 1) / CV_TH);
    k_corr2d<<<grid, 256, smem, st>>>(a.in, a.H, a.W, t.d_w, t.d_dy, t.d_dx, t.ntaps, R, a.out,
                                      a.mode, a.base, a.alpha, a.lo, a.hi, a.y0, a.y1 < 0 ? a.H : a.y1);
    SR_LAUNCH_CHECK("k_corr2d");
    return SR459_OK;
}

}  // namespace sr459
