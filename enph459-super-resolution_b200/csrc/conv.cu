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

// ---- PSFs up to 7 x 7: register-window form -------------------------------------------------------------------------------
// The generic kernel above spends one shared-memory load per multiply-add.  For reach <= 3 (every 7 x 7 PSF of the reference:
// Gaussian and measured, run_sr.py:104-152) a thread instead walks DOWN a 4-column strip with a sliding window of 7 staged
// rows x 12 columns in registers (three 128-bit loads per new row) and evaluates all 49 taps of its 4 outputs from registers,
// the weights coming from the kernel-parameter constant bank (196 FMA per 3 loads).  Tile = 32 rows x 128 columns per CTA of
// 128 threads (lane <-> 4 columns, warp <-> 8 output rows).  Same zero padding, same fused update epilogue.
// Rows per warp (RPW) is chosen per launch so that the grid fills whole waves (768 tiles of 32 rows on 740 resident CTA slots
// would run as two waves: 23 us instead of 12 at 1536 x 2048).
constexpr int C7_TW = 128, C7_SW = C7_TW + 8;
struct Corr7Args {
    const float* in; int H, W;
    float w[7][7];                  // correlation weights: out[y,x] = sum w[u][v] * in0[y + u - 3, x + v - 3]
    float* out;
    int mode; const float* base; float alpha, lo, hi;
    int sy0, sy1;
};

template <int C7_RPW>
__global__ void __launch_bounds__(128) k_corr7(const __grid_constant__ Corr7Args a) {
    constexpr int C7_TH = 4 * C7_RPW, C7_SH = C7_TH + 6;
    __shared__ __align__(16) float tile[C7_SH * C7_SW];            // rows y0-3 .. y0+TH+2, columns x0-4 .. x0+131
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int y0 = blockIdx.y * C7_TH, x0 = blockIdx.x * C7_TW;
    const bool vec = (a.W & 3) == 0;
    {   // staging: every load of the thread is in flight before the first shared-memory store (one DRAM latency, not one per element)
        constexpr int NV = C7_SH * (C7_SW / 4), PER = (NV + 127) / 128;
        float4 v[PER];
        #pragma unroll
        for (int i = 0; i < PER; ++i) {
            const int t = tid + 128 * i;
            const int ly = t / (C7_SW / 4), l4 = t - ly * (C7_SW / 4);
            const int gy = y0 - 3 + ly, gx = x0 - 4 + 4 * l4;
            v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (t < NV && gy >= 0 && gy < a.H) {
                const float* row = a.in + (size_t)gy * a.W;
                if (vec && gx >= 0 && gx + 3 < a.W) v[i] = __ldg(reinterpret_cast<const float4*>(row + gx));
                else {
                    if (gx >= 0 && gx < a.W) v[i].x = __ldg(row + gx);
                    if (gx + 1 >= 0 && gx + 1 < a.W) v[i].y = __ldg(row + gx + 1);
                    if (gx + 2 >= 0 && gx + 2 < a.W) v[i].z = __ldg(row + gx + 2);
                    if (gx + 3 >= 0 && gx + 3 < a.W) v[i].w = __ldg(row + gx + 3);
                }
            }
        }
        #pragma unroll
        for (int i = 0; i < PER; ++i) {
            const int t = tid + 128 * i;
            if (t < NV) *reinterpret_cast<float4*>(tile + 4 * t) = v[i];          // ly * C7_SW + 4 * l4 == 4 * t
        }
    }
    __syncthreads();
    // window slot r holds staged row (step + r) of this warp's strip: columns 4 lane .. 4 lane + 11 (= x - 4 .. x + 7)
    float win[7][12];
    const float* col = tile + (warp * C7_RPW) * C7_SW + 4 * lane;
    #pragma unroll
    for (int r = 0; r < 6; ++r) {
        #pragma unroll
        for (int q = 0; q < 3; ++q) {
            const float4 v = *reinterpret_cast<const float4*>(col + r * C7_SW + 4 * q);
            win[r][4 * q] = v.x; win[r][4 * q + 1] = v.y; win[r][4 * q + 2] = v.z; win[r][4 * q + 3] = v.w;
        }
    }
    const int gx = x0 + 4 * lane;
    #pragma unroll
    for (int step = 0; step < C7_RPW; ++step) {
        {   // the 7th row of this step's window
            #pragma unroll
            for (int q = 0; q < 3; ++q) {
                const float4 v = *reinterpret_cast<const float4*>(col + (step + 6) * C7_SW + 4 * q);
                win[(step + 6) % 7][4 * q] = v.x; win[(step + 6) % 7][4 * q + 1] = v.y; win[(step + 6) % 7][4 * q + 2] = v.z; win[(step + 6) % 7][4 * q + 3] = v.w;
            }
        }
        float o[4] = {0.f, 0.f, 0.f, 0.f};
        #pragma unroll
        for (int u = 0; u < 7; ++u) {
            #pragma unroll
            for (int v = 0; v < 7; ++v) {
                const float wv = a.w[u][v];
                #pragma unroll
                for (int c = 0; c < 4; ++c) o[c] = fmaf(wv, win[(step + u) % 7][c + v + 1], o[c]);   // column x + c + v - 3 = window index c + v + 1
            }
        }
        const int gy = y0 + warp * C7_RPW + step;
        if (gy < a.H && gy >= a.sy0 && gy < a.sy1 && gx < a.W) {
            const size_t off = (size_t)gy * a.W + gx;
            if (vec) {
                float4 r = make_float4(o[0], o[1], o[2], o[3]);
                if (a.mode == 1) {
                    const float4 b = __ldg(reinterpret_cast<const float4*>(a.base + off));
                    r.x = fminf(fmaxf(fmaf(a.alpha, r.x, b.x), a.lo), a.hi); r.y = fminf(fmaxf(fmaf(a.alpha, r.y, b.y), a.lo), a.hi);
                    r.z = fminf(fmaxf(fmaf(a.alpha, r.z, b.z), a.lo), a.hi); r.w = fminf(fmaxf(fmaf(a.alpha, r.w, b.w), a.lo), a.hi);
                }
                *reinterpret_cast<float4*>(a.out + off) = r;
            } else {
                #pragma unroll
                for (int c = 0; c < 4; ++c) {
                    if (gx + c >= a.W) break;
                    float val = o[c];
                    if (a.mode == 1) val = fminf(fmaxf(fmaf(a.alpha, val, __ldg(a.base + off + c)), a.lo), a.hi);
                    a.out[off + c] = val;
                }
            }
        }
    }
}

int launch_corr2d(const CorrArgs& a, cudaStream_t st) {
    const TapList& t = *a.taps;
    const int R = t.reach;
    static const bool no7 = getenv("SR459_NO_CORR7") != nullptr;
    if (R <= 3 && !no7) {
        Corr7Args k{};
        k.in = a.in; k.H = a.H; k.W = a.W; k.out = a.out; k.mode = a.mode; k.base = a.base; k.alpha = a.alpha; k.lo = a.lo; k.hi = a.hi;
        k.sy0 = a.y0; k.sy1 = a.y1 < 0 ? a.H : a.y1;
        for (int i = 0; i < t.ntaps; ++i) k.w[t.dy[i] + 3][t.dx[i] + 3] = t.w[i];      // out += w * in0[y + dy, x + dx]
        static int n_sm_memo[64] = {};
        int dev = 0, n_sm = 148;
        if (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < 64) {
            if (!n_sm_memo[dev]) cudaDeviceGetAttribute(&n_sm_memo[dev], cudaDevAttrMultiProcessorCount, dev);
            if (n_sm_memo[dev] > 0) n_sm = n_sm_memo[dev];
        }
        const int tx = (a.W + C7_TW - 1) / C7_TW;
        int best = 8; double best_cost = 1e30;
        for (int rpw : {4, 6, 8, 10, 12}) {
            const long long tiles = (long long)tx * ((a.H + 4 * rpw - 1) / (4 * rpw));
            const int per_sm = 5;                                                  // resident CTAs (92 registers x 128 threads)
            const double waves = (double)((tiles + (long long)n_sm * per_sm - 1) / ((long long)n_sm * per_sm));
            const double cost = waves * (rpw + 3.0);
            if (cost < best_cost - 1e-9) { best_cost = cost; best = rpw; }
        }
        dim3 grid(tx, (a.H + 4 * best - 1) / (4 * best));
        switch (best) {
            case 4:  k_corr7<4><<<grid, 128, 0, st>>>(k); break;
            case 6:  k_corr7<6><<<grid, 128, 0, st>>>(k); break;
            case 10: k_corr7<10><<<grid, 128, 0, st>>>(k); break;
            case 12: k_corr7<12><<<grid, 128, 0, st>>>(k); break;
            default: k_corr7<8><<<grid, 128, 0, st>>>(k); break;
        }
        SR_LAUNCH_CHECK("k_corr7");
        return SR459_OK;
    }
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
