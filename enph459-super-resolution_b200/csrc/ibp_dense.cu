// ibp_dense.cu -- one fused IBP iteration for INTEGER high-resolution shifts, dense form.
//
// Restates the body of ibp() (mono_cal_target/run_sr.py:190-209) for the case
// where every factor*shift is an integer (3 of the reference's 4 scripts:
// nominal +-0.5 LR px = +-1 HR px, quirk Q3).  scipy's order-3 shift with an
// integer offset returns the edge-clamped sample exactly, so with
//     B   = conv0(hr, psf)                               (blur, zero padded, :157-158)
//     C   : C(a, s)[y,x] = a[clamp(y - sy), clamp(x - sx)] (ndi_shift, mode='nearest')
//     U   : zero-insert at even/even sites                (:170-171)
// the iteration is
//     err_k = lr_k - C(B, +s_k)[::2, ::2]                 (:161-165, :201)
//     acc   = sum_k C(U(err_k), -s_k)                     (:176, summed before the blur: linearity)
//     hr'   = clip(hr + step/n * conv0(acc, flip(psf)), lo, hi)     (:178, :204-205)
//     mse   = 1/n sum_k mean(err_k^2)                     (:202, :206)
// One CTA owns an output tile, stages hr with a (2R+2m)-pixel halo in shared
// memory, evaluates B, the residual accumulation and the transposed blur
// entirely on chip and writes the updated tile: HBM traffic per iteration is
// hr once in, hr once out and the n LR frames once (plus halo re-reads that
// hit L2).  This "dense" variant evaluates every HR site and every boundary
// clamp literally; it is the always-correct form used for image borders, small
// images, unusual shifts and PSF sizes.  ibp_fast.cu is the polyphase interior
// specialisation that carries the throughput.
#include "internal.h"
#include <algorithm>

namespace sr459 {

struct DenseSmemLayout {
    int hR, bR, aR;          // halos of the hr / B / acc regions
    int hw, hh, bw, bh, aw, ah;
    size_t off_hr, off_b, off_acc, off_wf, off_of, off_wb, off_ob, total;
};

__host__ __device__ inline DenseSmemLayout dense_layout(int R, int m, int th, int tw, int nf, int nb) {
    DenseSmemLayout L;
    L.aR = R; L.bR = R + 2 * m; L.hR = 2 * R + 2 * m;
    L.hh = th + 2 * L.hR; L.hw = tw + 2 * L.hR;
    L.bh = th + 2 * L.bR; L.bw = tw + 2 * L.bR;
    L.ah = th + 2 * L.aR; L.aw = tw + 2 * L.aR;
    size_t o = 0;
    L.off_hr = o;  o += (size_t)L.hh * L.hw;
    L.off_b = o;   o += (size_t)L.bh * L.bw;
    L.off_acc = o; o += (size_t)L.ah * L.aw;
    L.off_wf = o;  o += nf;
    L.off_of = o;  o += nf;
    L.off_wb = o;  o += nb;
    L.off_ob = o;  o += nb;
    L.total = o * sizeof(float);
    return L;
}

size_t ibp_dense_smem(int R, int smax, int tile_h, int tile_w, int nf, int nb) {
    return dense_layout(R, smax, tile_h, tile_w, nf, nb).total;
}

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }

__global__ void __launch_bounds__(256) k_ibp_dense(const DenseArgs a) {
    extern __shared__ float sm[];
    const IbpGeom& g = a.g;
    const int tid = threadIdx.x;

    // ---- which tile ----------------------------------------------------------------
    int ri = 0;
    while (ri + 1 < a.nrect && (int)blockIdx.x >= a.tiles_before[ri + 1]) ++ri;
    const Rect rc = a.rect[ri];
    const int tl = blockIdx.x - a.tiles_before[ri];
    const int tyi = tl / a.tiles_x[ri], txi = tl - tyi * a.tiles_x[ri];
    const int y0 = rc.y0 + tyi * a.tile_h, x0 = rc.x0 + txi * a.tile_w;      // global, even
    const int y1 = min(y0 + a.tile_h, rc.y1), x1 = min(x0 + a.tile_w, rc.x1);

    const DenseSmemLayout L = dense_layout(a.R, g.smax, a.tile_h, a.tile_w, a.nf, a.nb);
    float* hr_s  = sm + L.off_hr;
    float* b_s   = sm + L.off_b;
    float* acc_s = sm + L.off_acc;
    float* wf_s  = sm + L.off_wf;
    int*   of_s  = reinterpret_cast<int*>(sm + L.off_of);
    float* wb_s  = sm + L.off_wb;
    int*   ob_s  = reinterpret_cast<int*>(sm + L.off_ob);

    // ---- stage 1: hr tile + halo (zero outside the image = conv0's padding) -----------
    for (int t = tid; t < L.hh * L.hw; t += 256) {
        int ly = t / L.hw, lx = t - ly * L.hw;
        int gy = y0 - L.hR + ly, gx = x0 - L.hR + lx;
        float v = 0.0f;
        if (gy >= 0 && gy < g.H && gx >= 0 && gx < g.W && gy >= g.hr_lo && gy < g.hr_hi)
            v = __ldg(a.hr_in + (size_t)(gy - g.hr_lo) * g.W + gx);
        hr_s[t] = v;
    }
    for (int t = tid; t < a.nf; t += 256) { wf_s[t] = a.wf[t]; of_s[t] = a.fdy[t] * L.hw + a.fdx[t]; }
    for (int t = tid; t < a.nb; t += 256) { wb_s[t] = a.wb[t]; ob_s[t] = a.bdy[t] * L.aw + a.bdx[t]; }
    __syncthreads();

    // ---- stage 2: B = conv0(hr, psf) on tile +- bR ------------------------------------
    for (int t = tid; t < L.bh * L.bw; t += 256) {
        int ly = t / L.bw, lx = t - ly * L.bw;
        int gy = y0 - L.bR + ly, gx = x0 - L.bR + lx;
        float v = 0.0f;
        if (gy >= 0 && gy < g.H && gx >= 0 && gx < g.W) {
            const float* c = hr_s + (ly + L.hR - L.bR) * L.hw + (lx + L.hR - L.bR);
            for (int k = 0; k < a.nf; ++k) v = fmaf(wf_s[k], c[of_s[k]], v);
        }
        b_s[t] = v;
    }
    __syncthreads();

    // ---- stage 3: acc = sum_k C(U(lr_k - sim_k), -s_k) on tile +- aR --------------------
    for (int t = tid; t < L.ah * L.aw; t += 256) {
        int ly = t / L.aw, lx = t - ly * L.aw;
        int gy = y0 - L.aR + ly, gx = x0 - L.aR + lx;
        float v = 0.0f;
        if (gy >= 0 && gy < g.H && gx >= 0 && gx < g.W) {
            for (int k = 0; k < g.n; ++k) {
                int Y = clampi(gy + g.sy[k], 0, g.H - 1);
                int X = clampi(gx + g.sx[k], 0, g.W - 1);
                if ((Y | X) & 1) continue;                    // zero-inserted sites
                int i = Y >> 1, j = X >> 1;
                if (i >= g.h || j >= g.w) continue;
                int by = clampi(2 * i - g.sy[k], 0, g.H - 1) - (y0 - L.bR);
                int bx = clampi(2 * j - g.sx[k], 0, g.W - 1) - (x0 - L.bR);
                float e = __ldg(a.lr[k] + (size_t)(i - g.lr_lo) * g.w + j) - b_s[by * L.bw + bx];
                v += e;
            }
        }
        acc_s[t] = v;
    }

    // ---- stage 3b: sum of squared residuals over the LR sites this tile owns ----------
    float e2 = 0.0f;
    {
        const int i0 = y0 >> 1, j0 = x0 >> 1;
        const int ni = (y1 - y0 + 1) >> 1, nj = (x1 - x0 + 1) >> 1;
        for (int t = tid; t < ni * nj; t += 256) {
            int i = i0 + t / nj, j = j0 + t % nj;
            if (i >= g.h || j >= g.w || i < g.own_i0 || i >= g.own_i1) continue;
            for (int k = 0; k < g.n; ++k) {
                int by = clampi(2 * i - g.sy[k], 0, g.H - 1) - (y0 - L.bR);
                int bx = clampi(2 * j - g.sx[k], 0, g.W - 1) - (x0 - L.bR);
                float e = __ldg(a.lr[k] + (size_t)(i - g.lr_lo) * g.w + j) - b_s[by * L.bw + bx];
                e2 = fmaf(e, e, e2);
            }
        }
    }
    __syncthreads();

    // ---- stage 4: hr' = clip(hr + alpha * corr(acc, psf)) ------------------------------
    const int th = y1 - y0, tw = x1 - x0;
    for (int t = tid; t < th * tw; t += 256) {
        int ly = t / tw, lx = t - ly * tw;
        int gy = y0 + ly, gx = x0 + lx;
        if (gy < g.upd_y0 || gy >= g.upd_y1) continue;
        const float* c = acc_s + (ly + L.aR) * L.aw + (lx + L.aR);
        float v = 0.0f;
        for (int k = 0; k < a.nb; ++k) v = fmaf(wb_s[k], c[ob_s[k]], v);
        float hv = hr_s[(ly + L.hR) * L.hw + lx + L.hR];
        a.hr_out[(size_t)(gy - g.hr_lo) * g.W + gx] = fminf(fmaxf(fmaf(g.alpha, v, hv), g.lo), g.hi);
    }

    // ---- block reduction of the residual energy -----------------------------------------
    __shared__ double red[8];
    double d = (double)e2;
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    if ((tid & 31) == 0) red[tid >> 5] = d;
    __syncthreads();
    if (tid == 0) {
        double s = 0.0;
        #pragma unroll
        for (int i = 0; i < 8; ++i) s += red[i];
        a.partial[blockIdx.x] = s;
    }
}

int launch_ibp_dense(const DenseArgs& a, cudaStream_t st) {
    int ntiles = a.tiles_before[a.nrect];
    if (ntiles <= 0) return SR459_OK;
    size_t smem = ibp_dense_smem(a.R, a.g.smax, a.tile_h, a.tile_w, a.nf, a.nb);
    if (smem > 220 * 1024) { set_error("dense IBP tile needs %zu B of shared memory", smem); return SR459_EUNSUPPORTED; }
    static PerDeviceMax configured;
    if (smem > 48 * 1024 && configured.raise(smem)) {
        SR_CUDA(cudaFuncSetAttribute(k_ibp_dense, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    k_ibp_dense<<<ntiles, 256, smem, st>>>(a);
    SR_LAUNCH_CHECK("k_ibp_dense");
    return SR459_OK;
}

__global__ void k_reduce_partials(const double* __restrict__ partial, int per_iter, double scale,
                                  const double* __restrict__ add, double* __restrict__ errors) {
    __shared__ double red[32];
    const double* p = partial + (size_t)blockIdx.x * per_iter;
    double s = 0.0;
    for (int i = threadIdx.x; i < per_iter; i += blockDim.x) s += p[i];
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += red[i];
        if (add) t += *add;
        errors[blockIdx.x] = t * scale;
    }
}

int launch_reduce_partials(const double* partial, int per_iter, int n_iter, double scale,
                           const double* d_add, double* d_errors, cudaStream_t st) {
    if (n_iter <= 0) return SR459_OK;
    k_reduce_partials<<<n_iter, 256, 0, st>>>(partial, per_iter, scale, d_add, d_errors);
    SR_LAUNCH_CHECK("k_reduce_partials");
    return SR459_OK;
}

}  // namespace sr459
