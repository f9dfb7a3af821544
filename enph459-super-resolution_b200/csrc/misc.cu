// misc.cu -- frame ingest / egress and measured-PSF reductions.
//
//   load_gray        uint8 -> float            mono_cal_target/run_sr.py:73-75
//   extract_red      img[0::2, 0::2]           rgb_cal_target/run_sr.py:73-75
//   np.mean(frames)  rep / frame averaging     run_sr.py:274, rgb_cal_target/run_sr.py:107-108
//   uint8 cast       clip(0,255) + truncation  run_sr.py:303
//   load_measured_psf: argmax (first maximum, row-major) and 19x19 patch mean
//                                              run_sr.py:127, :132, :137
#include "internal.h"

namespace sr459 {

// 16 elements per thread: one 16-byte load -> four float4 stores (and the reverse); the ragged / unaligned tail is scalar
__global__ void __launch_bounds__(256) k_u8_to_f32(const uint8_t* __restrict__ in, int64_t n, float* __restrict__ out, int aligned) {
    const int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 16;
    if (i0 >= n) return;
    if (aligned && i0 + 15 < n) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(in + i0));
        const unsigned w[4] = {v.x, v.y, v.z, v.w};
        float4* o = reinterpret_cast<float4*>(out + i0);
        #pragma unroll
        for (int q = 0; q < 4; ++q)
            o[q] = make_float4((float)(w[q] & 0xffu), (float)((w[q] >> 8) & 0xffu), (float)((w[q] >> 16) & 0xffu), (float)(w[q] >> 24));
    } else {
        for (int64_t i = i0; i < n && i < i0 + 16; ++i) out[i] = (float)in[i];
    }
}

int launch_u8_to_f32(const uint8_t* in, int64_t n, float* out, cudaStream_t st) {
    if (n <= 0) return SR459_OK;
    const int aligned = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
    const int64_t nt = (n + 15) / 16;
    k_u8_to_f32<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(in, n, out, aligned);
    SR_LAUNCH_CHECK("k_u8_to_f32");
    return SR459_OK;
}

__device__ __forceinline__ unsigned f32_to_u8_trunc(float x) {       // clip to [0, 255], truncation toward zero
    return (unsigned)(int)fminf(fmaxf(x, 0.0f), 255.0f);           //   like np.clip(...).astype(np.uint8) (run_sr.py:303)
}

__global__ void __launch_bounds__(256) k_f32_to_u8(const float* __restrict__ in, int64_t n, uint8_t* __restrict__ out, int aligned) {
    const int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 16;
    if (i0 >= n) return;
    if (aligned && i0 + 15 < n) {
        const float4* p = reinterpret_cast<const float4*>(in + i0);
        unsigned w[4];
        #pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float4 v = __ldg(p + q);
            w[q] = f32_to_u8_trunc(v.x) | (f32_to_u8_trunc(v.y) << 8) | (f32_to_u8_trunc(v.z) << 16) | (f32_to_u8_trunc(v.w) << 24);
        }
        *reinterpret_cast<uint4*>(out + i0) = make_uint4(w[0], w[1], w[2], w[3]);
    } else {
        for (int64_t i = i0; i < n && i < i0 + 16; ++i) out[i] = (uint8_t)f32_to_u8_trunc(in[i]);
    }
}

int launch_f32_to_u8(const float* in, int64_t n, uint8_t* out, cudaStream_t st) {
    if (n <= 0) return SR459_OK;
    const int aligned = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
    const int64_t nt = (n + 15) / 16;
    k_f32_to_u8<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(in, n, out, aligned);
    SR_LAUNCH_CHECK("k_f32_to_u8");
    return SR459_OK;
}

template <typename T>
__global__ void k_extract_red(const T* __restrict__ in, int H, int W, int ro, int co, int oh, int ow,
                              float* __restrict__ out) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    int i = blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= oh || j >= ow) return;
    out[(size_t)i * ow + j] = (float)in[(size_t)(2 * i + ro) * W + 2 * j + co];
}

int launch_extract_red(const void* in, int is_f32, int H, int W, int ro, int co, float* out, cudaStream_t st) {
    int oh = (H - ro + 1) / 2, ow = (W - co + 1) / 2;
    if (oh <= 0 || ow <= 0) return SR459_OK;
    dim3 block(32, 8), grid((ow + 31) / 32, (oh + 7) / 8);
    if (is_f32) k_extract_red<float><<<grid, block, 0, st>>>((const float*)in, H, W, ro, co, oh, ow, out);
    else        k_extract_red<uint8_t><<<grid, block, 0, st>>>((const uint8_t*)in, H, W, ro, co, oh, ow, out);
    SR_LAUNCH_CHECK("k_extract_red");
    return SR459_OK;
}

struct FramePtrs { const float* p[32]; };

__global__ void k_mean_frames(FramePtrs f, int n, int64_t npx, float fn, float* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npx) return;
    float s = 0.0f;
    for (int k = 0; k < n; ++k) s += __ldg(f.p[k] + i);
    out[i] = s / fn;
}

int launch_mean_frames(const float* const* h_frames, int n, int64_t npx, float* out, cudaStream_t st) {
    if (n < 1 || n > 32) { set_error("mean_frames supports 1..32 frames, got %d", n); return SR459_EINVAL; }
    FramePtrs f;
    for (int k = 0; k < n; ++k) f.p[k] = h_frames[k];
    k_mean_frames<<<(unsigned)((npx + 255) / 256), 256, 0, st>>>(f, n, npx, (float)n, out);
    SR_LAUNCH_CHECK("k_mean_frames");
    return SR459_OK;
}

// ---- fused frame-set ingest ----------------------------------------------------------------------------
// load_gray + extract_red + the per-corner repetition mean (rgb_cal_target/run_sr.py:107-108) + np.mean over
// the frames (run_sr.py:274) in ONE pass over the uint8 files, in float64 and in numpy's order of operations
// (sequential sums along axis 0, one division each), so LR_mean.png truncates exactly as the reference's does.
struct IngestPtrs { const uint8_t* const* img; float* frame[8]; int first[9]; };   // images of frame g: img[first[g] .. first[g+1])

__global__ void __launch_bounds__(256) k_ingest_groups(const __grid_constant__ IngestPtrs p, int n_groups, int W,
                                                       int step, int ro, int co, int oh, int ow,
                                                       float* __restrict__ mean_f32, uint8_t* __restrict__ mean_u8) {
    const int j = blockIdx.x * 32 + (threadIdx.x & 31);
    const int i = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (i >= oh || j >= ow) return;
    const size_t src = (size_t)(step * i + ro) * W + (size_t)(step * j + co), dst = (size_t)i * ow + j;
    double total = 0.0;
    for (int g = 0; g < n_groups; ++g) {
        const int reps = p.first[g + 1] - p.first[g];
        double s = 0.0;
        for (int r = p.first[g]; r < p.first[g + 1]; ++r) s += (double)p.img[r][src];
        const double m = (reps > 1) ? s / (double)reps : s;
        p.frame[g][dst] = (float)m;
        total += m;
    }
    const double mean = total / (double)n_groups;
    if (mean_f32) mean_f32[dst] = (float)mean;
    if (mean_u8) mean_u8[dst] = (uint8_t)(int)fmin(fmax(mean, 0.0), 255.0);
}

// h_reps[g] = repetitions of frame g (any count >= 1; the frames may differ, as np.stack(...).mean(axis=0) per corner
// allows, rgb_cal_target/run_sr.py:107-108); the image-pointer table travels through a stream-ordered device buffer
int launch_ingest_groups(const uint8_t* const* h_imgs, int n_groups, const int* h_reps, int H, int W, int step, int ro, int co,
                         float* const* h_frames, float* mean_f32, uint8_t* mean_u8, cudaStream_t st) {
    if (n_groups < 1 || n_groups > 8) { set_error("ingest_groups: 1..8 frames, got %d", n_groups); return SR459_EINVAL; }
    if (step < 1 || ro < 0 || co < 0 || ro >= H || co >= W) { set_error("ingest_groups: bad lattice"); return SR459_EINVAL; }
    const int oh = (H - ro + step - 1) / step, ow = (W - co + step - 1) / step;
    IngestPtrs p{};
    int total = 0;
    for (int g = 0; g < n_groups; ++g) {
        if (h_reps[g] < 1) { set_error("ingest_groups: frame %d has %d repetitions", g, h_reps[g]); return SR459_EINVAL; }
        p.first[g] = total; total += h_reps[g]; p.frame[g] = h_frames[g];
    }
    p.first[n_groups] = total;
    const uint8_t** d_tbl = nullptr;
    SR_CUDA(cudaMallocAsync((void**)&d_tbl, sizeof(void*) * total, st));
    cudaError_t e = cudaMemcpyAsync(d_tbl, h_imgs, sizeof(void*) * total, cudaMemcpyHostToDevice, st);   // pageable source: staged before return
    if (e != cudaSuccess) { cudaFreeAsync(d_tbl, st); set_error("ingest_groups: %s", cudaGetErrorString(e)); return SR459_ECUDA; }
    p.img = d_tbl;
    dim3 grid((ow + 31) / 32, (oh + 7) / 8);
    k_ingest_groups<<<grid, 256, 0, st>>>(p, n_groups, W, step, ro, co, oh, ow, mean_f32, mean_u8);
    e = cudaGetLastError();
    cudaFreeAsync(d_tbl, st);
    if (e != cudaSuccess) { set_error("launch of k_ingest_groups failed: %s", cudaGetErrorString(e)); return SR459_ECUDA; }
    count_launch();
    return SR459_OK;
}

// ---- argmax: first maximum in row-major order (np.argmax) -------------------------------------
// key = (value-bits << 32) | (0xffffffff - index): a max over keys picks the largest value and,
// among equals, the smallest index.  Pixel values are non-negative so float bits order correctly.
template <typename T>
__global__ void k_argmax(const T* __restrict__ imgs, int64_t npx, unsigned long long* __restrict__ out) {
    const T* img = imgs + (size_t)blockIdx.y * npx;
    unsigned long long best = 0ULL;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < npx; i += (int64_t)gridDim.x * blockDim.x) {
        float v = (float)img[i];
        unsigned long long key = ((unsigned long long)__float_as_uint(v) << 32) | (unsigned long long)(0xffffffffu - (unsigned)i);
        best = key > best ? key : best;
    }
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
        best = other > best ? other : best;
    }
    if ((threadIdx.x & 31) == 0) atomicMax(out + blockIdx.y, best);
}

int launch_argmax(const void* imgs, int is_f32, int n_img, int64_t npx, long long* d_out, cudaStream_t st) {
    if (npx >= (1LL << 32)) { set_error("image too large for argmax key"); return SR459_EUNSUPPORTED; }
    SR_CUDA(cudaMemsetAsync(d_out, 0, sizeof(long long) * n_img, st));
    dim3 grid(296, n_img);
    if (is_f32) k_argmax<float><<<grid, 256, 0, st>>>((const float*)imgs, npx, (unsigned long long*)d_out);
    else        k_argmax<uint8_t><<<grid, 256, 0, st>>>((const uint8_t*)imgs, npx, (unsigned long long*)d_out);
    SR_LAUNCH_CHECK("k_argmax");
    return SR459_OK;
}

// patch[py,px] = sum over accepted images of img[pr-m+py, pc-m+px]   (double accumulation)
template <typename T>
__global__ void k_patch_sum(const T* __restrict__ imgs, int n_img, int H, int W, const int* __restrict__ peaks,
                            int margin, double* __restrict__ patch) {
    int side = 2 * margin + 1;
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= side * side) return;
    int py = t / side, px = t - py * side;
    double s = 0.0;
    for (int k = 0; k < n_img; ++k) {
        int pr = peaks[2 * k], pc = peaks[2 * k + 1];
        if (pr < 0) continue;
        s += (double)imgs[(size_t)k * H * W + (size_t)(pr - margin + py) * W + (pc - margin + px)];
    }
    patch[t] = s;
}

int launch_patch_sum(const void* imgs, int is_f32, int n_img, int H, int W, const int* d_peaks, int margin,
                     double* d_patch, cudaStream_t st) {
    int side = 2 * margin + 1;
    int nt = side * side;
    if (is_f32) k_patch_sum<float><<<(nt + 127) / 128, 128, 0, st>>>((const float*)imgs, n_img, H, W, d_peaks, margin, d_patch);
    else        k_patch_sum<uint8_t><<<(nt + 127) / 128, 128, 0, st>>>((const uint8_t*)imgs, n_img, H, W, d_peaks, margin, d_patch);
    SR_LAUNCH_CHECK("k_patch_sum");
    return SR459_OK;
}

}  // namespace sr459
