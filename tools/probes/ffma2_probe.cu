// FFMA2 issue-rate probe: register vs uniform-register (constant bank) multiplicand, 1 / 2 / 4 warps per SMSP.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_probe ffma2_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
struct P { u64 w[16]; u64* out; int iters; };
template <int MODE>
__global__ void k(const __grid_constant__ P p, const u64* __restrict__ gw) {
    u64 acc[16], x[16], w[16];
    for (int i = 0; i < 16; ++i) { acc[i] = 0; x[i] = (u64)(threadIdx.x + i) * 0x3f8000013f800001ull; w[i] = MODE == 0 ? gw[i] : 0; }
    for (int it = 0; it < p.iters; ++it) {
        #pragma unroll
        for (int m = 0; m < 16; ++m) {
            #pragma unroll
            for (int b = 0; b < 16; ++b) {
                if (MODE == 0) acc[b] = fma2(x[(b + m) & 15], w[m], acc[b]);
                else if (MODE == 1) acc[b] = fma2(x[(b + m) & 15], p.w[m], acc[b]);
                else { float lo = __uint_as_float((unsigned)acc[b]); lo = fmaf(__uint_as_float((unsigned)x[(b + m) & 15]), __uint_as_float((unsigned)p.w[m]), lo); acc[b] = (acc[b] & 0xffffffff00000000ull) | __float_as_uint(lo); }
            }
        }
    }
    u64 s = 0; for (int i = 0; i < 16; ++i) s ^= acc[i];
    p.out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    P p; for (int i = 0; i < 16; ++i) p.w[i] = 0x3f8000003f800000ull;
    cudaMalloc(&p.out, 148 * 1024 * 8); p.iters = 2000;
    u64* gw; cudaMalloc(&gw, 128); cudaMemcpy(gw, p.w, 128, cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int mode = 0; mode < 3; ++mode)
        for (int threads = 128; threads <= 1024; threads *= 2) {
            float best = 1e9;
            for (int rep = 0; rep < 3; ++rep) {
                cudaEventRecord(e0);
                if (mode == 0) k<0><<<148, threads>>>(p, gw); else if (mode == 1) k<1><<<148, threads>>>(p, gw); else k<2><<<148, threads>>>(p, gw);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
            }
            double inst = 148.0 * (threads / 32) * p.iters * 256.0;
            printf("mode %d (%s) warps/SMSP %d: %.3f ms, %.3f warp-instr/clk/SMSP at 1.965 GHz\n", mode,
                   mode == 0 ? "FFMA2 R,R,R" : mode == 1 ? "FFMA2 R,UR,R" : "FFMA R,UR", threads / 128, best,
                   inst / (148 * 4) / (best * 1e-3 * 1.965e9));
        }
    return 0;
}
