// dev probe: which TMA box / tensor configurations work on this driver (run on the GPU box)
#include "../enph459-super-resolution_b200/csrc/tma.h"
#include <cstdio>
#include <vector>
#include <mutex>
namespace sr459 { void set_error(const char* fmt, ...) { } }
using namespace sr459;
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn getfn() {
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    printf("entry point: err=%d q=%d p=%p\n", (int)e, (int)q, p);
    return (EncodeTiledFn)p;
}
struct alignas(64) Args { CUtensorMap tm; float* out; int c0, c1, n; };
__global__ void k(const __grid_constant__ Args a) {
    extern __shared__ __align__(128) float sm[];
    __shared__ __align__(8) uint64_t mbar;
    if (threadIdx.x == 0) {
        mbar_init(&mbar, 1); mbar_fence_init();
        mbar_expect_tx(&mbar, a.n * 4);
        tma_load_2d(sm, &a.tm, a.c0, a.c1, &mbar);
    }
    __syncthreads();
    mbar_wait(&mbar, 0);
    for (int i = threadIdx.x; i < a.n; i += blockDim.x) a.out[i] = sm[i];
}
int main() {
    EncodeTiledFn fn = getfn();
    const int W = 200, H = 100;
    std::vector<float> h(W * H);
    for (int i = 0; i < W * H; ++i) h[i] = (float)i;
    float *d, *o; cudaMalloc(&d, W * H * 4); cudaMalloc(&o, 256 * 256 * 4);
    cudaMemcpy(d, h.data(), W * H * 4, cudaMemcpyHostToDevice);
    struct Case { int bc, br, c0, c1; } cases[] = {{64, 16, 0, 0}, {64, 16, 8, 4}, {64, 16, -8, -4}, {144, 80, -8, -8}, {144, 80, 100, 60}, {68, 36, -2, -2}, {256, 80, 0, 0}};
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (auto cs : cases) {
        Args a{};
        cuuint64_t gdim[2] = {W, H}; cuuint64_t gstr[1] = {W * 4}; cuuint32_t box[2] = {(cuuint32_t)cs.bc, (cuuint32_t)cs.br}; cuuint32_t es[2] = {1, 1};
        CUresult r = fn(&a.tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, d, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        a.out = o; a.c0 = cs.c0; a.c1 = cs.c1; a.n = cs.bc * cs.br;
        k<<<1, 128, a.n * 4>>>(a);
        cudaError_t e = cudaDeviceSynchronize();
        std::vector<float> res(a.n);
        cudaMemcpy(res.data(), o, a.n * 4, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int y = 0; y < cs.br; ++y) for (int x = 0; x < cs.bc; ++x) {
            int gy = cs.c1 + y, gx = cs.c0 + x;
            float want = (gy >= 0 && gy < H && gx >= 0 && gx < W) ? h[gy * W + gx] : 0.f;
            if (res[y * cs.bc + x] != want) ++bad;
        }
        printf("box %dx%d at (%d,%d): encode=%d run=%s bad=%d\n", cs.bc, cs.br, cs.c0, cs.c1, (int)r, cudaGetErrorString(e), bad);
        if (e != cudaSuccess) return 1;
    }
    return 0;
}
