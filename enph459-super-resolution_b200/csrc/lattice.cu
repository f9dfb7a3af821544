// lattice.cu -- what every fused separable-PSF iteration kernel (ibp_ws.cu, ibp_stream.cu, ibp_stream_ml.cu) shares:
// the TMA tensor-map encoder and the once-per-call pre-summation of co-sited frames.
//
// The mathematics is the body of ibp() (mono_cal_target/run_sr.py:190-209) in the closed form of
// ibp_dense.cu.  Three exact restructurings make one iteration cost ~9 bytes and ~11 FMA per HR pixel:
//
//  (1) polyphase lattices.  With factor 2 and |HR shift| <= 1, frame k samples the blurred image
//      B = conv0(hr, psf) only at the sites (y, x) = (2i - sy_k, 2j - sx_k), one of the four
//      parity lattices (P, Q) = (sy_k & 1, sx_k & 1), and its zero-inserted, back-shifted residual
//      (:170-176) lands on the same sites.  The nominal +-0.5 px corner shifts (quirk Q3) put all
//      four frames on lattice (1, 1).
//  (2) co-sited frames are pre-summed.  All frames of a lattice see the same B(site), so
//          sum_k (lr_k - B)   = S1 - m B                    S1 = sum_k lr_k(site), m = #frames
//          sum_k (lr_k - B)^2 = V + (S1 - m B)^2 / m        V  = sum_k (lr_k - S1/m)^2
//      S1 (one h x w plane per active lattice) and the scalar sum of V are constant over the
//      iterations: k_lattice_sums computes them once per ibp() call, and an iteration then reads one
//      S1 plane per lattice instead of n LR frames (1 instead of n bytes per HR pixel).
//  (3) the Gaussian PSF of make_gaussian_psf (:104-111) is exactly rank 1, psf = acol (x) brow, so
//      both blurs run as a column pass and a row pass that only touch lattice rows / columns.
//
// (The first two kernels built on this -- a four-phase CTA-wide one and a persistent pipelined one, plan paths 5 and 6 -- were
// superseded by k_ibp_ws and k_ibp_stream and have been removed.)
#include "internal.h"
#include "tma.h"

#include <algorithm>
#include <mutex>

namespace sr459 {

bool ibp_sep_supported(int kh, int kw, int smax, bool separable, int W) {
    return separable && kh == kw && (kh == 1 || kh == 3 || kh == 5 || kh == 7) && smax <= 1 && (W % 4) == 0;
}

// ---- tensor maps ---------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int encode_tile_map_f32(CUtensorMap* out, const void* base, uint64_t cols, uint64_t rows, uint64_t pitch_elems,
                        uint32_t box_c, uint32_t box_r) {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    if (!fn) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return SR459_ECUDA; }
    if ((reinterpret_cast<uintptr_t>(base) & 15) || ((pitch_elems * 4) & 15)) {
        set_error("TMA needs a 16-byte aligned base and row pitch (base %p, pitch %llu floats)", base, (unsigned long long)pitch_elems);
        return SR459_EINVAL;
    }
    cuuint64_t gdim[2] = {cols, rows};
    cuuint64_t gstr[1] = {pitch_elems * 4};
    cuuint32_t box[2] = {box_c, box_r};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return SR459_ECUDA; }
    return SR459_OK;
}

constexpr int LS_ROWS = 8;                             // plane rows per CTA: 8 x n independent loads per thread, one atomic per CTA
// NF = compile-time bound on the frames of a lattice (>= every slot_n): the per-frame code is fully unrolled and predicated,
// so a 4-frame set must not pay for kMaxFrames slots
template <int NF>
__global__ void __launch_bounds__(256) k_lattice_sums(const __grid_constant__ LatticeArgs a) {
    const int slot = blockIdx.z;
    const int b = blockIdx.x * 256 + threadIdx.x;
    double vd = 0.0;
    if (b < a.w) {
        #pragma unroll (NF <= 4 ? 4 : 2)
        for (int rr = 0; rr < LS_ROWS; ++rr) {
            const int ar = blockIdx.y * LS_ROWS + rr;      // plane row; site row index a = a.row0 + ar
            if (ar >= a.rows) break;
            const int ag = a.row0 + ar;
            float x[NF];
            int m = 0;
            float s = 0.0f;
            #pragma unroll
            for (int f = 0; f < NF; ++f) {
                x[f] = 0.0f;
                if (f < a.slot_n[slot]) {
                    const int k = a.slot_k[slot][f];
                    const int i = ag + a.slot_oy[slot][f], j = b + a.slot_ox[slot][f];
                    if (i >= a.lr_lo && i < a.lr_hi && i < a.h && j < a.w) {
                        x[f] = __ldg(a.lr[k] + (size_t)(i - a.lr_lo) * a.w + j);
                        m |= 1 << f;
                    }
                }
            }
            #pragma unroll
            for (int f = 0; f < NF; ++f) s += x[f];      // frame order, as before (absent frames add 0)
            a.s1[slot][(size_t)ar * a.pitch + b] = s;
            const int cnt = __popc(m);
            if (cnt > 0 && ag >= a.own_a0 && ag < a.own_a1) {
                const float mu = s / (float)cnt;
                float v = 0.0f;
                #pragma unroll
                for (int f = 0; f < NF; ++f)
                    if (m & (1 << f)) { const float d = x[f] - mu; v = fmaf(d, d, v); }
                vd += (double)v;                               // per-site f32, across sites f64 (as before)
            }
        }
    }
    __shared__ double red[8];
    double d = vd;
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = d;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        #pragma unroll
        for (int i = 0; i < 8; ++i) t += red[i];
        if (t != 0.0) atomicAdd(a.vsum, t);
    }
}

int launch_lattice_sums(const LatticeArgs& a, cudaStream_t st) {
    if (a.nslot <= 0 || a.rows <= 0) return SR459_OK;
    SR_CUDA(cudaMemsetAsync(a.vsum, 0, sizeof(double), st));
    dim3 grid((a.w + 255) / 256, (a.rows + LS_ROWS - 1) / LS_ROWS, a.nslot);
    int nmax = 1;
    for (int sl = 0; sl < a.nslot; ++sl) nmax = std::max(nmax, a.slot_n[sl]);
    if (nmax <= 1)      k_lattice_sums<1><<<grid, 256, 0, st>>>(a);
    else if (nmax <= 2) k_lattice_sums<2><<<grid, 256, 0, st>>>(a);
    else if (nmax <= 4) k_lattice_sums<4><<<grid, 256, 0, st>>>(a);
    else                k_lattice_sums<kMaxFrames><<<grid, 256, 0, st>>>(a);
    SR_LAUNCH_CHECK("k_lattice_sums");
    return SR459_OK;
}

}  // namespace sr459
