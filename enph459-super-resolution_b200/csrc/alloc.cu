// alloc.cu -- the one place device memory is allocated and released (plan scratch, tables, temporaries).
//
// Guard-band mode (sr459_debug_guard_bands(bytes), or SR459_GUARD=<bytes> in the environment when the library is
// loaded): every allocation is padded with `bytes` of a fixed pattern on both sides and remembered;
// sr459_debug_check_guards() reads all the bands back and reports the allocations whose bands were written to.
// compute-sanitizer is not available on the GPU pool this engine is developed on, so the parity suite is run once
// in this mode instead (tests/test_gpu_parity.py::test_guard_bands_*): an out-of-range WRITE trips a band, an
// out-of-range READ that matters pulls the pattern (0xC3C3C3C3 = -391.5f) into the result and fails parity.
#include "internal.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <map>
#include <mutex>
#include <vector>

namespace sr459 {

namespace {
struct Block { void* base; size_t bytes; size_t guard; int device; };
std::mutex g_mu;
std::map<void*, Block> g_blocks;                  // user pointer -> block (guard mode only)
size_t g_guard = []() -> size_t { const char* e = getenv("SR459_GUARD"); return e ? (size_t)atol(e) : 0; }();
constexpr unsigned char kPattern = 0xC3;
int g_bad_at_free = 0;                            // blocks released with a damaged band since the last check (guard mode)
char g_bad_msg[160] = "";

// both bands of one block, read back; true if any byte differs from the pattern (offset relative to the user pointer in *at)
bool bands_damaged(const Block& b, long* at) {
    const size_t user = (b.bytes + 255) & ~(size_t)255;
    std::vector<unsigned char> host(user + 2 * b.guard);
    if (cudaDeviceSynchronize() != cudaSuccess) return false;
    if (cudaMemcpy(host.data(), b.base, host.size(), cudaMemcpyDeviceToHost) != cudaSuccess) return false;
    for (size_t i = 0; i < b.guard; ++i) if (host[i] != kPattern) { *at = (long)i - (long)b.guard; return true; }
    for (size_t i = b.guard + b.bytes; i < host.size(); ++i) if (host[i] != kPattern) { *at = (long)(i - b.guard); return true; }
    return false;
}
}  // namespace

cudaError_t gmalloc(void** p, size_t bytes) {
    *p = nullptr;
    size_t g = g_guard;
    if (g == 0) return cudaMalloc(p, bytes ? bytes : 1);
    g = (g + 255) & ~(size_t)255;                 // keeps the user pointer 256-byte aligned like cudaMalloc's
    const size_t user = (bytes + 255) & ~(size_t)255;
    unsigned char* base = nullptr;
    cudaError_t e = cudaMalloc((void**)&base, user + 2 * g);
    if (e != cudaSuccess) return e;
    e = cudaMemset(base, kPattern, user + 2 * g);  // the tail between `bytes` and the band is patterned too
    if (e != cudaSuccess) { cudaFree(base); return e; }
    int dev = 0;
    cudaGetDevice(&dev);
    *p = base + g;
    std::lock_guard<std::mutex> lk(g_mu);
    g_blocks[*p] = Block{base, bytes, g, dev};
    return cudaSuccess;
}

void gfree(void* p) {
    if (!p) return;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        auto it = g_blocks.find(p);
        if (it != g_blocks.end()) {
            // entry points that free their scratch before returning (calib.cu, analysis.cu, the stand-alone operators) would
            // otherwise never have their bands looked at: check them on release (guard mode only: this synchronises the device)
            long at = 0;
            if (bands_damaged(it->second, &at)) {
                if (!g_bad_at_free) snprintf(g_bad_msg, sizeof(g_bad_msg), "guard band of a released %zu-byte allocation was written at byte offset %ld", it->second.bytes, at);
                ++g_bad_at_free;
            }
            p = it->second.base; g_blocks.erase(it);
        }
    }
    cudaFree(p);
}

struct DeviceGuardLite {
    int prev = -1;
    explicit DeviceGuardLite(int d) { cudaGetDevice(&prev); if (prev != d) cudaSetDevice(d); else prev = -1; }
    ~DeviceGuardLite() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace sr459

using namespace sr459;

extern "C" {

int sr459_debug_guard_bands(int bytes) {
    if (bytes < 0) { set_error("guard_bands: negative size"); return SR459_EINVAL; }
    std::lock_guard<std::mutex> lk(g_mu);
    g_guard = (size_t)bytes;
    return SR459_OK;
}

// returns the number of allocations whose guard bands were modified (0 = clean), or a negative error code;
// the first damaged one is described in sr459_last_error()
int sr459_debug_check_guards(void) {
    std::vector<std::pair<void*, Block>> blocks;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        blocks.assign(g_blocks.begin(), g_blocks.end());
    }
    int prev = 0, bad = 0;
    cudaGetDevice(&prev);
    SR_CUDA(cudaDeviceSynchronize());
    std::vector<unsigned char> host;
    for (auto& kv : blocks) {
        const Block& b = kv.second;
        cudaSetDevice(b.device);
        const size_t user = (b.bytes + 255) & ~(size_t)255;
        host.resize(user + 2 * b.guard);
        cudaError_t e = cudaMemcpy(host.data(), b.base, host.size(), cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { cudaSetDevice(prev); set_error("check_guards: %s", cudaGetErrorString(e)); return SR459_ECUDA; }
        bool hit = false; long at = 0;                                     // offset relative to the user pointer
        for (size_t i = 0; i < b.guard && !hit; ++i) if (host[i] != kPattern) { hit = true; at = (long)i - (long)b.guard; }
        for (size_t i = b.guard + b.bytes; i < host.size() && !hit; ++i) if (host[i] != kPattern) { hit = true; at = (long)(i - b.guard); }
        if (hit) {
            if (!bad) set_error("guard band of a %zu-byte allocation was written at byte offset %ld", b.bytes, at);
            ++bad;
        }
    }
    cudaSetDevice(prev);
    {
        std::lock_guard<std::mutex> lk(g_mu);
        if (g_bad_at_free) { if (!bad) set_error("%s", g_bad_msg); bad += g_bad_at_free; g_bad_at_free = 0; }
    }
    return bad;
}

// self-test of the checker: damage (restore == 0) or repair (restore != 0) one byte just past the end of the oldest
// live guarded allocation
int sr459_debug_poke_guard(int restore) {
    Block b{};
    {
        std::lock_guard<std::mutex> lk(g_mu);
        if (g_blocks.empty()) { set_error("poke_guard: no guarded allocation is live"); return SR459_EINVAL; }
        b = g_blocks.begin()->second;
    }
    DeviceGuardLite dg(b.device);
    SR_CUDA(cudaMemset((unsigned char*)b.base + b.guard + b.bytes, restore ? kPattern : 0x00, 1));
    SR_CUDA(cudaDeviceSynchronize());
    return SR459_OK;
}

int sr459_debug_live_allocations(void) {
    std::lock_guard<std::mutex> lk(g_mu);
    return (int)g_blocks.size();
}

}  // extern "C"
