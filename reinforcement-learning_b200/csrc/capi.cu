// Process-wide pieces of the C ABI: last-error text, launch counter, version.
#include "common.cuh"
#include <stdarg.h>
#include <string.h>

namespace ggl {
static thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace ggl

extern "C" const char* ggl_last_error(void) { return ggl::g_err; }
extern "C" uint64_t ggl_launch_count(void) { return ggl::g_launches.load(); }
extern "C" const char* ggl_version(void) { return "gigaenv-b200 0.1 (sm_100a)"; }

// Host-side stand-in for the policy in benchmarks and examples: one uniformly random mask-valid action per player from
// host-resident mask rows (the same k-th-set-bit rule and hash as the device helper ggl_envset_random_actions).
namespace {
// 0x01 in every byte of w that is non-zero
inline uint64_t nonzero_bytes(uint64_t w) {
    w |= w >> 4; w |= w >> 2; w |= w >> 1;
    return w & 0x0101010101010101ull;
}
}  // namespace

extern "C" int ggl_host_random_actions(const uint8_t* h_masks, int32_t* h_actions, int num_players, int num_actions, uint64_t seed) {
    if (!h_masks || !h_actions || num_players < 0 || num_actions <= 0) { ggl::set_error("ggl_host_random_actions: bad argument"); return GGL_ERR_INVALID; }
    const int nw = num_actions / 8;
    for (int i = 0; i < num_players; i++) {
        uint64_t z = seed + 0x9E3779B97F4A7C15ull * (uint64_t)(i + 1);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
        const uint8_t* m = h_masks + (size_t)i * num_actions;
        // count the valid actions 8 mask bytes at a time
        uint64_t words[32];
        int cnt = 0;
        const int nwc = nw < 32 ? nw : 0;       // wide rows fall back to the byte loop
        for (int w = 0; w < nwc; w++) { uint64_t x; memcpy(&x, m + 8 * w, 8); words[w] = nonzero_bytes(x); cnt += (int)((words[w] * 0x0101010101010101ull) >> 56); }
        for (int a = 8 * nwc; a < num_actions; a++) cnt += m[a] != 0;
        int sel = 0;
        if (cnt > 0) {
            int k = (int)(z % (uint64_t)cnt);
            int a = 0;
            for (int w = 0; w < nwc; w++) {
                const int c = (int)((words[w] * 0x0101010101010101ull) >> 56);     // bytes are 0/1: their sum lands in the top byte
                if (k >= c) { k -= c; a += 8; continue; }
                uint64_t x = words[w];
                while (k-- > 0) x &= x - 1;                       // drop the k lowest set bytes
                sel = a + (__builtin_ctzll(x) >> 3);
                a = -1;
                break;
            }
            if (a >= 0)
                for (a = 8 * nwc; a < num_actions; a++)
                    if (m[a]) { if (k-- == 0) { sel = a; break; } }
        }
        h_actions[i] = sel;
    }
    return GGL_OK;
}

// Device-memory plumbing for host code that sees the GPU through this C ABI alone (the g++-only C++ mirrors under host/).
extern "C" int ggl_dev_malloc(int device, size_t bytes, void** out) {
    if (!out) { ggl::set_error("ggl_dev_malloc: null output"); return GGL_ERR_INVALID; }
    GGL_CUDA(cudaSetDevice(device));
    cudaError_t e = cudaMalloc(out, bytes ? bytes : 1);
    if (e != cudaSuccess) { ggl::set_error("ggl_dev_malloc(%zu bytes) -> %s", bytes, cudaGetErrorString(e)); return GGL_ERR_NOMEM; }
    return GGL_OK;
}
extern "C" int ggl_dev_free(void* p) {
    if (p) GGL_CUDA(cudaFree(p));
    return GGL_OK;
}
extern "C" int ggl_dev_upload(void* d_dst, const void* h_src, size_t bytes, void* stream) {
    GGL_CUDA(cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return GGL_OK;
}
extern "C" int ggl_dev_download(void* h_dst, const void* d_src, size_t bytes, void* stream) {
    GGL_CUDA(cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    GGL_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return GGL_OK;
}
