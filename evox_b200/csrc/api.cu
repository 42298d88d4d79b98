// api.cu - error plumbing, launch accounting, host-entry arena, issue-rate probe.
#include <mutex>

#include "common.cuh"

namespace evx {

static thread_local char g_err[512] = "";
static thread_local int64_t g_launches = 0;

char* last_error_buf() { return g_err; }
int64_t& launch_counter() { return g_launches; }

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// Grow-only arena, one per device, for the entry points that take host buffers.
static std::recursive_mutex g_arena_mu;
std::recursive_mutex& arena_mutex() { return g_arena_mu; }
static void* g_arena_ptr[64] = {nullptr};
static size_t g_arena_bytes[64] = {0};

int arena_get(size_t bytes, void** ptr) {
    int dev = 0;
    EVX_CUDA(cudaGetDevice(&dev));
    EVX_REQUIRE(dev >= 0 && dev < 64, EVX_ERR_ARG, "arena: device index %d out of range", dev);
    std::lock_guard<std::recursive_mutex> lock(g_arena_mu);
    if (g_arena_bytes[dev] < bytes) {
        if (g_arena_ptr[dev]) {
            EVX_CUDA(cudaDeviceSynchronize());
            EVX_CUDA(cudaFree(g_arena_ptr[dev]));
            g_arena_ptr[dev] = nullptr;
            g_arena_bytes[dev] = 0;
        }
        const size_t want = bytes + bytes / 8;
        cudaError_t e = cudaMalloc(&g_arena_ptr[dev], want);
        if (e != cudaSuccess) {
            cudaGetLastError();
            set_error("arena: cudaMalloc(%zu bytes) failed: %s", want, cudaGetErrorString(e));
            return EVX_ERR_NOMEM;
        }
        g_arena_bytes[dev] = want;
    }
    *ptr = g_arena_ptr[dev];
    return EVX_OK;
}

// Register-only probe of the tile kernel's inner step: per 32 pair tests a warp issues (M-1)
// FSETP and one predicated bit-set.  No memory traffic; values come from the thread index so the
// compiler cannot fold the loop.  pair tests = threads * iters * 32 (columns per word) * RB.
template <int MM>
__global__ void k_probe(float seed, int64_t iters, uint32_t* sink) {
    float xi[4][MM];
    float cv[32][MM > 0 ? MM : 1];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int k = 0; k < MM; ++k) xi[q][k] = seed * (float)(threadIdx.x + 3 * q + k);
    uint32_t acc = 0;
    float drift = seed;
    for (int64_t it = 0; it < iters; ++it) {
        uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
        for (int b = 0; b < 32; ++b) {
#pragma unroll
            for (int k = 0; k < MM; ++k) cv[b][k] = drift + (float)(b * 7 + k);
        }
#pragma unroll
        for (int b = 0; b < 32; ++b) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                bool p = true;
#pragma unroll
                for (int k = 0; k < MM; ++k) p = p && (xi[q][k] <= cv[b][k]);
                if (p) w[q] |= 1u << b;
            }
        }
        acc ^= w[0] ^ (w[1] << 1) ^ (w[2] >> 1) ^ (w[3] << 3);
        drift += 1.0f;
    }
    if (acc == 0x12345678u) sink[0] = acc;
}

}  // namespace evx

using namespace evx;

extern "C" int evx_version(void) { return 100; }
extern "C" const char* evx_last_error(void) { return last_error_buf(); }
extern "C" int64_t evx_launch_count(void) { return launch_counter(); }
extern "C" void evx_launch_count_reset(void) { launch_counter() = 0; }

extern "C" int evx_probe_compare_rate(int32_t m, int64_t iters, double* pairs_per_s, void* stream) {
    EVX_REQUIRE(pairs_per_s && iters > 0, EVX_ERR_ARG, "evx_probe_compare_rate: bad argument");
    EVX_REQUIRE(m == 3 || m == 10, EVX_ERR_ARG, "evx_probe_compare_rate: m must be 3 or 10");
    cudaStream_t s = (cudaStream_t)stream;
    int dev = 0, sms = 0;
    EVX_CUDA(cudaGetDevice(&dev));
    EVX_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    std::lock_guard<std::recursive_mutex> hold(arena_mutex());
    uint32_t* sink = nullptr;
    int rc = arena_get(256, reinterpret_cast<void**>(&sink));
    if (rc) return rc;
    cudaEvent_t e0, e1;
    EVX_CUDA(cudaEventCreate(&e0));
    EVX_CUDA(cudaEventCreate(&e1));
    const int blocks = sms * 8, threads = 256;
    for (int rep = 0; rep < 2; ++rep) {
        EVX_CUDA(cudaEventRecord(e0, s));
        if (m == 3) k_probe<2><<<blocks, threads, 0, s>>>(1.0f, iters, sink);
        else k_probe<9><<<blocks, threads, 0, s>>>(1.0f, iters, sink);
        EVX_LAUNCH_CHECK();
        EVX_CUDA(cudaEventRecord(e1, s));
        EVX_CUDA(cudaEventSynchronize(e1));
    }
    float ms = 0.f;
    EVX_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *pairs_per_s = (double)blocks * threads * (double)iters * 32.0 * 4.0 / (ms * 1e-3);
    return EVX_OK;
}
