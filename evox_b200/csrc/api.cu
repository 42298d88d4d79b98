// api.cu - error plumbing, launch accounting, host-entry arena.
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

}  // namespace evx

using namespace evx;

extern "C" int evx_version(void) { return 100; }
extern "C" const char* evx_last_error(void) { return last_error_buf(); }
extern "C" int64_t evx_launch_count(void) { return launch_counter(); }
extern "C" void evx_launch_count_reset(void) { launch_counter() = 0; }
