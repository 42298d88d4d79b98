// common.cuh - shared helpers for libevoxmo (error plumbing, launch accounting, float keys).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include <mutex>

#include "../../include/evoxmo.h"

namespace evx {

// ---- error plumbing: thread-local message, integer status, nothing thrown across the ABI ------
char* last_error_buf();
void set_error(const char* fmt, ...);
int64_t& launch_counter();

#define EVX_CUDA(call)                                                                         \
    do {                                                                                       \
        cudaError_t e__ = (call);                                                              \
        if (e__ != cudaSuccess) {                                                              \
            evx::set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__,                \
                           cudaGetErrorString(e__));                                           \
            return EVX_ERR_CUDA;                                                               \
        }                                                                                      \
    } while (0)

#define EVX_LAUNCH_CHECK()                                                                     \
    do {                                                                                       \
        ++evx::launch_counter();                                                               \
        cudaError_t e__ = cudaGetLastError();                                                  \
        if (e__ != cudaSuccess) {                                                              \
            evx::set_error("kernel launch failed at %s:%d: %s", __FILE__, __LINE__,            \
                           cudaGetErrorString(e__));                                           \
            return EVX_ERR_CUDA;                                                               \
        }                                                                                      \
    } while (0)

#define EVX_REQUIRE(cond, code, ...)                                                           \
    do {                                                                                       \
        if (!(cond)) {                                                                         \
            evx::set_error(__VA_ARGS__);                                                       \
            return (code);                                                                     \
        }                                                                                      \
    } while (0)

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t a, size_t b) { return (a + b - 1) / b * b; }

// Order-preserving float32 -> uint32 key.  -0.0 is folded onto +0.0 first (IEEE compares treat
// them as equal), so key equality == float equality and key order == float order for non-NaN.
__host__ __device__ __forceinline__ uint32_t float_key(float v) {
    v = v + 0.0f;
#ifdef __CUDA_ARCH__
    uint32_t b = __float_as_uint(v);
#else
    union { float f; uint32_t u; } c; c.f = v; uint32_t b = c.u;
#endif
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
// Total-order key used by the sorts that must mirror jnp.sort: NaN (any sign) goes last.
__device__ __forceinline__ uint32_t float_key_nan_last(float v) {
    return (v != v) ? 0xFFFFFFFFu : float_key(v);
}

// Small-population non-dominated sort (nds_small.cu): n <= 4096 rows, scratch = n*words uint32 + n int.
int small_nds_launch(const float* x, int n, int m, int* rank, uint32_t* bits, int words, int* count, int* fronts_out,
                     cudaStream_t s);

// Grow-only device arena for the *_host_* entry points (one per device).  A host entry point holds
// arena_mutex() from arena_get until its last copy has finished: two host threads calling *_host_*
// functions are serialised instead of sharing (or freeing) each other's scratch.
int arena_get(size_t bytes, void** ptr);
std::recursive_mutex& arena_mutex();

}  // namespace evx
