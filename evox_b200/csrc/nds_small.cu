// nds_small.cu - non-dominated sort for small populations with NO device allocation.
//
// Who needs this: code that must rank a population from a place where it cannot obtain device
// scratch memory - a custom-call handler whose host framework gives it only input and output
// buffers, a callback running while the framework's allocator is locked, or a process that has
// handed all free device memory to another pool.  evx_nds_f32 needs a caller-supplied workspace and
// evx_nds_host_f32 allocates an arena; this entry point needs neither.  Its scratch (the n x n
// dominance bit matrix and the dominator counts) lives in a fixed array that is part of the library
// image on every GPU that loads it (about 2.1 MB, held until the process exits), so the population
// is limited to EVX_NDS_NOALLOC_MAX_N rows.
//
// There is no counterpart in the reference: its non_dominated_sort always materialises fresh
// arrays through XLA.  The ranks are the same as evx_nds_f32's.
#include <mutex>

#include "common.cuh"

namespace evx {

constexpr int SMALL_MAX_N = EVX_NDS_NOALLOC_MAX_N;
constexpr int SMALL_WORDS = SMALL_MAX_N / 32;

struct SmallScratch {
    uint32_t bits[(size_t)SMALL_MAX_N * SMALL_WORDS];   // bit j%32 of bits[i][j/32]: row i dominates row j
    int count[SMALL_MAX_N];                             // dominators of row j not yet ranked
};

// File-scope device array.  In host code this NAME is only a handle: the device address has to be
// asked for with cudaGetSymbolAddress (done once per device below) and that pointer is what the
// kernels receive.
static __device__ __align__(256) unsigned char g_small_scratch[sizeof(SmallScratch)];

static std::mutex g_small_mu;
static SmallScratch* g_small_ptr[64] = {nullptr};

// Once per device, with that device current: resolve the scratch array's device address.
static int small_scratch(SmallScratch** out) {
    int dev = 0;
    EVX_CUDA(cudaGetDevice(&dev));
    EVX_REQUIRE(dev >= 0 && dev < 64, EVX_ERR_ARG, "evx_nds_noalloc_f32: device index %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_small_mu);
    if (!g_small_ptr[dev]) {
        void* p = nullptr;
        EVX_CUDA(cudaGetSymbolAddress(&p, g_small_scratch));
        EVX_REQUIRE(p != nullptr, EVX_ERR_CUDA, "evx_nds_noalloc_f32: scratch symbol has no device address");
        g_small_ptr[dev] = reinterpret_cast<SmallScratch*>(p);
    }
    *out = g_small_ptr[dev];
    return EVX_OK;
}

// The counts are the only part of the scratch that must start at zero: every bit-matrix word that
// is read later is written first by k_small_bits.
__global__ void k_small_clear(int* __restrict__ count, int n) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) count[j] = 0;
}

// Lane <-> column j, warp <-> 8 rows: one ballot per (row, 32 columns).  Full definition of
// dominance (utils/common.py:88-91), no pre-sort: n is small.
__global__ void __launch_bounds__(256)
k_small_bits(const float* __restrict__ x, int n, int m, int words, uint32_t* __restrict__ bits, int* __restrict__ count) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int j = blockIdx.x * 32 + lane;
    const int r0 = blockIdx.y * 64 + warp * 8;
    const bool jv = j < n;
    int cnt = 0;
    for (int r = r0; r < r0 + 8 && r < n; ++r) {
        bool le = jv, lt = false;
        for (int k = 0; k < m && jv; ++k) {
            const float a = __ldg(x + (size_t)r * m + k), b = __ldg(x + (size_t)j * m + k);
            le = le && (a <= b);
            lt = lt || (a < b);
        }
        const bool d = le && lt;
        const uint32_t w = __ballot_sync(0xFFFFFFFFu, d);
        if (lane == 0) bits[(size_t)r * words + blockIdx.x] = w;
        cnt += d;
    }
    if (cnt) atomicAdd(&count[j], cnt);
}

// One CTA peels every front: no host round trip, no grid-wide barrier.
__global__ void __launch_bounds__(1024)
k_small_peel(const uint32_t* __restrict__ bits, int* __restrict__ count, int n, int words, int* __restrict__ rank,
             int* __restrict__ fronts_out) {
    __shared__ int front[SMALL_MAX_N];
    __shared__ int nf;
    constexpr int PER = SMALL_MAX_N / 1024;
    for (int k = 0;; ++k) {
        if (threadIdx.x == 0) nf = 0;
        __syncthreads();
#pragma unroll
        for (int t = 0; t < PER; ++t) {
            const int j = threadIdx.x + t * 1024;
            if (j < n && count[j] == 0) {
                rank[j] = k;
                count[j] = -1;
                front[atomicAdd(&nf, 1)] = j;
            }
        }
        __syncthreads();
        const int f = nf;
        if (f == 0) {
            if (threadIdx.x == 0 && fronts_out) *fronts_out = k;
            break;
        }
#pragma unroll
        for (int t = 0; t < PER; ++t) {
            const int j = threadIdx.x + t * 1024;
            if (j < n && count[j] > 0) {
                int dec = 0;
                for (int q = 0; q < f; ++q) dec += (bits[(size_t)front[q] * words + (j >> 5)] >> (j & 31)) & 1u;
                count[j] -= dec;
            }
        }
        __syncthreads();
    }
}

// Shared by evx_nds_noalloc_f32 (built-in scratch) and by evx_nds_f32 for populations of at most one
// panel (scratch inside the caller's workspace): three launches, no host round trip.
int small_nds_launch(const float* x, int n, int m, int* rank, uint32_t* bits, int words, int* count, int* fronts_out,
                     cudaStream_t s) {
    k_small_clear<<<(n + 255) / 256, 256, 0, s>>>(count, n);
    EVX_LAUNCH_CHECK();
    dim3 grid((n + 31) / 32, (n + 63) / 64);
    k_small_bits<<<grid, 256, 0, s>>>(x, n, m, words, bits, count);
    EVX_LAUNCH_CHECK();
    k_small_peel<<<1, 1024, 0, s>>>(bits, count, n, words, rank, fronts_out);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}

}  // namespace evx

using namespace evx;

extern "C" int evx_nds_noalloc_f32(const float* x, int64_t n, int32_t m, int32_t* rank, void* stream) {
    EVX_REQUIRE(n >= 0 && m >= 1 && m <= EVX_MAX_OBJECTIVES, EVX_ERR_ARG, "evx_nds_noalloc_f32: bad shape n=%lld m=%d",
                (long long)n, m);
    EVX_REQUIRE(n <= SMALL_MAX_N, EVX_ERR_ARG,
                "evx_nds_noalloc_f32: n=%lld exceeds the built-in scratch (at most %d rows); use evx_nds_f32 with a workspace",
                (long long)n, SMALL_MAX_N);
    if (n == 0) return EVX_OK;
    EVX_REQUIRE(x && rank, EVX_ERR_ARG, "evx_nds_noalloc_f32: null pointer");
    SmallScratch* sc = nullptr;
    int rc = small_scratch(&sc);
    if (rc) return rc;
    return small_nds_launch(x, (int)n, m, rank, sc->bits, SMALL_WORDS, sc->count, nullptr, (cudaStream_t)stream);
}
