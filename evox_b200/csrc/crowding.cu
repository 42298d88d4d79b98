// crowding.cu - crowding distance and its argsort.
//
// Replaces operators/selection/non_dominate.py:118-158 `crowding_distance(costs, mask)` and
// :161-180 `crowding_distance_sort`.  The reference runs, per objective, a stable two-key lexsort
// (valid rows first, then cost) over all n rows and three scatters (:146-155).  Only valid rows
// ever influence a valid row's distance, so here all M objectives are sorted in ONE segmented
// radix sort over 64-bit keys  (invalid | objective | order-preserving cost bits)  with the row
// index as payload: segment k of the sorted array is objective k's valid rows in ascending cost,
// ties in index order (radix sort is stable), NaN last, -0 == +0 - the order jnp.lexsort produces.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace evx {

struct CdScal { int nv; };

__global__ void k_cd_count(const uint8_t* __restrict__ mask, int64_t n, CdScal* sc) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool v = (i < n) && (mask == nullptr || mask[i] != 0);
    const uint32_t b = __ballot_sync(0xFFFFFFFFu, v);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(&sc->nv, __popc(b));
}

__global__ void k_cd_keys(const float* __restrict__ costs, const uint8_t* __restrict__ mask, int64_t n, int m,
                          int obj_bits, uint64_t* __restrict__ keys, int* __restrict__ vals) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * m) return;
    const int64_t i = t / m;
    const int k = (int)(t - i * m);
    const bool valid = (mask == nullptr) || mask[i] != 0;
    const uint64_t key = valid ? (((uint64_t)k << 32) | float_key_nan_last(costs[t]))
                               : ((uint64_t)1 << (32 + obj_bits));
    keys[(int64_t)k * n + i] = key;           // objective-major so equal keys keep index order
    vals[(int64_t)k * n + i] = (int)i;
}

// Sorted position q = k*nv + r  ->  per-objective distance of row vals[q].
__global__ void k_cd_scatter(const float* __restrict__ costs, const int* __restrict__ vals, int64_t n, int m,
                             const CdScal* __restrict__ sc, float* __restrict__ dk) {
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nv = sc->nv;
    if (q >= nv * m) return;
    const int k = (int)(q / nv);
    const int64_t r = q - (int64_t)k * nv;
    const int i = vals[q];
    float d;
    if (r == 0 || r == nv - 1) {
        d = __int_as_float(0x7F800000);                             // +inf (:152-153)
    } else {
        const float lo = costs[(int64_t)vals[(int64_t)k * nv] * m + k];
        const float hi = costs[(int64_t)vals[(int64_t)k * nv + nv - 1] * m + k];
        const float a = costs[(int64_t)vals[q + 1] * m + k], b = costs[(int64_t)vals[q - 1] * m + k];
        d = __fdiv_rn(__fsub_rn(a, b), __fsub_rn(hi, lo));          // (:149, :151) 0/0 -> NaN as in the reference
    }
    dk[(int64_t)k * n + i] = d;
}

__global__ void k_cd_sum(const float* __restrict__ dk, const uint8_t* __restrict__ mask, int64_t n, int m,
                         float* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (mask != nullptr && mask[i] == 0) { out[i] = __int_as_float(0xFF800000); return; }   // -inf (:155)
    float s = dk[i];
    for (int k = 1; k < m; ++k) s = __fadd_rn(s, dk[(int64_t)k * n + i]);                    // (:158)
    out[i] = s;
}

__global__ void k_cd_sortkeys(const float* __restrict__ d, int64_t n, uint32_t* __restrict__ keys, int* __restrict__ vals) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    keys[i] = float_key_nan_last(d[i]);
    vals[i] = (int)i;
}

struct CdLayout {
    size_t keysA, keysB, valsA, valsB, dk, dist, sc, tmp, tmp_bytes, total;
    int obj_bits;
};

static CdLayout cd_layout(int64_t n, int m) {
    CdLayout L{};
    size_t off = 0;
    auto take = [&](size_t b) { size_t o = off; off = align_up(off + b, 256); return o; };
    const size_t nm = (size_t)n * m;
    L.keysA = take(sizeof(uint64_t) * nm);
    L.keysB = take(sizeof(uint64_t) * nm);
    L.valsA = take(sizeof(int) * nm);
    L.valsB = take(sizeof(int) * nm);
    L.dk = take(sizeof(float) * nm);
    L.dist = take(sizeof(float) * (size_t)n);
    L.sc = take(sizeof(CdScal));
    size_t a = 0, b = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, a, (const uint64_t*)nullptr, (uint64_t*)nullptr, (const int*)nullptr,
                                    (int*)nullptr, (int)nm);
    cub::DeviceRadixSort::SortPairs(nullptr, b, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const int*)nullptr,
                                    (int*)nullptr, (int)n);
    L.tmp_bytes = a > b ? a : b;
    L.tmp = take(L.tmp_bytes);
    L.total = off;
    int ob = 0;
    while ((1 << ob) < m) ++ob;
    L.obj_bits = ob;
    return L;
}

// `nv_known`: device int holding the number of valid rows when the caller already has it (the fused
// survivor selection reads it off the rank histogram); nullptr = count the mask here.
static int crowding(const float* costs, const uint8_t* mask, int64_t n, int m, float* dist, char* ws,
                    const CdLayout& L, cudaStream_t s, const int* nv_known = nullptr) {
    const int T = 256;
    const int64_t nm = n * m;
    CdScal* sc = reinterpret_cast<CdScal*>(ws + L.sc);
    uint64_t* kA = reinterpret_cast<uint64_t*>(ws + L.keysA);
    uint64_t* kB = reinterpret_cast<uint64_t*>(ws + L.keysB);
    int* vA = reinterpret_cast<int*>(ws + L.valsA);
    int* vB = reinterpret_cast<int*>(ws + L.valsB);
    float* dk = reinterpret_cast<float*>(ws + L.dk);
    if (nv_known) {
        sc = reinterpret_cast<CdScal*>(const_cast<int*>(nv_known));
    } else {
        EVX_CUDA(cudaMemsetAsync(sc, 0, sizeof(CdScal), s));
        k_cd_count<<<(unsigned)ceil_div(ceil_div(n, 32) * 32, T), T, 0, s>>>(mask, n, sc);
        EVX_LAUNCH_CHECK();
    }
    k_cd_keys<<<(unsigned)ceil_div(nm, T), T, 0, s>>>(costs, mask, n, m, L.obj_bits, kA, vA);
    EVX_LAUNCH_CHECK();
    size_t tb = L.tmp_bytes;
    EVX_CUDA(cub::DeviceRadixSort::SortPairs(ws + L.tmp, tb, kA, kB, vA, vB, (int)nm, 0, 33 + L.obj_bits, s));
    k_cd_scatter<<<(unsigned)ceil_div(nm, T), T, 0, s>>>(costs, vB, n, m, sc, dk);
    EVX_LAUNCH_CHECK();
    k_cd_sum<<<(unsigned)ceil_div(n, T), T, 0, s>>>(dk, mask, n, m, dist);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}

static int crowding_sort(const float* costs, const uint8_t* mask, int64_t n, int m, int32_t* order, char* base,
                         cudaStream_t s, const int* nv_known) {
    const CdLayout L = cd_layout(n, m);
    float* dist = reinterpret_cast<float*>(base + L.dist);
    int rc = crowding(costs, mask, n, m, dist, base, L, s, nv_known);
    if (rc) return rc;
    uint32_t* kA = reinterpret_cast<uint32_t*>(base + L.keysA);
    uint32_t* kB = reinterpret_cast<uint32_t*>(base + L.keysB);
    int* vA = reinterpret_cast<int*>(base + L.valsA);
    k_cd_sortkeys<<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(dist, n, kA, vA);
    EVX_LAUNCH_CHECK();
    size_t tb = L.tmp_bytes;
    EVX_CUDA(cub::DeviceRadixSort::SortPairs(base + L.tmp, tb, kA, kB, vA, order, (int)n, 0, 32, s));
    return EVX_OK;
}

// For select.cu: crowding_distance_sort with the valid-row count already on the device.
int crowding_sort_known(const float* costs, const uint8_t* mask, const int* nv_dev, int64_t n, int m, int32_t* order,
                        void* ws, cudaStream_t s) {
    return crowding_sort(costs, mask, n, m, order, reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(ws), 256)), s,
                         nv_dev);
}

}  // namespace evx

using namespace evx;

extern "C" size_t evx_crowding_workspace_bytes(int64_t n, int32_t m) {
    if (n < 0 || m < 1 || m > EVX_MAX_OBJECTIVES || n * (int64_t)m > 0x7FFFFFFF) return 0;
    return cd_layout(n > 0 ? n : 1, m).total + 256;
}

static int cd_check(const char* who, const float* costs, int64_t n, int32_t m, const void* out, void* ws,
                    size_t ws_bytes) {
    EVX_REQUIRE(n >= 0 && m >= 1 && m <= EVX_MAX_OBJECTIVES && n * (int64_t)m <= 0x7FFFFFFF, EVX_ERR_ARG,
                "%s: bad shape n=%lld m=%d", who, (long long)n, m);
    if (n == 0) return EVX_OK;
    EVX_REQUIRE(costs && out, EVX_ERR_ARG, "%s: null pointer", who);
    EVX_REQUIRE(ws && ws_bytes >= evx_crowding_workspace_bytes(n, m), EVX_ERR_WORKSPACE,
                "%s: workspace %zu < required %zu", who, ws_bytes, evx_crowding_workspace_bytes(n, m));
    return EVX_OK;
}

extern "C" int evx_crowding_distance_f32(const float* costs, const uint8_t* mask, int64_t n, int32_t m, float* dist,
                                         void* ws, size_t ws_bytes, void* stream) {
    int rc = cd_check("evx_crowding_distance_f32", costs, n, m, dist, ws, ws_bytes);
    if (rc || n == 0) return rc;
    char* base = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(ws), 256));
    return crowding(costs, mask, n, m, dist, base, cd_layout(n, m), (cudaStream_t)stream);
}

extern "C" int evx_crowding_distance_sort_f32(const float* costs, const uint8_t* mask, int64_t n, int32_t m,
                                              int32_t* order, void* ws, size_t ws_bytes, void* stream) {
    int rc = cd_check("evx_crowding_distance_sort_f32", costs, n, m, order, ws, ws_bytes);
    if (rc || n == 0) return rc;
    char* base = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(ws), 256));
    return crowding_sort(costs, mask, n, m, order, base, (cudaStream_t)stream, nullptr);
}
