// select.cu - fused survivor selection.
//
// Replaces operators/selection/non_dominate.py:210-222 - what `non_dominate(population, fitness, topk,
// deduplicate)` does after the optional de-duplication (:188-208, which stays host glue):
//
//     rank   = non_dominated_sort(fitness)                       (:211)   evx_nds_f32, same stream
//     order  = argsort(rank); worst = rank[order[topk-1]]         (:213-214)
//     cd     = crowding_distance_sort(fitness, rank == worst)     (:215-216) an argsort PERMUTATION
//     chosen = lexsort((-cd, rank))[:topk]                        (:218)  secondary key = -cd[i], i.e.
//                                                                          entry i of that permutation
//     return population[chosen], fitness[chosen]                  (:219-222)
//
// The reference runs three full-length sorts plus the crowding sorts for this.  Here: `worst` comes
// from a rank histogram (no sort), the crowding permutation from the library's masked crowding sort
// with the valid-row count already known from the histogram, and the final order from ONE radix sort of
// the packed key (rank, n-1-cd[i]) - rows past the cut rank get the largest key - followed by one
// gather kernel.  The :216/:218 quirk (the permutation, not the distance, is the tie-breaker) is kept
// bit for bit: it decides both which rows of the cut front survive and the order of the output rows.
// Everything is enqueued on the caller's stream; nothing synchronises.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace evx {

int crowding_sort_known(const float* costs, const uint8_t* mask, const int* nv_dev, int64_t n, int m, int32_t* order,
                        void* ws, cudaStream_t s);

struct SelScal {
    int cut;        // rows with rank == worst: the valid-row count of the crowding step (must stay first)
    int worst;      // rank of the topk-th row in rank order
    int below;      // rows with rank < worst
    int maxrank;
};

// Rank histogram.  Ranks repeat heavily inside a warp (a converged population is one front), so equal
// ranks are merged with match_any before they reach the global counter.
__global__ void __launch_bounds__(256)
k_sel_hist(const int* __restrict__ rank, int64_t n, int* __restrict__ hist, SelScal* sc) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const bool ok = i < n;
    const int r = ok ? rank[i] : -1;
    const uint32_t peers = __match_any_sync(0xFFFFFFFFu, r);
    if (ok && lane == __ffs(peers) - 1) atomicAdd(&hist[r], __popc(peers));
    int mx = r;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) mx = max(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, o));
    if (lane == 0 && mx >= 0) atomicMax(&sc->maxrank, mx);
}

// One CTA: first rank whose cumulative count reaches topk (= rank[argsort(rank)[topk-1]]).
__global__ void __launch_bounds__(1024)
k_sel_cut(const int* __restrict__ hist, int64_t topk, SelScal* sc) {
    __shared__ int wsum[32];
    __shared__ int found;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nr = sc->maxrank + 1;
    if (threadIdx.x == 0) found = 0;
    __syncthreads();
    int64_t base = 0;
    for (int r0 = 0; r0 < nr; r0 += 1024) {
        const int r = r0 + threadIdx.x;
        const int c = (r < nr) ? hist[r] : 0;
        int incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        int before = 0;
        for (int w = 0; w < warp; ++w) before += wsum[w];
        int total = 0;
        for (int w = 0; w < 32; ++w) total += wsum[w];
        const int64_t cum = base + before + incl;                   // rows with rank <= r
        if (r < nr && cum >= topk && cum - c < topk) {              // exactly one r satisfies this
            sc->worst = r;
            sc->below = (int)(cum - c);
            sc->cut = c;
            found = 1;
        }
        __syncthreads();
        if (found) return;
        base += total;
    }
}

__global__ void k_sel_mask(const int* __restrict__ rank, int64_t n, const SelScal* __restrict__ sc, uint8_t* __restrict__ mask) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) mask[i] = rank[i] == sc->worst;
}

// key = (rank, n-1-cd[i]) for rows up to the cut rank; rows past it sort to the end.
__global__ void k_sel_keys(const int* __restrict__ rank, const int* __restrict__ cd, int64_t n, int nb,
                           const SelScal* __restrict__ sc, uint64_t* __restrict__ keys, int* __restrict__ vals) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int r = rank[i];
    keys[i] = (r > sc->worst) ? ~0ull >> (64 - 2 * nb) : (((uint64_t)r << nb) | (uint64_t)(n - 1 - cd[i]));
    vals[i] = (int)i;
}

// One thread per output element: out_pop[q, :] = pop[chosen[q], :], out_fit[q, :] = fit_src[chosen[q], :].
__global__ void k_sel_gather(const int* __restrict__ chosen, int64_t topk, const float* __restrict__ pop, int d,
                             const float* __restrict__ fit, int m, float* __restrict__ out_pop,
                             float* __restrict__ out_fit, int* __restrict__ out_index) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int w = d + m;
    if (t >= topk * w) return;
    const int64_t q = t / w;
    const int c = (int)(t - q * w);
    const int64_t src = chosen[q];
    if (c < d) {
        if (out_pop) out_pop[q * d + c] = pop[src * d + c];
    } else if (out_fit) {
        out_fit[q * m + (c - d)] = fit[src * m + (c - d)];
    }
    if (c == 0 && out_index) out_index[q] = (int)src;
}

struct SelLayout {
    size_t rank, hist, sc, mask, cd, keysA, keysB, valsA, valsB, tmp, tmp_bytes, inner, inner_bytes, total;
    int nb;
};

static SelLayout sel_layout(int64_t n, int m) {
    SelLayout L{};
    size_t off = 0;
    auto take = [&](size_t b) { size_t o = off; off = align_up(off + b, 256); return o; };
    L.rank = take(sizeof(int) * (size_t)n);
    L.hist = take(sizeof(int) * (size_t)(n + 1));
    L.sc = take(sizeof(SelScal));
    L.mask = take((size_t)n);
    L.cd = take(sizeof(int) * (size_t)n);
    L.keysA = take(sizeof(uint64_t) * (size_t)n);
    L.keysB = take(sizeof(uint64_t) * (size_t)n);
    L.valsA = take(sizeof(int) * (size_t)n);
    L.valsB = take(sizeof(int) * (size_t)n);
    cub::DeviceRadixSort::SortPairs(nullptr, L.tmp_bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr, (const int*)nullptr,
                                    (int*)nullptr, (int)n);
    L.tmp = take(L.tmp_bytes);
    // the non-dominated sort and the crowding sort run one after the other: they share one region
    const size_t a = evx_nds_workspace_bytes(n, m), b = evx_crowding_workspace_bytes(n, m);
    L.inner_bytes = a > b ? a : b;
    L.inner = take(L.inner_bytes);
    L.total = off;
    int nb = 1;
    while ((1ll << nb) < n) ++nb;                                   // rank < n and cd[i] < n fit in nb bits
    L.nb = nb;
    return L;
}

}  // namespace evx

using namespace evx;

extern "C" size_t evx_non_dominate_workspace_bytes(int64_t n, int32_t m) {
    if (n < 1 || m < 1 || m > EVX_MAX_OBJECTIVES || evx_nds_workspace_bytes(n, m) == 0 ||
        evx_crowding_workspace_bytes(n, m) == 0)
        return 0;
    return sel_layout(n, m).total + 256;
}

// Everything after the ranks: cut rank, crowding permutation, final order, gather.  `rank` may live
// anywhere; `inner` is scratch for the crowding sort.
static int select_by_rank(const float* pop, const float* fit, const float* fit_out_src, const int* rank, int64_t n, int d,
                          int m, int64_t topk, float* out_pop, float* out_fit, int32_t* out_index, char* base,
                          const SelLayout& L, cudaStream_t s) {
    int* hist = reinterpret_cast<int*>(base + L.hist);
    SelScal* sc = reinterpret_cast<SelScal*>(base + L.sc);
    uint8_t* mask = reinterpret_cast<uint8_t*>(base + L.mask);
    int* cd = reinterpret_cast<int*>(base + L.cd);
    uint64_t* kA = reinterpret_cast<uint64_t*>(base + L.keysA);
    uint64_t* kB = reinterpret_cast<uint64_t*>(base + L.keysB);
    int* vA = reinterpret_cast<int*>(base + L.valsA);
    int* vB = reinterpret_cast<int*>(base + L.valsB);
    const int T = 256;
    const unsigned gn = (unsigned)ceil_div(n, T);
    EVX_CUDA(cudaMemsetAsync(hist, 0, sizeof(int) * (size_t)(n + 1), s));
    EVX_CUDA(cudaMemsetAsync(sc, 0, sizeof(SelScal), s));
    k_sel_hist<<<gn, T, 0, s>>>(rank, n, hist, sc);
    EVX_LAUNCH_CHECK();
    k_sel_cut<<<1, 1024, 0, s>>>(hist, topk, sc);                                                       // :213-214
    EVX_LAUNCH_CHECK();
    k_sel_mask<<<gn, T, 0, s>>>(rank, n, sc, mask);                                                     // :215
    EVX_LAUNCH_CHECK();
    int rc = crowding_sort_known(fit, mask, &sc->cut, n, m, cd, base + L.inner, s);                     // :216
    if (rc) return rc;
    k_sel_keys<<<gn, T, 0, s>>>(rank, cd, n, L.nb, sc, kA, vA);
    EVX_LAUNCH_CHECK();
    size_t tb = L.tmp_bytes;
    EVX_CUDA(cub::DeviceRadixSort::SortPairs(base + L.tmp, tb, kA, kB, vA, vB, (int)n, 0, 2 * L.nb, s));   // :218
    const int64_t elems = topk * (int64_t)((out_pop ? d : 0) + m);
    k_sel_gather<<<(unsigned)ceil_div(elems, T), T, 0, s>>>(vB, topk, pop, out_pop ? d : 0, fit_out_src ? fit_out_src : fit,
                                                           m, out_pop, out_fit, out_index);            // :219-222
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}

static int select_check(const char* who, const float* pop, const float* fit, int64_t n, int32_t d, int32_t m, int64_t topk,
                        const float* out_pop, const float* out_fit, const int32_t* out_index, void* ws, size_t ws_bytes) {
    EVX_REQUIRE(n >= 1 && m >= 1 && m <= EVX_MAX_OBJECTIVES && d >= 0, EVX_ERR_ARG, "%s: bad shape n=%lld d=%d m=%d", who,
                (long long)n, d, m);
    EVX_REQUIRE(topk >= 1 && topk <= n, EVX_ERR_ARG, "%s: topk=%lld must be in [1, n=%lld]", who, (long long)topk,
                (long long)n);
    EVX_REQUIRE(fit && (out_pop == nullptr || (pop && d > 0)) && (out_pop || out_fit || out_index), EVX_ERR_ARG,
                "%s: null pointer", who);
    const size_t need = evx_non_dominate_workspace_bytes(n, m);
    EVX_REQUIRE(need != 0, EVX_ERR_ARG, "%s: n=%lld too large", who, (long long)n);
    EVX_REQUIRE(ws && ws_bytes >= need, EVX_ERR_WORKSPACE, "%s: workspace %zu < required %zu", who, ws_bytes, need);
    return EVX_OK;
}

extern "C" int evx_non_dominate_f32(const float* pop, const float* fit, const float* fit_out_src, int64_t n, int32_t d,
                                    int32_t m, int64_t topk, float* out_pop, float* out_fit, int32_t* out_index,
                                    void* ws, size_t ws_bytes, void* stream) {
    int rc = select_check("evx_non_dominate_f32", pop, fit, n, d, m, topk, out_pop, out_fit, out_index, ws, ws_bytes);
    if (rc) return rc;
    char* base = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(ws), 256));
    const SelLayout L = sel_layout(n, m);
    int* rank = reinterpret_cast<int*>(base + L.rank);
    rc = evx_nds_f32(fit, n, m, rank, base + L.inner, L.inner_bytes, stream);                           // :211
    if (rc) return rc;
    return select_by_rank(pop, fit, fit_out_src, rank, n, d, m, topk, out_pop, out_fit, out_index, base, L,
                          (cudaStream_t)stream);
}

extern "C" int evx_select_by_rank_f32(const float* pop, const float* fit, const float* fit_out_src, const int32_t* rank,
                                      int64_t n, int32_t d, int32_t m, int64_t topk, float* out_pop, float* out_fit,
                                      int32_t* out_index, void* ws, size_t ws_bytes, void* stream) {
    int rc = select_check("evx_select_by_rank_f32", pop, fit, n, d, m, topk, out_pop, out_fit, out_index, ws, ws_bytes);
    if (rc) return rc;
    EVX_REQUIRE(rank, EVX_ERR_ARG, "evx_select_by_rank_f32: null rank");
    char* base = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(ws), 256));
    return select_by_rank(pop, fit, fit_out_src, rank, n, d, m, topk, out_pop, out_fit, out_index, base, sel_layout(n, m),
                          (cudaStream_t)stream);
}
