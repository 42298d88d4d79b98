// spea2.cu - SPEA2's fitness assignment and archive truncation.
//
// Replaces algorithms/mo/spea2.py:21-37 `cal_fitness(obj)` and :40-67 `_truncation(obj, k, mask)`.
// The reference materialises the n x n bool dominance matrix, the n x n float distance matrix, sorts
// every row of the latter, and in the truncation loop takes a row-minimum of the whole distance matrix
// for every row it removes (k times n^2).  Here nothing n x n is stored:
//   fitness    - pass 1 (one thread per row i, columns staged in shared memory): strength s[i] = rows i
//                dominates, and the two smallest distances to other rows (the reference's sorted row,
//                index floor(sqrt(6)) - 1 = 1); pass 2: raw[j] = sum of s[i] over the rows i dominating j.
//   truncation - every valid row keeps its nearest valid neighbour (distance, index); one persistent CTA
//                removes k rows one after the other: arg-min (ties: lowest index, as the reference's stable
//                argsort), then only the rows whose nearest neighbour was the removed row search again.
// Distances are sqrt(sum_k (x_k - y_k)^2) in float32, summed in objective order without fused
// multiply-adds, compared AFTER the square root (the root can merge neighbouring values into a tie, and
// ties decide which row goes first).
#include "common.cuh"

namespace evx {

constexpr int SP_T = 256;          // threads per CTA = rows per CTA
constexpr int SP_TILE = 256;       // staged columns per step
constexpr int SP_MAXM = EVX_MAX_OBJECTIVES;

__device__ __forceinline__ float sp_dist(const float* __restrict__ a, const float* __restrict__ b, int m) {
    float acc = 0.0f;
    for (int k = 0; k < m; ++k) {
        const float d = __fsub_rn(a[k], b[k]);
        acc = __fadd_rn(acc, __fmul_rn(d, d));
    }
    return __fsqrt_rn(acc);
}
__device__ __forceinline__ bool sp_dominates(const float* __restrict__ a, const float* __restrict__ b, int m) {
    bool le = true, lt = false;                       // utils/common.py:88-91, IEEE compares
    for (int k = 0; k < m; ++k) {
        le = le && (a[k] <= b[k]);
        lt = lt || (a[k] < b[k]);
    }
    return le && lt;
}

// Pass 1: strength and the distance to the second nearest other row.
__global__ void __launch_bounds__(SP_T)
k_spea2_strength(const float* __restrict__ obj, int n, int m, int* __restrict__ strength, float* __restrict__ dens) {
    extern __shared__ float tile[];                   // [SP_TILE][m]
    const int i = blockIdx.x * SP_T + threadIdx.x;
    float xi[SP_MAXM];
    for (int k = 0; k < m; ++k) xi[k] = (i < n) ? obj[(size_t)i * m + k] : 0.0f;
    int s = 0;
    float d1 = INFINITY, d2 = INFINITY;               // smallest, second smallest (NaN distances sort last: ignored)
    for (int j0 = 0; j0 < n; j0 += SP_TILE) {
        const int nt = min(SP_TILE, n - j0);
        __syncthreads();
        for (int t = threadIdx.x; t < nt * m; t += SP_T) tile[t] = obj[(size_t)j0 * m + t];
        __syncthreads();
        if (i < n) {
            for (int j = 0; j < nt; ++j) {
                const float* y = tile + j * m;
                s += sp_dominates(xi, y, m);
                if (j0 + j != i) {
                    const float d = sp_dist(xi, y, m);
                    if (d < d1) { d2 = d1; d1 = d; }
                    else if (d < d2) d2 = d;
                }
            }
        }
    }
    if (i < n) {
        strength[i] = s;
        // n == 1: the row holds only the +inf diagonal -> 1 / (inf + 2) = 0; n == 2: second entry is the diagonal's inf
        dens[i] = __fdiv_rn(1.0f, __fadd_rn(d2, 2.0f));                   // :36
    }
}

// Pass 2: raw fitness of column j = sum of the strengths of its dominators (:27-29), then fitness (:37).
__global__ void __launch_bounds__(SP_T)
k_spea2_raw(const float* __restrict__ obj, int n, int m, const int* __restrict__ strength, const float* __restrict__ dens,
            float* __restrict__ fitness) {
    extern __shared__ float tile[];                   // [SP_TILE][m] rows i, then [SP_TILE] strengths
    int* st = reinterpret_cast<int*>(tile + SP_TILE * m);
    const int j = blockIdx.x * SP_T + threadIdx.x;
    float yj[SP_MAXM];
    for (int k = 0; k < m; ++k) yj[k] = (j < n) ? obj[(size_t)j * m + k] : 0.0f;
    unsigned int raw = 0;                             // int32 arithmetic, as jnp.sum over an int32 array
    for (int i0 = 0; i0 < n; i0 += SP_TILE) {
        const int nt = min(SP_TILE, n - i0);
        __syncthreads();
        for (int t = threadIdx.x; t < nt * m; t += SP_T) tile[t] = obj[(size_t)i0 * m + t];
        for (int t = threadIdx.x; t < nt; t += SP_T) st[t] = strength[i0 + t];
        __syncthreads();
        if (j < n)
            for (int i = 0; i < nt; ++i)
                if (sp_dominates(tile + i * m, yj, m)) raw += (unsigned int)st[i];
    }
    if (j < n) fitness[j] = __fadd_rn(dens[j], (float)(int)raw);
}

// Nearest valid neighbour of every valid row: (distance, index); ties in distance: lowest index.
__global__ void __launch_bounds__(SP_T)
k_spea2_nearest(const float* __restrict__ obj, const uint8_t* __restrict__ mask, int n, int m, float* __restrict__ nn,
                int* __restrict__ nni) {
    extern __shared__ float tile[];
    uint8_t* mk = reinterpret_cast<uint8_t*>(tile + SP_TILE * m);
    const int i = blockIdx.x * SP_T + threadIdx.x;
    float xi[SP_MAXM];
    for (int k = 0; k < m; ++k) xi[k] = (i < n) ? obj[(size_t)i * m + k] : 0.0f;
    const bool vi = i < n && mask[i] != 0;
    float best = INFINITY;
    int bi = -1;
    for (int j0 = 0; j0 < n; j0 += SP_TILE) {
        const int nt = min(SP_TILE, n - j0);
        __syncthreads();
        for (int t = threadIdx.x; t < nt * m; t += SP_T) tile[t] = obj[(size_t)j0 * m + t];
        for (int t = threadIdx.x; t < nt; t += SP_T) mk[t] = mask[j0 + t];
        __syncthreads();
        if (vi)
            for (int j = 0; j < nt; ++j)
                if (mk[j] && j0 + j != i) {
                    const float d = sp_dist(xi, tile + j * m, m);
                    if (d < best) { best = d; bi = j0 + j; }
                }
    }
    if (i < n) { nn[i] = vi ? best : INFINITY; nni[i] = vi ? bi : -1; }
}

// One CTA removes k rows (:50-63): arg-min of nn over all rows (ties: lowest index), mark it removed, then
// the rows that had it as nearest neighbour search the remaining valid rows again.
constexpr int TR_T = 1024;
__global__ void __launch_bounds__(TR_T)
k_spea2_truncate(const float* __restrict__ obj, const uint8_t* __restrict__ mask, int n, int m, int k, float* __restrict__ nn,
                 int* __restrict__ nni, uint8_t* __restrict__ alive, uint8_t* __restrict__ keep) {
    __shared__ float rv[TR_T / 32];
    __shared__ int ri[TR_T / 32];
    __shared__ int sel, naff;
    __shared__ int aff[TR_T];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    for (int i = t; i < n; i += TR_T) alive[i] = mask[i] != 0;
    __syncthreads();
    for (int it = 0; it < k; ++it) {
        float bv = INFINITY;
        int bi = 0x7FFFFFFF;
        for (int i = t; i < n; i += TR_T) {           // rows that are gone (or masked) carry +inf
            const float v = alive[i] ? nn[i] : INFINITY;
            if (v < bv || (v == bv && i < bi)) { bv = v; bi = i; }
        }
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            const float ov = __shfl_xor_sync(0xFFFFFFFFu, bv, o);
            const int oi = __shfl_xor_sync(0xFFFFFFFFu, bi, o);
            if (ov < bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { rv[warp] = bv; ri[warp] = bi; }
        __syncthreads();
        if (warp == 0) {
            bv = rv[lane]; bi = ri[lane];
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) {
                const float ov = __shfl_xor_sync(0xFFFFFFFFu, bv, o);
                const int oi = __shfl_xor_sync(0xFFFFFFFFu, bi, o);
                if (ov < bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
            }
            if (lane == 0) { sel = bi; naff = 0; }
        }
        __syncthreads();
        const int idx = sel;                          // all +inf: index 0, like argsort of a constant row
        if (idx >= n) break;
        if (t == 0) alive[idx] = 0;
        for (int i = t; i < n; i += TR_T)
            if (i != idx && alive[i] && nni[i] == idx) {
                const int slot = atomicAdd(&naff, 1);
                if (slot < TR_T) aff[slot] = i;
            }
        __syncthreads();
        const int na = naff;
        // (more than TR_T rows sharing one nearest neighbour cannot happen for distinct points in a few
        //  dimensions; identical points can: the overflow is handled by a serial sweep below)
        for (int a = 0; a < min(na, TR_T); ++a) {
            const int i = aff[a];
            float best = INFINITY;
            int bj = 0x7FFFFFFF;
            for (int j = t; j < n; j += TR_T)
                if (alive[j] && j != i) {
                    const float d = sp_dist(obj + (size_t)i * m, obj + (size_t)j * m, m);
                    if (d < best || (d == best && j < bj)) { best = d; bj = j; }
                }
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) {
                const float ov = __shfl_xor_sync(0xFFFFFFFFu, best, o);
                const int oj = __shfl_xor_sync(0xFFFFFFFFu, bj, o);
                if (ov < best || (ov == best && oj < bj)) { best = ov; bj = oj; }
            }
            if (lane == 0) { rv[warp] = best; ri[warp] = bj; }
            __syncthreads();
            if (warp == 0) {
                best = rv[lane]; bj = ri[lane];
#pragma unroll
                for (int o = 16; o >= 1; o >>= 1) {
                    const float ov = __shfl_xor_sync(0xFFFFFFFFu, best, o);
                    const int oj = __shfl_xor_sync(0xFFFFFFFFu, bj, o);
                    if (ov < best || (ov == best && oj < bj)) { best = ov; bj = oj; }
                }
                if (lane == 0) { nn[i] = best; nni[i] = (bj == 0x7FFFFFFF) ? -1 : bj; }
            }
            __syncthreads();
        }
        if (na > TR_T) {                              // pathological: every thread fixes whole rows on its own
            for (int i = t; i < n; i += TR_T)
                if (alive[i] && nni[i] == idx) {
                    float best = INFINITY;
                    int bj = -1;
                    for (int j = 0; j < n; ++j)
                        if (alive[j] && j != i) {
                            const float d = sp_dist(obj + (size_t)i * m, obj + (size_t)j * m, m);
                            if (d < best) { best = d; bj = j; }
                        }
                    nn[i] = best; nni[i] = bj;
                }
            __syncthreads();
        }
    }
    __syncthreads();
    for (int i = t; i < n; i += TR_T) keep[i] = alive[i];                 // = del_ind & mask (:67)
}

static size_t sp_layout(int64_t n, size_t* strength, size_t* dens, size_t* nn, size_t* nni, size_t* alive) {
    size_t off = 0;
    auto take = [&](size_t b) { size_t o = off; off = align_up(off + b, 256); return o; };
    *strength = take(sizeof(int) * (size_t)n);
    *dens = take(sizeof(float) * (size_t)n);
    *nn = take(sizeof(float) * (size_t)n);
    *nni = take(sizeof(int) * (size_t)n);
    *alive = take((size_t)n);
    return off;
}

}  // namespace evx

using namespace evx;

extern "C" size_t evx_spea2_workspace_bytes(int64_t n, int32_t m) {
    if (n < 0 || n > 0x7FFFFFFF / 2 || m < 1 || m > EVX_MAX_OBJECTIVES) return 0;
    size_t a, b, c, d, e;
    return sp_layout(n > 0 ? n : 1, &a, &b, &c, &d, &e) + 256;
}

extern "C" int evx_spea2_fitness_f32(const float* obj, int64_t n, int32_t m, float* fitness, void* ws, size_t ws_bytes,
                                     void* stream) {
    const size_t need = evx_spea2_workspace_bytes(n, m);
    EVX_REQUIRE(need != 0, EVX_ERR_ARG, "evx_spea2_fitness_f32: bad shape n=%lld m=%d", (long long)n, m);
    if (n == 0) return EVX_OK;
    EVX_REQUIRE(obj && fitness, EVX_ERR_ARG, "evx_spea2_fitness_f32: null pointer");
    EVX_REQUIRE(ws && ws_bytes >= need, EVX_ERR_WORKSPACE, "evx_spea2_fitness_f32: workspace %zu < required %zu", ws_bytes, need);
    char* base = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(ws), 256));
    size_t o_s, o_d, o_nn, o_ni, o_al;
    sp_layout(n, &o_s, &o_d, &o_nn, &o_ni, &o_al);
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned grid = (unsigned)ceil_div(n, SP_T);
    int* strength = reinterpret_cast<int*>(base + o_s);
    float* dens = reinterpret_cast<float*>(base + o_d);
    k_spea2_strength<<<grid, SP_T, sizeof(float) * SP_TILE * m, s>>>(obj, (int)n, m, strength, dens);
    EVX_LAUNCH_CHECK();
    k_spea2_raw<<<grid, SP_T, sizeof(float) * SP_TILE * m + sizeof(int) * SP_TILE, s>>>(obj, (int)n, m, strength, dens, fitness);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}

extern "C" int evx_spea2_truncation_f32(const float* obj, int64_t n, int32_t m, int64_t k, const uint8_t* mask, uint8_t* keep,
                                        void* ws, size_t ws_bytes, void* stream) {
    const size_t need = evx_spea2_workspace_bytes(n, m);
    EVX_REQUIRE(need != 0, EVX_ERR_ARG, "evx_spea2_truncation_f32: bad shape n=%lld m=%d", (long long)n, m);
    EVX_REQUIRE(k >= 0 && k <= n, EVX_ERR_ARG, "evx_spea2_truncation_f32: k=%lld out of range", (long long)k);
    if (n == 0) return EVX_OK;
    EVX_REQUIRE(obj && mask && keep, EVX_ERR_ARG, "evx_spea2_truncation_f32: null pointer");
    EVX_REQUIRE(ws && ws_bytes >= need, EVX_ERR_WORKSPACE, "evx_spea2_truncation_f32: workspace %zu < required %zu", ws_bytes, need);
    char* base = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(ws), 256));
    size_t o_s, o_d, o_nn, o_ni, o_al;
    sp_layout(n, &o_s, &o_d, &o_nn, &o_ni, &o_al);
    cudaStream_t s = (cudaStream_t)stream;
    float* nn = reinterpret_cast<float*>(base + o_nn);
    int* nni = reinterpret_cast<int*>(base + o_ni);
    uint8_t* alive = reinterpret_cast<uint8_t*>(base + o_al);
    k_spea2_nearest<<<(unsigned)ceil_div(n, SP_T), SP_T, sizeof(float) * SP_TILE * m + SP_TILE, s>>>(obj, mask, (int)n, m, nn, nni);
    EVX_LAUNCH_CHECK();
    k_spea2_truncate<<<1, TR_T, 0, s>>>(obj, mask, (int)n, m, (int)k, nn, nni, alive, keep);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}
