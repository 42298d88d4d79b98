// spea2.cu - SPEA2's fitness assignment and archive truncation.
//
// Replaces algorithms/mo/spea2.py:21-37 `cal_fitness(obj)` and :40-67 `_truncation(obj, k, mask)`.
// The reference materialises the n x n bool dominance matrix, the n x n float distance matrix, sorts
// every row of the latter, and in the truncation loop takes a row-minimum of the whole distance matrix
// for every row it removes (k times n^2).  Here nothing n x n is stored:
//   fitness    - pass 1: strength s[i] = rows i dominates, and the two smallest distances to other rows
//                (the reference's sorted row, index floor(sqrt(6)) - 1 = 1); pass 2: raw[j] = sum of s[i]
//                over the rows i dominating j.  A CTA takes 32 rows (columns in pass 2) and splits the
//                other side eight ways over its warps - a population of 20 000 still fills the chip -
//                with the other side staged in shared memory; the eight partial results are merged there.
//   truncation - every valid row gets the list of its EIGHT nearest valid neighbours, by (distance, index);
//                one persistent CTA then removes k rows one after the other: arg-min over the rows of the
//                distance to the nearest neighbour (ties: lowest index, as the reference's stable argsort)
//                from per-thread minima.  Lists are repaired lazily: a row whose nearest neighbour has been
//                removed still shows the old, smaller distance, so it reaches the arg-min no later than it
//                should; only then is it moved on to its next living neighbour (or, its list used up,
//                searched again by the whole CTA) and the arg-min repeated.  Rows that never become the
//                minimum are never repaired.
// Distances are sqrt(sum_k (x_k - y_k)^2) in float32, summed in objective order without fused
// multiply-adds, compared AFTER the square root (the root can merge neighbouring values into a tie, and
// ties decide which row goes first).  The root itself is only taken when the sum of squares is below the
// one behind the value it would have to beat (sqrt is monotone under round-to-nearest).
#include "common.cuh"

namespace evx {

constexpr int SP_T = 256;          // threads per CTA
constexpr int SP_ROWS = 32;        // rows (pass 2: columns) per CTA, one per lane
constexpr int SP_SPLIT = SP_T / SP_ROWS;   // warps: each takes 32 of every 256 staged rows of the other side
constexpr int SP_TILE = 256;       // staged rows per step
constexpr int SP_MAXM = EVX_MAX_OBJECTIVES;
constexpr int SP_KNN = 8;          // neighbours kept per row for the truncation

// MT = number of objectives at compile time (arrays in registers, loops unrolled), 0 = run-time m
template <int MT>
struct SpDim {
    static constexpr int CAP = MT ? MT : SP_MAXM;
    __device__ __forceinline__ static int get(int m) { return MT ? MT : m; }
};

template <int MT>
__device__ __forceinline__ float sp_sumsq(const float* a, const float* b, int m) {
    const int M = SpDim<MT>::get(m);
    float acc = 0.0f;
#pragma unroll
    for (int k = 0; k < M; ++k) {
        const float d = __fsub_rn(a[k], b[k]);
        acc = __fadd_rn(acc, __fmul_rn(d, d));
    }
    return acc;
}
template <int MT>
__device__ __forceinline__ bool sp_dominates(const float* a, const float* b, int m) {
    const int M = SpDim<MT>::get(m);
    bool le = true, lt = false;                       // utils/common.py:88-91, IEEE compares
#pragma unroll
    for (int k = 0; k < M; ++k) {
        le = le && (a[k] <= b[k]);
        lt = lt || (a[k] < b[k]);
    }
    return le && lt;
}

// Pass 1: strength and the distance to the second nearest other row.
template <int MT>
__global__ void __launch_bounds__(SP_T)
k_spea2_strength(const float* __restrict__ obj, int n, int m, int* __restrict__ strength, float* __restrict__ dens) {
    extern __shared__ float tile[];                   // [SP_TILE][M]
    __shared__ int cs[SP_SPLIT][SP_ROWS];
    __shared__ float c1[SP_SPLIT][SP_ROWS], c2[SP_SPLIT][SP_ROWS];
    const int M = SpDim<MT>::get(m);
    const int r = threadIdx.x & 31, sp = threadIdx.x >> 5;
    const int i = blockIdx.x * SP_ROWS + r;
    float xi[SpDim<MT>::CAP];
#pragma unroll
    for (int k = 0; k < M; ++k) xi[k] = (i < n) ? obj[(size_t)i * M + k] : __int_as_float(0x7FC00000);
    int s = 0;
    float d1 = INFINITY, d2 = INFINITY;               // smallest, second smallest (NaN distances sort last: ignored)
    float a2 = INFINITY;                              // a sum of squares whose root is d2
    float a1 = INFINITY;
    for (int j0 = 0; j0 < n; j0 += SP_TILE) {
        const int nt = min(SP_TILE, n - j0);
        __syncthreads();
        for (int t = threadIdx.x; t < nt * M; t += SP_T) tile[t] = obj[(size_t)j0 * M + t];
        __syncthreads();
        const int je = min(nt, (sp + 1) * SP_ROWS);
        for (int j = sp * SP_ROWS; j < je; ++j) {
            const float* y = tile + j * M;
            s += sp_dominates<MT>(xi, y, m) ? 1 : 0;
            if (j0 + j != i) {
                const float acc = sp_sumsq<MT>(xi, y, m);
                if (acc < a2) {                       // otherwise the root cannot be below d2 (false for NaN)
                    const float d = __fsqrt_rn(acc);
                    if (d < d1) { d2 = d1; a2 = a1; d1 = d; a1 = acc; }
                    else if (d < d2) { d2 = d; a2 = acc; }
                }
            }
        }
    }
    cs[sp][r] = s; c1[sp][r] = d1; c2[sp][r] = d2;
    __syncthreads();
    if (sp == 0 && i < n) {
        int st = 0;
        float m1 = INFINITY, m2 = INFINITY;           // the two smallest of the sixteen
#pragma unroll
        for (int q = 0; q < SP_SPLIT; ++q) {
            st += cs[q][r];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const float v = h ? c2[q][r] : c1[q][r];
                if (v < m1) { m2 = m1; m1 = v; }
                else if (v < m2) m2 = v;
            }
        }
        strength[i] = st;
        // n == 1: the row holds only the +inf diagonal -> 1 / (inf + 2) = 0; n == 2: second entry is the diagonal's inf
        dens[i] = __fdiv_rn(1.0f, __fadd_rn(m2, 2.0f));                   // :36
    }
}

// Pass 2: raw fitness of column j = sum of the strengths of its dominators (:27-29), then fitness (:37).
template <int MT>
__global__ void __launch_bounds__(SP_T)
k_spea2_raw(const float* __restrict__ obj, int n, int m, const int* __restrict__ strength, const float* __restrict__ dens,
            float* __restrict__ fitness) {
    extern __shared__ float tile[];                   // [SP_TILE][M] rows i, then [SP_TILE] strengths
    __shared__ unsigned int cr[SP_SPLIT][SP_ROWS];
    const int M = SpDim<MT>::get(m);
    int* st = reinterpret_cast<int*>(tile + SP_TILE * M);
    const int c = threadIdx.x & 31, sp = threadIdx.x >> 5;
    const int j = blockIdx.x * SP_ROWS + c;
    float yj[SpDim<MT>::CAP];
#pragma unroll
    for (int k = 0; k < M; ++k) yj[k] = (j < n) ? obj[(size_t)j * M + k] : __int_as_float(0x7FC00000);
    unsigned int raw = 0;                             // int32 arithmetic, as jnp.sum over an int32 array
    for (int i0 = 0; i0 < n; i0 += SP_TILE) {
        const int nt = min(SP_TILE, n - i0);
        __syncthreads();
        for (int t = threadIdx.x; t < nt * M; t += SP_T) tile[t] = obj[(size_t)i0 * M + t];
        for (int t = threadIdx.x; t < nt; t += SP_T) st[t] = strength[i0 + t];
        __syncthreads();
        const int ie = min(nt, (sp + 1) * SP_ROWS);
        for (int i = sp * SP_ROWS; i < ie; ++i)
            if (st[i] != 0 && sp_dominates<MT>(tile + i * M, yj, m)) raw += (unsigned int)st[i];
    }
    cr[sp][c] = raw;
    __syncthreads();
    if (sp == 0 && j < n) {
        unsigned int tot = 0;
#pragma unroll
        for (int q = 0; q < SP_SPLIT; ++q) tot += cr[q][c];
        fitness[j] = __fadd_rn(dens[j], (float)(int)tot);
    }
}

// The SP_KNN nearest valid neighbours of every valid row, ascending by (distance, index).  Thread (row, split)
// keeps its own sorted list over its share of the columns (visited in ascending index order, so equal
// distances keep their index order); the eight lists of a row are merged in shared memory.
template <int MT>
__global__ void __launch_bounds__(SP_T)
k_spea2_knn(const float* __restrict__ obj, const uint8_t* __restrict__ mask, int n, int m, float* __restrict__ knn_d,
            int* __restrict__ knn_j, int* __restrict__ cnt, int* __restrict__ ptr, float* __restrict__ nn,
            int* __restrict__ nni) {
    extern __shared__ float tile[];                   // [SP_TILE][M], then SP_TILE mask bytes; reused for the merge
    const int M = SpDim<MT>::get(m);
    uint8_t* mk = reinterpret_cast<uint8_t*>(tile + SP_TILE * M);
    const int r = threadIdx.x & 31, sp = threadIdx.x >> 5;
    const int i = blockIdx.x * SP_ROWS + r;
    float xi[SpDim<MT>::CAP];
#pragma unroll
    for (int k = 0; k < M; ++k) xi[k] = (i < n) ? obj[(size_t)i * M + k] : __int_as_float(0x7FC00000);
    const bool vi = i < n && mask[i] != 0;
    float ld[SP_KNN], la[SP_KNN];                     // distance and the sum of squares behind it
    int lj[SP_KNN];
#pragma unroll
    for (int p = 0; p < SP_KNN; ++p) { ld[p] = INFINITY; la[p] = INFINITY; lj[p] = -1; }
    for (int j0 = 0; j0 < n; j0 += SP_TILE) {
        const int nt = min(SP_TILE, n - j0);
        __syncthreads();
        for (int t = threadIdx.x; t < nt * M; t += SP_T) tile[t] = obj[(size_t)j0 * M + t];
        for (int t = threadIdx.x; t < nt; t += SP_T) mk[t] = mask[j0 + t];
        __syncthreads();
        if (!vi) continue;
        const int je = min(nt, (sp + 1) * SP_ROWS);
        for (int j = sp * SP_ROWS; j < je; ++j) {
            if (!mk[j] || j0 + j == i) continue;
            const float acc = sp_sumsq<MT>(xi, tile + j * M, m);
            if (!(acc < la[SP_KNN - 1])) continue;    // its root cannot be below the list's last (NaN: never)
            const float d = __fsqrt_rn(acc);
            if (!(d < ld[SP_KNN - 1])) continue;      // equal distance, larger index: behind it
            bool placed = false;
#pragma unroll
            for (int p = SP_KNN - 1; p >= 0; --p) {
                if (!placed) {
                    if (p > 0 && d < ld[p - 1]) { ld[p] = ld[p - 1]; la[p] = la[p - 1]; lj[p] = lj[p - 1]; }
                    else { ld[p] = d; la[p] = acc; lj[p] = j0 + j; placed = true; }
                }
            }
        }
    }
    // merge the eight lists of each row: sd / sj [row][split][entry]
    __syncthreads();
    float* sd = tile;
    int* sj = reinterpret_cast<int*>(tile + SP_ROWS * SP_SPLIT * SP_KNN);
#pragma unroll
    for (int p = 0; p < SP_KNN; ++p) {
        sd[(r * SP_SPLIT + sp) * SP_KNN + p] = ld[p];
        sj[(r * SP_SPLIT + sp) * SP_KNN + p] = lj[p];
    }
    __syncthreads();
    if (sp == 0 && i < n) {
        int h[SP_SPLIT];
#pragma unroll
        for (int q = 0; q < SP_SPLIT; ++q) h[q] = 0;
        int c = 0;
        for (int o = 0; o < SP_KNN; ++o) {
            float bd = INFINITY;
            int bj = 0x7FFFFFFF, bq = -1;
#pragma unroll
            for (int q = 0; q < SP_SPLIT; ++q) {
                if (h[q] < SP_KNN) {
                    const float d = sd[(r * SP_SPLIT + q) * SP_KNN + h[q]];
                    const int jj = sj[(r * SP_SPLIT + q) * SP_KNN + h[q]];
                    if (jj >= 0 && (d < bd || (d == bd && jj < bj))) { bd = d; bj = jj; bq = q; }
                }
            }
            if (bq < 0) break;
#pragma unroll
            for (int q = 0; q < SP_SPLIT; ++q) if (q == bq) ++h[q];
            knn_d[(size_t)i * SP_KNN + o] = bd;
            knn_j[(size_t)i * SP_KNN + o] = bj;
            ++c;
        }
        cnt[i] = c;
        ptr[i] = 0;
        nn[i] = c ? knn_d[(size_t)i * SP_KNN] : INFINITY;
        nni[i] = c ? knn_j[(size_t)i * SP_KNN] : -1;
    }
}

// One CTA removes k rows (:50-63).  Thread t owns the rows i = t (mod TR_T) and keeps the minimum of
// (nn, index) over them; see the header for the lazy repair of nn.
constexpr int TR_T = 1024;
enum { TR_REMOVE = 0, TR_MOVED = 1, TR_SEARCH = 2 };
template <int MT>
__global__ void __launch_bounds__(TR_T)
k_spea2_truncate(const float* __restrict__ obj, const uint8_t* __restrict__ mask, int n, int m, int k, float* nn, int* nni,
                 float* knn_d, int* knn_j, int* cnt, int* ptr, uint8_t* alive, uint8_t* __restrict__ keep) {
    __shared__ float rv[TR_T / 32];
    __shared__ int ri[TR_T / 32];
    __shared__ int sel, action, partner;
    const int M = SpDim<MT>::get(m);
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    for (int i = t; i < n; i += TR_T) alive[i] = mask[i] != 0;
    __syncthreads();
    float mv = INFINITY;                              // minimum over this thread's rows (removed or masked rows carry +inf)
    int mi = 0x7FFFFFFF;
    auto rescan_own = [&]() {                         // nn is +inf for masked rows (k_spea2_knn) and removed rows (below)
        mv = INFINITY; mi = 0x7FFFFFFF;
#pragma unroll 8
        for (int i = t; i < n; i += TR_T) {
            const float v = nn[i];
            if (v < mv || (v == mv && i < mi)) { mv = v; mi = i; }
        }
    };
    rescan_own();
    int removed = 0;
    while (removed < k) {
        float bv = mv;
        int bi = mi;
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
            if (lane == 0) sel = bi;                  // all +inf: index 0, like argsort of a constant row
        }
        __syncthreads();
        const int idx = sel;
        if (idx >= n) break;
        const bool owner = (idx % TR_T) == t;
        if (owner) {
            int act = TR_REMOVE;
            if (alive[idx] && nn[idx] < INFINITY && !alive[nni[idx]]) {   // the neighbour it shows is gone: not its turn yet
                const int c = cnt[idx];
                int p = ptr[idx] + 1;
                while (p < c && !alive[knn_j[(size_t)idx * SP_KNN + p]]) ++p;
                if (p < c) {
                    ptr[idx] = p;
                    nn[idx] = knn_d[(size_t)idx * SP_KNN + p];
                    nni[idx] = knn_j[(size_t)idx * SP_KNN + p];
                    act = TR_MOVED;
                } else {
                    act = TR_SEARCH;
                }
            }
            int pj = -1;
            if (act == TR_REMOVE) {
                // the removed row's nearest neighbour, when the two are each other's nearest (the closest pair:
                // the usual case), shows a dead row from now on and would be the very next arg-min: it is
                // moved on in this same turn instead of spending a turn of its own on it
                if (alive[idx] && nn[idx] < INFINITY) {
                    const int j = nni[idx];
                    if (alive[j] && nni[j] == idx) pj = j;
                }
                alive[idx] = 0;
                nn[idx] = INFINITY;
            }
            action = act;
            partner = pj;
        }
        __syncthreads();
        const int act = action;
        const int pj = partner;
        if (pj >= 0 && (pj % TR_T) == t) {            // (its owner; a used-up list is left to the lazy repair)
            const int c = cnt[pj];
            int p = ptr[pj] + 1;
            while (p < c && !alive[knn_j[(size_t)pj * SP_KNN + p]]) ++p;
            if (p < c) {
                ptr[pj] = p;
                nn[pj] = knn_d[(size_t)pj * SP_KNN + p];
                nni[pj] = knn_j[(size_t)pj * SP_KNN + p];
            }
        }
        if (act == TR_SEARCH) {                       // list used up: nearest living row, by the whole CTA
            float xi[SpDim<MT>::CAP];
#pragma unroll
            for (int q = 0; q < M; ++q) xi[q] = obj[(size_t)idx * M + q];
            float best = INFINITY;
            int bj = 0x7FFFFFFF;
            for (int j = t; j < n; j += TR_T)
                if (alive[j] && j != idx) {
                    float yj[SpDim<MT>::CAP];
#pragma unroll
                    for (int q = 0; q < M; ++q) yj[q] = obj[(size_t)j * M + q];
                    const float d = __fsqrt_rn(sp_sumsq<MT>(xi, yj, m));
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
                if (lane == 0) {
                    const bool found = bj != 0x7FFFFFFF;
                    nn[idx] = found ? best : INFINITY;
                    nni[idx] = found ? bj : -1;
                    knn_d[(size_t)idx * SP_KNN] = best;
                    knn_j[(size_t)idx * SP_KNN] = found ? bj : -1;
                    cnt[idx] = found ? 1 : 0;
                    ptr[idx] = 0;
                }
            }
            __syncthreads();
        }
        // the owner's minimum is stale: its whole warp looks through the owner's rows (n / 1024 of them), lane j
        // taking every 32nd, instead of the owner walking them alone
        for (int which = 0; which < 2; ++which) {
            const int x = which == 0 ? idx : pj;
            if (x < 0 || warp != (x % TR_T) / 32) continue;                // (warp-uniform)
            __syncwarp();                                                   // the partner's owner is a lane of this warp
            const int ot = x % TR_T;
            float bv2 = INFINITY;
            int bi2 = 0x7FFFFFFF;
            for (int i = ot + TR_T * lane; i < n; i += TR_T * 32) {
                const float v = nn[i];
                if (v < bv2 || (v == bv2 && i < bi2)) { bv2 = v; bi2 = i; }
            }
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) {
                const float ov = __shfl_xor_sync(0xFFFFFFFFu, bv2, o);
                const int oi = __shfl_xor_sync(0xFFFFFFFFu, bi2, o);
                if (ov < bv2 || (ov == bv2 && oi < bi2)) { bv2 = ov; bi2 = oi; }
            }
            if (t == ot) { mv = bv2; mi = bi2; }
        }
        if (act == TR_REMOVE) ++removed;
        // (the next turn's first barrier orders this turn's reads of sel / action before their next writes)
    }
    __syncthreads();
    for (int i = t; i < n; i += TR_T) keep[i] = alive[i];                 // = del_ind & mask (:67)
}

struct SpLayout {
    size_t strength, dens, nn, nni, alive, knn_d, knn_j, cnt, ptr, total;
};
static SpLayout sp_layout(int64_t n) {
    SpLayout L{};
    size_t off = 0;
    auto take = [&](size_t b) { size_t o = off; off = align_up(off + b, 256); return o; };
    L.strength = take(sizeof(int) * (size_t)n);
    L.dens = take(sizeof(float) * (size_t)n);
    L.nn = take(sizeof(float) * (size_t)n);
    L.nni = take(sizeof(int) * (size_t)n);
    L.alive = take((size_t)n);
    L.knn_d = take(sizeof(float) * (size_t)n * SP_KNN);
    L.knn_j = take(sizeof(int) * (size_t)n * SP_KNN);
    L.cnt = take(sizeof(int) * (size_t)n);
    L.ptr = take(sizeof(int) * (size_t)n);
    L.total = off;
    return L;
}

template <int MT>
static void launch_fitness(const float* obj, int n, int m, int* strength, float* dens, float* fitness, cudaStream_t s) {
    const unsigned grid = (unsigned)ceil_div((int64_t)n, SP_ROWS);
    k_spea2_strength<MT><<<grid, SP_T, sizeof(float) * SP_TILE * m, s>>>(obj, n, m, strength, dens);
    k_spea2_raw<MT><<<grid, SP_T, sizeof(float) * SP_TILE * m + sizeof(int) * SP_TILE, s>>>(obj, n, m, strength, dens, fitness);
}
template <int MT>
static void launch_knn(const float* obj, const uint8_t* mask, int n, int m, float* knn_d, int* knn_j, int* cnt, int* ptr,
                       float* nn, int* nni, cudaStream_t s) {
    const size_t stage = sizeof(float) * SP_TILE * m + SP_TILE;
    const size_t merge = (sizeof(float) + sizeof(int)) * SP_ROWS * SP_SPLIT * SP_KNN;
    k_spea2_knn<MT><<<(unsigned)ceil_div((int64_t)n, SP_ROWS), SP_T, stage > merge ? stage : merge, s>>>(obj, mask, n, m, knn_d, knn_j,
                                                                                                  cnt, ptr, nn, nni);
}

}  // namespace evx

using namespace evx;

// objectives counts with kernels of their own (registers, unrolled); everything else takes the run-time loop
#define EVX_SP_DISPATCH(m, CALL)                                                                   \
    switch (m) {                                                                                   \
        case 2: { constexpr int MT = 2; CALL; break; }                                             \
        case 3: { constexpr int MT = 3; CALL; break; }                                             \
        case 4: { constexpr int MT = 4; CALL; break; }                                             \
        case 5: { constexpr int MT = 5; CALL; break; }                                             \
        case 6: { constexpr int MT = 6; CALL; break; }                                             \
        case 8: { constexpr int MT = 8; CALL; break; }                                             \
        case 10: { constexpr int MT = 10; CALL; break; }                                           \
        default: { constexpr int MT = 0; CALL; break; }                                            \
    }

extern "C" size_t evx_spea2_workspace_bytes(int64_t n, int32_t m) {
    if (n < 0 || n > 0x7FFFFFFF / 2 || m < 1 || m > EVX_MAX_OBJECTIVES) return 0;
    return sp_layout(n > 0 ? n : 1).total + 256;
}

extern "C" int evx_spea2_fitness_f32(const float* obj, int64_t n, int32_t m, float* fitness, void* ws, size_t ws_bytes,
                                     void* stream) {
    const size_t need = evx_spea2_workspace_bytes(n, m);
    EVX_REQUIRE(need != 0, EVX_ERR_ARG, "evx_spea2_fitness_f32: bad shape n=%lld m=%d", (long long)n, m);
    if (n == 0) return EVX_OK;
    EVX_REQUIRE(obj && fitness, EVX_ERR_ARG, "evx_spea2_fitness_f32: null pointer");
    EVX_REQUIRE(ws && ws_bytes >= need, EVX_ERR_WORKSPACE, "evx_spea2_fitness_f32: workspace %zu < required %zu", ws_bytes, need);
    char* base = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(ws), 256));
    const SpLayout L = sp_layout(n);
    cudaStream_t s = (cudaStream_t)stream;
    int* strength = reinterpret_cast<int*>(base + L.strength);
    float* dens = reinterpret_cast<float*>(base + L.dens);
    EVX_SP_DISPATCH(m, launch_fitness<MT>(obj, (int)n, m, strength, dens, fitness, s));
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
    const SpLayout L = sp_layout(n);
    cudaStream_t s = (cudaStream_t)stream;
    float* nn = reinterpret_cast<float*>(base + L.nn);
    int* nni = reinterpret_cast<int*>(base + L.nni);
    uint8_t* alive = reinterpret_cast<uint8_t*>(base + L.alive);
    float* knn_d = reinterpret_cast<float*>(base + L.knn_d);
    int* knn_j = reinterpret_cast<int*>(base + L.knn_j);
    int* cnt = reinterpret_cast<int*>(base + L.cnt);
    int* ptr = reinterpret_cast<int*>(base + L.ptr);
    EVX_SP_DISPATCH(m, launch_knn<MT>(obj, mask, (int)n, m, knn_d, knn_j, cnt, ptr, nn, nni, s));
    EVX_LAUNCH_CHECK();
    k_spea2_truncate<0><<<1, TR_T, 0, s>>>(obj, mask, (int)n, m, (int)k, nn, nni, knn_d, knn_j, cnt, ptr, alive, keep);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}
