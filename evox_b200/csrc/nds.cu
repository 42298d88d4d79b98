// nds.cu - non-dominated sort on B200: lexicographic pre-sort, tiled dominance test into a
// panel-major triangular bit matrix, and front peeling with carry-save vertical popcounts.
//
// Replaces (EvoX source tree): utils/common.py:88-97 and
// operators/selection/non_dominate.py:32-115.  See DESIGN.md for the derivation; in short:
//
//   1. Rows holding a NaN never dominate and are never dominated (IEEE compares) -> rank 0.
//   2. The remaining rows are sorted lexicographically by (f0, f1, ..) on order-preserving integer
//      keys (-0 folded onto +0) and de-duplicated: equal rows dominate nothing among themselves and
//      have identical dominator sets, hence identical ranks.  U unique rows remain.
//   3. For unique rows i, j in sorted order: i dominates j  <=>  i < j and f_k[i] <= f_k[j] for all
//      k >= 1  (f0 is implied by the order; strictness by uniqueness).  Only the strict upper
//      triangle exists, and each pair costs M-1 compares.
//   4. The bit matrix is stored by column PANELS of TW = 1024 columns: panel J holds rows
//      0 .. (J+1)*TW-1, each row PW = 32 words (128 B).  Packed triangular: N^2/16 bytes.  A panel
//      is owned by one rank (J mod world) and, inside a rank, by whoever processes that panel, so
//      dominator counts of a panel's columns are private to its owner.
//   5. Peeling: front k = columns whose count is 0.  Every front row is read exactly once per
//      panel (128 B line), rows are summed bit-sliced (carry-save adders), and the per-column sums
//      are subtracted from the counts.  Total peel traffic = the matrix, once.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "common.cuh"

namespace evx {

constexpr int TW = 1024;            // columns per panel
constexpr int PW = TW / 32;         // words per panel row (128 B)
constexpr int RB = 4;               // rows per lane in the tile kernel
constexpr int TILE_WARPS = 4;
constexpr int SLAB = 32 * RB;       // rows per warp
constexpr int CTA_ROWS = SLAB * TILE_WARPS;
constexpr int APPLY_WARPS = 8;
constexpr int ANY_SLOTS = 64;

struct Scal {
    int U;          // unique non-NaN rows
    int n_nan;      // rows holding a NaN
    int nf[2];      // size of the current front list, double buffered by front parity
    int any[ANY_SLOTS];
};

struct Layout {
    int64_t n, npad;
    int m, world, rank, np, npl, npl_max;
    size_t keys, idxA, idxB, skeyA, skeyB, flags, incl, cubtmp, cubtmp_bytes, ucols, row2u, count,
        rank_u, flist, mask, scal, bits, total;
    int64_t bits_words;
};

static inline int64_t panel_words_before(int64_t l, int world, int rank) {
    // sum_{l' < l} (l'*world + rank + 1) * TW * PW
    return (int64_t)TW * PW * ((int64_t)world * l * (l - 1) / 2 + l * (int64_t)(rank + 1));
}
__device__ __forceinline__ int64_t d_panel_words_before(int64_t l, int world, int rank) {
    return (int64_t)TW * PW * ((int64_t)world * l * (l - 1) / 2 + l * (int64_t)(rank + 1));
}

static size_t cub_temp_bytes(int64_t n) {
    size_t a = 0, b = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, a, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                    (const int*)nullptr, (int*)nullptr, (int)n);
    cub::DeviceScan::InclusiveSum(nullptr, b, (const int*)nullptr, (int*)nullptr, (int)n);
    return a > b ? a : b;
}

static Layout make_layout(int64_t n, int m, int world, int rank) {
    Layout L{};
    L.n = n; L.m = m; L.world = world; L.rank = rank;
    L.npad = ceil_div(n > 0 ? n : 1, TW) * TW;
    L.np = (int)(L.npad / TW);
    L.npl = (L.np - rank + world - 1) / world;
    if (L.npl < 0) L.npl = 0;
    L.npl_max = (L.np + world - 1) / world;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
    L.keys = take(sizeof(uint32_t) * (size_t)m * n);
    L.idxA = take(sizeof(int) * n);
    L.idxB = take(sizeof(int) * n);
    L.skeyA = take(sizeof(uint32_t) * n);
    L.skeyB = take(sizeof(uint32_t) * n);
    L.flags = take(sizeof(int) * n);
    L.incl = take(sizeof(int) * n);
    L.cubtmp_bytes = cub_temp_bytes(n);
    L.cubtmp = take(L.cubtmp_bytes);
    L.ucols = take(sizeof(float) * (size_t)(m > 1 ? m - 1 : 1) * L.npad);
    L.row2u = take(sizeof(int) * n);
    L.count = take(sizeof(int) * (size_t)L.npl_max * TW);
    L.rank_u = take(sizeof(int) * (size_t)L.npl_max * TW);
    L.flist = take(sizeof(int) * L.npad);
    L.mask = take(sizeof(uint32_t) * (size_t)L.npl_max * PW * world);
    L.scal = take(sizeof(Scal));
    L.bits_words = panel_words_before(L.npl, world, rank);
    L.bits = take(sizeof(uint32_t) * (size_t)L.bits_words);
    L.total = off;
    return L;
}

// ------------------------------------------------------------------------------------------------
// stage 1: keys, lexicographic order, unique rows
// ------------------------------------------------------------------------------------------------
__global__ void k_make_keys(const float* __restrict__ x, int64_t n, int m, uint32_t* __restrict__ keys,
                            int* __restrict__ idx, Scal* scal) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* row = x + i * m;
    bool has_nan = false;
    for (int k = 0; k < m; ++k) has_nan |= (row[k] != row[k]);
    for (int k = 0; k < m; ++k) keys[(int64_t)k * n + i] = has_nan ? 0xFFFFFFFFu : float_key(row[k]);
    idx[i] = (int)i;
    if (has_nan) atomicAdd(&scal->n_nan, 1);
}

__global__ void k_gather_key(const uint32_t* __restrict__ key, const int* __restrict__ idx, int64_t n,
                             uint32_t* __restrict__ out) {
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) out[p] = key[idx[p]];
}

__global__ void k_mark_heads(const uint32_t* __restrict__ keys, const int* __restrict__ idx, int64_t n, int m,
                             const Scal* __restrict__ scal, int* __restrict__ flags) {
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int64_t nvalid = n - scal->n_nan;
    int head = 0;
    if (p < nvalid) {
        if (p == 0) head = 1;
        else {
            const int a = idx[p], b = idx[p - 1];
            for (int k = 0; k < m; ++k) head |= (keys[(int64_t)k * n + a] != keys[(int64_t)k * n + b]);
        }
    }
    flags[p] = head;
}

__global__ void k_scatter_unique(const float* __restrict__ x, const int* __restrict__ idx,
                                 const int* __restrict__ incl, const int* __restrict__ flags, int64_t n, int m,
                                 int64_t npad, Scal* scal, float* __restrict__ ucols, int* __restrict__ row2u) {
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int64_t nvalid = n - scal->n_nan;
    const int row = idx[p];
    if (p >= nvalid) { row2u[row] = -1; if (nvalid == 0 && p == 0) scal->U = 0; return; }
    const int u = incl[p] - 1;
    row2u[row] = u;
    if (flags[p]) {
        for (int k = 1; k < m; ++k) ucols[(int64_t)(k - 1) * npad + u] = x[(int64_t)row * m + k] + 0.0f;
    }
    if (p == nvalid - 1) scal->U = u + 1;
}

// ------------------------------------------------------------------------------------------------
// stage 2: dominance tiles
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void csa(uint32_t& hi, uint32_t& lo, uint32_t a, uint32_t b, uint32_t c) {
    const uint32_t u = a ^ b;
    hi = (a & b) | (u & c);
    lo = u ^ c;
}

template <int MM>
__device__ __forceinline__ void load_col(const float* p, float (&v)[MM > 0 ? MM : 1]) {
    if constexpr (MM == 0) {
        v[0] = 0.f;
    } else if constexpr (MM % 4 == 0) {
#pragma unroll
        for (int t = 0; t < MM / 4; ++t) {
            float4 q = reinterpret_cast<const float4*>(p)[t];
            v[4 * t] = q.x; v[4 * t + 1] = q.y; v[4 * t + 2] = q.z; v[4 * t + 3] = q.w;
        }
    } else if constexpr (MM % 2 == 0) {
#pragma unroll
        for (int t = 0; t < MM / 2; ++t) {
            float2 q = reinterpret_cast<const float2*>(p)[t];
            v[2 * t] = q.x; v[2 * t + 1] = q.y;
        }
    } else {
#pragma unroll
        for (int t = 0; t < MM; ++t) v[t] = p[t];
    }
}

// One CTA: CTA_ROWS dominator rows x one panel (TW columns).  Lane <-> row (RB rows per lane, values
// in registers); the panel's column values sit in shared memory and are broadcast one column per
// step; each lane assembles its own 32-bit words.  Words are staged through shared memory so the
// 128-byte panel rows leave as full coalesced lines, and the same pass over the staged words feeds
// carry-save adders that produce the column counts.
template <int MM>
__global__ void __launch_bounds__(TILE_WARPS * 32)
k_dominance_tiles(const float* __restrict__ ucols, int64_t npad, const Scal* __restrict__ scal,
                  uint32_t* __restrict__ bits, int* __restrict__ count, int world, int rank) {
    const int l = blockIdx.y;
    const int J = l * world + rank;
    const int U = scal->U;
    const int col0 = J * TW;
    if (col0 >= U) return;
    const int rowlimit = min((J + 1) * TW, U);
    const int row0 = blockIdx.x * CTA_ROWS;
    if (row0 >= rowlimit) return;

    extern __shared__ __align__(16) uint32_t smem[];
    float* colv = reinterpret_cast<float*>(smem);                       // [TW][MM]
    uint32_t* stage = smem + (MM > 0 ? MM : 0) * TW;                    // [TILE_WARPS][PW][SLAB+1]
    int* scount = reinterpret_cast<int*>(stage + TILE_WARPS * PW * (SLAB + 1));   // [TW]

    if constexpr (MM > 0) {
        for (int t = threadIdx.x; t < MM * TW; t += blockDim.x) {
            const int c = t / MM, k = t - c * MM;
            colv[t] = ucols[(int64_t)k * npad + col0 + c];
        }
    }
    for (int t = threadIdx.x; t < TW; t += blockDim.x) scount[t] = 0;
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rs = row0 + warp * SLAB;                                  // first row of this warp's slab
    if (rs < rowlimit) {
        float xi[RB][MM > 0 ? MM : 1];
#pragma unroll
        for (int q = 0; q < RB; ++q)
#pragma unroll
            for (int k = 0; k < MM; ++k) xi[q][k] = ucols[(int64_t)k * npad + rs + q * 32 + lane];

        uint32_t* st = stage + warp * PW * (SLAB + 1);
        for (int wc = 0; wc < PW; ++wc) {
            const int cb = col0 + wc * 32;                              // first column of this word
            uint32_t w[RB];
#pragma unroll
            for (int q = 0; q < RB; ++q) w[q] = 0u;
            if (cb + 31 > rs) {                                         // else wholly on/below the diagonal
                const float* cp = colv + wc * 32 * MM;
#pragma unroll
                for (int b = 0; b < 32; ++b) {
                    float cv[MM > 0 ? MM : 1];
                    load_col<MM>(cp + b * MM, cv);
#pragma unroll
                    for (int q = 0; q < RB; ++q) {
                        bool p = true;
#pragma unroll
                        for (int k = 0; k < MM; ++k) p = p && (xi[q][k] <= cv[k]);
                        if (p) w[q] |= (1u << b);
                    }
                }
                if (MM == 0) {                                          // M == 1: only the order matters
#pragma unroll
                    for (int q = 0; q < RB; ++q) {
                        const int i = rs + q * 32 + lane;
                        uint32_t v = 0u;
                        for (int b = 0; b < 32; ++b) v |= (uint32_t)(i < U && cb + b < U) << b;
                        w[q] = v;
                    }
                }
                if (cb <= rs + SLAB - 1) {                              // touches the diagonal: keep j > i only
#pragma unroll
                    for (int q = 0; q < RB; ++q) {
                        const int i = rs + q * 32 + lane;
                        const int sh = i - cb + 1;                      // bits below sh are j <= i
                        w[q] = (sh <= 0) ? w[q] : (sh >= 32 ? 0u : (w[q] & (0xFFFFFFFFu << sh)));
                    }
                }
            }
#pragma unroll
            for (int q = 0; q < RB; ++q) st[wc * (SLAB + 1) + q * 32 + lane] = w[q];
        }
        __syncwarp();

        // copy-out (lane <-> word column) fused with the vertical popcount of the slab
        uint32_t* gout = bits + d_panel_words_before(l, world, rank) + (int64_t)rs * PW;
        uint32_t ones = 0, twos = 0, fours = 0, up[5] = {0, 0, 0, 0, 0};   // planes 8,16,32,64,128
        const uint32_t* sl = st + lane * (SLAB + 1);
        for (int rl = 0; rl < SLAB; rl += 8) {
            uint32_t v[8];
#pragma unroll
            for (int t = 0; t < 8; ++t) v[t] = sl[rl + t];
#pragma unroll
            for (int t = 0; t < 8; ++t) gout[(int64_t)(rl + t) * PW + lane] = v[t];
            uint32_t ta, tb, fa, fb, e;
            csa(ta, ones, ones, v[0], v[1]);
            csa(tb, ones, ones, v[2], v[3]);
            csa(fa, twos, twos, ta, tb);
            csa(ta, ones, ones, v[4], v[5]);
            csa(tb, ones, ones, v[6], v[7]);
            csa(fb, twos, twos, ta, tb);
            csa(e, fours, fours, fa, fb);
#pragma unroll
            for (int p = 0; p < 5; ++p) { const uint32_t c = up[p] & e; up[p] ^= e; e = c; }
        }
#pragma unroll 4
        for (int bb = 0; bb < 32; ++bb) {
            const int bit = (bb + lane) & 31;
            int v = ((ones >> bit) & 1) | (((twos >> bit) & 1) << 1) | (((fours >> bit) & 1) << 2);
#pragma unroll
            for (int p = 0; p < 5; ++p) v |= ((up[p] >> bit) & 1) << (3 + p);
            if (v) atomicAdd(&scount[lane * 32 + bit], v);
        }
    }
    __syncthreads();
    for (int t = threadIdx.x; t < TW; t += blockDim.x) {
        const int v = scount[t];
        if (v) atomicAdd(&count[(int64_t)l * TW + t], v);
    }
}

// Fallback for M-1 > MAX_TEMPLATE_MM: same tiling, objective loop at run time, one row per lane.
__global__ void __launch_bounds__(TILE_WARPS * 32)
k_dominance_tiles_generic(const float* __restrict__ ucols, int64_t npad, int mm, const Scal* __restrict__ scal,
                          uint32_t* __restrict__ bits, int* __restrict__ count, int world, int rank) {
    const int l = blockIdx.y;
    const int J = l * world + rank;
    const int U = scal->U;
    const int col0 = J * TW;
    if (col0 >= U) return;
    const int rowlimit = min((J + 1) * TW, U);
    const int row0 = blockIdx.x * (TILE_WARPS * 32);
    if (row0 >= rowlimit) return;
    extern __shared__ __align__(16) uint32_t smem[];
    float* colv = reinterpret_cast<float*>(smem);                       // [TW][mm]
    for (int t = threadIdx.x; t < mm * TW; t += blockDim.x) {
        const int c = t / mm, k = t - c * mm;
        colv[t] = ucols[(int64_t)k * npad + col0 + c];
    }
    __syncthreads();
    const int i = row0 + threadIdx.x;
    uint32_t* gout = bits + d_panel_words_before(l, world, rank) + (int64_t)i * PW;
    for (int wc = 0; wc < PW; ++wc) {
        const int cb = col0 + wc * 32;
        uint32_t w = 0u;
        for (int b = 0; b < 32; ++b) {
            const int j = cb + b;
            bool p = (j > i);
            for (int k = 0; k < mm && p; ++k) p = ucols[(int64_t)k * npad + i] <= colv[(wc * 32 + b) * mm + k];
            w |= (uint32_t)p << b;
            if (p) atomicAdd(&count[(int64_t)l * TW + wc * 32 + b], 1);
        }
        gout[wc] = w;
    }
}

// ------------------------------------------------------------------------------------------------
// stage 3: front peeling
// ------------------------------------------------------------------------------------------------
// Front k of this rank's columns: count == 0.  One thread per local column, one mask word per warp.
__global__ void k_front_detect(int* __restrict__ count, int* __restrict__ rank_u, uint32_t* __restrict__ mask,
                               const Scal* __restrict__ scal, int k, int world, int rank, int64_t ncols) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;      // ncols is a multiple of 32
    if (t >= ncols) return;
    const int64_t l = t / TW;
    const int64_t j = (l * world + rank) * TW + (t - l * TW);
    bool z = false;
    if (j < scal->U && count[t] == 0) {
        z = true;
        rank_u[t] = k;
        count[t] = -1;                       // never selected again (non_dominate.py:105-106)
    }
    const uint32_t w = __ballot_sync(0xFFFFFFFFu, z);
    if ((threadIdx.x & 31) == 0) mask[t >> 5] = w;
}

// All ranks' membership masks -> list of unique-row ids of front k (unordered).
__global__ void k_mask_to_list(const uint32_t* __restrict__ masks, int64_t mask_words, int world,
                               int* __restrict__ flist, Scal* scal, int k) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t == 0) scal->nf[(k + 1) & 1] = 0;
    if (t >= mask_words * world) return;
    uint32_t w = masks[t];
    if (!w) return;
    const int r = (int)(t / mask_words);
    const int64_t wl = t - (int64_t)r * mask_words;
    const int64_t l = wl / PW;
    const int64_t base_col = (l * world + r) * TW + (wl - l * PW) * 32;
    int slot = atomicAdd(&scal->nf[k & 1], __popc(w));
    while (w) {
        const int b = __ffs(w) - 1;
        w &= w - 1;
        flist[slot++] = (int)(base_col + b);
    }
}

// Subtract the front's rows from the counts of one panel.  grid = (owned panels, slices).  A warp
// reads one 128-byte row segment per step (lane <-> word), eight rows in flight, and adds them
// bit-sliced; the slab sums are expanded once per warp and merged through shared memory.
__global__ void __launch_bounds__(APPLY_WARPS * 32)
k_front_apply(const uint32_t* __restrict__ bits, int* __restrict__ count, const int* __restrict__ flist,
              const Scal* __restrict__ scal, int k, int world, int rank, int* any_out) {
    const int nf = scal->nf[k & 1];
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0 && any_out) *any_out = nf > 0;
    if (nf == 0) return;
    const int l = blockIdx.x;
    const int J = l * world + rank;
    const int U = scal->U;
    if (J * TW >= U) return;
    const int rowlimit = min((J + 1) * TW, U);
    __shared__ int scount[TW];
    for (int t = threadIdx.x; t < TW; t += blockDim.x) scount[t] = 0;
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t* pbase = bits + d_panel_words_before(l, world, rank) + lane;
    const int stride = gridDim.y * APPLY_WARPS;
    constexpr int NP_UP = 9;                                        // planes 8 .. 2048
    uint32_t ones = 0, twos = 0, fours = 0, up[NP_UP];
#pragma unroll
    for (int p = 0; p < NP_UP; ++p) up[p] = 0;
    int groups = 0;

    auto flush = [&]() {
#pragma unroll 2
        for (int bb = 0; bb < 32; ++bb) {
            const int bit = (bb + lane) & 31;
            int v = ((ones >> bit) & 1) | (((twos >> bit) & 1) << 1) | (((fours >> bit) & 1) << 2);
#pragma unroll
            for (int p = 0; p < NP_UP; ++p) v |= ((up[p] >> bit) & 1) << (3 + p);
            if (v) atomicAdd(&scount[lane * 32 + bit], v);
        }
        ones = twos = fours = 0;
#pragma unroll
        for (int p = 0; p < NP_UP; ++p) up[p] = 0;
        groups = 0;
    };

    for (int t0 = blockIdx.y * APPLY_WARPS + warp; t0 < nf; t0 += 8 * stride) {
        uint32_t v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int t = t0 + u * stride;
            int i = (t < nf) ? flist[t] : 0x7FFFFFFF;
            v[u] = (i < rowlimit) ? __ldg(pbase + (int64_t)i * PW) : 0u;
        }
        uint32_t ta, tb, fa, fb, e;
        csa(ta, ones, ones, v[0], v[1]);
        csa(tb, ones, ones, v[2], v[3]);
        csa(fa, twos, twos, ta, tb);
        csa(ta, ones, ones, v[4], v[5]);
        csa(tb, ones, ones, v[6], v[7]);
        csa(fb, twos, twos, ta, tb);
        csa(e, fours, fours, fa, fb);
#pragma unroll
        for (int p = 0; p < NP_UP; ++p) { const uint32_t c = up[p] & e; up[p] ^= e; e = c; }
        if (++groups == 500) flush();                               // 4000 rows < 2^12 - 1
    }
    if (groups) flush();
    __syncthreads();
    for (int t = threadIdx.x; t < TW; t += blockDim.x) {
        const int v = scount[t];
        if (v) atomicSub(&count[(int64_t)l * TW + t], v);
    }
}

__global__ void k_write_ranks(const int* __restrict__ row2u, const int* __restrict__ rank_u, int64_t n, int world,
                              int rank, int* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int u = row2u[i];
    int r = 0;
    if (u >= 0) {
        const int J = u / TW;
        if (J % world == rank) r = rank_u[(int64_t)(J / world) * TW + (u - J * TW)];
    }
    out[i] = r;
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
struct Plan {
    Layout L;
    char* ws;
    template <class T> T* at(size_t off) const { return reinterpret_cast<T*>(ws + off); }
};

template <int MM>
static int launch_tiles(const Plan& P, cudaStream_t s) {
    const Layout& L = P.L;
    const size_t smem = sizeof(uint32_t) * ((size_t)MM * TW + (size_t)TILE_WARPS * PW * (SLAB + 1) + TW);
    EVX_CUDA(cudaFuncSetAttribute(k_dominance_tiles<MM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)ceil_div(L.npad, CTA_ROWS), (unsigned)L.npl);
    k_dominance_tiles<MM><<<grid, TILE_WARPS * 32, smem, s>>>(P.at<float>(L.ucols), L.npad, P.at<Scal>(L.scal),
                                                              P.at<uint32_t>(L.bits), P.at<int>(L.count), L.world,
                                                              L.rank);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}

static int build(const Plan& P, const float* x, cudaStream_t s, cudaEvent_t tiles_begin = nullptr,
                 cudaEvent_t tiles_end = nullptr) {
    const Layout& L = P.L;
    const int64_t n = L.n;
    const int m = L.m;
    const int T = 256;
    const unsigned gn = (unsigned)ceil_div(n, T);
    Scal* scal = P.at<Scal>(L.scal);
    EVX_CUDA(cudaMemsetAsync(scal, 0, sizeof(Scal), s));
    EVX_CUDA(cudaMemsetAsync(P.at<char>(L.ucols), 0xFF, sizeof(float) * (size_t)(m > 1 ? m - 1 : 1) * L.npad, s));
    EVX_CUDA(cudaMemsetAsync(P.at<char>(L.count), 0, sizeof(int) * (size_t)L.npl_max * TW, s));
    EVX_CUDA(cudaMemsetAsync(P.at<char>(L.rank_u), 0, sizeof(int) * (size_t)L.npl_max * TW, s));

    uint32_t* keys = P.at<uint32_t>(L.keys);
    int* idx_cur = P.at<int>(L.idxA);
    int* idx_nxt = P.at<int>(L.idxB);
    k_make_keys<<<gn, T, 0, s>>>(x, n, m, keys, idx_cur, scal);
    EVX_LAUNCH_CHECK();
    for (int k = m - 1; k >= 0; --k) {            // LSD passes of a stable sort = lexicographic order
        k_gather_key<<<gn, T, 0, s>>>(keys + (int64_t)k * n, idx_cur, n, P.at<uint32_t>(L.skeyA));
        EVX_LAUNCH_CHECK();
        size_t tb = L.cubtmp_bytes;
        EVX_CUDA(cub::DeviceRadixSort::SortPairs(P.at<void>(L.cubtmp), tb, P.at<uint32_t>(L.skeyA),
                                                 P.at<uint32_t>(L.skeyB), idx_cur, idx_nxt, (int)n, 0, 32, s));
        int* t = idx_cur; idx_cur = idx_nxt; idx_nxt = t;
    }
    k_mark_heads<<<gn, T, 0, s>>>(keys, idx_cur, n, m, scal, P.at<int>(L.flags));
    EVX_LAUNCH_CHECK();
    size_t tb = L.cubtmp_bytes;
    EVX_CUDA(cub::DeviceScan::InclusiveSum(P.at<void>(L.cubtmp), tb, P.at<int>(L.flags), P.at<int>(L.incl), (int)n, s));
    k_scatter_unique<<<gn, T, 0, s>>>(x, idx_cur, P.at<int>(L.incl), P.at<int>(L.flags), n, m, L.npad, scal,
                                      P.at<float>(L.ucols), P.at<int>(L.row2u));
    EVX_LAUNCH_CHECK();

    if (L.npl == 0) return EVX_OK;
    const int mm = m - 1;
    int rc = EVX_OK;
    if (tiles_begin) EVX_CUDA(cudaEventRecord(tiles_begin, s));
    switch (mm) {
        case 0: rc = launch_tiles<0>(P, s); break;
        case 1: rc = launch_tiles<1>(P, s); break;
        case 2: rc = launch_tiles<2>(P, s); break;
        case 3: rc = launch_tiles<3>(P, s); break;
        case 4: rc = launch_tiles<4>(P, s); break;
        case 5: rc = launch_tiles<5>(P, s); break;
        case 6: rc = launch_tiles<6>(P, s); break;
        case 7: rc = launch_tiles<7>(P, s); break;
        case 8: rc = launch_tiles<8>(P, s); break;
        case 9: rc = launch_tiles<9>(P, s); break;
        case 10: rc = launch_tiles<10>(P, s); break;
        case 11: rc = launch_tiles<11>(P, s); break;
        default: {
            const size_t smem = sizeof(float) * (size_t)mm * TW;
            EVX_CUDA(cudaFuncSetAttribute(k_dominance_tiles_generic, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)smem));
            dim3 grid((unsigned)ceil_div(L.npad, TILE_WARPS * 32), (unsigned)L.npl);
            k_dominance_tiles_generic<<<grid, TILE_WARPS * 32, smem, s>>>(
                P.at<float>(L.ucols), L.npad, mm, scal, P.at<uint32_t>(L.bits), P.at<int>(L.count), L.world, L.rank);
            EVX_LAUNCH_CHECK();
        }
    }
    if (tiles_end && rc == EVX_OK) EVX_CUDA(cudaEventRecord(tiles_end, s));
    return rc;
}

static int detect(const Plan& P, int k, uint32_t* mask, cudaStream_t s) {
    const Layout& L = P.L;
    const int64_t ncols = (int64_t)L.npl_max * TW;
    k_front_detect<<<(unsigned)ceil_div(ncols, 256), 256, 0, s>>>(P.at<int>(L.count), P.at<int>(L.rank_u), mask,
                                                                 P.at<Scal>(L.scal), k, L.world, L.rank, ncols);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}

static int apply_slices(const Layout& L) {
    int s = (int)ceil_div(4 * 148, L.npl > 0 ? L.npl : 1);
    return s < 1 ? 1 : (s > 64 ? 64 : s);
}

static int apply(const Plan& P, int k, const uint32_t* masks, int* any_out, cudaStream_t s) {
    const Layout& L = P.L;
    const int64_t mw = (int64_t)L.npl_max * PW;
    k_mask_to_list<<<(unsigned)ceil_div(mw * L.world, 256), 256, 0, s>>>(masks, mw, L.world, P.at<int>(L.flist),
                                                                        P.at<Scal>(L.scal), k);
    EVX_LAUNCH_CHECK();
    dim3 grid((unsigned)(L.npl > 0 ? L.npl : 1), (unsigned)apply_slices(L));
    k_front_apply<<<grid, APPLY_WARPS * 32, 0, s>>>(P.at<uint32_t>(L.bits), P.at<int>(L.count), P.at<int>(L.flist),
                                                    P.at<Scal>(L.scal), k, L.world, L.rank, any_out);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}

static thread_local int32_t g_last_fronts = 0;

// Optional per-stage timing (CUDA events on the caller's stream) for bench.py's roofline line.
struct Prof {
    bool on = false, have = false;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};   // start, tiles begin, tiles end, end
    float prep_ms = 0, tiles_ms = 0, peel_ms = 0;
};
static thread_local Prof g_prof;

}  // namespace evx

using namespace evx;

extern "C" size_t evx_nds_shard_workspace_bytes(int64_t n, int32_t m, int32_t world, int32_t rank) {
    if (n < 0 || m < 1 || m > EVX_MAX_OBJECTIVES || world < 1 || rank < 0 || rank >= world || n > 0x7FFFFFFF - 2 * TW)
        return 0;
    return make_layout(n, m, world, rank).total + 256;
}

extern "C" size_t evx_nds_workspace_bytes(int64_t n, int32_t m) { return evx_nds_shard_workspace_bytes(n, m, 1, 0); }

struct evx_nds_shard {
    Plan P;
};

extern "C" evx_nds_shard* evx_nds_shard_create(int64_t n, int32_t m, int32_t world, int32_t rank, void* ws,
                                               size_t ws_bytes) {
    const size_t need = evx_nds_shard_workspace_bytes(n, m, world, rank);
    if (need == 0) { set_error("evx_nds_shard_create: bad arguments n=%lld m=%d world=%d rank=%d", (long long)n, m, world, rank); return nullptr; }
    if (!ws || ws_bytes < need) { set_error("evx_nds_shard_create: workspace %zu < required %zu", ws_bytes, need); return nullptr; }
    evx_nds_shard* p = new (std::nothrow) evx_nds_shard();
    if (!p) { set_error("out of host memory"); return nullptr; }
    p->P.L = make_layout(n, m, world, rank);
    p->P.ws = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(ws), 256));
    return p;
}

extern "C" void evx_nds_shard_destroy(evx_nds_shard* p) { delete p; }

extern "C" int evx_nds_shard_build(evx_nds_shard* p, const float* x, void* stream) {
    EVX_REQUIRE(p && (x || p->P.L.n == 0), EVX_ERR_ARG, "evx_nds_shard_build: null argument");
    if (p->P.L.n == 0) return EVX_OK;
    return build(p->P, x, (cudaStream_t)stream);
}

extern "C" int64_t evx_nds_shard_mask_words(const evx_nds_shard* p) { return p ? (int64_t)p->P.L.npl_max * PW : 0; }

extern "C" int evx_nds_shard_detect(evx_nds_shard* p, int32_t k, uint32_t* local_mask, void* stream) {
    EVX_REQUIRE(p && local_mask, EVX_ERR_ARG, "evx_nds_shard_detect: null argument");
    return detect(p->P, k, local_mask, (cudaStream_t)stream);
}

extern "C" int evx_nds_shard_apply(evx_nds_shard* p, int32_t k, const uint32_t* all_masks, int32_t* any_out,
                                   void* stream) {
    EVX_REQUIRE(p && all_masks, EVX_ERR_ARG, "evx_nds_shard_apply: null argument");
    return apply(p->P, k, all_masks, any_out, (cudaStream_t)stream);
}

extern "C" int evx_nds_shard_ranks(evx_nds_shard* p, int32_t* rank_partial, void* stream) {
    EVX_REQUIRE(p && (rank_partial || p->P.L.n == 0), EVX_ERR_ARG, "evx_nds_shard_ranks: null argument");
    const Layout& L = p->P.L;
    if (L.n == 0) return EVX_OK;
    k_write_ranks<<<(unsigned)ceil_div(L.n, 256), 256, 0, (cudaStream_t)stream>>>(
        p->P.at<int>(L.row2u), p->P.at<int>(L.rank_u), L.n, L.world, L.rank, rank_partial);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}

extern "C" int32_t evx_nds_last_num_fronts(void) { return g_last_fronts; }

extern "C" int evx_nds_f32(const float* x, int64_t n, int32_t m, int32_t* rank, void* ws, size_t ws_bytes,
                           void* stream) {
    EVX_REQUIRE(n >= 0 && m >= 1 && m <= EVX_MAX_OBJECTIVES, EVX_ERR_ARG, "evx_nds_f32: bad shape n=%lld m=%d",
                (long long)n, m);
    g_last_fronts = 0;
    if (n == 0) return EVX_OK;
    EVX_REQUIRE(x && rank, EVX_ERR_ARG, "evx_nds_f32: null pointer");
    const size_t need = evx_nds_workspace_bytes(n, m);
    EVX_REQUIRE(need != 0, EVX_ERR_ARG, "evx_nds_f32: n=%lld too large", (long long)n);
    EVX_REQUIRE(ws && ws_bytes >= need, EVX_ERR_WORKSPACE, "evx_nds_f32: workspace %zu < required %zu", ws_bytes, need);
    cudaStream_t s = (cudaStream_t)stream;
    Plan P;
    P.L = make_layout(n, m, 1, 0);
    P.ws = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(ws), 256));
    Prof& pf = g_prof;
    pf.have = false;
    if (pf.on) {
        for (int i = 0; i < 4; ++i)
            if (!pf.ev[i]) EVX_CUDA(cudaEventCreate(&pf.ev[i]));
        EVX_CUDA(cudaEventRecord(pf.ev[0], s));
    }
    int rc = build(P, x, s, pf.on ? pf.ev[1] : nullptr, pf.on ? pf.ev[2] : nullptr);
    if (rc) return rc;

    // Peel.  The trip count is data dependent (non_dominate.py:78-80): fronts are enqueued in
    // batches and one flag per front is read back after each batch; steps past the last front are
    // no-ops (nothing has count 0 any more).
    Scal* scal = P.at<Scal>(P.L.scal);
    uint32_t* mask = P.at<uint32_t>(P.L.mask);
    int host_any[ANY_SLOTS];
    int k = 0, batch = 8;
    for (;;) {
        for (int b = 0; b < batch; ++b) {
            if ((rc = detect(P, k + b, mask, s))) return rc;
            if ((rc = apply(P, k + b, mask, &scal->any[b], s))) return rc;
        }
        EVX_CUDA(cudaMemcpyAsync(host_any, scal->any, sizeof(int) * batch, cudaMemcpyDeviceToHost, s));
        EVX_CUDA(cudaStreamSynchronize(s));
        int b = 0;
        while (b < batch && host_any[b]) ++b;
        if (b < batch) { g_last_fronts = k + b; break; }
        k += batch;
        if (batch < ANY_SLOTS) batch *= 2;
    }
    k_write_ranks<<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(P.at<int>(P.L.row2u), P.at<int>(P.L.rank_u), n, 1, 0, rank);
    EVX_LAUNCH_CHECK();
    if (pf.on) EVX_CUDA(cudaEventRecord(pf.ev[3], s));
    EVX_CUDA(cudaStreamSynchronize(s));
    if (pf.on) {
        EVX_CUDA(cudaEventElapsedTime(&pf.prep_ms, pf.ev[0], pf.ev[1]));
        EVX_CUDA(cudaEventElapsedTime(&pf.tiles_ms, pf.ev[1], pf.ev[2]));
        EVX_CUDA(cudaEventElapsedTime(&pf.peel_ms, pf.ev[2], pf.ev[3]));
        pf.have = true;
    }
    return EVX_OK;
}

extern "C" void evx_profile_enable(int32_t on) { g_prof.on = on != 0; }

extern "C" int evx_profile_last(float* prep_ms, float* tiles_ms, float* peel_ms) {
    EVX_REQUIRE(g_prof.have, EVX_ERR_ARG, "evx_profile_last: no profiled evx_nds_f32 call on this thread");
    if (prep_ms) *prep_ms = g_prof.prep_ms;
    if (tiles_ms) *tiles_ms = g_prof.tiles_ms;
    if (peel_ms) *peel_ms = g_prof.peel_ms;
    return EVX_OK;
}

extern "C" int evx_nds_host_f32(const float* x_host, int64_t n, int32_t m, int32_t* rank_host) {
    EVX_REQUIRE(n >= 0 && m >= 1 && m <= EVX_MAX_OBJECTIVES, EVX_ERR_ARG, "evx_nds_host_f32: bad shape");
    if (n == 0) { g_last_fronts = 0; return EVX_OK; }
    EVX_REQUIRE(x_host && rank_host, EVX_ERR_ARG, "evx_nds_host_f32: null pointer");
    const size_t wsb = evx_nds_workspace_bytes(n, m);
    EVX_REQUIRE(wsb != 0, EVX_ERR_ARG, "evx_nds_host_f32: n too large");
    const size_t xb = align_up(sizeof(float) * (size_t)n * m, 256), rb = align_up(sizeof(int) * (size_t)n, 256);
    void* base = nullptr;
    int rc = arena_get(xb + rb + wsb, &base);
    if (rc) return rc;
    float* dx = reinterpret_cast<float*>(base);
    int* dr = reinterpret_cast<int*>(reinterpret_cast<char*>(base) + xb);
    void* ws = reinterpret_cast<char*>(base) + xb + rb;
    EVX_CUDA(cudaMemcpyAsync(dx, x_host, sizeof(float) * (size_t)n * m, cudaMemcpyHostToDevice, 0));
    rc = evx_nds_f32(dx, n, m, dr, ws, wsb, nullptr);
    if (rc) return rc;
    EVX_CUDA(cudaMemcpy(rank_host, dr, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost));
    return EVX_OK;
}
