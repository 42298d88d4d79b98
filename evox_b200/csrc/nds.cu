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
//   3. Labels: blocks of TW = 1024 consecutive lexicographic positions; inside a block rows are
//      ordered by f1; inside each run of 32 (one bit-matrix word) by f2.
//   4. For labelled rows i, j: block(i) < block(j)  =>  i dominates j  <=>  f_k[i] <= f_k[j] for all
//      k >= 1 (f0 by the order, strictness by uniqueness); block(i) > block(j) => never; same block:
//      all M objectives.  Only the block-upper triangle exists.
//   5. Tiles, M = 3: per-panel tables (k_panel_tables) turn a row's f2 and f1 tests against a whole
//      32-column word into one table byte + shift and one selector + byte permute (tile_offdiag_m3,
//      lane <-> word).  Other M: a slab of rows (32, or 64 for M > 3) against a word is classified by
//      its f1 ranges (skip / f1 passes / mixed); f2 is a binary-searched suffix mask; what remains is
//      compared on the FMA pipe (sign of b - a).
//   6. The bit matrix is stored by column PANELS of TW columns, two panels side by side: pair P holds
//      rows 0 .. (2P+2)*TW-1 as 256-byte super-rows [panel 2P | panel 2P+1] (a random gather of
//      256-byte rows runs at full HBM bandwidth, one of 128-byte rows does not).  Packed: N^2/16
//      bytes.  A pair and the dominator counts of its columns are owned by one rank (P mod world).
//   7. Peeling: front k = columns whose count is 0.  Every front row is read exactly once per pair,
//      rows are summed bit-sliced (carry-save adders), subtracted from bit-sliced counts, and
//      "count == 0" is the NOR of the planes.  One persistent cooperative kernel (k_peel_dynamic) peels
//      all fronts and re-balances the work over the CTAs for every front; on several GPUs the same
//      kernel also exchanges front membership over NVLink peer memory.  Fronts are separated by data
//      dependence on epoch-tagged words, not by a barrier.
//   8. Up to 8192 rows on one GPU two single-CTA kernels replace the launch- and latency-bound parts:
//      k_small_sort_unique (step 2 in one launch) and k_peel_single (step 7 without flags between CTAs).
//
// File map: geometry and workspace layout - sort / unique (k_small_sort_unique) / labels / panel tables -
// tile kernels (word_fail, f2_mask, slab_out, counts_flush, tile_offdiag, tile_offdiag_m3, tile_diag) -
// one-front-per-launch peel (k_front_detect / k_mask_to_list / k_front_apply) -
// tagged front words and the persistent load-balanced peel (front_row, peel_commit_seg, publish_pending,
// merge_tree, k_peel_dynamic) - k_peel_single - host side (launch_tiles, build, detect / apply,
// peel_persistent) - C entry points.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <cstdlib>
#include <cstring>

#include <type_traits>

#include "common.cuh"

namespace evx {

constexpr int TW = 1024;            // columns per panel
constexpr int PW = TW / 32;         // words per panel row (128 B)
// Rows per lane in the tile kernel.  M <= 3: one row per lane (more CTAs per SM hide the latency of the
// f2 searches: 19.1 -> 17.7 ms at N = 1M); M > 3: two, so that every column value staged for the
// sign-bit compares serves two rows (N = 100k, M = 10: 1.64 vs 1.78 ms).  A warp always covers 128 rows.
__host__ __device__ constexpr int tile_rb(int m) { return m <= 3 ? 1 : 2; }
constexpr int TILE_WARPS = 8;
// f2 rank table of a panel: entry [g][w] = how many of the panel's g smallest-f2 columns lie in word w
// (one byte, 0..32), g = 0..TW.  See k_panel_tables.
constexpr int F2TAB_BYTES = (TW + 1) * PW;
// Per-panel table blob used by the M = 3 tile kernel (k_panel_tables), byte offsets:
constexpr int PT_SF2 = 0;                                  // float [TW]      the panel's f2 values, sorted, in level order (panel_rank)
constexpr int PT_S1 = PT_SF2 + 4 * TW;                     // float [PW][32]  per word: its columns' f1 values, ascending
constexpr int PT_TAB = PT_S1 + 4 * TW;                     // uint8 [TW+1][PW] f2 rank -> columns of word w below that rank
constexpr int PT_M1 = PT_TAB + F2TAB_BYTES;                // uint32 [PW][33] per word: columns whose f1 rank is >= k
constexpr int PT_BYTES = PT_M1 + 4 * PW * 33;
static_assert(PT_BYTES % 16 == 0 && PT_TAB % 16 == 0, "table blob is copied in 16-byte pieces");
constexpr int TILE_ROWS_PER_WARP = 128;                 // slabs per warp = 128 / (32 * rows per lane)
constexpr int APPLY_WARPS = 8;
constexpr int SMALL_PATH_MAX_N = 1024;   // populations of at most one panel take the direct small path
constexpr int PEEL_MAX_GRID = 1024;  // upper bound on the CTAs of the persistent peel kernel
constexpr int PEEL_KMAX = 24;      // count planes the load-balanced peel kernel supports (N < 16M)
constexpr int ANY_SLOTS = 64;

struct Scal {
    int U;          // unique non-NaN rows
    int n_nan;      // rows holding a NaN
    int nf[2];      // one-front-per-launch peel: size of the current front list, double buffered by front parity
    int any[ANY_SLOTS];
    int fronts;     // persistent peel: number of fronts found (-1: a peer timed out)
    unsigned int arrived;   // persistent peel, one GPU: panels published so far in this launch (see PeerState::arrived)
    unsigned int pad0;
    unsigned long long dbg[16];  // persistent peel: ns spent by CTA 0 / last CTA in accumulate, merge+commit, waiting (see evx_debug_peel_times)
};

struct Layout {
    int64_t n, npad;
    int m, world, rank, np, npl, npl_max;
    size_t keys, idxA, idxB, skeyA, skeyB, flags, incl, cubtmp, cubtmp_bytes, ucols, row2u, count,
        rank_u, flist, flist2, cplanes, mask, scal, bits, total, tcols, label, partials, pdone, pcnt, ptab, ctadbg;
    int kplanes;
    int64_t bits_words;
};

// ---- geometry -------------------------------------------------------------------------------------
// Panels are stored, owned and peeled in PAIRS: the two panels 2P, 2P+1 share 256-byte "super-rows"
// [panel 2P: 32 words | panel 2P+1: 32 words] for rows 0 .. (2P+2)*TW-1.  A random gather of 256-byte
// rows runs at full HBM bandwidth on B200 (~6.9 TB/s) where 128-byte rows top out at 4.0-4.7 TB/s
// (profiles/r1_gather_probe.txt), and the peel is exactly such a gather.  Pair P belongs to rank
// P mod world; local panel index l = 2*lp + h (lp = local pair, h = half).
constexpr int PW2 = 2 * PW;          // words per super-row

__host__ __device__ __forceinline__ int panel_of_local(int l, int world, int rank) {
    return 2 * ((l >> 1) * world + rank) + (l & 1);
}
__host__ __device__ __forceinline__ int64_t pair_words_before(int64_t lp, int world, int rank) {
    // sum_{l' < lp} (2*(l'*world + rank) + 2) * TW * PW2
    return (int64_t)TW * PW2 * ((int64_t)world * lp * (lp - 1) + 2 * lp * (int64_t)(rank + 1));
}
// word offset of (local panel l, row i, word 0)
__host__ __device__ __forceinline__ int64_t bits_offset(int l, int64_t row, int world, int rank) {
    return pair_words_before(l >> 1, world, rank) + row * PW2 + (l & 1) * PW;
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
    const int npairs = (L.np + 1) / 2;
    int nlp = (npairs - rank + world - 1) / world;                 // pairs owned by this rank
    if (nlp < 0) nlp = 0;
    L.npl = 2 * nlp;                                               // local panel slots (the last may not exist)
    L.npl_max = 2 * ((npairs + world - 1) / world);
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
    L.ucols = take(sizeof(float) * (size_t)m * L.npad);
    L.tcols = take(sizeof(float) * (size_t)m * L.npad);
    L.label = take(sizeof(int) * n);
    L.ptab = take(m == 3 ? (size_t)PT_BYTES * (size_t)L.np : 0);              // per-panel tables of the M = 3 tile kernel
    L.row2u = take(sizeof(int) * n);
    L.count = take(sizeof(int) * (size_t)L.npl_max * TW);
    L.rank_u = take(sizeof(int) * (size_t)L.npl_max * TW);
    L.flist = take(sizeof(uint64_t) * L.npad);                   // front segments (64-bit tagged words), double buffered;
    L.flist2 = take(sizeof(uint64_t) * L.npad);                  // the one-front-per-launch peel uses the first as an int list
    L.kplanes = 1;
    while ((1ll << L.kplanes) <= L.npad) ++L.kplanes;          // counts < npad fit in kplanes bits
    L.cplanes = take(sizeof(uint32_t) * (size_t)L.npl_max * (L.kplanes + 1) * PW);
    L.mask = take(sizeof(uint32_t) * (size_t)L.npl_max * PW * world);
    L.partials = take(sizeof(uint32_t) * (size_t)(PEEL_MAX_GRID + L.npl_max / 2 + 1) * PEEL_KMAX * 64);
    L.pcnt = take(sizeof(uint64_t) * 2 * (size_t)(L.np + 2));          // per-panel front sizes (tagged words), double buffered
    L.pdone = take(sizeof(unsigned int) * 2 * (size_t)L.npl_max);      // arrivals, commits
    L.ctadbg = take(sizeof(unsigned long long) * 4 * PEEL_MAX_GRID);   // per-CTA phase clocks of the peel (evx_debug_peel_cta_times)
    L.scal = take(sizeof(Scal));
    L.bits_words = pair_words_before(L.npl / 2, world, rank);
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
                                 int64_t npad, Scal* scal, float* __restrict__ tcols, int* __restrict__ row2u) {
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int64_t nvalid = n - scal->n_nan;
    const int row = idx[p];
    if (p >= nvalid) { row2u[row] = -1; if (nvalid == 0 && p == 0) scal->U = 0; return; }
    const int u = incl[p] - 1;
    row2u[row] = u;
    if (flags[p]) {
        for (int k = 0; k < m; ++k) tcols[(int64_t)k * npad + u] = x[(int64_t)row * m + k] + 0.0f;
    }
    if (p == nvalid - 1) scal->U = u + 1;
}

// ---- small populations: sort + unique in ONE launch ----------------------------------------------------
// Up to 8192 rows, while the keys fit into one CTA's shared memory (M <= 6 at 8192 rows, M <= 13 at 4096, ...), this replaces
// the radix path: there the M stable radix sorts cost 7 launches each and every onesweep pass
// ~10 us of latency on a handful of tiles (12 of them at M = 3: 0.13 ms at N = 6000, a quarter of the
// whole sort).  Here one CTA holds all keys in shared memory and bitonic-sorts the row indices by
// (key_0, .., key_{M-1}, index) - a total order, so the result is the stable lexicographic order the
// radix passes give - then marks the heads of runs of equal keys, scans, and scatters the unique rows
// (k_mark_heads + InclusiveSum + k_scatter_unique).
constexpr int PREP_SMALL_MAX_N = 8192;            // (16 384 rows fit at M = 3, but 105 bitonic stages are no faster than the radix path)
constexpr size_t PREP_SMALL_MAX_SMEM = 227 * 1024 - 256;   // keys [m][n] + row indices: 8192 rows x 6 objectives, 4096 x 13, ...
constexpr int SU_T = 1024, SU_ITEMS = PREP_SMALL_MAX_N / SU_T;
static size_t small_prep_smem(int64_t n, int m) {
    int64_t pw2 = 1;
    while (pw2 < n) pw2 <<= 1;
    return sizeof(uint32_t) * (size_t)m * (size_t)n + sizeof(uint16_t) * (size_t)pw2;
}

__global__ void __launch_bounds__(SU_T, 1)
k_small_sort_unique(const float* __restrict__ x, const uint32_t* __restrict__ keys, int n, int m, int64_t npad, Scal* scal,
                    float* __restrict__ tcols, int* __restrict__ row2u) {
    extern __shared__ uint32_t ss_smem[];             // [m][n] keys, then P row indices (uint16)
    __shared__ int wsum[SU_T / 32];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    int P = 1;
    while (P < n) P <<= 1;                            // sorted length: rows + sentinels (index >= n: after every row)
    uint32_t* ks = ss_smem;
    uint16_t* idx = reinterpret_cast<uint16_t*>(ss_smem + (size_t)m * n);
    for (int i = t; i < m * n; i += SU_T) ks[i] = keys[i];
    for (int i = t; i < P; i += SU_T) idx[i] = (uint16_t)i;
    __syncthreads();
    auto before = [&](int a, int b) -> bool {         // does row a come before row b?
        if (a >= n || b >= n) return a < b;
        for (int k = 0; k < m; ++k) {
            const uint32_t ka = ks[k * n + a], kb = ks[k * n + b];
            if (ka != kb) return ka < kb;
        }
        return a < b;
    };
    for (int k = 2; k <= P; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int q = t; q < (P >> 1); q += SU_T) {
                const int i = ((q & ~(j - 1)) << 1) | (q & (j - 1));      // q-th index with bit j clear
                const int a = idx[i], b = idx[i | j];
                const bool up = (i & k) == 0;
                if (before(b, a) == up) { idx[i] = (uint16_t)b; idx[i | j] = (uint16_t)a; }
            }
            __syncthreads();
        }
    }
    // ---- heads of runs, scan, unique rows out ----------------------------------------------------------
    const int nvalid = n - scal->n_nan;
    int cnt = 0;
    uint32_t heads = 0;
#pragma unroll
    for (int e = 0; e < SU_ITEMS; ++e) {
        const int p = t * SU_ITEMS + e;
        int head = 0;
        if (p < nvalid) {
            if (p == 0) head = 1;
            else {
                const int a = idx[p], b = idx[p - 1];
                for (int k = 0; k < m; ++k) head |= (ks[k * n + a] != ks[k * n + b]) ? 1 : 0;
            }
        }
        heads |= (uint32_t)head << e;
        cnt += head;
    }
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    int run = incl - cnt, total = 0;
    for (int w = 0; w < SU_T / 32; ++w) {
        const int v = wsum[w];
        if (w < warp) run += v;
        total += v;
    }
#pragma unroll
    for (int e = 0; e < SU_ITEMS; ++e) {
        const int p = t * SU_ITEMS + e;
        if (p < n) {
            const int row = idx[p];
            if (p >= nvalid) {
                row2u[row] = -1;
            } else {
                const int head = (heads >> e) & 1;
                run += head;
                const int u = run - 1;
                row2u[row] = u;
                if (head)
                    for (int k = 0; k < m; ++k) tcols[(size_t)k * npad + u] = x[(size_t)row * m + k] + 0.0f;
            }
        }
    }
    if (t == 0) scal->U = total;
}

// ------------------------------------------------------------------------------------------------
// stage 1b: block-local re-labelling
// ------------------------------------------------------------------------------------------------
// Unique rows are cut into blocks of TW consecutive lexicographic positions, and inside each block
// they are re-ordered by f1.  Between different blocks the lexicographic order still decides f0
// (lower block => f0 <=), so off-diagonal tiles test objectives 1..M-1 only; and because both the
// rows of a warp's slab and the 32 columns of a word are now contiguous in f1, whole (slab, word)
// sub-tiles are decided by their f1 ranges.  Inside a block either direction may dominate: diagonal
// tiles test all M.  Second level: inside a block, every run of 32 consecutive (f1-ordered) positions -
// one bit-matrix word - is re-ordered by f2, so that "f2_row <= f2_col" over a word is a suffix mask.
// One CTA per block: bitonic sort of (f1 key, position) in shared memory, then warp w ranks the 32 rows
// of word w by f2 with shuffles; rows are written to their label (ucols) and label_of_u is filled.
// (Any order with these two properties gives the same ranks; ties are broken by position.)
__global__ void __launch_bounds__(TW)
k_block_labels(const float* __restrict__ tcols, int64_t npad, int m, const Scal* __restrict__ scal,
               float* __restrict__ ucols, int* __restrict__ label_of_u) {
    __shared__ unsigned long long key[TW];
    const int b = blockIdx.x, t = threadIdx.x, lane = t & 31;
    const int U = scal->U;
    const int64_t base = (int64_t)b * TW;
    if (base >= U) return;
    const bool valid = base + t < U;
    key[t] = valid ? (((unsigned long long)(m > 1 ? float_key(tcols[npad + base + t]) : 0u) << 32) | (unsigned int)t) : ~0ull;
    __syncthreads();
    for (int k = 2; k <= TW; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            const int o = t ^ j;
            if (o > t) {
                const unsigned long long a = key[t], c = key[o];
                const bool up = (t & k) == 0;
                if ((a > c) == up) { key[t] = c; key[o] = a; }
            }
            __syncthreads();
        }
    }
    const unsigned long long kk = key[t];                                  // position t of the f1 order holds row `src`
    const bool have = kk != ~0ull;
    const int src = (int)(unsigned int)kk;
    int q = t;                                                              // final position inside the block
    if (m >= 3) {
        const uint32_t k2 = have ? float_key(tcols[2 * npad + base + src]) : 0xFFFFFFFFu;
        int rk = 0;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            const uint32_t o = __shfl_sync(0xFFFFFFFFu, k2, i);
            rk += (o < k2) || (o == k2 && i < lane);
        }
        q = (t & ~31) + rk;                                                 // missing rows (largest key) stay at the end
    }
    if (have) {
        label_of_u[base + src] = (int)(base + q);
        for (int k = 0; k < m; ++k) ucols[(int64_t)k * npad + base + q] = tcols[(int64_t)k * npad + base + src];
    }
}

__global__ void k_relabel_rows(int* __restrict__ row2u, const int* __restrict__ label_of_u, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int u = row2u[i];
    if (u >= 0) row2u[i] = label_of_u[u];
}

// Level-order (Eytzinger) slot of sorted position p (0-based) among TW values: the first TW - 1 form a
// perfect tree (node k has children 2k, 2k + 1), the largest value sits in slot 0.  See panel_rank.
__device__ __forceinline__ int eytzinger_slot(int p) {
    if (p == TW - 1) return 0;
    const int x = p + 1, tz = __ffs(x) - 1;              // in-order number 1 .. TW - 1
    return (1 << (9 - tz)) + (x >> (tz + 1));
}

// ------------------------------------------------------------------------------------------------
// stage 1c: per-panel tables (M = 3)
// ------------------------------------------------------------------------------------------------
// f2: for a row r and a word w of panel J (32 columns, ascending f2), the columns with f2_r <= f2_c form
// a suffix of the word; its start is the number of the word's columns with f2_c < f2_r.  Instead of one
// 5-step binary search per (row, word) - 32 per row and panel - the row is searched ONCE in the panel's
// 1024 f2 values (sf2, ascending): g = #{c in panel : f2_c < f2_r}, and tab[g][w] says how many of those
// g columns belong to word w.  (g always falls on a boundary between distinct values, so ties in the
// sorted order do not matter.)
// f1: the words of a panel are consecutive runs of the block's f1 order, so for a row only ONE word
// needs an f1 test per column; for it the word's f1 values are kept sorted (s1) with, for every rank k,
// the mask m1[w][k] of the word's columns whose f1 rank is >= k: a 5-step search and one load.
// One CTA per panel: bitonic sort of (f2 key, column) in shared memory, warp w walks the sorted order
// and counts its own word's columns; warp w also ranks its word's 32 f1 values.
__global__ void __launch_bounds__(TW)
k_panel_tables(const float* __restrict__ ucols, int64_t npad, const Scal* __restrict__ scal, uint8_t* __restrict__ ptab) {
    __shared__ unsigned long long key[TW];
    __shared__ uint8_t wordof[TW];
    __shared__ __align__(16) uint8_t tab[F2TAB_BYTES];
    const int J = blockIdx.x, t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int U = scal->U;
    const int c = J * TW + t;
    const float* f1 = ucols + npad;
    const float* f2 = ucols + 2 * npad;
    uint8_t* blob = ptab + (size_t)J * PT_BYTES;
    // ---- f1: rank of my column inside my word (ties by column), sorted values and masks ----
    {
        const float v = (c < U) ? f1[c] : __int_as_float(0x7FC00000);
        const uint32_t k1 = (c < U) ? float_key(v) : 0xFFFFFFFFu;
        int rk = 0;
#pragma unroll
        for (int b = 0; b < 32; ++b) {
            const uint32_t o = __shfl_sync(0xFFFFFFFFu, k1, b);
            rk += (o < k1) || (o == k1 && b < lane);
        }
        reinterpret_cast<float*>(blob + PT_S1)[warp * 32 + rk] = v;         // NaN for missing columns: never "< r"
        uint32_t* m1 = reinterpret_cast<uint32_t*>(blob + PT_M1) + warp * 33;
        for (int k = 0; k <= 32; ++k) {
            const uint32_t mk = __ballot_sync(0xFFFFFFFFu, rk >= k);
            if (lane == 0) m1[k] = mk;
        }
    }
    // ---- f2 ----
    key[t] = (c < U) ? (((unsigned long long)float_key(f2[c]) << 32) | (unsigned int)t) : ~0ull;
    __syncthreads();
    for (int k = 2; k <= TW; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            const int o = t ^ j;
            if (o > t) {
                const unsigned long long a = key[t], b = key[o];
                const bool up = (t & k) == 0;
                if ((a > b) == up) { key[t] = b; key[o] = a; }
            }
            __syncthreads();
        }
    }
    const unsigned long long kk = key[t];
    const bool valid = kk != ~0ull;
    const int col = (int)(unsigned int)kk;                                  // column within the panel
    reinterpret_cast<float*>(blob + PT_SF2)[eytzinger_slot(t)] = valid ? f2[J * TW + col] : __int_as_float(0x7FC00000);   // NaN: never "< r"
    wordof[t] = valid ? (uint8_t)(col >> 5) : (uint8_t)255;
    __syncthreads();
    int base = 0;                                                           // warp <-> word
    if (lane == 0) tab[warp] = 0;                                           // g = 0
    for (int ch = 0; ch < TW / 32; ++ch) {
        const int p = ch * 32 + lane;
        const uint32_t b = __ballot_sync(0xFFFFFFFFu, wordof[p] == warp);
        tab[(p + 1) * PW + warp] = (uint8_t)(base + __popc(b & (0xFFFFFFFFu >> (31 - lane))));
        base += __popc(b);
    }
    __syncthreads();
    uint4* dst = reinterpret_cast<uint4*>(blob + PT_TAB);
    const uint4* src = reinterpret_cast<const uint4*>(tab);
    for (int i = t; i < F2TAB_BYTES / 16; i += TW) dst[i] = src[i];
}

// ------------------------------------------------------------------------------------------------
// stage 2: dominance tiles
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void csa(uint32_t& hi, uint32_t& lo, uint32_t a, uint32_t b, uint32_t c) {
    const uint32_t u = a ^ b;
    hi = (a & b) | (u & c);
    lo = u ^ c;
}

template <int C>
__device__ __forceinline__ void load_col(const float* p, float (&v)[C]) {
    if constexpr (C % 4 == 0) {
#pragma unroll
        for (int t = 0; t < C / 4; ++t) {
            float4 q = reinterpret_cast<const float4*>(p)[t];
            v[4 * t] = q.x; v[4 * t + 1] = q.y; v[4 * t + 2] = q.z; v[4 * t + 3] = q.w;
        }
    } else if constexpr (C % 2 == 0) {
#pragma unroll
        for (int t = 0; t < C / 2; ++t) {
            float2 q = reinterpret_cast<const float2*>(p)[t];
            v[2 * t] = q.x; v[2 * t + 1] = q.y;
        }
    } else {
#pragma unroll
        for (int t = 0; t < C; ++t) v[t] = p[t];
    }
}

// The compare runs on the FMA pipe: for finite or infinite a, b (no NaN reaches here)
//     a <= b   <=>   sign(b - a) == 0
// (b - a is exactly signed: it rounds to zero only when a == b, and inf - inf gives the canonical
// positive NaN, i.e. "pass", which is what inf <= inf says).  The fail bits of the objectives are
// OR-ed (LOP3) and shifted into the word with one funnel shift (SHF), instead of FSETP chains on the
// half-rate ALU pipe.
//
// Fail bits of 32 columns x RB rows for objective f1 (if F1) and the CR objectives 3 .. M-1.
// f1w: the word's 32 f1 values (SoA); restw: its [32][CR] remaining values.
template <int CR, bool F1, int RB>
__device__ __forceinline__ void word_fail(const float* f1w, const float* restw, const float (&xf1)[RB],
                                          const float (&xr)[RB][CR > 0 ? CR : 1], uint32_t (&w)[RB]) {
#pragma unroll
    for (int q = 0; q < RB; ++q) w[q] = 0u;
    if constexpr (!F1 && CR == 0) return;
    if constexpr (CR == 0) {                          // f1 only: four columns per 16-byte broadcast load
        // four independent shift chains per row (8 columns each) instead of one of 32
        uint32_t a[RB][4];
#pragma unroll
        for (int q = 0; q < RB; ++q)
#pragma unroll
            for (int g = 0; g < 4; ++g) a[q][g] = 0u;
#pragma unroll
        for (int h = 1; h >= 0; --h) {
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const float4 v = *reinterpret_cast<const float4*>(f1w + g * 8 + h * 4);
#pragma unroll
                for (int q = 0; q < RB; ++q) {
                    a[q][g] = __funnelshift_l(__float_as_uint(v.w - xf1[q]), a[q][g], 1);
                    a[q][g] = __funnelshift_l(__float_as_uint(v.z - xf1[q]), a[q][g], 1);
                    a[q][g] = __funnelshift_l(__float_as_uint(v.y - xf1[q]), a[q][g], 1);
                    a[q][g] = __funnelshift_l(__float_as_uint(v.x - xf1[q]), a[q][g], 1);
                }
            }
        }
#pragma unroll
        for (int q = 0; q < RB; ++q) w[q] = a[q][0] | (a[q][1] << 8) | (a[q][2] << 16) | (a[q][3] << 24);
    } else {
#pragma unroll
        for (int b = 31; b >= 0; --b) {
            float cv[CR];
            load_col<CR>(restw + b * CR, cv);
            const float c1 = F1 ? f1w[b] : 0.0f;
#pragma unroll
            for (int q = 0; q < RB; ++q) {
                uint32_t f = F1 ? __float_as_uint(c1 - xf1[q]) : 0u;
#pragma unroll
                for (int k = 0; k < CR; ++k) f |= __float_as_uint(cv[k] - xr[q][k]);
                w[q] = __funnelshift_l(f, w[q], 1);   // (w << 1) | sign(f)
            }
        }
    }
}

// g = #{c : f2_c < r} over a panel's TW f2 values.  The first TW - 1 sorted values are stored in level
// order (Eytzinger: node k has children 2k, 2k + 1; slot 0 holds the largest value), so that the 32 lanes'
// probes of one step fall into a contiguous range and spread over the banks - the same search over the
// plainly sorted array has every lane probe an index = step - 1 (mod 2 step): one bank, up to 32 wavefronts
// per step (measured: 95 of the 180 shared-memory wavefronts per slab).  NaN padding never compares below.
__device__ __forceinline__ int panel_rank(uint32_t saddr, const float* e, float r) {
    uint32_t a = saddr + 4u;                             // node 1
#pragma unroll
    for (int level = 0; level < 10; ++level) {
        float v;
        asm("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
        a = 2u * a - saddr;                              // left child: base + 4 * 2k
        if (v < r) a += 4u;                              // right child
    }
    int g = (int)((a - saddr) >> 2) - TW;                // leaf number = values below r among the first TW - 1
    if (g == TW - 1 && e[0] < r) g = TW;
    return g;
}
// Columns of a word are in ascending f2: the pairs with f2_row <= f2_col form a suffix of the word.
// Lower bound among the first 31 columns by walking a shared-memory BYTE address (load, compare,
// predicated add: three instructions per step - the compiler's version kept an element index and
// rebuilt the address every step); the compare with the largest column is off the dependent chain.
// `caddr` must be a multiple of 128.
// `c15` and `c31` are the word's 16th and last value, loaded by the caller: their addresses are the
// same for every lane and for both rows of a lane, so one load serves two searches and the first step
// is a compare and a select.
__device__ __forceinline__ uint32_t f2_mask(uint32_t caddr, float c15, float c31, float r) {
    uint32_t a;
    asm("{\n\t.reg .pred p;\n\t.reg .f32 v;\n\t"
        "mov.u32 %0, %2;\n\t"
        "setp.lt.f32 p, %3, %1;\n\t@p add.u32 %0, %0, 64;\n\t"
        "ld.shared.f32 v, [%0+28];\n\tsetp.lt.f32 p, v, %1;\n\t@p add.u32 %0, %0, 32;\n\t"
        "ld.shared.f32 v, [%0+12];\n\tsetp.lt.f32 p, v, %1;\n\t@p add.u32 %0, %0, 16;\n\t"
        "ld.shared.f32 v, [%0+4];\n\tsetp.lt.f32 p, v, %1;\n\t@p add.u32 %0, %0, 8;\n\t"
        "ld.shared.f32 v, [%0];\n\tsetp.lt.f32 p, v, %1;\n\t@p add.u32 %0, %0, 4;\n\t}"
        : "=&r"(a)
        : "f"(r), "r"(caddr), "f"(c15));
    // columns with f2 < r among the first 31 = (a - caddr) / 4 = (a / 4) mod 32: the word's values
    // start on a 128-byte boundary, and the funnel shift takes its count mod 32
    return (c31 < r) ? 0u : __funnelshift_l(0u, 0xFFFFFFFFu, a >> 2);
}

constexpr int TILE_NUP = 6;                            // count planes 8 .. 256 (a warp's rows: 128, or 256 for M = 3)

struct CountPlanes {
    uint32_t ones = 0, twos = 0, fours = 0, up[TILE_NUP] = {};
};

// Copy one staged slab out (lane <-> word column: full 128-byte lines, one half of each 256-byte
// super-row) and fold it into the planes.  `zero_sibling`: also clear the other half of the super-rows
// (rows of an odd panel's own block have no bits in the even panel of the pair).
template <int SLAB>
__device__ __forceinline__ void slab_out(const uint32_t* sl, uint32_t* gout, int lane, CountPlanes& cp,
                                         bool zero_sibling = false, uint32_t skipm = 0u) {
    const bool dead = (skipm >> lane) & 1u;            // this lane's word was skipped: all zero, never staged
    for (int rl = 0; rl < SLAB; rl += 8) {
        uint32_t v[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) v[t] = dead ? 0u : sl[rl + t];
#pragma unroll
        for (int t = 0; t < 8; ++t) gout[(int64_t)(rl + t) * PW2 + lane] = v[t];
        if (zero_sibling) {
#pragma unroll
            for (int t = 0; t < 8; ++t) gout[(int64_t)(rl + t) * PW2 - PW + lane] = 0u;
        }
        uint32_t ta, tb, fa, fb, e;
        csa(ta, cp.ones, cp.ones, v[0], v[1]);
        csa(tb, cp.ones, cp.ones, v[2], v[3]);
        csa(fa, cp.twos, cp.twos, ta, tb);
        csa(ta, cp.ones, cp.ones, v[4], v[5]);
        csa(tb, cp.ones, cp.ones, v[6], v[7]);
        csa(fb, cp.twos, cp.twos, ta, tb);
        csa(e, cp.fours, cp.fours, fa, fb);
#pragma unroll
        for (int p = 0; p < TILE_NUP; ++p) { const uint32_t c = cp.up[p] & e; cp.up[p] ^= e; e = c; }
    }
}

// Column counts of the CTA's tile: the eight warps' planes are added bit-sliced in shared memory
// (three rounds, a plane more each round), and only the sum is turned into integers - one global
// atomic per non-zero column.  `pl` is scratch of TILE_WARPS * FLUSH_PLANES * 32 words that no warp
// uses any more (the caller has synchronised).
constexpr int FLUSH_PLANES = 3 + TILE_NUP + 3;         // 8 warps x 256 rows: counts up to 2048
__device__ __forceinline__ void counts_flush(const CountPlanes& cp, int lane, int warp, uint32_t* pl,
                                             int* __restrict__ count, int l) {
    uint32_t* mine = pl + warp * FLUSH_PLANES * 32 + lane;
    mine[0] = cp.ones; mine[32] = cp.twos; mine[64] = cp.fours;
#pragma unroll
    for (int p = 0; p < TILE_NUP; ++p) mine[(3 + p) * 32] = cp.up[p];
#pragma unroll
    for (int p = 3 + TILE_NUP; p < FLUSH_PLANES; ++p) mine[p * 32] = 0u;
    __syncthreads();
#pragma unroll
    for (int half = TILE_WARPS / 2; half >= 1; half >>= 1) {
        if (warp < half) {
            const uint32_t* other = mine + half * FLUSH_PLANES * 32;
            uint32_t carry = 0;
#pragma unroll
            for (int p = 0; p < FLUSH_PLANES; ++p) {
                const uint32_t a = mine[p * 32], b = other[p * 32];
                const uint32_t x = a ^ b;
                mine[p * 32] = x ^ carry;
                carry = (a & b) | (x & carry);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < PW / TILE_WARPS; ++i) {              // warp w: words w, w + 8, ...; lane <-> column in the word
        const int wd = warp + TILE_WARPS * i;
        int v = 0;
#pragma unroll
        for (int p = 0; p < FLUSH_PLANES; ++p) v |= (int)((pl[p * 32 + wd] >> lane) & 1u) << p;
        if (v) atomicAdd(&count[(int64_t)l * TW + wd * 32 + lane], v);
    }
}

// ---- off-diagonal CTA: rows of a lower block against one panel ------------------------------------
// Lane <-> row (RB rows per lane, values in registers); the panel's column values sit in shared
// memory (f1 and f2 as separate arrays, objectives 3.. interleaved) and are broadcast.  Two passes
// per 64-row slab.  Pass 1, every word: classify by f1 range (skip / f1 passes / mixed), and for the
// words not skipped find the f2 suffix mask by binary search - the iterations are independent, so
// four searches are in flight per lane.  Pass 2, only the words that still need compares: fail bits
// on the FMA pipe, AND-ed into the staged word.  Then the slab leaves as full 128-byte lines while
// carry-save adders accumulate the column counts.
template <int M>
__device__ __forceinline__ void tile_offdiag(const float* __restrict__ ucols, int64_t npad, int U, int l, int col0,
                                             int row0, int rowlimit, int slabs, uint32_t* __restrict__ bits,
                                             int* __restrict__ count, int world, int rank, uint32_t* smem) {
    constexpr int RB = tile_rb(M), SLAB = 32 * RB;
    constexpr int CR = (M > 3) ? M - 3 : 0;
    constexpr int NF = (M > 1) ? M - 1 : 1;                                // staged objective arrays
    float* f1s = reinterpret_cast<float*>(smem);                           // [TW]
    float* f2s = f1s + ((M >= 3) ? TW : 0);                                // [TW]
    float* rest = f2s + TW;                                                // [TW][CR]
    uint32_t* stage = smem + NF * TW;                                      // [TILE_WARPS][PW][SLAB+1]
    float* wmin = reinterpret_cast<float*>(stage + TILE_WARPS * PW * (SLAB + 1));   // [PW] f1 range of each word
    float* wmax = wmin + PW;
    uint32_t* wcm = reinterpret_cast<uint32_t*>(wmax + PW);                // [PW] valid-column mask of each word

    for (int t = threadIdx.x; t < TW; t += blockDim.x) {
        if constexpr (M >= 2) f1s[t] = ucols[npad + col0 + t];
        if constexpr (M >= 3) f2s[t] = ucols[2 * npad + col0 + t];
#pragma unroll
        for (int k = 0; k < CR; ++k) rest[t * CR + k] = ucols[(int64_t)(3 + k) * npad + col0 + t];
    }
    __syncthreads();
    if (threadIdx.x < PW) {
        const int ncol = min(32, U - (col0 + (int)threadIdx.x * 32));      // valid columns in this word
        wcm[threadIdx.x] = (ncol >= 32) ? 0xFFFFFFFFu : (ncol <= 0 ? 0u : ((1u << ncol) - 1u));
        if constexpr (M >= 2) {
            float lo = INFINITY, hi = -INFINITY;
            for (int c = 0; c < 32; ++c) {                                  // fminf / fmaxf skip the NaN padding
                const float v = f1s[threadIdx.x * 32 + c];
                lo = fminf(lo, v);
                hi = fmaxf(hi, v);
            }
            wmin[threadIdx.x] = lo;
            wmax[threadIdx.x] = hi;
        }
    }
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* st = stage + warp * PW * (SLAB + 1);
    const uint32_t f2addr = (uint32_t)__cvta_generic_to_shared(f2s);
    CountPlanes cp;

    for (int sidx = 0; sidx < slabs; ++sidx) {
        const int rs = row0 + (sidx * TILE_WARPS + warp) * SLAB;            // first row of this slab
        if (rs >= rowlimit) break;                                          // warp-uniform
        float xf1[RB], xf2[RB], xr[RB][CR > 0 ? CR : 1];
#pragma unroll
        for (int q = 0; q < RB; ++q) {
            const int64_t i = rs + q * 32 + lane;                           // rows of a lower block: all valid
            xf1[q] = (M >= 2) ? ucols[npad + i] : 0.0f;
            xf2[q] = (M >= 3) ? ucols[2 * npad + i] : 0.0f;
#pragma unroll
            for (int k = 0; k < CR; ++k) xr[q][k] = ucols[(int64_t)(3 + k) * npad + i];
        }
        float rmin = xf1[0], rmax = xf1[0];                                 // f1 range of the slab
#pragma unroll
        for (int q = 1; q < RB; ++q) { rmin = fminf(rmin, xf1[q]); rmax = fmaxf(rmax, xf1[q]); }
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            rmin = fminf(rmin, __shfl_xor_sync(0xFFFFFFFFu, rmin, o));
            rmax = fmaxf(rmax, __shfl_xor_sync(0xFFFFFFFFu, rmax, o));
        }

        // Phase A: classify the 32 words (warp-uniform masks).  Lane w tests word w against the slab.
        uint32_t skipm, need_full, need_rest = 0u;
        {
            bool skip = wcm[lane] == 0u, f1pass = true;
            if constexpr (M >= 2) {
                skip = skip || (rmin > wmax[lane]);                         // every row's f1 > every column's
                f1pass = rmax <= wmin[lane];                                // f1 passes for every pair
            }
            skipm = __ballot_sync(0xFFFFFFFFu, skip);
            need_full = __ballot_sync(0xFFFFFFFFu, !skip && !f1pass);
            if constexpr (CR > 0) need_rest = ~skipm & ~need_full;
        }
        // Phase B: f2 suffix masks of the words that are not skipped, two words (four searches) in flight.
        // (words are in ascending f1 inside the panel, so the skipped ones are a prefix and the live ones a
        // range: a counted loop.  A skipped word inside the range would only cost work: slab_out ignores it.)
        const uint32_t live = ~skipm;
        const int wlo = live ? __ffs(live) - 1 : 0, whi = live ? 32 - __clz(live) : 0;
        auto phase_b = [&](auto full_tag) {
            constexpr bool FULLP = decltype(full_tag)::value;          // every column of the panel exists: no validity mask
            for (int wa = wlo; wa < whi; wa += 2) {
                const int wb = min(wa + 1, whi - 1);
                uint32_t ma[RB], mb[RB];
                if constexpr (M >= 3) {
                    const uint32_t aa = f2addr + wa * 128, ab = f2addr + wb * 128;
                    const float a15 = f2s[wa * 32 + 15], a31 = f2s[wa * 32 + 31];
                    const float b15 = f2s[wb * 32 + 15], b31 = f2s[wb * 32 + 31];
#pragma unroll
                    for (int q = 0; q < RB; ++q) {
                        ma[q] = f2_mask(aa, a15, a31, xf2[q]);
                        mb[q] = f2_mask(ab, b15, b31, xf2[q]);
                    }
                } else {
#pragma unroll
                    for (int q = 0; q < RB; ++q) ma[q] = mb[q] = 0xFFFFFFFFu;
                }
                const uint32_t ca = FULLP ? 0xFFFFFFFFu : wcm[wa], cb2 = FULLP ? 0xFFFFFFFFu : wcm[wb];
#pragma unroll
                for (int q = 0; q < RB; ++q) {
                    st[wa * (SLAB + 1) + q * 32 + lane] = ma[q] & ca;
                    st[wb * (SLAB + 1) + q * 32 + lane] = mb[q] & cb2;
                }
            }
        };
        if (col0 + TW <= U) phase_b(std::true_type{});
        else phase_b(std::false_type{});
        while (need_full) {
            const int wc = __ffs(need_full) - 1;
            need_full &= need_full - 1;
            uint32_t f[RB];
            word_fail<CR, true, RB>(f1s + wc * 32, rest + wc * 32 * CR, xf1, xr, f);
#pragma unroll
            for (int q = 0; q < RB; ++q) st[wc * (SLAB + 1) + q * 32 + lane] &= ~f[q];
        }
        if constexpr (CR > 0) {
            while (need_rest) {
                const int wc = __ffs(need_rest) - 1;
                need_rest &= need_rest - 1;
                uint32_t f[RB];
                word_fail<CR, false, RB>(f1s + wc * 32, rest + wc * 32 * CR, xf1, xr, f);
#pragma unroll
                for (int q = 0; q < RB; ++q) st[wc * (SLAB + 1) + q * 32 + lane] &= ~f[q];
            }
        }
        __syncwarp();
        slab_out<SLAB>(st + lane * (SLAB + 1), bits + bits_offset(l, rs, world, rank), lane, cp, false, skipm);
        __syncwarp();
    }
    __syncthreads();                                                       // every warp is done with its stage
    counts_flush(cp, lane, warp, stage, count, l);
}

// ---- off-diagonal CTA, M = 3: lane <-> word, no transposition --------------------------------------
// The panel's columns are ordered by f1 at word granularity (word w holds the w-th run of 32 columns of
// the block's f1 order) and by f2 inside a word.  For a row r of a lower block (f0 decided by the
// order) that leaves, per word:
//   words entirely below r in f1            -> no pair passes              (w < ws_r)
//   the ONE word whose f1 range contains r  -> mask m1[ws_r][k_r]          (k_r = rank of f1_r in that word)
//   words entirely at or above r in f1      -> f1 passes for every pair    (w > ws_r)
// and f2_r <= f2_c is a suffix of the word whose start is a table byte: tab[g_r][w], g_r = rank of f2_r
// among the panel's f2 values (k_panel_tables).  So a slab of 32 rows takes one pass with lane <-> row
// (g_r by a 10-step search, ws_r by a 5-step search over the words' smallest f1, k_r by a 5-step search
// in word ws_r's sorted f1 values), which leaves one 16-byte record per row in shared memory, and then
// one pass with lane <-> WORD over the 32 rows: the record arrives by one broadcast load, the lane reads
// ONE table byte, shifts, masks, stores its word of row r (the warp stores the row's full 128-byte
// line) and feeds it to the carry-save column counters - no shared-memory staging of results at all.
__device__ __forceinline__ void tile_offdiag_m3(const uint8_t* __restrict__ ptab, const float* __restrict__ ucols,
                                                int64_t npad, int U, int l, int col0, int row0, int rowlimit, int slabs,
                                                uint32_t* __restrict__ bits, int* __restrict__ count, int world, int rank,
                                                uint32_t* smem) {
    unsigned char* sb = reinterpret_cast<unsigned char*>(smem);
    const float* f2s = reinterpret_cast<const float*>(sb + PT_SF2);        // [TW]
    const float* s1 = reinterpret_cast<const float*>(sb + PT_S1);          // [PW][32]
    // the f1 rank masks are read once per row: they stay in global memory (L1) so that five CTAs fit on an SM
    const uint32_t* m1 = reinterpret_cast<const uint32_t*>(ptab + (size_t)(col0 / TW) * PT_BYTES + PT_M1);   // [PW][33]
    uint32_t* wcm = reinterpret_cast<uint32_t*>(sb + PT_M1);               // [PW] valid-column mask of each word
    uint32_t* seltab = wcm + PW;                                           // [64] byte selectors by (word - ws_r + 32), see below
    uint2* rec = reinterpret_cast<uint2*>(sb + PT_M1 + 4 * PW + 256);      // [TILE_WARPS][32] per-row records
    uint32_t* flush = smem;                                                // counts_flush scratch: reuses the (then dead) tables
    {
        const uint4* src = reinterpret_cast<const uint4*>(ptab + (size_t)(col0 / TW) * PT_BYTES);
        uint4* dst = reinterpret_cast<uint4*>(sb);
        for (int t = threadIdx.x; t < PT_M1 / 16; t += blockDim.x) dst[t] = src[t];
    }
    if (threadIdx.x < PW) {
        const int ncol = min(32, U - (col0 + (int)threadIdx.x * 32));      // valid columns in this word
        wcm[threadIdx.x] = (ncol >= 32) ? 0xFFFFFFFFu : (ncol <= 0 ? 0u : ((1u << ncol) - 1u));
    }
    // The f1 decision of word w for a row whose straddle word is ws_r: all pairs fail (w < ws_r), the row's
    // mask m1 applies (w == ws_r), all pairs pass (w > ws_r).  As a byte selector for PRMT over {m1, 0x000000FF}:
    // 0x5555 picks the zero byte four times, 0x3210 passes m1 through, 0x4444 picks 0xFF four times - one
    // shared-memory load (index w - ws_r + 32) and one PRMT instead of two compares and two selects on the
    // ALU pipe, which is the pipe this kernel saturates.
    if (threadIdx.x < 64) seltab[threadIdx.x] = threadIdx.x < 32 ? 0x5555u : (threadIdx.x == 32 ? 0x3210u : 0x4444u);
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t f2addr = (uint32_t)__cvta_generic_to_shared(f2s);
    const uint32_t tabaddr = (uint32_t)__cvta_generic_to_shared(sb + PT_TAB);
    const uint32_t seladdr = (uint32_t)__cvta_generic_to_shared(seltab);
    const uint32_t mywcm = wcm[lane];
    const uint32_t lane4 = 4u * (uint32_t)lane;
    uint2* myrec = rec + warp * 32;
    const uint32_t recaddr = (uint32_t)__cvta_generic_to_shared(myrec);
    CountPlanes cp;

    for (int sidx = 0; sidx < slabs; ++sidx) {
        const int rs = row0 + (sidx * TILE_WARPS + warp) * 32;              // first row of this slab
        if (rs >= rowlimit) break;                                          // warp-uniform
        // ---- lane <-> row ----
        const float xf1 = ucols[npad + rs + lane], xf2 = ucols[2 * npad + rs + lane];   // rows of a lower block: all valid
        const int g = panel_rank(f2addr, f2s, xf2);
        int ws = 0;                                                         // last word whose smallest f1 is < xf1 (0 if none)
#pragma unroll
        for (int step = PW / 2; step >= 1; step >>= 1)
            if (s1[(ws + step) * 32] < xf1) ws += step;                     // (NaN = word without columns: never)
        // words below ws: every f1 < xf1 (word ranges ascend); words above: every f1 >= xf1; word ws: by rank
        const float* sw = s1 + ws * 32;
        int k = 0;                                                          // columns of word ws with f1 < xf1
#pragma unroll
        for (int step = 16; step >= 1; step >>= 1)
            if (sw[k + step - 1] < xf1) k += step;
        if (k == 31 && sw[31] < xf1) k = 32;
        __syncwarp();                                                       // the previous slab's records have been read
        // (both shared-memory addresses are below 64 K: the CTA has 44 KB)
        myrec[lane] = make_uint2((tabaddr + (uint32_t)g * PW) | ((seladdr + 4u * (uint32_t)(32 - ws)) << 16), __ldg(m1 + ws * 33 + k));
        __syncwarp();
        // ---- lane <-> word ----
        uint32_t* gout = bits + bits_offset(l, rs, world, rank) + lane;
        uint32_t e8[4];                                                     // carries out of the fours plane, one per 8 rows
#pragma unroll
        for (int r0 = 0; r0 < 32; r0 += 8) {
            uint32_t v[8];
#pragma unroll
            for (int t = 0; t < 8; ++t) {
                uint32_t rp, rm;                                            // table row address | selector address of word 0 << 16, m1 mask
                asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(rp), "=r"(rm) : "r"(recaddr + 8u * (uint32_t)(r0 + t)));
                const uint32_t ra = rp & 0xFFFFu, rs2 = rp >> 16;
                uint32_t cnt, sel;                                          // columns of my word with f2 < the row's; f1 selector
                asm volatile("ld.shared.u8 %0, [%1];" : "=r"(cnt) : "r"(ra + (uint32_t)lane));
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(sel) : "r"(rs2 + lane4));
                uint32_t mk;
                asm("shf.l.clamp.b32 %0, %1, %2, %3;" : "=r"(mk) : "r"(0u), "r"(0xFFFFFFFFu), "r"(cnt));   // ~0 << cnt, 0 at 32
                v[t] = mk & mywcm & __byte_perm(rm, 0xFFu, sel);
            }
#pragma unroll
            for (int t = 0; t < 8; ++t) gout[(int64_t)(r0 + t) * PW2] = v[t];
            uint32_t ta, tb, fa, fb;
            csa(ta, cp.ones, cp.ones, v[0], v[1]);
            csa(tb, cp.ones, cp.ones, v[2], v[3]);
            csa(fa, cp.twos, cp.twos, ta, tb);
            csa(ta, cp.ones, cp.ones, v[4], v[5]);
            csa(tb, cp.ones, cp.ones, v[6], v[7]);
            csa(fb, cp.twos, cp.twos, ta, tb);
            csa(e8[r0 >> 3], cp.fours, cp.fours, fa, fb);
        }
        {                                                                   // the slab's four eights-carries: one ripple per slab
            uint32_t sa, sb2, c32;
            csa(sa, cp.up[0], cp.up[0], e8[0], e8[1]);
            csa(sb2, cp.up[0], cp.up[0], e8[2], e8[3]);
            csa(c32, cp.up[1], cp.up[1], sa, sb2);
#pragma unroll
            for (int p = 2; p < TILE_NUP; ++p) { const uint32_t c = cp.up[p] & c32; cp.up[p] ^= c32; c32 = c; }
        }
    }
    __syncthreads();
    counts_flush(cp, lane, warp, flush, count, l);
}

// ---- diagonal CTA: a block against its own panel ---------------------------------------------------
// Either direction may dominate inside a block, so all M objectives are tested (FSETP chain) and
// only i == j is excluded (rows are unique).  1/NP of the work.
template <int M>
__device__ __forceinline__ void tile_diag(const float* __restrict__ ucols, int64_t npad, int U, int l, int col0,
                                          int row0, int rowlimit, int slabs, uint32_t* __restrict__ bits,
                                          int* __restrict__ count, int world, int rank, uint32_t* smem) {
    constexpr int RB = tile_rb(M), SLAB = 32 * RB;
    float* colv = reinterpret_cast<float*>(smem);                          // [TW][M]
    uint32_t* stage = smem + M * TW;
    for (int t = threadIdx.x; t < M * TW; t += blockDim.x) {
        const int c = t / M, k = t - c * M;
        colv[t] = ucols[(int64_t)k * npad + col0 + c];
    }
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* st = stage + warp * PW * (SLAB + 1);
    CountPlanes cp;
    for (int sidx = 0; sidx < slabs; ++sidx) {
        const int rs = row0 + (sidx * TILE_WARPS + warp) * SLAB;
        if (rs >= rowlimit) break;
        float xi[RB][M];
#pragma unroll
        for (int q = 0; q < RB; ++q)
#pragma unroll
            for (int k = 0; k < M; ++k) xi[q][k] = ucols[(int64_t)k * npad + rs + q * 32 + lane];
        for (int wc = 0; wc < PW; ++wc) {
            const int cb = col0 + wc * 32;
            const int ncol = min(32, U - cb);
            uint32_t w[RB];
#pragma unroll
            for (int q = 0; q < RB; ++q) w[q] = 0u;
            if (ncol > 0) {
                const float* cpw = colv + wc * 32 * M;
#pragma unroll
                for (int b = 0; b < 32; ++b) {
                    float cv[M];
                    load_col<M>(cpw + b * M, cv);
#pragma unroll
                    for (int q = 0; q < RB; ++q) {
                        bool p = true;
#pragma unroll
                        for (int k = 0; k < M; ++k) p = p && (xi[q][k] <= cv[k]);   // NaN padding never passes
                        if (p) w[q] |= (1u << b);
                    }
                }
#pragma unroll
                for (int q = 0; q < RB; ++q) {                          // unique rows: only i == j is "equal"
                    const int d = rs + q * 32 + lane - cb;
                    if (d >= 0 && d < 32) w[q] &= ~(1u << d);
                }
            }
#pragma unroll
            for (int q = 0; q < RB; ++q) st[wc * (SLAB + 1) + q * 32 + lane] = w[q];
        }
        __syncwarp();
        slab_out<SLAB>(st + lane * (SLAB + 1), bits + bits_offset(l, rs, world, rank), lane, cp, (l & 1) != 0);
        __syncwarp();
    }
    __syncthreads();
    counts_flush(cp, lane, warp, stage, count, l);
}

// Off-diagonal tiles: grid = (row chunks, owned panels); chunks at or past the panel's own block leave.
template <int M>
__global__ void __launch_bounds__(TILE_WARPS * 32)
k_dominance_tiles(const float* __restrict__ ucols, const uint8_t* __restrict__ ptab, int64_t npad,
                  const Scal* __restrict__ scal, uint32_t* __restrict__ bits, int* __restrict__ count, int world, int rank,
                  int slabs) {
    const int l = blockIdx.y;
    const int J = panel_of_local(l, world, rank);
    const int U = scal->U;
    const int col0 = J * TW;
    if (col0 >= U) return;
    const int row0 = blockIdx.x * (slabs * TILE_WARPS * 32 * tile_rb(M));
    if (row0 >= col0) return;
    extern __shared__ __align__(128) uint32_t smem[];
    if constexpr (M == 3) tile_offdiag_m3(ptab, ucols, npad, U, l, col0, row0, col0, slabs, bits, count, world, rank, smem);
    else tile_offdiag<M>(ucols, npad, U, l, col0, row0, col0, slabs, bits, count, world, rank, smem);
}

// Diagonal tiles: grid = (row chunks of one block, owned panels).
template <int M>
__global__ void __launch_bounds__(TILE_WARPS * 32)
k_dominance_diag(const float* __restrict__ ucols, int64_t npad, const Scal* __restrict__ scal,
                 uint32_t* __restrict__ bits, int* __restrict__ count, int world, int rank, int slabs) {
    const int l = blockIdx.y;
    const int J = panel_of_local(l, world, rank);
    const int U = scal->U;
    const int col0 = J * TW;
    if (col0 >= U) return;
    const int rowlimit = min((J + 1) * TW, U);
    const int row0 = col0 + blockIdx.x * (slabs * TILE_WARPS * 32 * tile_rb(M));
    if (row0 >= rowlimit) return;
    extern __shared__ __align__(128) uint32_t smem[];
    tile_diag<M>(ucols, npad, U, l, col0, row0, rowlimit, slabs, bits, count, world, rank, smem);
}

// Fallback for M > MAX_TEMPLATE_M: same labelling and layout, objective loop at run time.
__global__ void __launch_bounds__(TILE_WARPS * 32)
k_dominance_tiles_generic(const float* __restrict__ ucols, int64_t npad, int m, const Scal* __restrict__ scal,
                          uint32_t* __restrict__ bits, int* __restrict__ count, int world, int rank) {
    const int l = blockIdx.y;
    const int J = panel_of_local(l, world, rank);
    const int U = scal->U;
    const int col0 = J * TW;
    if (col0 >= U) return;
    const int rowlimit = min((J + 1) * TW, U);
    const int row0 = blockIdx.x * (TILE_WARPS * 32);
    if (row0 >= rowlimit) return;
    const bool diag = row0 >= col0;
    const int k0 = diag ? 0 : 1;
    const int i = row0 + threadIdx.x;
    uint32_t* gout = bits + bits_offset(l, i, world, rank);
    if (diag && (l & 1))                                                // sibling half of this super-row
        for (int wc = 0; wc < PW; ++wc) gout[wc - PW] = 0u;
    for (int wc = 0; wc < PW; ++wc) {
        const int cb = col0 + wc * 32;
        uint32_t w = 0u;
        for (int b = 0; b < 32; ++b) {
            const int j = cb + b;
            bool p = (i < U) && (j < U) && (i != j);
            for (int k = k0; k < m && p; ++k) p = ucols[(int64_t)k * npad + i] <= ucols[(int64_t)k * npad + j];
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
    const int64_t j = (int64_t)panel_of_local((int)l, world, rank) * TW + (t - l * TW);
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
    const int64_t base_col = (int64_t)panel_of_local((int)l, world, r) * TW + (wl - l * PW) * 32;
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
    const int J = panel_of_local(l, world, rank);
    const int U = scal->U;
    if (J * TW >= U) return;
    const int rowlimit = min((J + 1) * TW, U);
    __shared__ int scount[TW];
    for (int t = threadIdx.x; t < TW; t += blockDim.x) scount[t] = 0;
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t* pbase = bits + bits_offset(l, 0, world, rank) + lane;
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
            v[u] = (i < rowlimit) ? __ldg(pbase + (int64_t)i * PW2) : 0u;
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


// ------------------------------------------------------------------------------------------------
// stage 3b: the whole peel as ONE persistent cooperative kernel
// ------------------------------------------------------------------------------------------------
// The dominator counts of a panel live BIT-SLICED in global memory (L2): plane p, word w holds bit p of
// the counts of columns 32w..32w+31.  Per front and panel pair the front's 256-byte super-rows are
// gathered into carry-save adders (lane <-> word), the per-warp sums are merged bit-sliced through
// shared memory, subtracted from the count planes with a bit-sliced borrow chain, and "count == 0" is
// simply the NOR of the planes - which IS the next front's membership mask for these 32 x 32 columns.
// Nothing is ever expanded to integers.  One barrier per front: over the grid on one GPU, over all
// CTAs of all ranks in the fused multi-GPU peel.
constexpr int PEEL_WARPS = 8;
constexpr int PEEL_LOG_WARPS = 3;
constexpr int PEEL_UP = 9;                          // register planes 8 .. 2^11 (+ ones, twos, fours)
constexpr int PEEL_RP = 3 + PEEL_UP;                // 12 planes per warp: up to 4095 rows per pass
constexpr int PEEL_PM = PEEL_RP + PEEL_LOG_WARPS;   // merged over 8 warps: 15 planes (8*4095 < 2^15)
constexpr int PEEL_ROWS_PER_WARP = (1 << PEEL_RP) - 32;

struct Planes {
    uint32_t ones = 0, twos = 0, fours = 0, up[PEEL_UP] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
};

// ---- peer-visible state of the fused multi-rank peel ------------------------------------------------
// Every rank provides one buffer in peer-mapped memory (torch symmetric memory / CUDA IPC):
// [PeerState | per-panel front sizes x 2 | front segments A | front segments B], a full replica of the
// front (all panels of all ranks), made of 64-bit words (value, epoch).  The warp that commits a panel
// writes the panel's new members straight into EVERY rank's replica, and the CTA publishes the sizes of
// the panels it committed when it is done with the front (plain 8-byte stores, posted over NVLink: no
// cross-GPU atomics on data, no system-scope fence).  There is NO barrier between fronts, on one GPU or
// on eight: a CTA starts front f when all panels' size words carry f's epoch, and reads a member word
// only once it carries it too.  Reuse of the double buffers is safe by data dependence: a rank writes
// front f+2 only after it has seen every rank's sizes of front f+1, and a rank publishes its last size
// of front f+1 only after its last read of front f.
// So that waiting CTAs do not hammer the size words, every CTA adds the number of panels it published
// to a counter on its own GPU, and the CTA that finishes a front last on a GPU stores the front's epoch
// into its rank's slot on every peer (a plain store: atomics across NVLink cost several microseconds
// each).  One warp per CTA polls the local counter and the peers' slots, and the CTA looks at the size
// words when those say they should all be there.  Counter and slots are hints only: validity is decided
// by the words themselves.
// Epochs and the counter run on across launches (PeerState::epoch_base / arrived_base; every rank
// passes the same number of fronts), so stale words never validate and the next launch needs neither a
// reset nor a host-level hand-shake: the buffer is cleared once, by evx_nds_shard_peer_init.
constexpr int MAX_WORLD = 8;
struct PeerState {
    int abort;                              // a rank gave up waiting (4 s): everybody leaves
    unsigned int epoch_base;                // epoch of the last front of the previous launch (local use only)
    unsigned int arrived;                   // panels this rank's CTAs have published so far, all launches (a hint, local)
    unsigned int arrived_base;              // its value when the previous launch was complete (local use only)
    unsigned int done_epoch[MAX_WORLD];     // [r]: last front epoch rank r has published completely (a hint, written by rank r)
    unsigned int pad[52];
};
static_assert(sizeof(PeerState) == 256, "PeerState must stay 256 bytes");
struct PeerTable {
    char* base[MAX_WORLD];
    int world, rank;
};
__device__ __forceinline__ PeerState* peer_state(const PeerTable& pt, int r) {
    return reinterpret_cast<PeerState*>(pt.base[r]);
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
constexpr unsigned long long PEER_TIMEOUT_NS = 4000000000ull;      // a peer that is 4 s late is gone

__device__ __forceinline__ void csa8(uint32_t& ones, uint32_t& twos, uint32_t& fours, uint32_t (&up)[PEEL_UP],
                                     const uint32_t* x) {
    uint32_t ta, tb, fa, fb, e;
    csa(ta, ones, ones, x[0], x[1]);
    csa(tb, ones, ones, x[2], x[3]);
    csa(fa, twos, twos, ta, tb);
    csa(ta, ones, ones, x[4], x[5]);
    csa(tb, ones, ones, x[6], x[7]);
    csa(fb, twos, twos, ta, tb);
    csa(e, fours, fours, fa, fb);
#pragma unroll
    for (int p = 0; p < PEEL_UP; ++p) {
        if (!__any_sync(0xFFFFFFFFu, e)) break;
        const uint32_t c = up[p] & e;
        up[p] ^= e;
        e = c;
    }
}
__device__ __forceinline__ void csa8(Planes& pl, const uint32_t* x) { csa8(pl.ones, pl.twos, pl.fours, pl.up, x); }

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ------------------------------------------------------------------------------------------------
// Single-GPU peel with per-front load balancing.
//
// The static item table above gives every CTA a fixed set of pairs, weighted by how many rows a pair
// CAN see.  What a pair really has to fetch for a front is the number of front members stored for
// it, and that distribution moves during the peel: early fronts sit at small row indices (every
// pair sees nearly the whole list), late fronts at large ones (only the last pairs have work), so
// with fixed ownership the slowest CTA was 25-30 us behind the mean on every one of ~230 fronts.
//
// Here the front is kept as one SEGMENT PER PANEL (seg[J*TW ..], cnt[J]): the committer of a panel
// writes its new members in column order with no atomics.  After the grid barrier every CTA scans
// the panel counts into pre[] (members in panels < J), which makes the front a virtual list sorted
// by row, and the rows pair Q has to fetch are exactly the first pre[2Q+2] of it.  Work is counted in
// groups of 32 list positions; the groups of all pairs are laid end to end and cut into gridDim.x
// equal ranges, one per CTA.  A range covers whole pairs (committed directly) and at most two pairs
// shared with neighbours: their partial sums are parked in global memory and the last CTA to arrive
// adds them up and commits - nobody waits.  Count planes live in global memory (L2).
// ------------------------------------------------------------------------------------------------

// Front data words are 64 bits: low half = value, high half = the epoch it was written for.  A word is
// valid for front epoch e iff its high half equals e, so a reader needs no fence and no barrier with a
// writer on another GPU (8-byte stores arrive whole): it polls the word itself.
typedef unsigned long long FWord;
__device__ __forceinline__ FWord fword(int value, unsigned int epoch) {
    return ((FWord)epoch << 32) | (unsigned int)value;
}

// Position t of the virtual front list restricted to panels [0, npre): the member's row index.
__device__ __forceinline__ int front_row(const int* pre, int npre, const FWord* seg, int t, unsigned int fe) {
    int lo = 0, hi = npre;                            // pre[lo] <= t < pre[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (pre[mid] <= t) lo = mid; else hi = mid;
    }
    const FWord* w = seg + (size_t)lo * TW + (t - pre[lo]);
    FWord v = __ldcg(w);
    while ((unsigned int)(v >> 32) != fe)             // a peer's size word overtook its members on the way: rare
        asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(w) : "memory");
    return (int)(unsigned int)v;
}

// ---- cp.async gather ring (load-balanced kernel) ----------------------------------------------------
// A 256-byte super-row is fetched by 16 lanes x 16 bytes of cp.async.cg (LDGSTS): a real SIMT
// instruction, so one warp instruction starts TWO row fetches and nothing is serialised per lane.
// (cp.async.bulk is a warp-uniform instruction: with a different row per lane the hardware issues
// the lanes one after another, ~170 cycles per row per warp; tools/probes/gather_probe4.cu measures
// 6.0 TB/s for that ring with this consumer against 6.76 TB/s - the random-gather ceiling - for this
// one.)  Each warp owns DYN_STAGES stages of 16 rows; a stage is one commit group.
constexpr int DYN_STAGES = 2;
constexpr int DYN_STAGE_ROWS = 16;
constexpr int DYN_STAGE_WORDS = DYN_STAGE_ROWS * PW2;
struct Ring2 {
    uint32_t* slots;             // [DYN_STAGES][DYN_STAGE_ROWS][PW2]
    int issued = 0, consumed = 0;
    uint32_t rowsp = 0;          // rows in each stage, one byte per stage
};

// Fold the oldest stage into the planes (lane l: word l of panel 0 -> pa, of panel 1 -> pb).  The
// caller has waited for that stage's commit group.
__device__ __forceinline__ void ring2_consume(Ring2& rg, int lane, Planes& pa, Planes& pb) {
    const int st = rg.consumed & (DYN_STAGES - 1);
    const int n = (rg.rowsp >> (8 * st)) & 0xFF;
    const uint32_t* base = rg.slots + st * DYN_STAGE_WORDS + lane;
    for (int r0 = 0; r0 < n; r0 += 8) {
        uint32_t xa[8], xb[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const bool ok = r0 + r < n;
            xa[r] = ok ? base[(r0 + r) * PW2] : 0u;
            xb[r] = ok ? base[(r0 + r) * PW2 + PW] : 0u;
        }
        csa8(pa, xa);
        csa8(pb, xb);
    }
    ++rg.consumed;
    __syncwarp();                                     // every lane has read the stage before it is refilled
}

__device__ __forceinline__ void ring2_drain(Ring2& rg, int lane, Planes& pa, Planes& pb) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    while (rg.consumed < rg.issued) ring2_consume(rg, lane, pa, pb);
}

// Groups g0 + first, g0 + first + stride, ... < g1 of pair Q's list prefix.
// (first, stride) = (warp, PEEL_WARPS) when the CTA shares the groups, (0, 1) when one warp does them all.
__device__ __forceinline__ void peel_accumulate_groups(const int* pre, int npre, int rows, const FWord* seg, unsigned int fe,
                                                       int g0, int g1, int first, int stride, int lane,
                                                       const uint32_t* pair, Ring2& rg, Planes& pa, Planes& pb) {
    const int sub = lane >> 4, part = (lane & 15) * 4;              // which of the two rows / which 16 bytes of it
    int g = g0 + first;
    int mine = (g < g1 && g * 32 + lane < rows) ? front_row(pre, npre, seg, g * 32 + lane, fe) : -1;
    for (; g < g1; g += stride) {
        const int gn = g + stride;
        const int nxt = (gn < g1 && gn * 32 + lane < rows) ? front_row(pre, npre, seg, gn * 32 + lane, fe) : -1;
        const uint32_t valid = __ballot_sync(0xFFFFFFFFu, mine >= 0);     // a prefix of the lanes
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int total = __popc(valid & (0xFFFFu << (16 * half)));
            if (total) {
                if (rg.issued - rg.consumed == DYN_STAGES) {
                    asm volatile("cp.async.wait_group %0;" ::"n"(DYN_STAGES - 1) : "memory");
                    __syncwarp();
                    ring2_consume(rg, lane, pa, pb);
                }
                const int st = rg.issued & (DYN_STAGES - 1);
                uint32_t* dst = rg.slots + st * DYN_STAGE_WORDS + part;
#pragma unroll
                for (int i = 0; i < DYN_STAGE_ROWS / 2; ++i) {
                    const int r = 2 * i + sub;
                    const int idx = __shfl_sync(0xFFFFFFFFu, mine, 16 * half + r);
                    if (idx >= 0)
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst + r * PW2)),
                                     "l"(pair + (size_t)idx * PW2 + part)
                                     : "memory");
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
                rg.rowsp = (rg.rowsp & ~(0xFFu << (8 * st))) | ((uint32_t)total << (8 * st));
                ++rg.issued;
            }
        }
        mine = nxt;
    }
}

// What a CTA has committed for the coming front but not yet published: the panels' size words go out
// together when the CTA is done with the front (publish_pending), behind ONE fence, instead of one fence
// per panel.
constexpr int PUB_CAP = 64;
struct PubList {
    int n;                  // panels committed so far for the coming front (may exceed PUB_CAP)
    int extra;              // panels already published directly (list full, or untouched panels)
    int J[PUB_CAP], cnt[PUB_CAP];
};

// Store one size word into the local replica and, fused multi-rank, into every peer's.
template <bool MULTI>
__device__ __forceinline__ void store_count(FWord* cnt_next_J, int cnt, unsigned int fe_next, const PeerTable& pt) {
    const FWord w = fword(cnt, fe_next);
    __stcg(cnt_next_J, w);
    if constexpr (MULTI) {
        if (pt.world > 1) {
            const size_t coff = reinterpret_cast<const char*>(cnt_next_J) - pt.base[pt.rank];
            for (int r = 0; r < pt.world; ++r)
                if (r != pt.rank) *reinterpret_cast<volatile FWord*>(pt.base[r] + coff) = w;
        }
    }
}

// One warp: subtract a bit-sliced delta (plane p of this lane's word = delta(p), p < pm; pm = 0: none)
// from panel J's count planes in global memory, retire the columns that reached zero, write their
// rank and append them, in column order, as panel J's segment of the next front (epoch fe_next).
// MULTI: the segment goes into every rank's replica of the front (plain 8-byte stores to peer-mapped
// memory, posted over NVLink).  The segment's size is queued in `pub`.
template <bool MULTI, class Delta>
__device__ __forceinline__ void peel_commit_seg(uint32_t* cp, int K, Delta delta, int pm, int lane, int J, FWord* seg_next,
                                                FWord* cnt_next, unsigned int fe_next, int* rank_panel, int rank_value,
                                                const PeerTable& pt, PubList* pub) {
    uint32_t c[PEEL_KMAX];
#pragma unroll
    for (int p = 0; p < PEEL_KMAX; ++p) c[p] = (p < K) ? __ldcg(cp + p * PW + lane) : 0u;   // all planes in flight at once
    const uint32_t alive = __ldcg(cp + K * PW + lane);
    uint32_t borrow = 0, nz = 0;
#pragma unroll
    for (int p = 0; p < PEEL_KMAX; ++p) {
        if (p < K) {
            const uint32_t d = (p < pm) ? delta(p) : 0u;
            const uint32_t x = c[p] ^ d;
            const uint32_t r = x ^ borrow;
            borrow = (~c[p] & d) | (~x & borrow);
            if (r != c[p]) __stcg(cp + p * PW + lane, r);
            nz |= r;
        }
    }
    const uint32_t nw = alive & ~nz;
    if (nw) __stcg(cp + K * PW + lane, alive & ~nw);
    const int cnt = __popc(nw);
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += t;
    }
    FWord* dst = seg_next + (size_t)J * TW + (incl - cnt);
    uint32_t w = nw;
    const int c0 = lane * 32;
    int q = 0;
    while (w) {
        const int b = __ffs(w) - 1;
        w &= w - 1;
        const FWord v = fword(J * TW + c0 + b, fe_next);
        __stcg(dst + q, v);
        if constexpr (MULTI) {
            if (pt.world > 1) {
                const size_t soff = reinterpret_cast<const char*>(dst + q) - pt.base[pt.rank];
                for (int r = 0; r < pt.world; ++r)
                    if (r != pt.rank) *reinterpret_cast<volatile FWord*>(pt.base[r] + soff) = v;
            }
        }
        rank_panel[c0 + b] = rank_value;
        ++q;
    }
    __syncwarp();
    if (lane == 31) {
        const int slot = atomicAdd(&pub->n, 1);
        if (slot < PUB_CAP) {
            pub->J[slot] = J;
            pub->cnt[slot] = incl;
        } else {                                      // list full: publish now, behind a fence of its own
            __threadfence();
            store_count<MULTI>(cnt_next + J, incl, fe_next, pt);
        }
    }
}

// End of a CTA's work on a front: publish the size words of the panels it committed (after one fence
// that covers every warp's planes, ranks and members), then tell every rank how many panels this CTA
// has published for the coming front.
// `call` = how many times this CTA has been here in this launch (1 = after the init).  Fused multi-rank:
// every CTA adds its panels to the LOCAL counter; the CTA whose panels complete the front on this GPU
// (counter = abase + nlocal * call) stores the front's epoch into the rank's slot on every peer.
template <bool MULTI>
__device__ __forceinline__ void publish_pending(PubList* pub, FWord* cnt_next, unsigned int fe_next, Scal* scal,
                                                const PeerTable& pt, unsigned int call, unsigned int nlocal,
                                                unsigned int abase) {
    __syncthreads();                                  // every warp's commits are issued
    if (threadIdx.x == 0) {
        const int n = pub->n, listed = n < PUB_CAP ? n : PUB_CAP;
        const unsigned int total = (unsigned int)(n + pub->extra);
        __threadfence();                              // cumulative over the CTA barrier: planes, ranks, members
        for (int i = 0; i < listed; ++i) store_count<MULTI>(cnt_next + pub->J[i], pub->cnt[i], fe_next, pt);
        pub->n = 0;                                   // (the next commits are several CTA barriers away)
        pub->extra = 0;
        if constexpr (MULTI) {
            if (total && pt.world > 1) {
                // the CTA whose panels complete the front on this GPU tells the peers (the counter only moves on
                // to the next front after every rank has seen this one complete)
                unsigned int old;
                asm volatile("atom.relaxed.gpu.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(&peer_state(pt, pt.rank)->arrived), "r"(total) : "memory");
                if (old + total == abase + nlocal * call)
                    for (int r = 0; r < pt.world; ++r)
                        if (r != pt.rank)
                            asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(&peer_state(pt, r)->done_epoch[pt.rank]), "r"(fe_next) : "memory");
            } else if (total) {
                asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(&peer_state(pt, pt.rank)->arrived), "r"(total) : "memory");
            }
        } else {
            if (total) asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(&scal->arrived), "r"(total) : "memory");
        }
    }
}

struct NoDelta {
    __device__ __forceinline__ uint32_t operator()(int) const { return 0u; }
};
struct SmemDelta {                                            // planes in shared memory, stride PW2
    const uint32_t* base;                                     // already offset to this lane's word
    __device__ __forceinline__ uint32_t operator()(int p) const { return base[p * PW2]; }
};
constexpr int DYN_MAX_PPT = 4;              // pairs per thread in the per-front scan: up to 1024 pairs (N <= 2M)
#ifndef EVX_WPG
#define EVX_WPG 1
#endif
#ifndef EVX_CW
#define EVX_CW 2
#endif
#ifndef EVX_CC
#define EVX_CC 12
#endif
constexpr int PEEL_WARP_PAIR_GROUPS = EVX_WPG;    // a pair with at most this many groups for a front is done by one warp alone
// Fixed cost of visiting a pair, in units of one group's gather time (~0.4 us for a CTA): a pair done by
// one warp costs its CTA about one unit (eight of them run side by side), a pair done by the whole CTA
// pays for the merge tree, the commit and four CTA-wide barriers.
#ifndef EVX_WIDE
#define EVX_WIDE 3
#endif
constexpr int PEEL_WIDE_PARTS = EVX_WIDE;         // from this many CTAs sharing a pair on, their sums are added by all warps
constexpr int PEEL_COST_WARP_PAIR = EVX_CW;
constexpr int PEEL_COST_CTA_PAIR = EVX_CC;
static_assert(PEEL_WARP_PAIR_GROUPS + PEEL_COST_WARP_PAIR < 1 + PEEL_COST_CTA_PAIR, "the two kinds must be told apart by total cost");
__device__ __forceinline__ int pair_fixed_cost(int total) {      // total = groups + fixed cost (> 0)
    return total <= PEEL_WARP_PAIR_GROUPS + PEEL_COST_WARP_PAIR ? PEEL_COST_WARP_PAIR : PEEL_COST_CTA_PAIR;
}

struct DynShared {
    alignas(128) uint32_t merge[PEEL_WARPS][PEEL_PM][PW2];    // per-warp count planes of the pair being merged
    alignas(128) uint32_t acc[PEEL_KMAX][PW2];                // this CTA's sums for the pair (all passes)
    alignas(128) uint32_t ring[PEEL_WARPS][DYN_STAGES * DYN_STAGE_ROWS][PW2];   // per-warp cp.async gather ring
    int nf, dead;
    int cnt[8], cnt2[8];
    PubList pub;
};

// A warp's register planes -> its shared-memory region, planes [0, pm).  Lane l: word l of panel 0 (pa) and of panel 1 (pb).
__device__ __forceinline__ void store_planes(uint32_t (*m)[PW2], const Planes& pa, const Planes& pb, int pm, int lane) {
    m[0][lane] = pa.ones;   m[0][PW + lane] = pb.ones;
    m[1][lane] = pa.twos;   m[1][PW + lane] = pb.twos;
    m[2][lane] = pa.fours;  m[2][PW + lane] = pb.fours;
#pragma unroll
    for (int p = 0; p < PEEL_UP; ++p)
        if (3 + p < pm) { m[3 + p][lane] = pa.up[p]; m[3 + p][PW + lane] = pb.up[p]; }
    for (int p = PEEL_RP; p < pm; ++p) { m[p][lane] = 0u; m[p][PW + lane] = 0u; }
}

// Sum the eight warps' plane sets (planes [0, pm)) into sh.merge[0]: three levels, warps [0, half) take
// panel 0's words and warps [half, 2*half) panel 1's.  Called by the whole CTA, after a barrier.
__device__ __forceinline__ void merge_tree(uint32_t (*merge)[PEEL_PM][PW2], int pm, int warp, int lane) {
    for (int half = PEEL_WARPS / 2; half >= 1; half >>= 1) {
        if (warp < 2 * half) {
            const int wl = warp % half, off = (warp / half) * PW + lane;
            uint32_t carry = 0;
            for (int p = 0; p < pm; ++p) {
                const uint32_t a = merge[wl][p][off], b2 = merge[wl + half][p][off];
                const uint32_t x = a ^ b2;
                merge[wl][p][off] = x ^ carry;
                carry = (a & b2) | (x & carry);
            }
        }
        __syncthreads();
    }
}

template <bool MULTI>
__global__ void __launch_bounds__(PEEL_WARPS * 32, 2)
k_peel_dynamic(const uint32_t* __restrict__ bits, const int* __restrict__ count, int* __restrict__ rank_u, FWord* segA,
               FWord* segB, Scal* scal, uint32_t* __restrict__ cplanes, FWord* pcnt, int pstride, uint32_t* partials,
               unsigned int* pdone, int K, int world, int rank, PeerTable pt, unsigned long long* ctadbg) {
    extern __shared__ __align__(128) unsigned char peel_smem_raw[];
    DynShared& sh = *reinterpret_cast<DynShared*>(peel_smem_raw);
    uint32_t (*acc)[PW2] = sh.acc;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int G = gridDim.x, B = blockIdx.x;
    const int U = scal->U;
    const int npan = (U + TW - 1) / TW;                             // live panels, all ranks
    const int nqg = (npan + 1) / 2;                                 // live pairs, all ranks
    // pairs of this rank: local pair lp is global pair lp * world + rank (one GPU: the same thing)
    const int nq = MULTI ? max(0, (nqg - rank + world - 1) / world) : nqg;
    auto gpair = [&](int lp) { return MULTI ? lp * world + rank : lp; };
    int* pre = reinterpret_cast<int*>(peel_smem_raw + sizeof(DynShared));   // [npan + 1]
    int* S = pre + npan + 1;                                        // [nq + 1] cost units before local pair lp
    const size_t cp_stride = (size_t)(K + 1) * PW;
    // Front f of this launch carries epoch e0 + f.  One GPU: the host cleared the size words and the
    // counter, epochs start at 1.  Fused multi-rank: both run on from launch to launch (every rank
    // passes the same number of fronts), so stale words never validate and nothing is cleared.
    unsigned int e0 = 1, abase = 0;
    const unsigned int* arrived = &scal->arrived;
    if constexpr (MULTI) {
        e0 = peer_state(pt, pt.rank)->epoch_base + 1u;
        abase = peer_state(pt, pt.rank)->arrived_base;
        arrived = &peer_state(pt, pt.rank)->arrived;
    }
    const unsigned long long t_start = global_ns();
    unsigned int nlocal = 0;                                        // live panels owned by this rank
    for (int lp = 0; lp < nq; ++lp) nlocal += (2 * gpair(lp) + 1 < npan) ? 2u : 1u;
    if (threadIdx.x == 0) { sh.pub.n = 0; sh.pub.extra = 0; }
    __syncthreads();

    // ---- init: int counts -> planes, alive mask, front 0 ------------------------------------------
    {
        FWord* seg0 = (e0 & 1u) ? segB : segA;
        FWord* cnt0 = pcnt + (size_t)(e0 & 1u) * pstride;
        for (int lp = B; lp < nq; lp += G) {
            const int Q = gpair(lp);
            for (int c2 = threadIdx.x; c2 < 2 * TW; c2 += blockDim.x) {
                const int h = c2 >> 10, c = c2 & (TW - 1), J = 2 * Q + h, l = 2 * lp + h;
                const int j = J * TW + c;
                uint32_t* cp = cplanes + (size_t)l * cp_stride;
                const int v = (j < U) ? count[(size_t)l * TW + c] : 0;
                for (int p = 0; p < K; ++p) {
                    const uint32_t w = __ballot_sync(0xFFFFFFFFu, (v >> p) & 1);
                    if (lane == 0) __stcg(cp + p * PW + (c >> 5), w);
                }
                const uint32_t a = __ballot_sync(0xFFFFFFFFu, j < U);
                if (lane == 0) __stcg(cp + K * PW + (c >> 5), a);
            }
            __syncthreads();
            if (warp < 2) {
                const int J = 2 * Q + warp, l = 2 * lp + warp;
                if (J < npan)
                    peel_commit_seg<MULTI>(cplanes + (size_t)l * cp_stride, K, NoDelta{}, 0, lane, J, seg0, cnt0, e0,
                                           rank_u + (size_t)l * TW, 0, pt, &sh.pub);
            }
            __syncthreads();
        }
        publish_pending<MULTI>(&sh.pub, cnt0, e0, scal, pt, 1u, nlocal, abase);
    }

    Ring2 rg;
    rg.slots = &sh.ring[warp][0][0];
    int k = 0;
    unsigned long long t_acc = 0, t_mc = 0, t_bar = 0, t_hint = 0, t_scan = 0, t_mark = global_ns();
    for (;; ++k) {
        // ---- the front as a virtual list: pre[J] = members in panels < J; S[q] = cost units before pair q ----
        const unsigned int fe = e0 + (unsigned int)k;              // this front's epoch
        const FWord* cnt = pcnt + (size_t)(fe & 1u) * pstride;
        FWord* cnt_next = pcnt + (size_t)((fe + 1u) & 1u) * pstride;
        // Scan 1: thread t owns the global pairs [t*ppt, (t+1)*ppt) - their panel sizes in registers, one L2
        // round trip - and writes pre[].  Scan 2: thread t owns local pairs and writes S[] (cost units).
        // Loading the sizes IS the synchronisation between fronts: a size word is valid once its owner has
        // published the panel for this front (on whatever GPU).  Thread 0 first waits for the arrival
        // counter (one word to poll), then the CTA re-reads the words until all of them are valid.
        {
            const int ppt = (nqg + PEEL_WARPS * 32 - 1) / (PEEL_WARPS * 32);    // <= DYN_MAX_PPT (host checks)
            const int q0 = threadIdx.x * ppt;
            int c[2 * DYN_MAX_PPT];
            int psum = 0;
            bool done = false;
            if (warp == 0) {                                        // hints: my GPU's panels of fronts 0..k, and every peer done with front k
                const unsigned int target = abase + nlocal * (unsigned int)(k + 1);
                const bool peer = MULTI && lane < pt.world && lane != pt.rank && lane < nqg;   // ranks >= nqg own no live panel and send nothing
                const unsigned int* word = peer ? &peer_state(pt, pt.rank)->done_epoch[lane] : arrived;
                const unsigned int want = peer ? fe : target;
                for (unsigned int spin = 0;; ++spin) {
                    unsigned int v;
                    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(word) : "memory");
                    if (__all_sync(0xFFFFFFFFu, (int)(v - want) >= 0)) break;
                    __nanosleep(32);                                // 300 pollers on one line: leave the L2 slice some air
                    if ((spin & 1023u) == 1023u && global_ns() - t_start > PEER_TIMEOUT_NS * (MULTI ? 1 : 8)) break;
                }
                if (lane == 0) {
                    asm volatile("fence.acq_rel.gpu;" ::: "memory");    // acquire: the committers' planes and members
                    t_hint += global_ns() - t_mark;
                }
            }
            __syncthreads();
            for (unsigned int round = 0;; ++round) {
                bool ok = true;
                psum = 0;
#pragma unroll
                for (int i = 0; i < 2 * DYN_MAX_PPT; ++i) {
                    const int J = 2 * q0 + i;
                    c[i] = 0;
                    if (i < 2 * ppt && J < npan) {
                        FWord w;
                        asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(w) : "l"(cnt + J) : "memory");
                        ok = ok && (unsigned int)(w >> 32) == fe;
                        c[i] = (int)(unsigned int)w;
                    }
                    psum += c[i];
                }
                if (threadIdx.x == 0) {
                    int dead = 0;
                    if ((round & 255u) == 255u) {                   // a front never takes seconds: somebody is gone
                        dead = global_ns() - t_start > PEER_TIMEOUT_NS * (MULTI ? 1 : 8);
                        if constexpr (MULTI) {
                            int ab;
                            asm volatile("ld.relaxed.sys.global.s32 %0, [%1];" : "=r"(ab) : "l"(&peer_state(pt, pt.rank)->abort) : "memory");
                            dead = dead || ab != 0;
                        }
                    }
                    sh.dead = dead;
                }
                if (__syncthreads_and(ok)) { done = true; break; }
                if (sh.dead) break;
                __syncthreads();                                    // sh.dead is rewritten next round
            }
            if (!done) {
                if constexpr (MULTI) {
                    if ((int)threadIdx.x < pt.world)
                        asm volatile("st.relaxed.sys.global.s32 [%0], %1;" ::"l"(&peer_state(pt, threadIdx.x)->abort), "r"(1) : "memory");
                }
                k = -1;
                break;
            }
            if (threadIdx.x == 0) { const unsigned long long t = global_ns(); t_bar += t - t_mark; t_mark = t; }
            int incl = psum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
                if (lane >= o) incl += t;
            }
            if (lane == 31) sh.cnt[warp] = incl;
            __syncthreads();
            int run = incl - psum;
            for (int w = 0; w < warp; ++w) run += sh.cnt[w];
            if (threadIdx.x == 0) { pre[0] = 0; S[0] = 0; }
#pragma unroll
            for (int j = 0; j < DYN_MAX_PPT; ++j) {
                const int q = q0 + j;
                if (j < ppt && q < nqg) {
                    run += c[2 * j];
                    pre[2 * q + 1] = run;                           // 2q + 1 <= npan always
                    run += c[2 * j + 1];
                    if (2 * q + 2 <= npan) pre[2 * q + 2] = run;
                }
            }
            __syncthreads();
            const int ppl = (nq + PEEL_WARPS * 32 - 1) / (PEEL_WARPS * 32);
            const int l0 = threadIdx.x * ppl;
            int cost[DYN_MAX_PPT];
            int csum = 0;
#pragma unroll
            for (int j = 0; j < DYN_MAX_PPT; ++j) {
                const int lp = l0 + j;
                cost[j] = 0;
                if (j < ppl && lp < nq) {
                    const int r = pre[min(2 * gpair(lp) + 2, npan)];
                    const int gr = (r + 31) >> 5;                   // [fixed cost | one unit per group]
                    cost[j] = r ? gr + (gr <= PEEL_WARP_PAIR_GROUPS ? PEEL_COST_WARP_PAIR : PEEL_COST_CTA_PAIR) : 0;
                }
                csum += cost[j];
            }
            int incl2 = csum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xFFFFFFFFu, incl2, o);
                if (lane >= o) incl2 += t;
            }
            if (lane == 31) sh.cnt2[warp] = incl2;
            __syncthreads();
            int run2 = incl2 - csum;
            for (int w = 0; w < warp; ++w) run2 += sh.cnt2[w];
#pragma unroll
            for (int j = 0; j < DYN_MAX_PPT; ++j) {
                const int lp = l0 + j;
                if (j < ppl && lp < nq) { run2 += cost[j]; S[1 + lp] = run2; }
            }
            __syncthreads();
        }
        const int nf = pre[npan];
        if (threadIdx.x == 0) { const unsigned long long t = global_ns(); t_scan += t - t_mark; t_acc -= 0; t_mark = t; }
        if (nf <= 0) break;
        const long long T = S[nq];
        const int lo = (int)((long long)B * T / G), hi = (int)((long long)(B + 1) * T / G);
        const FWord* seg = (fe & 1u) ? segB : segA;
        FWord* seg_next = (fe & 1u) ? segA : segB;
        // Pairs the front does not reach (no member with a row stored for them) lose nothing and gain no
        // member, but their panels' size words must still turn valid for the next front.  Nothing they
        // depend on changed, so they go out at once, without a fence.
        for (int t = B * blockDim.x + threadIdx.x; t < 2 * nq; t += G * blockDim.x) {
            const int Q = gpair(t >> 1), J = 2 * Q + (t & 1);
            if (J < npan && pre[min(2 * Q + 2, npan)] == 0) {
                store_count<MULTI>(cnt_next + J, 0, fe + 1u, pt);
                atomicAdd(&sh.pub.extra, 1);
            }
        }
        auto owner = [&](int u) {                                   // CTA whose range holds cost unit u
            return (int)((((long long)(u + 1) * G) + T - 1) / T) - 1;
        };

        if (lo < hi) {
            int Qa = 0;
            {                                                       // first pair with S[Q+1] > lo
                int a = 0, b = nq;                                  // S[a] <= lo < S[b]
                while (b - a > 1) {
                    const int mid = (a + b) >> 1;
                    if (S[mid] <= lo) a = mid; else b = mid;
                }
                Qa = a;
            }
            // ---- pairs with only a few groups that lie wholly in this range: one warp each, no CTA-wide step ----
            int ord = 0;
            for (int lp = Qa; lp < nq && S[lp] < hi; ++lp) {
                const int Q = gpair(lp);
                const int tot = S[lp + 1] - S[lp];
                if (tot <= 0 || tot > PEEL_WARP_PAIR_GROUPS + PEEL_COST_WARP_PAIR) continue;
                const int gq = tot - PEEL_COST_WARP_PAIR;
                const int g0 = S[lp] + PEEL_COST_WARP_PAIR;          // cost unit of its first group
                if (g0 < lo || g0 + gq > hi) continue;              // shared with a neighbour: below
                if ((ord++ & (PEEL_WARPS - 1)) != warp) continue;
                const int npre = min(2 * Q + 2, npan);
                const int rows = pre[npre];
                Planes pa, pb;
                peel_accumulate_groups(pre, npre, rows, seg, fe, 0, gq, 0, 1, lane, bits + pair_words_before(lp, world, rank), rg, pa, pb);
                ring2_drain(rg, lane, pa, pb);
                int pm = 1;
                while ((1 << pm) <= rows) ++pm;                     // <= 8 planes: at most 128 rows
                store_planes(sh.merge[warp], pa, pb, pm, lane);     // this warp's own region: lane <-> word remap
                __syncwarp();
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int J = 2 * Q + h, l = 2 * lp + h;
                    if (J < npan)
                        peel_commit_seg<MULTI>(cplanes + (size_t)l * cp_stride, K, SmemDelta{&sh.merge[warp][0][h * PW + lane]},
                                               pm, lane, J, seg_next, cnt_next, fe + 1u, rank_u + (size_t)l * TW, k + 1, pt, &sh.pub);
                }
                __syncwarp();
            }
            if (threadIdx.x == 0) { const unsigned long long t = global_ns(); t_acc += t - t_mark; t_mark = t; }
            // ---- everything else: the whole CTA per pair (or per shared part of a pair) --------------------
            for (int lp = Qa; lp < nq && S[lp] < hi; ++lp) {
                const int Q = gpair(lp);
                const int tot = S[lp + 1] - S[lp];
                if (tot <= 0) continue;
                const int fixed = pair_fixed_cost(tot);
                const int gq = tot - fixed;
                const int g0 = S[lp] + fixed;
                const int ga = max(lo, g0) - g0, gb = min(hi, g0 + gq) - g0;
                if (ga >= gb) continue;                             // only the fixed-cost units fell in this range
                if (gq <= PEEL_WARP_PAIR_GROUPS && ga == 0 && gb == gq) continue;   // done above
                const int npre = min(2 * Q + 2, npan);
                const int rows = pre[npre];
                int pq = 1;
                while ((1 << pq) <= rows) ++pq;                     // planes that hold any partial sum of this pair
                // CTAs sharing this pair: owners of its first and last group; with fewer cost units than
                // CTAs some ranges are empty, and then every unit has its own CTA
                const int bf = owner(g0), bl = owner(g0 + gq - 1);
                const int parts = (T >= G) ? bl - bf + 1 : gq;
                const uint32_t* pbase = bits + pair_words_before(lp, world, rank);

                constexpr int PASS_GROUPS = PEEL_WARPS * (PEEL_ROWS_PER_WARP / 32);
                // nobody to share with and one pass: the merged planes are the pair's delta as they stand
                const bool direct = parts == 1 && gb - ga <= PASS_GROUPS;
                int pm_last = 0;
                if (!direct)
                    for (int i = threadIdx.x; i < pq * PW2; i += blockDim.x) acc[0][i] = 0u;
                for (int gp = ga; gp < gb; gp += PASS_GROUPS) {
                    const int ge = min(gb, gp + PASS_GROUPS);
                    Planes pa, pb;
                    peel_accumulate_groups(pre, npre, rows, seg, fe, gp, ge, warp, PEEL_WARPS, lane, pbase, rg, pa, pb);
                    ring2_drain(rg, lane, pa, pb);
                    const int per = ((ge - gp + PEEL_WARPS - 1) / PEEL_WARPS) << 5;   // rows a warp can have seen
                    int rp = 1;
                    while ((1 << rp) <= per) ++rp;
                    rp = min(rp, PEEL_RP);
                    const int pm = min(rp + PEEL_LOG_WARPS, PEEL_PM);
                    store_planes(sh.merge[warp], pa, pb, pm, lane);
                    __syncthreads();
                    if (threadIdx.x == 0) { const unsigned long long t = global_ns(); t_acc += t - t_mark; t_mark = t; }
                    merge_tree(sh.merge, pm, warp, lane);
                    pm_last = pm;
                    if (direct) break;
                    if (warp < 2) {                                 // acc += this pass (warp h: panel h)
                        const int off = warp * PW + lane;
                        uint32_t carry = 0;
                        for (int p = 0; p < pq; ++p) {
                            const uint32_t a = acc[p][off], b2 = (p < pm) ? sh.merge[0][p][off] : 0u;
                            const uint32_t x = a ^ b2;
                            acc[p][off] = x ^ carry;
                            carry = (a & b2) | (x & carry);
                        }
                    }
                    __syncthreads();
                }
                bool commit = true, wide = false;
                if (parts > 1) {
                    // Shared pair: park this CTA's sums (slot lp + B is unique per front), count
                    // arrivals; the last one adds the others' sums to its own and commits.
                    uint32_t* mine = partials + (size_t)(lp + B) * PEEL_KMAX * PW2;
                    for (int i = threadIdx.x; i < pq * PW2; i += blockDim.x) __stcg(mine + i, acc[0][i]);
                    __syncthreads();
                    if (threadIdx.x == 0) {
                        __threadfence();
                        const unsigned int old = atomicAdd(&pdone[lp], 1u);
                        const bool last = old + 1u == (unsigned int)parts;
                        if (last) {
                            pdone[lp] = 0u;
                            __threadfence();                        // acquire: the others' parked sums
                        }
                        sh.nf = last ? 1 : 0;
                    }
                    __syncthreads();
                    commit = sh.nf != 0;
                    wide = commit && parts >= PEEL_WIDE_PARTS && pq <= PEEL_PM;
                    if (wide) {
                        // Many parts: the L2 round trips of a serial sum would line up on the front's critical
                        // path.  Warp w sums parts w, w + 8, ... (this CTA's own from shared memory), the merge
                        // tree adds the eight results: one round trip for up to eight parts.
                        uint32_t sa[PEEL_PM], sb[PEEL_PM];
#pragma unroll
                        for (int p = 0; p < PEEL_PM; ++p) { sa[p] = 0u; sb[p] = 0u; }
                        for (int i = warp; i < parts; i += PEEL_WARPS) {
                            const int b2 = (T >= G) ? bf + i : owner(g0 + i);
                            uint32_t oa[PEEL_PM], ob[PEEL_PM];
                            if (b2 == B) {
#pragma unroll
                                for (int p = 0; p < PEEL_PM; ++p) {
                                    oa[p] = (p < pq) ? acc[p][lane] : 0u;
                                    ob[p] = (p < pq) ? acc[p][PW + lane] : 0u;
                                }
                            } else {
                                const uint32_t* other = partials + (size_t)(lp + b2) * PEEL_KMAX * PW2 + lane;
#pragma unroll
                                for (int p = 0; p < PEEL_PM; ++p) {
                                    oa[p] = (p < pq) ? __ldcg(other + p * PW2) : 0u;
                                    ob[p] = (p < pq) ? __ldcg(other + p * PW2 + PW) : 0u;
                                }
                            }
                            uint32_t ca = 0, cb = 0;
#pragma unroll
                            for (int p = 0; p < PEEL_PM; ++p) {
                                const uint32_t xa = sa[p] ^ oa[p], xb = sb[p] ^ ob[p];
                                const uint32_t na = (sa[p] & oa[p]) | (xa & ca), nb = (sb[p] & ob[p]) | (xb & cb);
                                sa[p] = xa ^ ca; sb[p] = xb ^ cb;
                                ca = na; cb = nb;
                            }
                        }
#pragma unroll
                        for (int p = 0; p < PEEL_PM; ++p)
                            if (p < pq) { sh.merge[warp][p][lane] = sa[p]; sh.merge[warp][p][PW + lane] = sb[p]; }
                        __syncthreads();
                        merge_tree(sh.merge, pq, warp, lane);       // the pair's total is below 2^pq: no carry out
                    }
                    if (commit && !wide && warp < 2) {
                        const int off = warp * PW + lane;
                        for (int i = 0; i < parts; ++i) {
                            const int b2 = (T >= G) ? bf + i : owner(g0 + i);
                            if (b2 == B) continue;
                            const uint32_t* other = partials + (size_t)(lp + b2) * PEEL_KMAX * PW2 + off;
                            uint32_t o[PEEL_KMAX];
#pragma unroll
                            for (int p = 0; p < PEEL_KMAX; ++p) o[p] = (p < pq) ? __ldcg(other + p * PW2) : 0u;
                            uint32_t carry = 0;
#pragma unroll
                            for (int p = 0; p < PEEL_KMAX; ++p) {
                                if (p < pq) {
                                    const uint32_t a = acc[p][off];
                                    const uint32_t x = a ^ o[p];
                                    acc[p][off] = x ^ carry;
                                    carry = (a & o[p]) | (x & carry);
                                }
                            }
                        }
                    }
                }
                if (commit && warp < 2) {
                    const int J = 2 * Q + warp, l = 2 * lp + warp;
                    if (J < npan)
                        peel_commit_seg<MULTI>(cplanes + (size_t)l * cp_stride, K,
                                               SmemDelta{direct || wide ? &sh.merge[0][0][warp * PW + lane] : &acc[0][warp * PW + lane]},
                                               direct ? pm_last : pq, lane, J, seg_next, cnt_next, fe + 1u,
                                               rank_u + (size_t)l * TW, k + 1, pt, &sh.pub);
                }
                __syncthreads();
                if (threadIdx.x == 0) { const unsigned long long t = global_ns(); t_mc += t - t_mark; t_mark = t; }
            }
        }
        publish_pending<MULTI>(&sh.pub, cnt_next, fe + 1u, scal, pt, (unsigned int)k + 2u, nlocal, abase);
        if (threadIdx.x == 0) { const unsigned long long t = global_ns(); t_mc += t - t_mark; t_mark = t; }
    }
    if constexpr (MULTI) {
        // k + 1 fronts (the last one empty) were published by every rank: the next launch continues from there
        if (B == 0 && threadIdx.x == 0 && k >= 0) {
            peer_state(pt, pt.rank)->epoch_base = e0 + (unsigned int)k;
            peer_state(pt, pt.rank)->arrived_base = abase + nlocal * (unsigned int)(k + 1);
        }
    }
    if (B == 0 && threadIdx.x == 0) scal->fronts = k;
    if (threadIdx.x == 0) { ctadbg[4 * B] = t_acc; ctadbg[4 * B + 1] = t_mc; ctadbg[4 * B + 2] = t_bar; ctadbg[4 * B + 3] = t_scan; }
    if (threadIdx.x == 0 && (B == 0 || B == G - 1)) {
        unsigned long long* d = scal->dbg + (B == 0 ? 0 : 4);
        d[0] = t_acc; d[1] = t_mc; d[2] = t_bar; d[3] = (unsigned long long)(k > 0 ? k : 1);
        unsigned long long* e = scal->dbg + (B == 0 ? 8 : 12);
        e[0] = t_hint; e[1] = t_scan;
    }
}

// ------------------------------------------------------------------------------------------------
// Small populations (at most PS_MAX_PAN panels, N <= 8192) on one GPU: ONE CTA peels all fronts.
// The persistent kernel above pays about 12 us per front in flags, fences and polling between its CTAs
// whatever the front's size; here a front costs three CTA barriers and a scan.  Count planes, alive
// masks and the front list (row indices, ascending) live in shared memory.
// The 32 warps are dealt to the pairs, W = 32, 16 or 8 each, and inside a pair as RS row shares x 8
// column shares: a warp covers 8 of the super-row's 64 words (32 bytes = one sector), its lanes are
// 8 row slots x 4 lanes of 8 bytes, so one load instruction fetches full sectors of eight different
// rows and a lane always sees the same two columns' words.  A lane counts its rows (at most 127 per
// pass) in carry-save counters, the eight row slots are added up with shuffles - no shared memory, no
// barrier - and slot 0's lanes add the result to the pair's sums.  After one barrier the same lanes
// subtract the sums from the panel's planes; the zero columns - the next front - are appended in row
// order by a scan over the new-member words.
// ------------------------------------------------------------------------------------------------
constexpr int PS_T = 1024, PS_WARPS = PS_T / 32;
constexpr int PS_MAX_PAN = 8;                     // N <= 8192: above, the multi-CTA kernel's bandwidth wins (measured)
constexpr int PS_MAX_PAIRS = PS_MAX_PAN / 2;
constexpr int PS_K = 14;                          // count planes: counts < 2^14 > 8192
constexpr int PS_CS = 8;                          // column shares of a pair: 8 words each
constexpr int PS_LANE_ROWS = 127;                 // rows per lane and pass: 7 counter planes
constexpr int PS_MP = 10;                         // planes after adding the 8 row slots: 8 * 127 < 2^10
struct SingleShared {
    uint32_t planes[PS_MAX_PAN][PS_K + 1][PW];      // [.][PS_K] = alive mask
    uint32_t acc[PS_MAX_PAIRS][PS_K][PW2];          // sums over the front's rows: [pair * RS + row share] (pairs * RS <= 4)
    uint32_t neww[PS_MAX_PAN * PW];                 // new members, one word per 32 columns
    uint16_t list[2][PS_MAX_PAN * TW];              // the front (row indices, ascending), double buffered
    int pre[PS_MAX_PAN + 1];                        // members in panels < J
    int wsum[PS_WARPS];
};

// eight more rows into carry-save counters c[0..6] (ones .. sixty-fours): at most 127 rows in all
__device__ __forceinline__ void csa8_small(uint32_t (&c)[PS_MP], const uint32_t* x) {
    uint32_t ta, tb, fa, fb, e;
    csa(ta, c[0], c[0], x[0], x[1]);
    csa(tb, c[0], c[0], x[2], x[3]);
    csa(fa, c[1], c[1], ta, tb);
    csa(ta, c[0], c[0], x[4], x[5]);
    csa(tb, c[0], c[0], x[6], x[7]);
    csa(fb, c[1], c[1], ta, tb);
    csa(e, c[2], c[2], fa, fb);
#pragma unroll
    for (int p = 3; p < 7; ++p) {
        const uint32_t carry = c[p] & e;
        c[p] ^= e;
        e = carry;
    }
}

__global__ void __launch_bounds__(PS_T, 1)
k_peel_single(const uint32_t* __restrict__ bits, const int* __restrict__ count, int* __restrict__ rank_u, Scal* scal) {
    extern __shared__ __align__(128) unsigned char ps_smem_raw[];
    SingleShared& sh = *reinterpret_cast<SingleShared*>(ps_smem_raw);
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int U = scal->U;
    const int npan = (U + TW - 1) / TW;
    const int nq = (npan + 1) / 2;
    int K = 1;
    while ((1 << K) <= U) ++K;                      // counts are below U
    // ---- init: int counts -> planes, alive masks ------------------------------------------------------
    for (int w = warp; w < npan * PW; w += PS_WARPS) {
        const int J = w / PW, wc = w % PW;
        const int j = J * TW + wc * 32 + lane;
        const int v = (j < U) ? count[(size_t)J * TW + wc * 32 + lane] : 0;
        for (int p = 0; p < K; ++p) {
            const uint32_t b = __ballot_sync(0xFFFFFFFFu, (v >> p) & 1);
            if (lane == 0) sh.planes[J][p][wc] = b;
        }
        const uint32_t a = __ballot_sync(0xFFFFFFFFu, j < U);
        const uint32_t z = __ballot_sync(0xFFFFFFFFu, j < U && v == 0);
        if (lane == 0) {
            sh.planes[J][PS_K][wc] = a & ~z;
            sh.neww[w] = z;                         // front 0
        }
    }
    __syncthreads();
    // new-member words -> list (ascending rows), members per panel, ranks; returns the front's size
    auto emit = [&](uint16_t* list, int front) -> int {
        const int words = npan * PW;                // <= 256
        const uint32_t w = t < words ? sh.neww[t] : 0u;
        const int c = __popc(w);
        int incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) sh.wsum[warp] = incl;
        __syncthreads();
        int before = incl - c, total = 0;
        for (int q = 0; q < PS_WARPS; ++q) {
            const int v = sh.wsum[q];
            if (q < warp) before += v;
            total += v;
        }
        if (t < words && (t % PW) == 0) sh.pre[t / PW] = before;
        if (t == 0) sh.pre[npan] = total;
        uint32_t x = w;
        while (x) {
            const int b = __ffs(x) - 1;
            x &= x - 1;
            const int row = t * 32 + b;             // word t covers columns [32 t, 32 t + 32)
            list[before++] = (uint16_t)row;
            rank_u[row] = front;
        }
        __syncthreads();                            // list and pre complete, wsum free again
        return total;
    };
    uint16_t* cur = sh.list[0];
    uint16_t* nxt = sh.list[1];
    int nf = emit(cur, 0);
    // warps per pair, and the warp's place: pair, row share, column share; lane: row slot, 8-byte piece
    const int W = nq <= 1 ? 32 : nq == 2 ? 16 : 8;
    const int RS = W / PS_CS;
    const int grp = warp / W, wl = warp % W;
    const int rs = wl / PS_CS, cs = wl % PS_CS;
    const int slot = lane >> 2, q = lane & 3;
    const bool live = grp < nq;
    const int word0 = cs * 8 + 2 * q;               // this lane's two words of the super-row: word0, word0 + 1
    const int J = 2 * grp + (word0 >> 5), wp = word0 & 31;    // their panel and place in it
    const uint32_t* pbase = bits + pair_words_before(live ? grp : 0, 1, 0) + word0;
    uint32_t* accw = &sh.acc[live ? grp * RS + rs : 0][0][word0];
    const int stride = 8 * RS;                      // list positions between a lane's rows
    int k = 0;
    while (nf > 0) {
        ++k;
        // rows pair Q needs: the members in panels < 2Q + 2, a prefix of the ascending list
        const int rows = live ? sh.pre[min(2 * grp + 2, npan)] : 0;
        const int passes = (nf + stride * PS_LANE_ROWS - 1) / (stride * PS_LANE_ROWS);   // pair nq - 1 sees every member
        if (live && slot == 0)
            for (int p = 0; p < K; ++p) { accw[p * PW2] = 0u; accw[p * PW2 + 1] = 0u; }
        for (int pass = 0; pass < passes; ++pass) {
            const int base = pass * stride * PS_LANE_ROWS;
            if (!live || rows <= base) continue;    // (warp-uniform)
            // ---- gather: list positions base + rs * 8 + slot + stride * u ----
            uint32_t ca[PS_MP], cb[PS_MP];
#pragma unroll
            for (int p = 0; p < PS_MP; ++p) { ca[p] = 0u; cb[p] = 0u; }
            const int most = min(PS_LANE_ROWS, (rows - base + stride - 1) / stride);    // rows of the pass's first lane
            const int first = base + rs * 8 + slot;
            for (int u0 = 0; u0 < most; u0 += 8) {
                uint32_t xa[8], xb[8];
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    const int pos = first + (u0 + r) * stride;
                    uint2 v = make_uint2(0u, 0u);
                    if (u0 + r < most && pos < rows)
                        v = __ldcg(reinterpret_cast<const uint2*>(pbase + (size_t)cur[pos] * PW2));
                    xa[r] = v.x; xb[r] = v.y;
                }
                csa8_small(ca, xa);
                csa8_small(cb, xb);
            }
            // ---- the eight row slots: bit-sliced adds across lanes 4, 8, 16 apart ----
#pragma unroll
            for (int m = 4; m <= 16; m <<= 1) {
                uint32_t cya = 0, cyb = 0;
#pragma unroll
                for (int p = 0; p < PS_MP; ++p) {
                    const uint32_t oa = __shfl_xor_sync(0xFFFFFFFFu, ca[p], m), ob = __shfl_xor_sync(0xFFFFFFFFu, cb[p], m);
                    const uint32_t sa = ca[p] ^ oa, sb = cb[p] ^ ob;
                    const uint32_t na = (ca[p] & oa) | (sa & cya), nb = (cb[p] & ob) | (sb & cyb);
                    ca[p] = sa ^ cya; cb[p] = sb ^ cyb;
                    cya = na; cyb = nb;
                }
            }
            // ---- into the sums of this row share (slot 0's lanes own their two words) ----
            if (slot == 0) {
                uint32_t cya = 0, cyb = 0;
#pragma unroll
                for (int p = 0; p < PS_K; ++p) {
                    if (p < K) {
                        const uint32_t a = accw[p * PW2], b = accw[p * PW2 + 1];
                        const uint32_t da = p < PS_MP ? ca[p] : 0u, db = p < PS_MP ? cb[p] : 0u;
                        const uint32_t sa = a ^ da, sb = b ^ db;
                        accw[p * PW2] = sa ^ cya; accw[p * PW2 + 1] = sb ^ cyb;
                        cya = (a & da) | (sa & cya); cyb = (b & db) | (sb & cyb);
                    }
                }
            }
        }
        __syncthreads();
        // ---- row shares together, subtract, retire the zero columns: slot 0's lanes of row share 0 ----
        if (live && rs == 0 && slot == 0 && J < npan) {
            uint32_t nwa = 0u, nwb = 0u;
            if (rows > 0) {
                for (int o = 1; o < RS; ++o) {
                    const uint32_t* other = &sh.acc[grp * RS + o][0][word0];
                    uint32_t cya = 0, cyb = 0;
                    for (int p = 0; p < K; ++p) {
                        const uint32_t a = accw[p * PW2], b = accw[p * PW2 + 1], da = other[p * PW2], db = other[p * PW2 + 1];
                        const uint32_t sa = a ^ da, sb = b ^ db;
                        accw[p * PW2] = sa ^ cya; accw[p * PW2 + 1] = sb ^ cyb;
                        cya = (a & da) | (sa & cya); cyb = (b & db) | (sb & cyb);
                    }
                }
                uint32_t bwa = 0, bwb = 0, nza = 0, nzb = 0;
                for (int p = 0; p < K; ++p) {
                    const uint32_t c0 = sh.planes[J][p][wp], c1 = sh.planes[J][p][wp + 1];
                    const uint32_t d0 = accw[p * PW2], d1 = accw[p * PW2 + 1];
                    const uint32_t x0 = c0 ^ d0, x1 = c1 ^ d1;
                    const uint32_t r0 = x0 ^ bwa, r1 = x1 ^ bwb;
                    bwa = (~c0 & d0) | (~x0 & bwa); bwb = (~c1 & d1) | (~x1 & bwb);
                    sh.planes[J][p][wp] = r0; sh.planes[J][p][wp + 1] = r1;
                    nza |= r0; nzb |= r1;
                }
                const uint32_t al0 = sh.planes[J][PS_K][wp], al1 = sh.planes[J][PS_K][wp + 1];
                nwa = al0 & ~nza; nwb = al1 & ~nzb;
                sh.planes[J][PS_K][wp] = al0 & ~nwa; sh.planes[J][PS_K][wp + 1] = al1 & ~nwb;
            }
            sh.neww[J * PW + wp] = nwa; sh.neww[J * PW + wp + 1] = nwb;
        }
        __syncthreads();
        nf = emit(nxt, k);
        uint16_t* tmp = cur; cur = nxt; nxt = tmp;
    }
    if (t == 0) scal->fronts = k;                  // front k was the first empty one
}

__global__ void k_write_ranks(const int* __restrict__ row2u, const int* __restrict__ rank_u, int64_t n, int world,
                              int rank, int* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int u = row2u[i];
    int r = 0;
    if (u >= 0) {
        const int J = u / TW, P = J >> 1;
        if (P % world == rank) r = rank_u[(int64_t)(2 * (P / world) + (J & 1)) * TW + (u - J * TW)];
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

template <int M>
static int launch_tiles(const Plan& P, cudaStream_t s) {
    const Layout& L = P.L;
    constexpr int SLAB = 32 * tile_rb(M), MAX_SLABS = TILE_ROWS_PER_WARP / SLAB;
    static_assert(TW % (MAX_SLABS * TILE_WARPS * SLAB) == 0, "a CTA's rows must tile a block");
    const size_t stage_words = (size_t)TILE_WARPS * PW * (SLAB + 1);
    const size_t smem = (M == 3) ? (size_t)PT_M1 + 4 * PW + 256 + 8 * 32 * TILE_WARPS
                                 : sizeof(uint32_t) * ((size_t)(M > 1 ? M - 1 : 1) * TW + stage_words + 3 * PW);
    const size_t smem_diag = sizeof(uint32_t) * ((size_t)M * TW + stage_words + TW);
    EVX_CUDA(cudaFuncSetAttribute(k_dominance_tiles<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    EVX_CUDA(cudaFuncSetAttribute(k_dominance_diag<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_diag));
    // rows per CTA = slabs * TILE_WARPS * SLAB (256 .. 1024, always dividing TW): as many passes per
    // warp as still leave a few waves of CTAs on the chip
    int slabs = MAX_SLABS;
    while (slabs > 1 && (int64_t)L.npl * L.np * (TW / (slabs * TILE_WARPS * SLAB)) / 2 < 148 * 5 * 4) slabs >>= 1;
    // M = 3: a CTA stages 45 KB of panel tables, so it takes twice the rows when there are CTAs to spare
    const int slabs_off = (M == 3 && slabs == MAX_SLABS && (int64_t)L.npl * L.np / 4 >= 148 * 4 * 8) ? 2 * MAX_SLABS : slabs;
    dim3 grid((unsigned)ceil_div(L.npad, slabs_off * TILE_WARPS * SLAB), (unsigned)L.npl);
    if (L.np > 1) {
        k_dominance_tiles<M><<<grid, TILE_WARPS * 32, smem, s>>>(P.at<float>(L.ucols), P.at<uint8_t>(L.ptab), L.npad,
                                                                  P.at<Scal>(L.scal), P.at<uint32_t>(L.bits),
                                                                  P.at<int>(L.count), L.world, L.rank, slabs_off);
        EVX_LAUNCH_CHECK();
    }
    // diagonal blocks: one slab per warp, i.e. TW / 256 CTAs per block - there are only NP blocks, and a
    // CTA with four slabs per warp left one CTA per SM at N = 200k (0.127 ms; the full compare is the slow path)
    const int slabs_diag = 1;
    dim3 gdiag((unsigned)ceil_div(TW, slabs_diag * TILE_WARPS * SLAB), (unsigned)L.npl);
    k_dominance_diag<M><<<gdiag, TILE_WARPS * 32, smem_diag, s>>>(P.at<float>(L.ucols), L.npad, P.at<Scal>(L.scal),
                                                                   P.at<uint32_t>(L.bits), P.at<int>(L.count), L.world,
                                                                   L.rank, slabs_diag);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}

static bool force_radix_prep() {                                   // EVX_PREP=radix: the radix-sort path at every size
    const char* e = getenv("EVX_PREP");
    return e && strcmp(e, "radix") == 0;
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
    EVX_CUDA(cudaMemsetAsync(P.at<char>(L.ucols), 0xFF, sizeof(float) * (size_t)m * L.npad, s));
    EVX_CUDA(cudaMemsetAsync(P.at<char>(L.count), 0, sizeof(int) * (size_t)L.npl_max * TW, s));
    EVX_CUDA(cudaMemsetAsync(P.at<char>(L.rank_u), 0, sizeof(int) * (size_t)L.npl_max * TW, s));

    uint32_t* keys = P.at<uint32_t>(L.keys);
    int* idx_cur = P.at<int>(L.idxA);
    int* idx_nxt = P.at<int>(L.idxB);
    k_make_keys<<<gn, T, 0, s>>>(x, n, m, keys, idx_cur, scal);
    EVX_LAUNCH_CHECK();
    if (n <= PREP_SMALL_MAX_N && small_prep_smem(n, m) <= PREP_SMALL_MAX_SMEM && !force_radix_prep()) {
        const size_t sm = small_prep_smem(n, m);
        EVX_CUDA(cudaFuncSetAttribute(k_small_sort_unique, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        k_small_sort_unique<<<1, SU_T, sm, s>>>(x, keys, (int)n, m, L.npad, scal, P.at<float>(L.tcols), P.at<int>(L.row2u));
        EVX_LAUNCH_CHECK();
    } else {
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
                                          P.at<float>(L.tcols), P.at<int>(L.row2u));
        EVX_LAUNCH_CHECK();
    }
    // labels: block-local f1 order, word-local f2 order (one CTA per block)
    k_block_labels<<<(unsigned)L.np, TW, 0, s>>>(P.at<float>(L.tcols), L.npad, m, scal, P.at<float>(L.ucols),
                                                 P.at<int>(L.label));
    EVX_LAUNCH_CHECK();
    k_relabel_rows<<<gn, T, 0, s>>>(P.at<int>(L.row2u), P.at<int>(L.label), n);
    EVX_LAUNCH_CHECK();

    if (L.npl == 0) return EVX_OK;
    if (m == 3 && L.np > 1) {                          // f2 rank tables of the panels (used by the off-diagonal tiles)
        k_panel_tables<<<(unsigned)L.np, TW, 0, s>>>(P.at<float>(L.ucols), L.npad, scal, P.at<uint8_t>(L.ptab));
        EVX_LAUNCH_CHECK();
    }
    int rc = EVX_OK;
    if (tiles_begin) EVX_CUDA(cudaEventRecord(tiles_begin, s));
    switch (m) {
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
        case 12: rc = launch_tiles<12>(P, s); break;
        default: {
            dim3 grid((unsigned)ceil_div(L.npad, TILE_WARPS * 32), (unsigned)L.npl);
            k_dominance_tiles_generic<<<grid, TILE_WARPS * 32, 0, s>>>(
                P.at<float>(L.ucols), L.npad, m, scal, P.at<uint32_t>(L.bits), P.at<int>(L.count), L.world, L.rank);
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


// Peer-mapped buffer of the fused multi-rank peel: [PeerState | counts 3 x stride | pad | segments A | segments B]
static int peer_cnt_stride(const Layout& L) { return L.np + 2; }
static size_t peer_seg_offset(const Layout& L) {
    return align_up(sizeof(PeerState) + sizeof(FWord) * 2 * (size_t)(L.np + 2), 256);
}

static size_t peel_smem_bytes(const Layout& L) {
    return sizeof(DynShared) + sizeof(int) * (size_t)(L.np + 1 + L.npl / 2 + 2);
}
// The persistent kernel keeps two per-front tables in shared memory (one entry per panel / per local
// pair) and scans the panel counts with DYN_MAX_PPT pairs per thread: populations beyond that (N > 2M)
// take the one-front-per-launch path.
static bool persistent_peel_fits(const Layout& L) {
    return L.kplanes <= PEEL_KMAX && peel_smem_bytes(L) <= 227 * 1024 &&
           (L.np + 1) / 2 <= DYN_MAX_PPT * PEEL_WARPS * 32;
}

static bool force_dynamic_peel() {                                 // EVX_PEEL=dynamic: the multi-CTA kernel at every size
    const char* e = getenv("EVX_PEEL");
    return e && strcmp(e, "dynamic") == 0;
}

// Launch the persistent peel.  peers == nullptr: single rank.  Otherwise the fused multi-rank kernel
// with the peer buffers.  Leaves the front count in scal->fronts (-1 after a time-out).
static int peel_persistent(const Plan& P, cudaStream_t s, const PeerTable* peers = nullptr) {
    const Layout& L = P.L;
    EVX_REQUIRE(persistent_peel_fits(L), EVX_ERR_ARG, "persistent peel: population of %lld rows exceeds the kernel's tables",
                (long long)L.n);
    if (!peers && L.world == 1 && L.np <= PS_MAX_PAN && !force_dynamic_peel()) {
        // small population on one GPU: one CTA, no flags between CTAs (k_peel_single)
        EVX_CUDA(cudaFuncSetAttribute(k_peel_single, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SingleShared)));
        k_peel_single<<<1, PS_T, sizeof(SingleShared), s>>>(P.at<uint32_t>(L.bits), P.at<int>(L.count), P.at<int>(L.rank_u),
                                                           P.at<Scal>(L.scal));
        EVX_LAUNCH_CHECK();
        ++evx::launch_counter();
        return EVX_OK;
    }
    int dev = 0, sms = 0, per_sm = 0;
    EVX_CUDA(cudaGetDevice(&dev));
    EVX_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int nlp = L.npl / 2;                                     // local pairs
    const size_t dyn_smem = peel_smem_bytes(L);
    const void* dk = peers ? (const void*)k_peel_dynamic<true> : (const void*)k_peel_dynamic<false>;
    EVX_CUDA(cudaFuncSetAttribute(dk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn_smem));
    EVX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, dk, PEEL_WARPS * 32, dyn_smem));
    EVX_REQUIRE(per_sm >= 1, EVX_ERR_CUDA, "peel kernel does not fit on an SM");
    const int resident = sms * per_sm;                             // co-resident CTAs: the cooperative launch limit
    int dgrid = resident < PEEL_MAX_GRID ? resident : PEEL_MAX_GRID;
    int useful = 16 * (nlp > 0 ? nlp : 1);                          // a small population cannot feed every CTA:
    if (useful < 32) useful = 32;                                   // fewer arrivals per grid barrier
    if (dgrid > useful) dgrid = useful;                             // (never above `resident`)
    EVX_CUDA(cudaMemsetAsync(P.at<unsigned int>(L.pdone), 0, sizeof(unsigned int) * (size_t)L.npl_max, s));
    const uint32_t* bits = P.at<uint32_t>(L.bits);
    const int* count = P.at<int>(L.count);
    int* rank_u = P.at<int>(L.rank_u);
    Scal* scal = P.at<Scal>(L.scal);
    uint32_t* cpl = P.at<uint32_t>(L.cplanes);
    uint32_t* partials = P.at<uint32_t>(L.partials);
    unsigned int* pdone = P.at<unsigned int>(L.pdone);
    int K = L.kplanes, ps = peer_cnt_stride(L), world = L.world, rank = L.rank;
    FWord *pcnt, *fa, *fb;
    PeerTable pt{};
    if (peers) {
        // front segments and per-panel counts live in the peer-mapped buffer (every rank holds all panels);
        // evx_nds_shard_peer_init cleared it once and every launch leaves the counts cleared
        pt = *peers;
        char* mine = pt.base[L.rank];
        pcnt = reinterpret_cast<FWord*>(mine + sizeof(PeerState));
        fa = reinterpret_cast<FWord*>(mine + peer_seg_offset(L));
        fb = fa + L.npad;
    } else {
        pcnt = P.at<FWord>(L.pcnt);
        fa = P.at<FWord>(L.flist);
        fb = P.at<FWord>(L.flist2);
        EVX_CUDA(cudaMemsetAsync(pcnt, 0, sizeof(FWord) * 2 * (size_t)ps, s));     // epoch 0 = never valid
    }
    unsigned long long* ctadbg = P.at<unsigned long long>(L.ctadbg);
    EVX_CUDA(cudaMemsetAsync(ctadbg, 0, sizeof(unsigned long long) * 4 * PEEL_MAX_GRID, s));
    void* args[] = {&bits, &count, &rank_u, &fa, &fb, &scal, &cpl, &pcnt, &ps, &partials, &pdone, &K, &world, &rank, &pt, &ctadbg};
    EVX_CUDA(cudaLaunchCooperativeKernel(dk, dim3(dgrid), dim3(PEEL_WARPS * 32), args, dyn_smem, s));
    ++evx::launch_counter();
    return EVX_OK;
}

static bool use_stepwise_peel() {
    const char* e = getenv("EVX_PEEL");
    return e && strcmp(e, "stepwise") == 0;
}

// Number of fronts of the last call, fetched lazily: the call enqueues a copy of the device-side front
// count into a pinned host word and records an event; evx_nds_last_num_fronts() waits for that event.
// The call itself never blocks the host.
struct FrontCount {
    int32_t value = 0;          // last known value
    int32_t* pinned = nullptr;  // one pinned word per host thread (allocated on first use)
    cudaEvent_t ev = nullptr;
    bool pending = false;
};
static thread_local FrontCount g_fc;

static int front_count_post(const int* dev_word, cudaStream_t s) {
    FrontCount& fc = g_fc;
    fc.pending = false;
    fc.value = -1;
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &cap) != cudaSuccess) { cudaGetLastError(); return EVX_OK; }
    if (cap != cudaStreamCaptureStatusNone) return EVX_OK;           // inside a graph capture: count not reported
    if (!fc.pinned) {
        EVX_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&fc.pinned), sizeof(int32_t), cudaHostAllocDefault));
        EVX_CUDA(cudaEventCreateWithFlags(&fc.ev, cudaEventDisableTiming));
    }
    EVX_CUDA(cudaMemcpyAsync(fc.pinned, dev_word, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    EVX_CUDA(cudaEventRecord(fc.ev, s));
    fc.pending = true;
    return EVX_OK;
}
static void front_count_set(int32_t v) { g_fc.pending = false; g_fc.value = v; }

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

extern "C" size_t evx_nds_shard_peer_bytes(const evx_nds_shard* p) {
    if (!p) return 0;
    const Layout& L = p->P.L;
    return peer_seg_offset(L) + 2 * sizeof(FWord) * (size_t)L.npad;
}

extern "C" int evx_nds_shard_peer_init(evx_nds_shard* p, void* local_peer_buf, void* stream) {
    EVX_REQUIRE(p && local_peer_buf, EVX_ERR_ARG, "evx_nds_shard_peer_init: null argument");
    // epoch base, counter, abort flag and every front word: epoch 0 never validates
    EVX_CUDA(cudaMemsetAsync(local_peer_buf, 0, evx_nds_shard_peer_bytes(p), (cudaStream_t)stream));
    return EVX_OK;
}

extern "C" int evx_nds_shard_peel_fused(evx_nds_shard* p, void* const* peer_bufs, int32_t* num_fronts, void* stream) {
    EVX_REQUIRE(p && peer_bufs, EVX_ERR_ARG, "evx_nds_shard_peel_fused: null argument");
    const Layout& L = p->P.L;
    EVX_REQUIRE(L.world <= MAX_WORLD, EVX_ERR_ARG, "evx_nds_shard_peel_fused: world %d > %d", L.world, MAX_WORLD);
    if (num_fronts) *num_fronts = 0;
    if (L.n == 0) return EVX_OK;
    PeerTable pt{};
    for (int r = 0; r < L.world; ++r) {
        EVX_REQUIRE(peer_bufs[r], EVX_ERR_ARG, "evx_nds_shard_peel_fused: peer buffer %d is null", r);
        pt.base[r] = reinterpret_cast<char*>(peer_bufs[r]);
    }
    pt.world = L.world; pt.rank = L.rank;
    cudaStream_t s = (cudaStream_t)stream;
    EVX_REQUIRE(persistent_peel_fits(L), EVX_ERR_ARG,
                "evx_nds_shard_peel_fused: n=%lld exceeds the fused kernel's tables; use the detect / all-gather / apply loop",
                (long long)L.n);
    int rc = peel_persistent(p->P, s, &pt);
    if (rc) return rc;
    Scal* scal = p->P.at<Scal>(L.scal);
    if (!num_fronts) return front_count_post(&scal->fronts, s);      // asynchronous: evx_nds_last_num_fronts() reads it lazily
    int fronts = 0;
    EVX_CUDA(cudaMemcpyAsync(&fronts, &scal->fronts, sizeof(int), cudaMemcpyDeviceToHost, s));
    EVX_CUDA(cudaStreamSynchronize(s));
    EVX_REQUIRE(fronts >= 0, EVX_ERR_CUDA,
                "evx_nds_shard_peel_fused: a peer did not reach the barrier within 4 s (call evx_nds_shard_peer_init on every "
                "rank again before the next attempt)");
    *num_fronts = fronts;
    front_count_set(fronts);
    return EVX_OK;
}

static thread_local uint64_t g_dbg_times[16];

extern "C" int evx_debug_peel_times(evx_nds_shard* p, uint64_t* out8) {
    EVX_REQUIRE(out8, EVX_ERR_ARG, "evx_debug_peel_times: null argument");
    if (!p) {                                             // last profiled evx_nds_f32 call of this thread
        memcpy(out8, g_dbg_times, sizeof(g_dbg_times));
        return EVX_OK;
    }
    Scal* scal = p->P.at<Scal>(p->P.L.scal);
    EVX_CUDA(cudaMemcpy(out8, scal->dbg, sizeof(uint64_t) * 16, cudaMemcpyDeviceToHost));
    return EVX_OK;
}

// Per-CTA phase clocks of the plan's last peel: out[4 * b + {0,1,2,3}] = ns CTA b spent gathering, in merge +
// commit + publish, waiting for the next front, in the scans (zero for CTAs that did not exist).
extern "C" int evx_debug_peel_cta_times(evx_nds_shard* p, uint64_t* out, int32_t max_ctas) {
    EVX_REQUIRE(p && out && max_ctas > 0, EVX_ERR_ARG, "evx_debug_peel_cta_times: bad argument");
    const int nb = max_ctas < PEEL_MAX_GRID ? max_ctas : PEEL_MAX_GRID;
    EVX_CUDA(cudaMemcpy(out, p->P.at<unsigned long long>(p->P.L.ctadbg), sizeof(uint64_t) * 4 * nb, cudaMemcpyDeviceToHost));
    return EVX_OK;
}

extern "C" int32_t evx_nds_last_num_fronts(void) {
    FrontCount& fc = g_fc;
    if (fc.pending) {
        fc.pending = false;
        if (cudaEventSynchronize(fc.ev) == cudaSuccess) fc.value = *fc.pinned;
        else cudaGetLastError();
    }
    return fc.value;
}

extern "C" int evx_nds_f32(const float* x, int64_t n, int32_t m, int32_t* rank, void* ws, size_t ws_bytes,
                           void* stream) {
    EVX_REQUIRE(n >= 0 && m >= 1 && m <= EVX_MAX_OBJECTIVES, EVX_ERR_ARG, "evx_nds_f32: bad shape n=%lld m=%d",
                (long long)n, m);
    front_count_set(0);
    if (n == 0) return EVX_OK;
    EVX_REQUIRE(x && rank, EVX_ERR_ARG, "evx_nds_f32: null pointer");
    const size_t need = evx_nds_workspace_bytes(n, m);
    EVX_REQUIRE(need != 0, EVX_ERR_ARG, "evx_nds_f32: n=%lld too large", (long long)n);
    EVX_REQUIRE(ws && ws_bytes >= need, EVX_ERR_WORKSPACE, "evx_nds_f32: workspace %zu < required %zu", ws_bytes, need);
    cudaStream_t s = (cudaStream_t)stream;
    Plan P;
    P.L = make_layout(n, m, 1, 0);
    P.ws = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(ws), 256));
    Scal* scal = P.at<Scal>(P.L.scal);
    if (n <= SMALL_PATH_MAX_N && !use_stepwise_peel()) {
        // At most one panel: the sort / tile / persistent-peel pipeline is all launch latency here.
        // Direct n x n test + single-CTA peel: three launches (measured 30 us at n = 200 against 235 us).
        const int words = (int)ceil_div(n, 32);
        int rc = small_nds_launch(x, (int)n, m, rank, P.at<uint32_t>(P.L.bits), words, P.at<int>(P.L.count),
                                  &scal->fronts, s);
        if (rc) return rc;
        g_prof.have = false;
        return front_count_post(&scal->fronts, s);
    }
    Prof& pf = g_prof;
    pf.have = false;
    if (pf.on) {
        for (int i = 0; i < 4; ++i)
            if (!pf.ev[i]) EVX_CUDA(cudaEventCreate(&pf.ev[i]));
        EVX_CUDA(cudaEventRecord(pf.ev[0], s));
    }
    int rc = build(P, x, s, pf.on ? pf.ev[1] : nullptr, pf.on ? pf.ev[2] : nullptr);
    if (rc) return rc;

    const bool stepwise = use_stepwise_peel() || !persistent_peel_fits(P.L);
    if (!stepwise) {
        // one persistent cooperative kernel peels every front: no host round trip, nothing blocks
        if ((rc = peel_persistent(P, s))) return rc;
    } else {
        // Fallback (EVX_PEEL=stepwise, or a population beyond the persistent kernel's tables): one
        // detect + apply pair of launches per front.  The trip count is data dependent
        // (non_dominate.py:78-80), so fronts are enqueued in batches and one flag per front is read back
        // after each batch - this path blocks the host; steps past the last front are no-ops.
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
            if (b < batch) { front_count_set(k + b); break; }
            k += batch;
            if (batch < ANY_SLOTS) batch *= 2;
        }
    }
    k_write_ranks<<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(P.at<int>(P.L.row2u), P.at<int>(P.L.rank_u), n, 1, 0, rank);
    EVX_LAUNCH_CHECK();
    if (!stepwise && (rc = front_count_post(&scal->fronts, s))) return rc;
    if (pf.on) {                                                    // profiling mode only: blocks to read the events
        EVX_CUDA(cudaEventRecord(pf.ev[3], s));
        EVX_CUDA(cudaEventSynchronize(pf.ev[3]));
        EVX_CUDA(cudaEventElapsedTime(&pf.prep_ms, pf.ev[0], pf.ev[1]));
        EVX_CUDA(cudaEventElapsedTime(&pf.tiles_ms, pf.ev[1], pf.ev[2]));
        EVX_CUDA(cudaEventElapsedTime(&pf.peel_ms, pf.ev[2], pf.ev[3]));
        EVX_CUDA(cudaMemcpy(g_dbg_times, scal->dbg, sizeof(uint64_t) * 16, cudaMemcpyDeviceToHost));
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
    if (n == 0) { front_count_set(0); return EVX_OK; }
    EVX_REQUIRE(x_host && rank_host, EVX_ERR_ARG, "evx_nds_host_f32: null pointer");
    const size_t wsb = evx_nds_workspace_bytes(n, m);
    EVX_REQUIRE(wsb != 0, EVX_ERR_ARG, "evx_nds_host_f32: n too large");
    const size_t xb = align_up(sizeof(float) * (size_t)n * m, 256), rb = align_up(sizeof(int) * (size_t)n, 256);
    std::lock_guard<std::recursive_mutex> hold(arena_mutex());      // the arena is this call's until the D2H copy is done
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
