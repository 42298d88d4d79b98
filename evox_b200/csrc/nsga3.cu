// nsga3.cu - NSGA-III's association and niching.
//
// Replaces algorithms/mo/nsga3.py:140-147 (association of every solution with its nearest reference
// direction: `cos_dist` N x R, perpendicular distance, nanargmin / nanmin) and :148-205 (niche counts and
// the one-survivor-per-iteration `while_loop`, up to pop_size sequential steps, each of which touches
// several arrays of N elements in the reference).  The normalisation before it (:103-139: ideal point,
// extreme points, a tiny M x M solve) stays host glue.
//   associate - a thread takes up to four solutions, the reference directions are staged in shared memory
//               already divided by their norm and split eight ways over the CTA's warps; nothing N x R is
//               stored.  The distance is monotone in 1 - c^2, so the square root is only taken when a
//               direction improves on the best so far.
//   niche     - the last front's members are sorted by (group, index) once (radix sort); one persistent
//               CTA then runs the reference's loop LEVEL BY LEVEL instead of pick by pick: the loop always
//               serves the first niche with the smallest count, so all niches at the smallest count c are
//               served in index order before any niche at c + 1 - within a level the picks are independent,
//               and the loop index of each (which selects its random words) is its position in the level.
//               One step per level: minimum, flags, one scan, then every thread makes the picks of its own
//               niches; a niche with many candidates left is served by the whole CTA.
// The random words are an INPUT (two 32-bit words per loop index, what randint(keys[i], ...) consumes),
// so the host framework's PRNG stream stays the reference's.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace evx {

constexpr int AS_T = 256;
constexpr int AS_TILE = 128;

// group_id[i] = nanargmin_r dist(i, r) (R if every distance is NaN), group_dist[i] = nanmin (NaN then).
// dist = ||x|| * sqrt(1 - c^2), c = (x / ||x||) . (y_r / ||y_r||)   (utils/common.py:75-80, nsga3.py:140-143)
__global__ void __launch_bounds__(AS_T)
k_nsga3_associate(const float* __restrict__ x, const float* __restrict__ ref, int n, int R, int m, int* __restrict__ gid,
                  float* __restrict__ gdist) {
    extern __shared__ float tile[];                   // [AS_TILE][m] normalised reference directions
    const int i = blockIdx.x * AS_T + threadIdx.x;
    float xn[EVX_MAX_OBJECTIVES];
    float nx = 0.0f;
    if (i < n) {
        float acc = 0.0f;
        for (int k = 0; k < m; ++k) { const float v = x[(size_t)i * m + k]; acc = __fadd_rn(acc, __fmul_rn(v, v)); }
        nx = __fsqrt_rn(acc);
        for (int k = 0; k < m; ++k) xn[k] = __fdiv_rn(x[(size_t)i * m + k], nx);
    }
    float best = __int_as_float(0x7FC00000);          // NaN until a number is seen
    int bi = R;
    for (int r0 = 0; r0 < R; r0 += AS_TILE) {
        const int nt = min(AS_TILE, R - r0);
        __syncthreads();
        for (int t = threadIdx.x; t < nt; t += AS_T) {
            const float* y = ref + (size_t)(r0 + t) * m;
            float acc = 0.0f;
            for (int k = 0; k < m; ++k) acc = __fadd_rn(acc, __fmul_rn(y[k], y[k]));
            const float ny = __fsqrt_rn(acc);
            for (int k = 0; k < m; ++k) tile[t * m + k] = __fdiv_rn(y[k], ny);
        }
        __syncthreads();
        if (i < n) {
            for (int t = 0; t < nt; ++t) {
                float c = 0.0f;
                for (int k = 0; k < m; ++k) c = __fadd_rn(c, __fmul_rn(xn[k], tile[t * m + k]));
                const float d = __fmul_rn(nx, __fsqrt_rn(__fsub_rn(1.0f, __fmul_rn(c, c))));
                if (d == d && !(best <= d)) { best = d; bi = r0 + t; }      // first minimum, NaN skipped
            }
        }
    }
    if (i < n) { gid[i] = bi; gdist[i] = best; }
}

// The same for a compile-time number of objectives.  A CTA takes 32 * RPT solutions (lane r: solutions
// q * 32 + r) and splits the reference directions eight ways over its warps (warp sp: 16 of every 128 staged
// directions), so that 100 000 solutions still fill the chip and each staged direction is read once for RPT
// solutions; FMA dot product; d = |x| sqrt(1 - c^2) only when 1 - c^2 is below the split's best so far (sqrt
// and the product are monotone under round-to-nearest, so a direction with a larger or equal 1 - c^2 can
// never be strictly closer: the first minimum is unchanged).  The eight splits' results are merged by
// (distance, direction index): the first minimum over all directions.
constexpr int AS_SPLIT = AS_T / 32;
template <int M, int RPT>
__global__ void __launch_bounds__(AS_T)
k_nsga3_associate_t(const float* __restrict__ x, const float* __restrict__ ref, int n, int R, int* __restrict__ gid,
                    float* __restrict__ gdist) {
    extern __shared__ float tile[];                   // [AS_TILE][M]
    __shared__ float md[AS_SPLIT][32 * RPT];
    __shared__ int mi[AS_SPLIT][32 * RPT];
    const int r = threadIdx.x & 31, sp = threadIdx.x >> 5;
    float xn[RPT][M], nx[RPT], best[RPT], best_s[RPT];
    int bi[RPT];
#pragma unroll
    for (int q = 0; q < RPT; ++q) {
        const int i = blockIdx.x * (32 * RPT) + q * 32 + r;
        nx[q] = 0.0f;
        best[q] = __int_as_float(0x7FC00000);         // NaN until a number is seen
        best_s[q] = INFINITY;
        bi[q] = R;
        if (i < n) {
            float acc = 0.0f;
#pragma unroll
            for (int k = 0; k < M; ++k) { xn[q][k] = x[(size_t)i * M + k]; acc = __fadd_rn(acc, __fmul_rn(xn[q][k], xn[q][k])); }
            nx[q] = __fsqrt_rn(acc);
#pragma unroll
            for (int k = 0; k < M; ++k) xn[q][k] = __fdiv_rn(xn[q][k], nx[q]);
        } else {
#pragma unroll
            for (int k = 0; k < M; ++k) xn[q][k] = __int_as_float(0x7FC00000);      // never improves
        }
    }
    constexpr int PER = AS_TILE / AS_SPLIT;           // directions of a tile per split
    for (int r0 = 0; r0 < R; r0 += AS_TILE) {
        const int nt = min(AS_TILE, R - r0);
        __syncthreads();
        for (int t = threadIdx.x; t < nt; t += AS_T) {
            const float* y = ref + (size_t)(r0 + t) * M;
            float acc = 0.0f;
#pragma unroll
            for (int k = 0; k < M; ++k) acc = __fadd_rn(acc, __fmul_rn(y[k], y[k]));
            const float ny = __fsqrt_rn(acc);
#pragma unroll
            for (int k = 0; k < M; ++k) tile[t * M + k] = __fdiv_rn(y[k], ny);
        }
        __syncthreads();
        const int te = min(nt, (sp + 1) * PER);
        for (int t = sp * PER; t < te; ++t) {
            float y[M];
#pragma unroll
            for (int k = 0; k < M; ++k) y[k] = tile[t * M + k];
#pragma unroll
            for (int q = 0; q < RPT; ++q) {
                float c = __fmul_rn(xn[q][0], y[0]);
#pragma unroll
                for (int k = 1; k < M; ++k) c = __fmaf_rn(xn[q][k], y[k], c);
                const float sq = __fmaf_rn(-c, c, 1.0f);
                if (sq < best_s[q]) {                                           // false for NaN
                    const float d = __fmul_rn(nx[q], __fsqrt_rn(sq));
                    if (d == d && !(best[q] <= d)) { best[q] = d; best_s[q] = sq; bi[q] = r0 + t; }
                }
            }
        }
    }
#pragma unroll
    for (int q = 0; q < RPT; ++q) { md[sp][q * 32 + r] = best[q]; mi[sp][q * 32 + r] = bi[q]; }
    __syncthreads();
    if (sp == 0) {
#pragma unroll
        for (int q = 0; q < RPT; ++q) {
            const int i = blockIdx.x * (32 * RPT) + q * 32 + r;
            float bd = __int_as_float(0x7FC00000);
            int bj = R;
#pragma unroll
            for (int w = 0; w < AS_SPLIT; ++w) {
                const float d = md[w][q * 32 + r];
                const int j = mi[w][q * 32 + r];
                if (j < R && (bj == R || d < bd || (d == bd && j < bj))) { bd = d; bj = j; }
            }
            if (i < n) { gid[i] = bj; gdist[i] = bd; }
        }
    }
}

template <int M>
static void launch_associate_t(const float* x, const float* ref, int n, int R, int* gid, float* gd, cudaStream_t s) {
    constexpr int RPT = M <= 12 ? 4 : 2;
    k_nsga3_associate_t<M, RPT><<<(unsigned)ceil_div(n, 32 * RPT), AS_T, sizeof(float) * AS_TILE * M, s>>>(x, ref, n, R, gid, gd);
}

// ---- niching ------------------------------------------------------------------------------------------
struct NicheScal { int selected; };

__global__ void k_niche_counts(const int* __restrict__ rank, const int* __restrict__ gid, int n, int R, int last_rank,
                               int* __restrict__ rho, int* __restrict__ rho_last, uint32_t* __restrict__ keys,
                               int* __restrict__ vals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int g = gid[i], r = rank[i];
    if (g >= 0 && g < R) {
        if (r < last_rank) atomicAdd(&rho[g], 1);                           // :148-150
        else if (r == last_rank) atomicAdd(&rho_last[g], 1);                // :151-153
    }
    keys[i] = (r == last_rank && g >= 0 && g < R) ? (uint32_t)g : (uint32_t)R;   // candidates first, by group; stable = by index
    vals[i] = i;
}

constexpr int NI_T = 1024;
constexpr int NI_HEAVY = 64;          // a niche with more entries than this is served by a warp, not by its owner thread
constexpr int RHO_INF = 0x7FFFFFFF;

// One CTA.  cand[] = last-front members sorted by (group, index); off[g] = first entry of group g;
// taken = one bit per entry; live[] = the niches that have candidates; heavy[] = (niche, loop index) of
// the level's picks that are left to a warp each.
__global__ void __launch_bounds__(NI_T)
k_niche_loop(int* __restrict__ rank, const float* __restrict__ gdist, const int* __restrict__ cand, int n, int R,
             int last_rank, int pop_size, int* __restrict__ rho, int* __restrict__ rho_last, int* __restrict__ off,
             uint32_t* taken, const uint32_t* __restrict__ rbits, int* __restrict__ selected_idx,
             int* __restrict__ nsel_out, int* __restrict__ heavy, int* __restrict__ live) {
    __shared__ int wv[32], wi[32], wc[32];
    __shared__ int s_sel, s_nheavy, s_nlive;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int chunk = (R + NI_T - 1) / NI_T;
    // offsets (exclusive scan of rho_last), selected = sum(rho), rho = inf where the group has no candidate (:158-160)
    {
        int part = 0, lsum = 0;
        for (int g = t * chunk; g < min(R, (t + 1) * chunk); ++g) { part += rho_last[g]; lsum += rho[g]; }
        int incl = part, tot = lsum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += v;
        }
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) tot += __shfl_xor_sync(0xFFFFFFFFu, tot, o);
        if (lane == 31) wv[warp] = incl;
        if (lane == 0) wi[warp] = tot;
        __syncthreads();
        int run = incl - part;
        for (int w = 0; w < warp; ++w) run += wv[w];
        if (t == 0) { int s = 0; for (int w = 0; w < 32; ++w) s += wi[w]; s_sel = s; }
        for (int g = t * chunk; g < min(R, (t + 1) * chunk); ++g) {
            off[g] = run;
            run += rho_last[g];
            if (rho_last[g] == 0) rho[g] = RHO_INF;
        }
        if (t * chunk < R && min(R, (t + 1) * chunk) == R) off[R] = run;    // end of the last group's entries
        __syncthreads();
    }
    // ---- the niches that have candidates, ascending, compact: thread t owns live[t * lchunk ..) ------------
    {
        int c = 0;
        for (int g = t * chunk; g < min(R, (t + 1) * chunk); ++g) c += rho_last[g] > 0 ? 1 : 0;
        int incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) wv[warp] = incl;
        __syncthreads();
        int run = incl - c, tot = 0;
        for (int w = 0; w < 32; ++w) { if (w < warp) run += wv[w]; tot += wv[w]; }
        for (int g = t * chunk; g < min(R, (t + 1) * chunk); ++g)
            if (rho_last[g] > 0) live[run++] = g;
        if (t == 0) s_nlive = tot;
        __syncthreads();
    }
    const int nlive = s_nlive;
    const int lchunk = (nlive + NI_T - 1) / NI_T;
    const int l_lo = min(nlive, t * lchunk), l_hi = min(nlive, (t + 1) * lchunk);
    // jax.random.randint(key, (), 0, span) from the two words it consumes (oracle/jax_prng.randint_from_bits)
    auto draw = [&](int loop_index, int left) -> int {
        const uint32_t span = (uint32_t)left;
        const uint32_t hi = rbits[2 * loop_index], lo = rbits[2 * loop_index + 1];
        uint32_t mult = (65536u % span) * (65536u % span);
        mult %= span;
        uint32_t offv = (hi % span) * mult + (lo % span);
        return (int)(offv % span);
    };
    // The group's entries are cand[off[g] .. off[g + 1]) in index order; bit q of `taken` = entry q is gone.
    // Remaining entries of word w that belong to [first, end):
    auto avail = [&](int w, int first, int end) -> uint32_t {
        uint32_t m = ~__ldcg(&taken[w]);                                     // (bits are set by atomics: read them where those land)
        const int lo = w * 32;
        if (first > lo) m &= 0xFFFFFFFFu << (first - lo);
        if (end < lo + 32) m &= (end <= lo) ? 0u : (0xFFFFFFFFu >> (lo + 32 - end));
        return m;
    };
    auto retire = [&](int group, int level, int pick, int left) {
        atomicOr(&taken[pick >> 5], 1u << (pick & 31));                     // neighbouring groups share boundary words
        rank[cand[pick]] = last_rank - 1;                                   // :189
        rho_last[group] = left - 1;                                         // :191
        rho[group] = (left - 1 == 0) ? RHO_INF : level + 1;                 // :193-201
    };
    // One pick by ONE thread (group with at most NI_HEAVY entries: three words at most).
    auto light_pick = [&](int group, int level, int loop_index) {
        const int first = off[group], end = off[group + 1];
        const int left = rho_last[group];
        int pick = -1;
        if (level == 0 && left > 1) {                                       // :187: closest remaining candidate
            float bd = INFINITY;
            for (int w = first >> 5; w <= (end - 1) >> 5; ++w) {
                uint32_t m = avail(w, first, end);
                while (m) {
                    const int q = w * 32 + __ffs(m) - 1;
                    m &= m - 1;
                    const float d = gdist[cand[q]];
                    if (pick < 0 || d < bd) { bd = d; pick = q; }           // ascending index: first minimum
                }
            }
        } else {                                                            // :173-183: the rand_index-th remaining, by index
            int k = left == 1 ? 0 : draw(loop_index, left);                 // (one left: it is also the closest)
            for (int w = first >> 5; w <= (end - 1) >> 5; ++w) {
                const uint32_t m = avail(w, first, end);
                const int c = __popc(m);
                if (k < c) { pick = w * 32 + (int)__fns(m, 0, k + 1); break; }
                k -= c;
            }
        }
        retire(group, level, pick, left);
    };
    // The same pick by ONE WARP, for a group with many entries: lane l takes a contiguous run of words.
    auto heavy_pick = [&](int group, int level, int loop_index) {
        const int first = off[group], end = off[group + 1], left = rho_last[group];
        const int w0 = first >> 5, nw = ((end - 1) >> 5) - w0 + 1;
        const int per = (nw + 31) / 32;
        const int wa = w0 + min(nw, lane * per), wb = w0 + min(nw, (lane + 1) * per);
        if (level == 0 && left > 1) {
            float bd = INFINITY;
            int bq = 0x7FFFFFFF;
            for (int w = wa; w < wb; ++w) {
                uint32_t m = avail(w, first, end);
                while (m) {
                    const int q = w * 32 + __ffs(m) - 1;
                    m &= m - 1;
                    const float d = gdist[cand[q]];
                    if (bq == 0x7FFFFFFF || d < bd) { bd = d; bq = q; }
                }
            }
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) {
                const float od = __shfl_xor_sync(0xFFFFFFFFu, bd, o);
                const int oq = __shfl_xor_sync(0xFFFFFFFFu, bq, o);
                if (oq != 0x7FFFFFFF && (bq == 0x7FFFFFFF || od < bd || (od == bd && oq < bq))) { bd = od; bq = oq; }
            }
            if (lane == 0) retire(group, level, bq, left);
        } else {
            const int k = left == 1 ? 0 : draw(loop_index, left);
            int c = 0;
            for (int w = wa; w < wb; ++w) c += __popc(avail(w, first, end));
            int incl = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
                if (lane >= o) incl += v;
            }
            int kk = k - (incl - c);
            if (kk >= 0 && kk < c) {                                        // exactly one lane
                for (int w = wa; w < wb; ++w) {
                    const uint32_t m = avail(w, first, end);
                    const int cw = __popc(m);
                    if (kk < cw) { retire(group, level, w * 32 + (int)__fns(m, 0, kk + 1), left); break; }
                    kk -= cw;
                }
            }
        }
        __syncwarp();
    };

    int selected = s_sel;
    while (selected < pop_size) {                                           // :196-200, one level per turn
        // The smallest count: every niche that has it is served once, in index order (:169 takes the first).
        // One barrier for both the minimum and the positions: per thread (its minimum, how many of its niches
        // have it), per warp the same, and every thread folds the 32 warp entries.
        int bv = RHO_INF, cnt = 0;
        for (int i = l_lo; i < l_hi; ++i) {
            const int v = rho[live[i]];
            if (v < bv) { bv = v; cnt = 1; } else if (v == bv && v != RHO_INF) ++cnt;
        }
        int wmin = bv;
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) wmin = min(wmin, __shfl_xor_sync(0xFFFFFFFFu, wmin, o));
        const int mine = (bv == wmin && bv != RHO_INF) ? cnt : 0;          // my niches at the warp's minimum
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) { wv[warp] = wmin; wc[warp] = incl; }
        if (t == 0) s_nheavy = 0;
        __syncthreads();
        int level = RHO_INF;
        for (int w = 0; w < 32; ++w) level = min(level, wv[w]);
        if (level == RHO_INF) break;                                        // no candidate anywhere (cannot happen for pop_size < n)
        int before = 0, total = 0;
        for (int w = 0; w < 32; ++w) {
            const int v = wv[w] == level ? wc[w] : 0;
            if (w < warp) before += v;
            total += v;
        }
        if (wmin == level) before += incl - mine;
        // the niche at position p of the level is loop index selected + p (its random words); positions
        // past pop_size are never reached by the reference's loop
        if (bv == level) {
            int p = selected + before;
            for (int i = l_lo; i < l_hi; ++i) {
                const int g = live[i];
                if (rho[g] != level) continue;
                if (p < pop_size) {
                    if (off[g + 1] - off[g] > NI_HEAVY) {
                        const int slot = atomicAdd(&s_nheavy, 1);
                        heavy[2 * slot] = g;
                        heavy[2 * slot + 1] = p;
                    } else {
                        light_pick(g, level, p);
                    }
                }
                ++p;
            }
        }
        __syncthreads();
        const int nh = s_nheavy;
        for (int h = warp; h < nh; h += NI_T / 32) heavy_pick(heavy[2 * h], level, heavy[2 * h + 1]);
        selected += min(total, pop_size - selected);
        if (nh) __syncthreads();                                            // (uniform) the warps' updates of rho before the next minimum
    }
    __syncthreads();
    // survivors: rank < last_rank, ascending index, first pop_size (:202-204)
    __shared__ int base;
    if (t == 0) base = 0;
    __syncthreads();
    for (int i0 = 0; i0 < n; i0 += NI_T) {
        const int i = i0 + t;
        const bool keep = i < n && rank[i] < last_rank;
        const uint32_t b = __ballot_sync(0xFFFFFFFFu, keep);
        if (lane == 0) wv[warp] = __popc(b);
        __syncthreads();
        int before = base;
        for (int w = 0; w < warp; ++w) before += wv[w];
        const int pos = before + __popc(b & ((1u << lane) - 1u));
        if (keep && pos < pop_size) selected_idx[pos] = i;
        __syncthreads();
        if (t == 0) { int s = 0; for (int w = 0; w < 32; ++w) s += wv[w]; base += s; }
        __syncthreads();
    }
    if (t == 0 && nsel_out) *nsel_out = base;
}

struct NicheLayout {
    size_t rank, rho, rho_last, off, heavy, live, keysA, keysB, valsA, valsB, taken, tmp, tmp_bytes, total;
};
static NicheLayout niche_layout(int64_t n, int64_t R) {
    NicheLayout L{};
    size_t o = 0;
    auto take = [&](size_t b) { size_t r = o; o = align_up(o + b, 256); return r; };
    L.rank = take(sizeof(int) * (size_t)n);
    L.rho = take(sizeof(int) * (size_t)(R + 1));
    L.rho_last = take(sizeof(int) * (size_t)(R + 1));
    L.off = take(sizeof(int) * (size_t)(R + 1));
    L.heavy = take(sizeof(int) * 2 * (size_t)(R + 1));
    L.live = take(sizeof(int) * (size_t)(R + 1));
    L.keysA = take(sizeof(uint32_t) * (size_t)n);
    L.keysB = take(sizeof(uint32_t) * (size_t)n);
    L.valsA = take(sizeof(int) * (size_t)n);
    L.valsB = take(sizeof(int) * (size_t)n);
    L.taken = take(sizeof(uint32_t) * ((size_t)n / 32 + 2));
    cub::DeviceRadixSort::SortPairs(nullptr, L.tmp_bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const int*)nullptr,
                                    (int*)nullptr, (int)n);
    L.tmp = take(L.tmp_bytes);
    L.total = o;
    return L;
}

}  // namespace evx

using namespace evx;

extern "C" int evx_nsga3_associate_f32(const float* normalized, const float* ref, int64_t n, int64_t nref, int32_t m,
                                       int32_t* group_id, float* group_dist, void* stream) {
    EVX_REQUIRE(n >= 0 && nref >= 1 && nref < 0x7FFFFFFF && n < 0x7FFFFFFF && m >= 1 && m <= EVX_MAX_OBJECTIVES, EVX_ERR_ARG,
                "evx_nsga3_associate_f32: bad shape n=%lld nref=%lld m=%d", (long long)n, (long long)nref, m);
    if (n == 0) return EVX_OK;
    EVX_REQUIRE(normalized && ref && group_id && group_dist, EVX_ERR_ARG, "evx_nsga3_associate_f32: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    switch (m) {
#define EVX_ASSOC_CASE(MM) case MM: launch_associate_t<MM>(normalized, ref, (int)n, (int)nref, group_id, group_dist, s); break;
        EVX_ASSOC_CASE(2) EVX_ASSOC_CASE(3) EVX_ASSOC_CASE(4) EVX_ASSOC_CASE(5) EVX_ASSOC_CASE(6) EVX_ASSOC_CASE(7) EVX_ASSOC_CASE(8)
        EVX_ASSOC_CASE(9) EVX_ASSOC_CASE(10) EVX_ASSOC_CASE(11) EVX_ASSOC_CASE(12) EVX_ASSOC_CASE(13) EVX_ASSOC_CASE(14)
        EVX_ASSOC_CASE(15) EVX_ASSOC_CASE(16)
#undef EVX_ASSOC_CASE
        default:
            k_nsga3_associate<<<(unsigned)ceil_div(n, AS_T), AS_T, sizeof(float) * AS_TILE * m, s>>>(normalized, ref, (int)n, (int)nref, m,
                                                                                               group_id, group_dist);
    }
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}

extern "C" size_t evx_nsga3_niche_workspace_bytes(int64_t n, int64_t nref) {
    if (n < 1 || nref < 1 || n >= 0x7FFFFFFF || nref >= 0x7FFFFFFF) return 0;
    return niche_layout(n, nref).total + 256;
}

extern "C" int evx_nsga3_niche(const int32_t* rank, int32_t last_rank, const int32_t* group_id, const float* group_dist,
                               int64_t n, int64_t nref, int64_t pop_size, const uint32_t* rand_words, int32_t* selected_idx,
                               void* ws, size_t ws_bytes, void* stream) {
    const size_t need = evx_nsga3_niche_workspace_bytes(n, nref);
    EVX_REQUIRE(need != 0, EVX_ERR_ARG, "evx_nsga3_niche: bad shape n=%lld nref=%lld", (long long)n, (long long)nref);
    EVX_REQUIRE(pop_size >= 1 && pop_size < n, EVX_ERR_ARG, "evx_nsga3_niche: pop_size=%lld must be in [1, n)", (long long)pop_size);
    EVX_REQUIRE(rank && group_id && group_dist && rand_words && selected_idx, EVX_ERR_ARG, "evx_nsga3_niche: null pointer");
    EVX_REQUIRE(ws && ws_bytes >= need, EVX_ERR_WORKSPACE, "evx_nsga3_niche: workspace %zu < required %zu", ws_bytes, need);
    cudaStream_t s = (cudaStream_t)stream;
    char* base = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(ws), 256));
    const NicheLayout L = niche_layout(n, nref);
    int* rk = reinterpret_cast<int*>(base + L.rank);
    int* rho = reinterpret_cast<int*>(base + L.rho);
    int* rho_last = reinterpret_cast<int*>(base + L.rho_last);
    int* off = reinterpret_cast<int*>(base + L.off);
    int* heavy = reinterpret_cast<int*>(base + L.heavy);
    uint32_t* kA = reinterpret_cast<uint32_t*>(base + L.keysA);
    uint32_t* kB = reinterpret_cast<uint32_t*>(base + L.keysB);
    int* vA = reinterpret_cast<int*>(base + L.valsA);
    int* vB = reinterpret_cast<int*>(base + L.valsB);
    uint32_t* taken = reinterpret_cast<uint32_t*>(base + L.taken);
    int* live = reinterpret_cast<int*>(base + L.live);
    EVX_CUDA(cudaMemcpyAsync(rk, rank, sizeof(int) * (size_t)n, cudaMemcpyDeviceToDevice, s));
    EVX_CUDA(cudaMemsetAsync(rho, 0, sizeof(int) * (size_t)(nref + 1), s));
    EVX_CUDA(cudaMemsetAsync(rho_last, 0, sizeof(int) * (size_t)(nref + 1), s));
    EVX_CUDA(cudaMemsetAsync(taken, 0, sizeof(uint32_t) * ((size_t)n / 32 + 2), s));
    k_niche_counts<<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(rk, group_id, (int)n, (int)nref, last_rank, rho, rho_last, kA, vA);
    EVX_LAUNCH_CHECK();
    int bits = 1;
    while ((1ll << bits) <= nref) ++bits;
    size_t tb = L.tmp_bytes;
    EVX_CUDA(cub::DeviceRadixSort::SortPairs(base + L.tmp, tb, kA, kB, vA, vB, (int)n, 0, bits, s));
    k_niche_loop<<<1, NI_T, 0, s>>>(rk, group_dist, vB, (int)n, (int)nref, last_rank, (int)pop_size, rho, rho_last, off, taken,
                                    rand_words, selected_idx, nullptr, heavy, live);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}
