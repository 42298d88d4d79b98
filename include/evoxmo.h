/* evoxmo.h - C ABI of libevoxmo.so: EvoX's multi-objective environmental-selection hot path
 * (dominance, non-dominated sort, crowding distance, DTLZ/LSMOP/MaF1 evaluate) as hand-written
 * CUDA for sm_100a (B200).
 *
 * Conventions
 *   - Every pointer marked "device" is a CUDA device pointer owned by the caller (XLA, torch, ...);
 *     the library never frees or retains it.  Inputs are read-only.  Arrays are row-major,
 *     float32 / int32 / uint8 exactly as the reference holds them (x64 is never enabled there).
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  All work is
 *     enqueued on it.  Functions documented "synchronises" block the host on that stream.
 *   - Scratch memory comes from the caller: ask `*_workspace_bytes`, pass `ws`/`ws_bytes`
 *     (device, 256-byte aligned).  The `*_host_*` entry points take HOST buffers and use a
 *     grow-only per-device arena owned by the library instead.
 *   - Return value: 0 on success, non-zero on failure; `evx_last_error()` returns a
 *     thread-local message.  No C++ exception, abort or exit crosses this boundary.
 *   - Reference citations are paths under the EvoX source tree (src/evox/...).
 */
#ifndef EVOXMO_H_
#define EVOXMO_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EVX_OK 0
#define EVX_ERR_ARG 1        /* bad shape / null pointer / unsupported m                 */
#define EVX_ERR_WORKSPACE 2  /* workspace missing, misaligned or too small               */
#define EVX_ERR_CUDA 3       /* a CUDA runtime call or kernel failed                     */
#define EVX_ERR_NOMEM 4      /* arena allocation failed (host entry points)              */

#define EVX_MAX_OBJECTIVES 32

int evx_version(void);
const char* evx_last_error(void);

/* ---- (a1) dominance relation ------------------------------------------------------------------
 * replaces utils/common.py:88-97  `dominate_relation(x, y)`:
 *   out[i, j] = all_k(x[i,k] <= y[j,k]) && any_k(x[i,k] < y[j,k]),  IEEE compares (NaN -> false).
 * x: device float32 [nx, m], y: device float32 [ny, m].
 * _u8  : out device uint8  [nx, ny]           (the reference's bool matrix, 1 byte per pair)
 * _bits: out device uint32 [nx, ceil(ny/32)]  (bit j%32 of word j/32 = pair (i,j); tail bits 0)   */
int evx_dominate_relation_u8(const float* x, int64_t nx, const float* y, int64_t ny, int32_t m,
                             uint8_t* out, void* stream);
int evx_dominate_relation_bits(const float* x, int64_t nx, const float* y, int64_t ny, int32_t m,
                               uint32_t* out, void* stream);

/* ---- (a2) non-dominated sort ------------------------------------------------------------------
 * replaces operators/selection/non_dominate.py:32-115  `non_dominated_sort(x, method)`
 * (all three `method`s compute the same ranks; this is the one implementation behind them).
 * x: device float32 [n, m];  rank: device int32 [n], front index starting at 0, bit-exact.
 * Asynchronous: the data-dependent front loop (non_dominate.py:78-80) runs inside one persistent kernel,
 * the host never waits.  (Two documented exceptions block the host: n > 2M rows, and EVX_PEEL=stepwise in
 * the environment - both take a one-front-per-launch path that reads one flag per front back.)         */
size_t evx_nds_workspace_bytes(int64_t n, int32_t m);
int evx_nds_f32(const float* x, int64_t n, int32_t m, int32_t* rank, void* ws, size_t ws_bytes,
                void* stream);
/* Same, host buffers in and out (H2D + D2H inside; library arena for scratch).                    */
int evx_nds_host_f32(const float* x_host, int64_t n, int32_t m, int32_t* rank_host);
/* Number of fronts found by the last evx_nds_* / evx_non_dominate_f32 call on this host thread.  The call
 * leaves a copy of the device-side count in flight; this function waits for it (it is the only place
 * that blocks).  -1: not available (call made during a stream capture, or a peer timed out).        */
int32_t evx_nds_last_num_fronts(void);

/* ---- non-dominated sort without any device allocation -------------------------------------------
 * No counterpart in the reference (its non_dominated_sort always gets fresh arrays from XLA).
 * For callers that cannot obtain device scratch memory at the point of the call: a custom-call
 * handler that is handed only its input and output buffers, a callback that runs while the host
 * framework's allocator is locked, a process whose free device memory all belongs to another pool.
 * The scratch (n x n dominance bits + counts, about 2.1 MB) is a fixed array inside the library
 * image, present on every GPU that loads the library until the process exits; hence the limit
 * n <= EVX_NDS_NOALLOC_MAX_N, beyond which the call fails with EVX_ERR_ARG and a message.
 * The call allocates nothing (no cudaMalloc / cudaFree / library temp storage), is asynchronous on
 * `stream`, and gives the same ranks as evx_nds_f32.  x: device float32 [n, m]; rank: device int32 [n].
 * - The first call on each device resolves the array's device address (cudaGetSymbolAddress, with
 *   that device current); make that first call before starting any stream capture.
 * - All calls on a device share the one array, which keeps the previous call's contents; what has to
 *   start from zero is cleared by the call itself.  Two calls that overlap on the same GPU (different
 *   streams or threads) race on it: serialise them.                                                 */
#define EVX_NDS_NOALLOC_MAX_N 4096
int evx_nds_noalloc_f32(const float* x, int64_t n, int32_t m, int32_t* rank, void* stream);

/* ---- (a3) rank from a dense domination matrix ------------------------------------------------
 * replaces operators/selection/non_dominate.py:13-29  `host_rank_from_domination_matrix`
 * (the `method="host"` pure_callback, :65-72), on the device.
 * dom: device uint8 [n, n] (dom[i,j] != 0: i dominates j); count: device int32 [n] column sums, or
 * NULL to have them computed.  rank: device int32 [n].  Synchronises.                              */
size_t evx_rank_from_matrix_workspace_bytes(int64_t n);
int evx_rank_from_matrix_u8(const uint8_t* dom, const int32_t* count, int64_t n, int32_t* rank,
                            void* ws, size_t ws_bytes, void* stream);

/* ---- (a4, a5) crowding distance ---------------------------------------------------------------
 * replaces operators/selection/non_dominate.py:118-158 `crowding_distance(costs, mask)` and
 * :161-180 `crowding_distance_sort(x, mask)`.
 * costs: device float32 [n, m]; mask: device uint8 [n] (non-zero = participates) or NULL (all).
 * dist: device float32 [n] (+inf boundaries, -inf masked rows, exact);
 * order: device int32 [n] = stable ascending argsort of dist (NaN last).  Asynchronous.            */
size_t evx_crowding_workspace_bytes(int64_t n, int32_t m);
int evx_crowding_distance_f32(const float* costs, const uint8_t* mask, int64_t n, int32_t m,
                              float* dist, void* ws, size_t ws_bytes, void* stream);
int evx_crowding_distance_sort_f32(const float* costs, const uint8_t* mask, int64_t n, int32_t m,
                                   int32_t* order, void* ws, size_t ws_bytes, void* stream);

/* ---- (f1) survivor selection --------------------------------------------------------------------
 * replaces operators/selection/non_dominate.py:210-222, the body of `non_dominate(population, fitness,
 * topk, deduplicate)` after its optional de-duplication (:188-208 stays with the caller): non-dominated
 * sort, cut rank, crowding-distance permutation of the cut front, lexsort((-permutation, rank))[:topk]
 * - including the reference's choice of the argsort PERMUTATION as tie-breaker (:216/:218) - and the
 * gather of the chosen rows, in the reference's output order.
 * pop: device float32 [n, d] (may be NULL with out_pop NULL); fit: device float32 [n, m], what is ranked;
 * fit_out_src: device float32 [n, m] gathered into out_fit (NULL = fit; the de-duplicating caller passes
 * the fitness backup of :201); 1 <= topk <= n.  Outputs (each may be NULL): out_pop [topk, d],
 * out_fit [topk, m], out_index int32 [topk] (row numbers chosen, for populations that are not float32).
 * Asynchronous on `stream`.                                                                          */
size_t evx_non_dominate_workspace_bytes(int64_t n, int32_t m);
int evx_non_dominate_f32(const float* pop, const float* fit, const float* fit_out_src, int64_t n, int32_t d,
                         int32_t m, int64_t topk, float* out_pop, float* out_fit, int32_t* out_index,
                         void* ws, size_t ws_bytes, void* stream);

/* The same selection on ranks the caller already holds (device int32 [n], values in [0, n)): what the
 * sharded driver uses after the multi-GPU sort has produced the ranks (non_dominate.py:213-222).     */
int evx_select_by_rank_f32(const float* pop, const float* fit, const float* fit_out_src, const int32_t* rank,
                           int64_t n, int32_t d, int32_t m, int64_t topk, float* out_pop, float* out_fit,
                           int32_t* out_index, void* ws, size_t ws_bytes, void* stream);

/* ---- (f4) SPEA2 fitness and archive truncation ---------------------------------------------------
 * replaces algorithms/mo/spea2.py:21-37 `cal_fitness(obj)`: fitness[j] = 1 / (d2[j] + 2) + sum of the
 * strengths s[i] (= rows i dominates) over the rows i dominating j, d2 = distance to the second nearest
 * other row (sorted-row index floor(sqrt(6)) - 1); and :40-67 `_truncation(obj, k, mask)`: k times
 * remove the valid row whose nearest valid neighbour is closest (ties: lowest index), keep[i] = row i is
 * valid and was not removed.  Nothing n x n is stored.  obj: device float32 [n, m]; fitness: device
 * float32 [n]; mask, keep: device uint8 [n]; k < number of valid rows, as the reference requires.
 * Asynchronous.                                                                                      */
size_t evx_spea2_workspace_bytes(int64_t n, int32_t m);
int evx_spea2_fitness_f32(const float* obj, int64_t n, int32_t m, float* fitness, void* ws, size_t ws_bytes,
                          void* stream);
int evx_spea2_truncation_f32(const float* obj, int64_t n, int32_t m, int64_t k, const uint8_t* mask, uint8_t* keep,
                             void* ws, size_t ws_bytes, void* stream);

/* ---- (f3) NSGA-III association and niching -------------------------------------------------------
 * replaces algorithms/mo/nsga3.py:140-147: group_id[i] = nearest reference direction of row i by
 * perpendicular distance ||x|| * sqrt(1 - cos^2) (`cos_dist`, utils/common.py:75-80), group_dist[i] = that
 * distance; a row of NaN (beyond the last front) gets group nref and distance NaN.
 * normalized: device float32 [n, m]; ref: device float32 [nref, m].  Asynchronous.                   */
int evx_nsga3_associate_f32(const float* normalized, const float* ref, int64_t n, int64_t nref, int32_t m,
                            int32_t* group_id, float* group_dist, void* stream);
/* replaces nsga3.py:148-205: niche counts of the fronts before `last_rank`, then the niche-preserving
 * fill from the last front - the reference's one-survivor-per-iteration while_loop, served here one level
 * of niche counts per step with the same picks and the same use of the random words - and the final index list.
 * rank: device int32 [n] (not modified); rand_words: device uint32 [2 * (pop_size + 1)], words (2i, 2i+1)
 * = the two 32-bit draws jax.random.randint(keys[i], (), 0, span) consumes, keys = split(state.key,
 * pop_size + 1) (:165) - the caller supplies them so that the PRNG stream stays the host framework's;
 * selected_idx: device int32 [pop_size], ascending row numbers of the survivors.  1 <= pop_size < n.
 * Asynchronous.                                                                                      */
size_t evx_nsga3_niche_workspace_bytes(int64_t n, int64_t nref);
int evx_nsga3_niche(const int32_t* rank, int32_t last_rank, const int32_t* group_id, const float* group_dist,
                    int64_t n, int64_t nref, int64_t pop_size, const uint32_t* rand_words, int32_t* selected_idx,
                    void* ws, size_t ws_bytes, void* stream);

/* ---- (a7) DTLZ1..7 evaluate -------------------------------------------------------------------
 * replaces problems/numerical/dtlz.py:48-69, 84-103, 115-145, 152-172, 188-208, 244-265, 303-322
 * `DTLZk.evaluate(state, X)`.  X: device float32 [n, d]; F: device float32 [n, m].  Asynchronous. */
int evx_eval_dtlz_f32(int32_t which, const float* X, int64_t n, int32_t d, int32_t m, float* F,
                      void* stream);

/* ---- (a9) MaF1 evaluate -----------------------------------------------------------------------
 * replaces problems/numerical/maf.py:95-108 `MaF1.evaluate`.                                      */
int evx_eval_maf1_f32(const float* X, int64_t n, int32_t d, int32_t m, float* F, void* stream);

/* ---- (a8) LSMOP1..9 evaluate ------------------------------------------------------------------
 * replaces problems/numerical/lsmop.py:127-141 ... 400-426 `LSMOPk.evaluate` with `_calc_g`
 * (:88-111).  sublen: HOST int32 [m] and off: HOST int32 [m+1] are the constructor's group lengths
 * and offsets (lsmop.py:46-54; `evx_lsmop_groups` computes them the same way, in float32).        */
int evx_lsmop_groups(int32_t d, int32_t m, int32_t* sublen_host, int32_t* off_host);
int evx_eval_lsmop_f32(int32_t which, const float* X, int64_t n, int32_t d, int32_t m,
                       const int32_t* sublen_host, const int32_t* off_host, float* F, void* stream);

/* ---- (f2) NSGA-II variation -------------------------------------------------------------------
 * replaces operators/crossover/sbx.py:8-40 (type 1), operators/mutation/pm_mutation.py:7-54 and the
 * clip of algorithms/mo/nsga2.py:85 in one kernel.  pop, out: device float32 [n, d] (rows p and
 * p + n/2 are mated; an odd last row passes through); lb, ub: device float32 [d].  Random numbers
 * come from Philox4x32-10 keyed by (seed, generation): reproducible, statistically - not bitwise -
 * equivalent to the reference's jax.random draws.  Asynchronous.                                   */
int evx_variation_sbx_pm_f32(const float* pop, int64_t n, int32_t d, const float* lb, const float* ub,
                             float pro_c, float dis_c, float pro_m, float dis_m, uint32_t seed,
                             uint32_t generation, float* out, void* stream);

/* ---- sharded non-dominated sort (one process per GPU; column panels dealt round-robin) --------
 * The exchange step between ranks - all-gather of each front's membership - is done by the caller
 * (torch.distributed / NCCL) between these calls; see INTEGRATION.md.  No reference counterpart:
 * the reference replicates selection on every device (workflows/std_workflow.py:160,187-200).     */
typedef struct evx_nds_shard evx_nds_shard;   /* opaque plan living in host memory                 */
size_t evx_nds_shard_workspace_bytes(int64_t n, int32_t m, int32_t world, int32_t rank);
evx_nds_shard* evx_nds_shard_create(int64_t n, int32_t m, int32_t world, int32_t rank, void* ws,
                                    size_t ws_bytes);
void evx_nds_shard_destroy(evx_nds_shard* p);
/* sort + unique + this rank's panels of the bit matrix + their column counts (asynchronous)       */
int evx_nds_shard_build(evx_nds_shard* p, const float* x, void* stream);
/* number of uint32 words each rank contributes to the per-front membership mask                   */
int64_t evx_nds_shard_mask_words(const evx_nds_shard* p);
/* detect this rank's members of front k -> local_mask (device, mask_words); asynchronous          */
int evx_nds_shard_detect(evx_nds_shard* p, int32_t k, uint32_t* local_mask, void* stream);
/* apply the all-gathered masks (device, world*mask_words, rank-major) of front k to this rank's
 * counts; `any_out` (device int32) is set non-zero if the front was non-empty; asynchronous       */
int evx_nds_shard_apply(evx_nds_shard* p, int32_t k, const uint32_t* all_masks, int32_t* any_out,
                        void* stream);
/* Fused alternative to the detect / all-gather / apply loop: ONE persistent kernel per rank peels
 * every front.  Every rank holds a replica of the current front (per-panel member lists and sizes, as
 * 64-bit words tagged with the front's epoch) in peer-mapped memory; the warp that finds a panel's new
 * members stores them straight into every rank's replica over NVLink (plain 8-byte peer stores: no
 * cross-GPU atomics on data, no system-scope fences).  There is no barrier between fronts: a CTA starts
 * the next front when every panel's size word carries that front's epoch.  Each rank provides
 * `peer_bytes` of peer-mapped memory (e.g. torch symmetric memory); peer_bufs[r] (HOST array of world
 * device pointers) is rank r's buffer as mapped in this process.
 * Sequence: peer_init ONCE per plan -> host-level barrier across ranks -> then, per population,
 * build -> peel_fused -> ranks, with no further hand-shake between calls (epochs continue across
 * launches).  num_fronts == NULL: asynchronous, the front count is available from
 * evx_nds_last_num_fronts(); otherwise the call waits and fails with a message (never hangs) if a rank
 * was more than 4 s late - after which peer_init + host barrier are needed again.  world <= 8, n <= 2M. */
size_t evx_nds_shard_peer_bytes(const evx_nds_shard* p);
int evx_nds_shard_peer_init(evx_nds_shard* p, void* local_peer_buf, void* stream);
int evx_nds_shard_peel_fused(evx_nds_shard* p, void* const* peer_bufs, int32_t* num_fronts, void* stream);
/* ranks of the columns this rank owns, scattered to original row order into rank_partial (device
 * int32 [n], zero elsewhere: a SUM all-reduce over ranks yields the final ranks); asynchronous    */
int evx_nds_shard_ranks(evx_nds_shard* p, int32_t* rank_partial, void* stream);

/* ---- instrumentation --------------------------------------------------------------------------
 * kernels launched by this library on this host thread since the last reset (bench.py reports it) */
int64_t evx_launch_count(void);
void evx_launch_count_reset(void);
/* per-stage device times (CUDA events on the call's stream) of the last evx_nds_f32 call made on
 * this host thread while enabled: sort+unique, dominance tile kernel, front peeling.              */
void evx_profile_enable(int32_t on);
int evx_profile_last(float* prep_ms, float* tiles_ms, float* peel_ms);
/* nanoseconds the persistent peel kernel of the plan's last run spent in {gathering, merge + commit +
 * publish, waiting for the next front's sizes, number of fronts} as seen by its first CTA (out16[0..3])
 * and its last CTA (out16[4..7]); out16[8..9] / out16[12..13]: of the waiting time, the part until the arrival counter
 * said "complete", and the time of the two scans that follow (first / last CTA).  p == NULL: the values of this host thread's
 * last evx_nds_f32 call made with evx_profile_enable(1).                                            */
int evx_debug_peel_times(evx_nds_shard* p, uint64_t* out16);
/* the same four clocks for every CTA of the plan's last (fused or single-rank shard) peel: out[4*b + {0,1,2,3}] =
 * ns gathering, merge + commit + publish, waiting, scans; max_ctas <= 1024 entries are written.     */
int evx_debug_peel_cta_times(evx_nds_shard* p, uint64_t* out, int32_t max_ctas);
#ifdef __cplusplus
}
#endif
#endif /* EVOXMO_H_ */
