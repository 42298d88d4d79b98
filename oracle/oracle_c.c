/* CPU restatement in plain C of the reference's dominance / peel / crowding path.
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT: only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs load this.  It exists so that parity checks and the
 * timed CPU baseline finish in seconds at N = 10k..50k where the NumPy oracle takes minutes.
 * It follows the same reference lines as oracle/oracle_np.py (paths under /root/reference):
 *   orc_dominate_relation  src/evox/utils/common.py:88-97
 *   orc_nds                src/evox/operators/selection/non_dominate.py:62-63 (matrix, column
 *                          sums) and :13-29 / :74-113 (front peeling on the dense byte matrix)
 *   orc_crowding           src/evox/operators/selection/non_dominate.py:139-158
 * and is pinned against oracle_np.py and the reference's known-answer vectors in
 * tests/test_oracle_c.py.  Like the reference it materialises the N x N byte matrix.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

/* ---- tiny pthread parallel-for (libgomp is not in the image) ---------------------------- */
static int g_threads = 0;
int orc_num_threads(void) {
    if (g_threads <= 0) {
        long c = sysconf(_SC_NPROCESSORS_ONLN);
        g_threads = c < 1 ? 1 : (c > 256 ? 256 : (int)c);
    }
    return g_threads;
}
void orc_set_threads(int t) { g_threads = t < 1 ? 1 : (t > 256 ? 256 : t); }

typedef void (*range_fn)(int64_t lo, int64_t hi, int tid, void* ctx);
typedef struct { range_fn fn; int64_t lo, hi; int tid; void* ctx; } job_t;
static void* job_main(void* p) { job_t* j = (job_t*)p; j->fn(j->lo, j->hi, j->tid, j->ctx); return 0; }
static void parallel_for(int64_t n, range_fn fn, void* ctx) {
    int t = orc_num_threads();
    if (n < t) t = n > 0 ? (int)n : 1;
    pthread_t th[256]; job_t jobs[256];
    for (int i = 0; i < t; ++i) {
        jobs[i].fn = fn; jobs[i].ctx = ctx; jobs[i].tid = i;
        jobs[i].lo = n * i / t; jobs[i].hi = n * (i + 1) / t;
        if (i) pthread_create(&th[i], 0, job_main, &jobs[i]);
    }
    job_main(&jobs[0]);
    for (int i = 1; i < t; ++i) pthread_join(th[i], 0);
}

/* out[i*ny + j] = all_k(x[i,k] <= y[j,k]) && any_k(x[i,k] < y[j,k]) */
typedef struct { const float* x; const float* y; int64_t ny; int m; uint8_t* out; } rel_ctx;
static void rel_rows(int64_t lo, int64_t hi, int tid, void* p) {
    (void)tid;
    rel_ctx* c = (rel_ctx*)p;
    for (int64_t i = lo; i < hi; ++i) {
        const float* xi = c->x + i * c->m;
        uint8_t* o = c->out + i * c->ny;
        for (int64_t j = 0; j < c->ny; ++j) {
            const float* yj = c->y + j * c->m;
            int all_le = 1, any_lt = 0;
            for (int k = 0; k < c->m; ++k) {
                all_le &= (xi[k] <= yj[k]);
                any_lt |= (xi[k] < yj[k]);
            }
            o[j] = (uint8_t)(all_le & any_lt);
        }
    }
}
void orc_dominate_relation(const float* x, int64_t nx, const float* y, int64_t ny, int m, uint8_t* out) {
    rel_ctx c = { x, y, ny, m, out };
    parallel_for(nx, rel_rows, &c);
}

/* column-block workers for the peel: each thread owns a contiguous range of columns */
typedef struct { const uint8_t* dom; const uint8_t* front; int64_t* count; int64_t n; int all_rows; } col_ctx;
static void col_sub(int64_t lo, int64_t hi, int tid, void* p) {
    (void)tid;
    col_ctx* c = (col_ctx*)p;
    for (int64_t jb = lo; jb < hi; jb += 4096) {
        int64_t je = jb + 4096 < hi ? jb + 4096 : hi;
        for (int64_t i = 0; i < c->n; ++i) {
            if (!c->all_rows && !c->front[i]) continue;
            const uint8_t* row = c->dom + i * c->n;
            if (c->all_rows) for (int64_t j = jb; j < je; ++j) c->count[j] += row[j];
            else             for (int64_t j = jb; j < je; ++j) c->count[j] -= row[j];
        }
    }
}

/* rank[n]; returns number of fronts, or -1 if the matrix cannot be allocated. */
int orc_nds(const float* x, int64_t n, int m, int32_t* rank) {
    if (n == 0) return 0;
    uint8_t* dom = (uint8_t*)malloc((size_t)n * (size_t)n);
    int64_t* count = (int64_t*)calloc((size_t)n, sizeof(int64_t));
    uint8_t* front = (uint8_t*)malloc((size_t)n);
    if (!dom || !count || !front) { free(dom); free(count); free(front); return -1; }
    orc_dominate_relation(x, n, x, n, m, dom);
    /* column sums (:63) */
    col_ctx cc = { dom, front, count, n, 1 };
    parallel_for(n, col_sub, &cc);
    cc.all_rows = 0;
    for (int64_t i = 0; i < n; ++i) rank[i] = 0;
    int k = 0;
    for (;;) {
        int64_t nf = 0;
        for (int64_t i = 0; i < n; ++i) { front[i] = (count[i] == 0); nf += front[i]; }
        if (!nf) break;
        for (int64_t i = 0; i < n; ++i) if (front[i]) rank[i] = k;
        /* count -= sum of the front's rows (:24 / :94-102), column-parallel */
        parallel_for(n, col_sub, &cc);
        for (int64_t i = 0; i < n; ++i) count[i] -= front[i];
        ++k;
    }
    free(dom); free(count); free(front);
    return k;
}

typedef struct { float c; uint8_t inv; int32_t idx; } key_t_;

static int key_cmp(const void* a, const void* b) {
    const key_t_* p = (const key_t_*)a; const key_t_* q = (const key_t_*)b;
    if (p->inv != q->inv) return (int)p->inv - (int)q->inv;          /* valid rows first      */
    int pn = isnan(p->c), qn = isnan(q->c);
    if (pn != qn) return pn - qn;                                     /* NaN last              */
    if (!pn) { if (p->c < q->c) return -1; if (p->c > q->c) return 1; }   /* -0 == +0           */
    return (p->idx > q->idx) - (p->idx < q->idx);                     /* stable                */
}

/* mask may be NULL (all valid).  out[n] float32. */
void orc_crowding(const float* costs, const uint8_t* mask, int64_t n, int m, float* out) {
    if (n == 0) return;
    int64_t nv = 0;
    for (int64_t i = 0; i < n; ++i) nv += mask ? (mask[i] != 0) : 1;
    key_t_* keys = (key_t_*)malloc((size_t)n * sizeof(key_t_));
    float* d = (float*)malloc((size_t)n * sizeof(float));
    int64_t last = (nv - 1 + n) % n;                                  /* nv==0 wraps to n-1    */
    for (int k = 0; k < m; ++k) {
        for (int64_t i = 0; i < n; ++i) {
            keys[i].c = costs[i * m + k];
            keys[i].inv = mask ? (mask[i] == 0) : 0;
            keys[i].idx = (int32_t)i;
        }
        qsort(keys, (size_t)n, sizeof(key_t_), key_cmp);
        float range = keys[last].c - keys[0].c;
        for (int64_t i = 0; i < n; ++i) d[i] = 0.0f;
        for (int64_t i = 1; i + 1 < n; ++i) d[keys[i].idx] = (keys[i + 1].c - keys[i - 1].c) / range;
        d[keys[0].idx] = INFINITY;
        d[keys[last].idx] = INFINITY;
        for (int64_t i = 0; i < n; ++i) {
            float v = (mask && !mask[i]) ? -INFINITY : d[i];
            out[i] = k ? out[i] + v : v;
        }
    }
    free(keys); free(d);
}
