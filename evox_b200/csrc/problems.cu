// problems.cu - fused DTLZ1-7 / MaF1 / LSMOP1-9 evaluation over a population.
//
// Replaces problems/numerical/dtlz.py (evaluate methods :48-69, 84-103, 115-145, 152-172, 188-208,
// 244-265, 303-322), problems/numerical/maf.py:95-108 and problems/numerical/lsmop.py:88-111,
// 127-426 with the inner functions sphere.py:5-6, griewank.py:6-12, rosenbrock.py:6-12,
// ackley.py:7-13.  The reference issues ~10 small XLA fusions per call; here each problem is one
// kernel: a CTA stages a block of rows of X in shared memory with coalesced loads, and all
// reductions, running products and the shape function happen on chip in float32, in the
// reference's left-to-right order.
#include <math.h>

#include "common.cuh"

namespace evx {

constexpr int MAXM = EVX_MAX_OBJECTIVES;
#define EVX_PI_D 3.14159265358979323846
__device__ constexpr float kHalfPi = (float)(EVX_PI_D / 2.0);
__device__ constexpr float k20Pi = (float)(20.0 * EVX_PI_D);
__device__ constexpr float k3Pi = (float)(3.0 * EVX_PI_D);
__device__ constexpr float k2Pi = (float)(2.0 * EVX_PI_D);
__device__ constexpr float kE = 2.718281828459045f;

// lax.integer_pow(x, 100): square-and-multiply, every product rounded to float32 (dtlz.py:154)
__device__ __forceinline__ float pow100(float x) {
    const float x2 = x * x, x4 = x2 * x2, x8 = x4 * x4, x16 = x8 * x8, x32 = x16 * x16, x64 = x32 * x32;
    return (x4 * x32) * x64;
}

enum Shape { SH_LINEAR = 0, SH_SPHERE_CLAMP = 1, SH_SPHERE = 2 };

// f_i = (scale_i * cumprod_rev_i) * last_i   (dtlz.py:62-67, :87-101; lsmop.py:137-141, :246-256)
template <int SHAPE>
__device__ __forceinline__ void shape_write(const float* xp, int m, const float* scale, bool scale_per_obj,
                                            float* f) {
    float cp[MAXM];
    cp[0] = 1.0f;
    for (int t = 1; t < m; ++t) {
        float v = xp[t - 1];
        if (SHAPE == SH_SPHERE_CLAMP) v = fmaxf(cosf(v * kHalfPi), 0.0f);
        if (SHAPE == SH_SPHERE) v = cosf(v * kHalfPi);
        cp[t] = cp[t - 1] * v;
    }
    for (int i = 0; i < m; ++i) {
        const float sc = scale_per_obj ? scale[i] : scale[0];
        float last = 1.0f;
        if (i > 0) {
            const float v = xp[m - 1 - i];
            last = (SHAPE == SH_LINEAR) ? (1.0f - v) : sinf(v * kHalfPi);
        }
        f[i] = (sc * cp[m - 1 - i]) * last;
    }
}

// DTLZ7 / LSMOP9 tail (dtlz.py:306-321, lsmop.py:409-425)
__device__ __forceinline__ void shape_dtlz7(const float* xp, int m, float g, float* f) {
    float s = 0.0f;
    const float og = 1.0f + g;
    for (int i = 0; i < m - 1; ++i) {
        f[i] = xp[i];
        s += xp[i] / og * (1.0f + sinf(k3Pi * xp[i]));
    }
    f[m - 1] = og * ((float)m - s);
}

// One thread per individual.  X rows are short (d = m + 4 .. m + 19): a warp's 32 rows are one
// contiguous run of 32*d floats, so the strided per-thread walk is served from L1 lines that are
// fully used.  which: 1..7 = DTLZ, 8 = MaF1.
__global__ void __launch_bounds__(128)
k_eval_dtlz(int which, const float* __restrict__ X, int64_t n, int d, int m, float* __restrict__ F) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n) return;
    const float* x = X + row * d;
    float xp[MAXM];                                  // position variables x_0 .. x_{m-2}
    for (int t = 0; t < m - 1; ++t) xp[t] = (t < d) ? __ldg(x + t) : 0.0f;
    float f[MAXM];
    float g = 0.0f;
    if (which == 1 || which == 3) {
        for (int j = m - 1; j < d; ++j) {
            const float t = __ldg(x + j) - 0.5f;
            g += t * t - cosf(k20Pi * t);
        }
        g = 100.0f * ((float)(d - m + 1) + g);
    } else if (which == 6) {
        for (int j = m - 1; j < d; ++j) g += powf(__ldg(x + j), 0.1f);
    } else if (which == 7) {
        for (int j = m - 1; j < d; ++j) g += __ldg(x + j);
        g = 1.0f + 9.0f * (g / (float)(d - m + 1));
    } else {
        for (int j = m - 1; j < d; ++j) {
            const float t = __ldg(x + j) - 0.5f;
            g += t * t;
        }
    }
    if (which == 4) for (int t = 0; t < m - 1; ++t) xp[t] = pow100(xp[t]);
    if (which == 5 || which == 6) {
        for (int t = 1; t < m - 1; ++t) xp[t] = (1.0f + 2.0f * g * xp[t]) / (2.0f + 2.0f * g);
    }
    float scale;
    switch (which) {
        case 1: scale = 0.5f * (1.0f + g); shape_write<SH_LINEAR>(xp, m, &scale, false, f); break;
        case 7: shape_dtlz7(xp, m, g, f); break;
        case 8: {
            scale = 1.0f + g;
            shape_write<SH_LINEAR>(xp, m, &scale, false, f);
            for (int i = 0; i < m; ++i) f[i] = scale - f[i];               // maf.py:107
            break;
        }
        default: scale = 1.0f + g; shape_write<SH_SPHERE_CLAMP>(xp, m, &scale, false, f); break;
    }
    float* o = F + row * m;
    for (int i = 0; i < m; ++i) o[i] = f[i];
}

// ---- LSMOP -----------------------------------------------------------------------------------------
struct LsmopGroups { int sublen[MAXM]; int off[MAXM + 1]; };
enum Inner { IN_SPHERE = 0, IN_GRIEWANK, IN_ROSENBROCK, IN_ACKLEY, IN_SCHWEFEL, IN_RASTRIGIN };

__device__ __forceinline__ float inner_eval(int fn, const float* x, int len) {
    float r = 0.0f;
    switch (fn) {
        case IN_SPHERE:
            for (int t = 0; t < len; ++t) r += x[t] * x[t];
            return r;
        case IN_GRIEWANK: {
            float p = 1.0f;
            for (int t = 0; t < len; ++t) {
                r += x[t] * x[t];
                p *= cosf(x[t] / sqrtf((float)(t + 1)));
            }
            return 0.00025f * r - p + 1.0f;
        }
        case IN_ROSENBROCK:
            for (int t = 0; t + 1 < len; ++t) {
                const float a = x[t], b = x[t + 1];
                const float u = b - a * a, v = a - 1.0f;
                r += 100.0f * (u * u) + v * v;
            }
            return r;
        case IN_ACKLEY: {
            float c = 0.0f;
            for (int t = 0; t < len; ++t) {
                r += x[t] * x[t];
                c += cosf(k2Pi * x[t]);
            }
            const float ms = r / (float)len, mc = c / (float)len;
            return -20.0f * expf(-0.2f * sqrtf(ms)) - expf(mc) + 20.0f + kE;
        }
        case IN_SCHWEFEL: {
            float mx = -INFINITY;
            for (int t = 0; t < len; ++t) mx = fmaxf(mx, fabsf(x[t]));
            return mx;
        }
        default:   // IN_RASTRIGIN
            for (int t = 0; t < len; ++t) r += x[t] * x[t] - 10.0f * cosf(k2Pi * x[t]) + 10.0f;
            return r;
    }
}

// A CTA stages R rows of X in shared memory (coalesced), applies the variable linkage in place
// (lsmop.py:130-133 / :243-247), evaluates the m*nk sub-block functions one per thread-task, then
// one thread per row finishes g and the shape function.
__global__ void __launch_bounds__(256)
k_eval_lsmop(int which, const float* __restrict__ X, int64_t n, int d, int m, int R, LsmopGroups G, int fn0, int fn1,
             int nfn, float* __restrict__ F) {
    constexpr int NK = 5;
    extern __shared__ float sm[];
    float* xs = sm;                                   // [R][d]
    float* res = sm + (size_t)R * d;                  // [R][m][NK]
    float* gs = res + (size_t)R * m * NK;             // [R][m]
    const int64_t row0 = (int64_t)blockIdx.x * R;
    const int rows = (int)min((int64_t)R, n - row0);
    const float* src = X + row0 * d;
    for (int t = threadIdx.x; t < rows * d; t += blockDim.x) xs[t] = src[t];
    __syncthreads();
    for (int t = threadIdx.x; t < rows * d; t += blockDim.x) {
        const int r = t / d, c = t - r * d;
        if (c >= m - 1) {
            const float jd = (float)(c + 1) / (float)d;
            const float w = (which <= 4) ? (1.0f + jd) : (1.0f + cosf(jd * kHalfPi));
            xs[t] = w * xs[t] - src[(int64_t)r * d] * 10.0f;       // x_0 of the untouched input
        }
    }
    __syncthreads();
    const int tasks = rows * m * NK;
    for (int t = threadIdx.x; t < tasks; t += blockDim.x) {
        const int r = t / (m * NK), rem = t - r * (m * NK);
        const int i = rem / NK, b = rem - i * NK;
        const int sl = G.sublen[i];
        int start = G.off[i] + m - 1 + b * sl;
        int end = start + sl;
        start = min(start, d);
        end = min(end, d);
        const int fn = (nfn == 1 || (i & 1) == 0) ? fn0 : fn1;
        res[t] = inner_eval(fn, xs + (size_t)r * d + start, end - start);
    }
    __syncthreads();
    for (int t = threadIdx.x; t < rows * m; t += blockDim.x) {
        const int i = t % m;
        const float* rr = res + (size_t)t * NK;
        float s = rr[0];
        for (int b = 1; b < NK; ++b) s += rr[b];
        gs[t] = s / (float)G.sublen[i] / (float)NK;               // lsmop.py:110
    }
    __syncthreads();
    for (int r = threadIdx.x; r < rows; r += blockDim.x) {
        const float* xp = xs + (size_t)r * d;                      // columns < m-1 are untouched
        const float* g = gs + (size_t)r * m;
        float f[MAXM], sc[MAXM];
        if (which <= 4) {
            for (int i = 0; i < m; ++i) sc[i] = 1.0f + g[i];
            shape_write<SH_LINEAR>(xp, m, sc, true, f);
        } else if (which <= 8) {
            for (int i = 0; i < m; ++i) sc[i] = 1.0f + g[i] + (i + 1 < m ? g[i + 1] : 0.0f);
            shape_write<SH_SPHERE>(xp, m, sc, true, f);
        } else {
            float s = g[0];
            for (int i = 1; i < m; ++i) s += g[i];
            shape_dtlz7(xp, m, 1.0f + s, f);
        }
        float* o = F + (row0 + r) * m;
        for (int i = 0; i < m; ++i) o[i] = f[i];
    }
}

// ---- pipelined variant ---------------------------------------------------------------------------------
// LSMOP is the one evaluate that is bound by reading X (d = 100 m columns per row).  Persistent CTAs walk
// over blocks of R consecutive rows - one contiguous piece of X - and fetch block k+1 with ONE bulk copy
// (cp.async.bulk, completion on an mbarrier) while block k is evaluated, so the loads of an SM are always
// in flight.  The 5 m sub-block functions of a row are cut into LS_SPLIT partial reductions each (sums, a
// product, a maximum): 64 task slots per row whose column ranges are tabulated once per CTA, as are the
// linkage weights of the columns - the linkage is applied while the values are read, not in a pass of
// its own, and no index arithmetic with divisions is left in the per-block loop.
#ifndef EVX_LS_STAGE
#define EVX_LS_STAGE 20480
#endif
constexpr int LS_STAGE_BYTES = EVX_LS_STAGE;
constexpr int LS_SPLIT = 1;                           // partial reductions per sub-block (1: a thread takes a whole sub-block)
constexpr int LS_SLOTS = 64;                          // at most this many task slots per row (m * 5 * LS_SPLIT <= 64), else the simple kernel
struct Partial { float a, b; };                      // griewank: a = sum, b = product; ackley: two sums; others: a

template <class Get>
__device__ __forceinline__ Partial inner_partial(int fn, Get x, int len, int t0, int t1) {
    Partial p{0.0f, (fn == IN_GRIEWANK) ? 1.0f : 0.0f};
    switch (fn) {
        case IN_SPHERE:
#pragma unroll 4
            for (int t = t0; t < t1; ++t) { const float v = x(t); p.a += v * v; }
            break;
        case IN_GRIEWANK:
            for (int t = t0; t < t1; ++t) { const float v = x(t); p.a += v * v; p.b *= cosf(v / sqrtf((float)(t + 1))); }
            break;
        case IN_ROSENBROCK:
            for (int t = t0; t < min(t1, len - 1); ++t) {
                const float a = x(t), b = x(t + 1);
                const float u = b - a * a, v = a - 1.0f;
                p.a += 100.0f * (u * u) + v * v;
            }
            break;
        case IN_ACKLEY:
            for (int t = t0; t < t1; ++t) { const float v = x(t); p.a += v * v; p.b += cosf(k2Pi * v); }
            break;
        case IN_SCHWEFEL:
            p.a = -INFINITY;
            for (int t = t0; t < t1; ++t) p.a = fmaxf(p.a, fabsf(x(t)));
            break;
        default:   // IN_RASTRIGIN
            for (int t = t0; t < t1; ++t) { const float v = x(t); p.a += v * v - 10.0f * cosf(k2Pi * v) + 10.0f; }
            break;
    }
    return p;
}
__device__ __forceinline__ float inner_finish(int fn, const Partial* ps, int len) {
    float a = ps[0].a, b = ps[0].b;
    for (int q = 1; q < LS_SPLIT; ++q) {
        if (fn == IN_SCHWEFEL) a = fmaxf(a, ps[q].a); else a += ps[q].a;
        if (fn == IN_GRIEWANK) b *= ps[q].b; else b += ps[q].b;
    }
    switch (fn) {
        case IN_GRIEWANK: return 0.00025f * a - b + 1.0f;
        case IN_ACKLEY: return -20.0f * expf(-0.2f * sqrtf(a / (float)len)) - expf(b / (float)len) + 20.0f + kE;
        default: return a;
    }
}

__device__ __forceinline__ uint32_t ls_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(256)
k_eval_lsmop_pipe(int which, const float* __restrict__ X, int64_t nblocks, int d, int m, int R, int slots, LsmopGroups G,
                  int fn0, int fn1, int nfn, float* __restrict__ F) {
    constexpr int NK = 5;
    extern __shared__ __align__(128) unsigned char ls_raw[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(ls_raw);                          // [2]
    int4* ttab = reinterpret_cast<int4*>(ls_raw + 128);                            // [LS_SLOTS] (start, t0, t1, len | fn << 24)
    float* stage0 = reinterpret_cast<float*>(ls_raw + 128 + sizeof(int4) * LS_SLOTS);
    const size_t stage_floats = (size_t)R * d;
    float* wtab = stage0 + 2 * stage_floats;                                       // [d] linkage weight of each column
    Partial* part = reinterpret_cast<Partial*>(wtab + ((d + 3) & ~3));             // [R][slots]
    float* gs = reinterpret_cast<float*>(part + (size_t)R * slots);                // [R][m]
    const uint32_t bytes = (uint32_t)(stage_floats * sizeof(float));
    const int ntask = m * NK * LS_SPLIT;                                           // <= slots (a power of two <= LS_SLOTS)
    if (threadIdx.x == 0) {
        for (int q = 0; q < 2; ++q) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(ls_smem_u32(bars + q)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int c = threadIdx.x; c < d; c += blockDim.x) {                            // lsmop.py:130-133 / :243-247
        const float jd = (float)(c + 1) / (float)d;
        wtab[c] = (which <= 4) ? (1.0f + jd) : (1.0f + cosf(jd * kHalfPi));
    }
    if (threadIdx.x < LS_SLOTS) {
        int4 e = make_int4(0, 0, 0, 0);
        if ((int)threadIdx.x < ntask) {
            const int q = threadIdx.x % LS_SPLIT, u = threadIdx.x / LS_SPLIT, i = u / NK, bk = u - i * NK;
            const int sl = G.sublen[i];
            const int start = min(G.off[i] + m - 1 + bk * sl, d), end = min(G.off[i] + m - 1 + bk * sl + sl, d);
            const int len = end - start, cl = (len + LS_SPLIT - 1) / LS_SPLIT;
            const int fn = (nfn == 1 || (i & 1) == 0) ? fn0 : fn1;
            e = make_int4(start, min(q * cl, len), min((q + 1) * cl, len), len | (fn << 24));
        }
        ttab[threadIdx.x] = e;
    }
    __syncthreads();
    auto issue = [&](int64_t b, int st) {              // thread 0: fetch block b into stage st
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // the stage's earlier generic-proxy reads are done
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(ls_smem_u32(bars + st)), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         ls_smem_u32(stage0 + (size_t)st * stage_floats)),
                     "l"(X + b * (int64_t)stage_floats), "r"(bytes), "r"(ls_smem_u32(bars + st))
                     : "memory");
    };
    const int slot = threadIdx.x & (slots - 1), rsub = threadIdx.x / slots, rstep = blockDim.x / slots;
    const int4 my = ttab[slot];
    const int my_len = my.w & 0xFFFFFF, my_fn = my.w >> 24;
    int it = 0;
    if (threadIdx.x == 0 && (int64_t)blockIdx.x < nblocks) issue(blockIdx.x, 0);
    for (int64_t b = blockIdx.x; b < nblocks; b += gridDim.x, ++it) {
        const int st = it & 1;
        if (threadIdx.x == 0 && b + gridDim.x < nblocks) issue(b + gridDim.x, st ^ 1);
        {
            const uint32_t parity = (uint32_t)(it >> 1) & 1u;
            asm volatile(
                "{\n\t.reg .pred p;\n\tLSW_%=:\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
                "@p bra LSD_%=;\n\tbra LSW_%=;\n\tLSD_%=:\n\t}" ::"r"(ls_smem_u32(bars + st)),
                "r"(parity)
                : "memory");
        }
        const float* xs = stage0 + (size_t)st * stage_floats;
        if (slot < ntask) {
            const float* wt = wtab + my.x;
            for (int r = rsub; r < R; r += rstep) {
                const float* xr = xs + (size_t)r * d;
                const float x0 = xr[0] * 10.0f;
                const float* xv = xr + my.x;
                part[r * slots + slot] = inner_partial(my_fn, [&](int t) { return wt[t] * xv[t] - x0; }, my_len, my.y, my.z);
            }
        }
        __syncthreads();
        for (int t = threadIdx.x; t < R * m; t += blockDim.x) {
            const int i = t % m, r = t / m;
            float ssum = 0.0f;
            for (int bk = 0; bk < NK; ++bk) {
                const int4 e = ttab[(i * NK + bk) * LS_SPLIT];
                const float v = inner_finish(e.w >> 24, part + r * slots + (i * NK + bk) * LS_SPLIT, e.w & 0xFFFFFF);
                ssum = (bk == 0) ? v : ssum + v;
            }
            gs[t] = ssum / (float)G.sublen[i] / (float)NK;                // lsmop.py:110
        }
        __syncthreads();
        for (int r = threadIdx.x; r < R; r += blockDim.x) {
            const float* xp = xs + (size_t)r * d;                          // the position variables (columns < m-1) carry no linkage
            const float* g = gs + (size_t)r * m;
            float f[MAXM], sc[MAXM];
            if (which <= 4) {
                for (int i = 0; i < m; ++i) sc[i] = 1.0f + g[i];
                shape_write<SH_LINEAR>(xp, m, sc, true, f);
            } else if (which <= 8) {
                for (int i = 0; i < m; ++i) sc[i] = 1.0f + g[i] + (i + 1 < m ? g[i + 1] : 0.0f);
                shape_write<SH_SPHERE>(xp, m, sc, true, f);
            } else {
                float ssum = g[0];
                for (int i = 1; i < m; ++i) ssum += g[i];
                shape_dtlz7(xp, m, 1.0f + ssum, f);
            }
            float* o = F + (b * R + r) * m;
            for (int i = 0; i < m; ++i) o[i] = f[i];
        }
        __syncthreads();                                                    // the stage may be refilled from now on
    }
}

}  // namespace evx

using namespace evx;

static int eval_check(const char* who, const float* X, int64_t n, int32_t d, int32_t m, const float* F) {
    EVX_REQUIRE(n >= 0 && d >= 1 && m >= 2 && m <= EVX_MAX_OBJECTIVES && d >= m - 1, EVX_ERR_ARG,
                "%s: bad shape n=%lld d=%d m=%d (need 2 <= m <= %d, d >= m-1)", who, (long long)n, d, m,
                EVX_MAX_OBJECTIVES);
    if (n > 0) EVX_REQUIRE(X && F, EVX_ERR_ARG, "%s: null pointer", who);
    return EVX_OK;
}

extern "C" int evx_eval_dtlz_f32(int32_t which, const float* X, int64_t n, int32_t d, int32_t m, float* F,
                                 void* stream) {
    EVX_REQUIRE(which >= 1 && which <= 7, EVX_ERR_ARG, "evx_eval_dtlz_f32: which=%d not in 1..7", which);
    int rc = eval_check("evx_eval_dtlz_f32", X, n, d, m, F);
    if (rc || n == 0) return rc;
    k_eval_dtlz<<<(unsigned)ceil_div(n, 128), 128, 0, (cudaStream_t)stream>>>(which, X, n, d, m, F);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}

extern "C" int evx_eval_maf1_f32(const float* X, int64_t n, int32_t d, int32_t m, float* F, void* stream) {
    int rc = eval_check("evx_eval_maf1_f32", X, n, d, m, F);
    if (rc || n == 0) return rc;
    k_eval_dtlz<<<(unsigned)ceil_div(n, 128), 128, 0, (cudaStream_t)stream>>>(8, X, n, d, m, F);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}

// lsmop.py:46-54: c_i by the logistic map in double (Python floats), then float32 arithmetic for
// floor(c / sum(c) * d / nk); offsets are the running sum of sublen*nk.
extern "C" int evx_lsmop_groups(int32_t d, int32_t m, int32_t* sublen, int32_t* off) {
    EVX_REQUIRE(d >= 1 && m >= 1 && m <= EVX_MAX_OBJECTIVES && sublen && off, EVX_ERR_ARG, "evx_lsmop_groups: bad argument");
    double c64[MAXM];
    c64[0] = 3.8 * 0.1 * (1 - 0.1);
    for (int i = 1; i < m; ++i) c64[i] = 3.8 * c64[i - 1] * (1 - c64[i - 1]);
    float c[MAXM], sum = 0.0f;
    for (int i = 0; i < m; ++i) { c[i] = (float)c64[i]; }
    // jnp.sum over a short float32 vector: sequential accumulation
    for (int i = 0; i < m; ++i) sum += c[i];
    float acc = 0.0f;
    off[0] = 0;
    for (int i = 0; i < m; ++i) {
        const float v = floorf(c[i] / sum * (float)d / 5.0f);
        sublen[i] = (int)v;
        acc += v * 5.0f;
        off[i + 1] = (int)acc;
    }
    return EVX_OK;
}

extern "C" int evx_eval_lsmop_f32(int32_t which, const float* X, int64_t n, int32_t d, int32_t m,
                                  const int32_t* sublen, const int32_t* off, float* F, void* stream) {
    EVX_REQUIRE(which >= 1 && which <= 9, EVX_ERR_ARG, "evx_eval_lsmop_f32: which=%d not in 1..9", which);
    int rc = eval_check("evx_eval_lsmop_f32", X, n, d, m, F);
    if (rc || n == 0) return rc;
    EVX_REQUIRE(sublen && off, EVX_ERR_ARG, "evx_eval_lsmop_f32: null group arrays");
    static const int table[10][3] = {{0, 0, 0},
                                     {IN_SPHERE, IN_SPHERE, 1},
                                     {IN_GRIEWANK, IN_SCHWEFEL, 2},
                                     {IN_RASTRIGIN, IN_ROSENBROCK, 2},
                                     {IN_ACKLEY, IN_GRIEWANK, 2},
                                     {IN_SPHERE, IN_SPHERE, 1},
                                     {IN_ROSENBROCK, IN_SCHWEFEL, 2},
                                     {IN_ACKLEY, IN_ROSENBROCK, 2},
                                     {IN_GRIEWANK, IN_SPHERE, 2},
                                     {IN_SPHERE, IN_ACKLEY, 2}};
    LsmopGroups G;
    for (int i = 0; i < m; ++i) { G.sublen[i] = sublen[i]; G.off[i] = off[i]; }
    G.off[m] = off[m];
    cudaStream_t s = (cudaStream_t)stream;
    // Full blocks of Rp rows go through the pipelined kernel (bulk copies need 16-byte pieces: Rp is a
    // multiple of 4 and X must be 16-byte aligned); what is left - fewer than Rp rows, or everything when
    // a row is too long for two stages - takes the simple kernel.
    int64_t done = 0;
    {
        const size_t row_bytes = sizeof(float) * (size_t)d;
        int Rp = (int)((size_t)LS_STAGE_BYTES / row_bytes) & ~3;               // small stages: many CTAs per SM hide the short serial phases
        if (Rp > 32) Rp = 32;
        if (Rp >= 4 && (reinterpret_cast<size_t>(X) & 15) == 0 && n >= Rp && m * 5 * LS_SPLIT <= LS_SLOTS) {
            const int64_t nblocks = n / Rp;
            int slots = 1;
            while (slots < m * 5 * LS_SPLIT) slots <<= 1;
            const size_t smem = 128 + sizeof(int4) * LS_SLOTS + 2 * row_bytes * Rp + sizeof(float) * (size_t)((d + 3) & ~3) +
                                sizeof(Partial) * (size_t)Rp * slots + sizeof(float) * (size_t)Rp * m;
            // device properties and occupancy are looked up once per (device, shared-memory size), not per call
            static thread_local int c_dev = -1, c_sms = 0, c_per_sm = 0;
            static thread_local size_t c_smem = 0;
            int dev = 0;
            EVX_CUDA(cudaGetDevice(&dev));
            if (dev != c_dev || smem != c_smem) {
                EVX_CUDA(cudaDeviceGetAttribute(&c_sms, cudaDevAttrMultiProcessorCount, dev));
                EVX_CUDA(cudaFuncSetAttribute(k_eval_lsmop_pipe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                EVX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c_per_sm, k_eval_lsmop_pipe, 256, smem));
                c_dev = dev;
                c_smem = smem;
            }
            const int sms = c_sms, per_sm = c_per_sm;
            if (per_sm >= 1) {
                int64_t grid = (int64_t)sms * per_sm;
                if (grid > nblocks) grid = nblocks;
                k_eval_lsmop_pipe<<<(unsigned)grid, 256, smem, s>>>(which, X, nblocks, d, m, Rp, slots, G, table[which][0],
                                                                    table[which][1], table[which][2], F);
                EVX_LAUNCH_CHECK();
                done = nblocks * Rp;
            }
        }
    }
    if (done == n) return EVX_OK;
    X += done * d;
    F += done * m;
    n -= done;
    const size_t per_row = sizeof(float) * ((size_t)d + (size_t)m * 5 + m);
    const size_t budget = 200 * 1024;
    EVX_REQUIRE(per_row <= budget, EVX_ERR_ARG, "evx_eval_lsmop_f32: d=%d does not fit one row in shared memory", d);
    int R = (int)(budget / per_row);
    if (R > 32) R = 32;
    // keep at least ~2 CTAs per SM's worth of blocks when n is small
    while (R > 1 && ceil_div(n, R) < 296) R /= 2;
    const size_t smem = per_row * R;
    EVX_CUDA(cudaFuncSetAttribute(k_eval_lsmop, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
    k_eval_lsmop<<<(unsigned)ceil_div(n, R), 256, smem, s>>>(which, X, n, d, m, R, G, table[which][0],
                                                             table[which][1], table[which][2], F);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}
