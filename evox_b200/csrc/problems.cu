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
    const size_t per_row = sizeof(float) * ((size_t)d + (size_t)m * 5 + m);
    const size_t budget = 200 * 1024;
    EVX_REQUIRE(per_row <= budget, EVX_ERR_ARG, "evx_eval_lsmop_f32: d=%d does not fit one row in shared memory", d);
    int R = (int)(budget / per_row);
    if (R > 32) R = 32;
    // keep at least ~2 CTAs per SM's worth of blocks when n is small
    while (R > 1 && ceil_div(n, R) < 296) R /= 2;
    const size_t smem = per_row * R;
    EVX_CUDA(cudaFuncSetAttribute(k_eval_lsmop, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
    k_eval_lsmop<<<(unsigned)ceil_div(n, R), 256, smem, (cudaStream_t)stream>>>(which, X, n, d, m, R, G, table[which][0],
                                                                             table[which][1], table[which][2], F);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}
