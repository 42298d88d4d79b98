// variation.cu - NSGA-II's variation step as one kernel: simulated binary crossover + polynomial
// mutation + clip.
//
// Replaces operators/crossover/sbx.py:8-40 (`simulated_binary`, type 1), operators/mutation/
// pm_mutation.py:7-54 (`polynomial`) and the clip of algorithms/mo/nsga2.py:85, which the
// reference runs as ~40 XLA ops over [pop, d].  The formulas are the reference's; the random
// numbers are not: the reference draws from jax.random (Threefry, one array per draw), here each
// (pair, gene) pulls its eight uniforms from a counter-based Philox4x32-10 keyed by
// (seed, generation), so results are reproducible for a given seed but only *statistically*
// equivalent to the reference's (tests check the distributional and algebraic properties).
#include "common.cuh"

namespace evx {

struct Philox {
    uint32_t c[4];
};

__device__ __forceinline__ Philox philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                uint32_t k1) {
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        const uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += W0; k1 += W1;
    }
    return Philox{{c0, c1, c2, c3}};
}

__device__ __forceinline__ float u01(uint32_t b) { return (float)(b >> 8) * (1.0f / 16777216.0f); }   // [0, 1)

// pm_mutation.py:20-51 for one gene already clamped to [lo, up]
__device__ __forceinline__ float poly_mutate(float x, float lo, float up, bool site, float mu, float dis_m) {
    x = fmaxf(fminf(x, up), lo);
    if (!site) return x;
    const float span = up - lo, e = 1.0f / (dis_m + 1.0f);
    if (mu <= 0.5f) {
        const float norm = (x - lo) / span;
        return x + span * (powf(2.0f * mu + (1.0f - 2.0f * mu) * powf(1.0f - norm, dis_m + 1.0f), e) - 1.0f);
    }
    const float norm = (up - x) / span;
    return x + span * (1.0f - powf(2.0f * (1.0f - mu) + 2.0f * (mu - 0.5f) * powf(1.0f - norm, dis_m + 1.0f), e));
}

// One thread per (pair p, gene j): parents p and p + n/2 -> offspring rows p and p + n/2.
__global__ void __launch_bounds__(256)
k_sbx_pm(const float* __restrict__ pop, int64_t n, int d, const float* __restrict__ lb, const float* __restrict__ ub,
         float pro_c, float dis_c, float pro_m, float dis_m, uint32_t seed, uint32_t gen, float* __restrict__ out) {
    const int64_t half = n / 2;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= half * d) {
        const int64_t r = t - half * d;                       // odd n: last row passes through (sbx.py:36-37)
        if ((n & 1) && r < d) out[(n - 1) * d + r] = fmaxf(fminf(pop[(n - 1) * d + r], ub[r]), lb[r]);
        return;
    }
    const int64_t p = t / d;
    const int j = (int)(t - p * d);
    const Philox a = philox4x32_10((uint32_t)t, (uint32_t)(t >> 32), 0u, gen, seed, 0x5BD1E995u);
    const Philox b = philox4x32_10((uint32_t)t, (uint32_t)(t >> 32), 1u, gen, seed, 0x5BD1E995u);
    const Philox pr = philox4x32_10((uint32_t)p, (uint32_t)(p >> 32), 2u, gen, seed, 0x5BD1E995u);   // per pair

    const float p1 = pop[p * d + j], p2 = pop[(p + half) * d + j];
    // sbx.py:15-25
    const float mu = u01(a.c[0]);
    float beta = (mu <= 0.5f) ? powf(2.0f * mu, 1.0f / (dis_c + 1.0f)) : powf(2.0f - 2.0f * mu, -1.0f / (dis_c + 1.0f));
    if (a.c[1] & 1u) beta = -beta;
    if (u01(a.c[2]) < 0.5f) beta = 1.0f;
    if (u01(pr.c[0]) > pro_c) beta = 1.0f;
    const float mid = (p1 + p2) / 2.0f, dev = beta * (p1 - p2) / 2.0f;
    float c1 = mid + dev, c2 = mid - dev;
    // pm_mutation.py:16-51, then nsga2.py:85
    const float lo = lb[j], up = ub[j], rate = pro_m / (float)d;
    c1 = poly_mutate(c1, lo, up, u01(a.c[3]) < rate, u01(b.c[0]), dis_m);
    c2 = poly_mutate(c2, lo, up, u01(b.c[1]) < rate, u01(b.c[2]), dis_m);
    out[p * d + j] = fmaxf(fminf(c1, up), lo);
    out[(p + half) * d + j] = fmaxf(fminf(c2, up), lo);
}

}  // namespace evx

using namespace evx;

extern "C" int evx_variation_sbx_pm_f32(const float* pop, int64_t n, int32_t d, const float* lb, const float* ub,
                                        float pro_c, float dis_c, float pro_m, float dis_m, uint32_t seed,
                                        uint32_t generation, float* out, void* stream) {
    EVX_REQUIRE(n >= 0 && d >= 1, EVX_ERR_ARG, "evx_variation_sbx_pm_f32: bad shape n=%lld d=%d", (long long)n, d);
    if (n == 0) return EVX_OK;
    EVX_REQUIRE(pop && lb && ub && out, EVX_ERR_ARG, "evx_variation_sbx_pm_f32: null pointer");
    const int64_t threads = (n / 2) * d + ((n & 1) ? d : 0);
    if (threads == 0) return EVX_OK;
    k_sbx_pm<<<(unsigned)ceil_div(threads, 256), 256, 0, (cudaStream_t)stream>>>(pop, n, d, lb, ub, pro_c, dis_c, pro_m,
                                                                                dis_m, seed, generation, out);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}
