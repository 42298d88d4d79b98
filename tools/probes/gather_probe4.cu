// gather_probe4.cu - random row gather through SHARED MEMORY rings, as the peel kernel does it:
//   mode 0: cp.async.bulk (TMA 1-D copy), one row per lane, mbarrier per stage
//   mode 1: cp.async.cg 16 B per lane (LDGSTS), ROW/16 lanes per row, commit/wait groups
// 8 warps per CTA, CTAS CTAs per SM; each warp owns NST stages of SR rows of ROW bytes.  Footprint 48 GB.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

constexpr int UP = 9;
struct Planes { uint32_t ones = 0, twos = 0, fours = 0, up[UP] = {0,0,0,0,0,0,0,0,0}; };
__device__ __forceinline__ void csa(uint32_t& hi, uint32_t& lo, uint32_t a, uint32_t b, uint32_t c) {
    const uint32_t u = a ^ b; hi = (a & b) | (u & c); lo = u ^ c;
}
__device__ __forceinline__ void csa8(Planes& pl, const uint32_t* x) {
    uint32_t ta, tb, fa, fb, e;
    csa(ta, pl.ones, pl.ones, x[0], x[1]); csa(tb, pl.ones, pl.ones, x[2], x[3]); csa(fa, pl.twos, pl.twos, ta, tb);
    csa(ta, pl.ones, pl.ones, x[4], x[5]); csa(tb, pl.ones, pl.ones, x[6], x[7]); csa(fb, pl.twos, pl.twos, ta, tb);
    csa(e, pl.fours, pl.fours, fa, fb);
#pragma unroll
    for (int p = 0; p < UP; ++p) { if (!__any_sync(0xFFFFFFFFu, e)) break; const uint32_t c = pl.up[p] & e; pl.up[p] ^= e; e = c; }
}

template <int MODE, int ROW, int SR, int NST, int CSA = 0>
__global__ void __launch_bounds__(256, 2) k_ring(const unsigned char* __restrict__ buf, const uint32_t* __restrict__ rnd,
                                              int per_warp, uint64_t rows, uint32_t* out) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char* ring = smem + (size_t)warp * NST * SR * ROW;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)8 * NST * SR * ROW) + warp * NST;
    if (MODE == 0 && lane == 0) {
        for (int s = 0; s < NST; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bars + s)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const int gw = blockIdx.x * 8 + warp;
    const uint32_t* my = rnd + (size_t)gw * per_warp;
    uint32_t acc = 0, phase = 0;
    Planes pa, pb;
    int issued = 0, consumed = 0;
    const int nstage = per_warp / SR;
    auto consume = [&]() {
        const int st = consumed % NST;
        if (MODE == 0) {
            asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(smem_u32(bars + st)), "r"((phase >> st) & 1u) : "memory");
            phase ^= 1u << st;
        } else {
            asm volatile("cp.async.wait_group %0;" ::"n"(NST - 1) : "memory");
            __syncwarp();
        }
        const uint32_t* base = reinterpret_cast<const uint32_t*>(ring + (size_t)st * SR * ROW);
        if (CSA) {
            for (int r0 = 0; r0 < SR; r0 += 8) {
                uint32_t xa[8], xb[8];
#pragma unroll
                for (int r = 0; r < 8; ++r) { xa[r] = base[(r0 + r) * (ROW / 4) + lane]; xb[r] = base[(r0 + r) * (ROW / 4) + 32 + lane]; }
                csa8(pa, xa); csa8(pb, xb);
            }
        } else {
#pragma unroll
        for (int r = 0; r < SR; ++r)
#pragma unroll
            for (int w = 0; w < ROW / 128; ++w) acc ^= base[r * (ROW / 4) + w * 32 + lane];
        }
        ++consumed;
    };
    for (int s = 0; s < nstage; ++s) {
        if (issued - consumed == NST) consume();
        const int st = issued % NST;
        unsigned char* dst = ring + (size_t)st * SR * ROW;
        if (MODE == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bars + st)), "r"(SR * ROW) : "memory");
            __syncwarp();
            if (lane < SR) {
                const uint64_t row = ((uint64_t)__ldg(my + s * SR + lane) * rows) >> 32;
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst + lane * ROW)), "l"(buf + row * ROW), "n"(ROW), "r"(smem_u32(bars + st)) : "memory");
            }
        } else {
            constexpr int LPR = ROW / 16;                 // lanes per row
            constexpr int RPI = 32 / LPR;                 // rows per instruction
#pragma unroll
            for (int i = 0; i < SR / RPI; ++i) {
                const int r = i * RPI + lane / LPR;
                const uint64_t row = ((uint64_t)__ldg(my + s * SR + r) * rows) >> 32;
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst + r * ROW + (lane % LPR) * 16)), "l"(buf + row * ROW + (lane % LPR) * 16) : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
        ++issued;
    }
    while (consumed < issued) {
        if (MODE == 1) { asm volatile("cp.async.wait_group 0;" ::: "memory"); __syncwarp(); }
        consume();
    }
    if (CSA) { acc ^= pa.ones ^ pa.twos ^ pa.fours ^ pb.ones ^ pb.twos ^ pb.fours; for (int p = 0; p < UP; ++p) acc ^= pa.up[p] ^ pb.up[p]; }
    if (acc == 0x12345u) *out = acc;
}

template <int MODE, int ROW, int SR, int NST, int CSA = 0>
int run(const unsigned char* buf, const uint32_t* rnd, size_t total, uint32_t* out, int ctas) {
    const size_t smem = (size_t)8 * NST * SR * ROW + 8 * NST * 8;
    CK(cudaFuncSetAttribute(k_ring<MODE, ROW, SR, NST, CSA>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_ring<MODE, ROW, SR, NST, CSA>, 256, smem));
    if (occ < ctas) { printf("mode %d row %d SR %d NST %d: only %d CTAs/SM fit\n", MODE, ROW, SR, NST, occ); return 0; }
    const uint64_t rows = total / ROW;
    const int per_warp = 4096 * 256 / ROW;              // same bytes per warp for every row size
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e9f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        k_ring<MODE, ROW, SR, NST, CSA><<<148 * ctas, 256, smem>>>(buf, rnd, per_warp, rows, out);
        cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    const double nrows = (double)148 * ctas * 8 * per_warp;
    printf("%s%s row %4d B, %d CTA/SM x 8 warps, %d stages x %2d rows (%3d KB in flight/SM): %7.0f GB/s %7.1f rows/us/SM\n",
           MODE ? "ldgsts" : "bulk  ", CSA ? "+csa" : "    ", ROW, ctas, NST, SR, (int)(ctas * 8 * NST * SR * ROW / 1024), nrows * ROW / (best * 1e-3) / 1e9,
           nrows / 148 / (best * 1e3));
    return 0;
}

int main() {
    const size_t total = (size_t)48 << 30;
    unsigned char* buf; uint32_t* rnd; uint32_t* out;
    CK(cudaMalloc(&buf, total)); CK(cudaMemset(buf, 1, total));
    const size_t nidx = (size_t)148 * 4 * 8 * 8192;
    uint32_t* h = (uint32_t*)malloc(nidx * sizeof(uint32_t));
    uint64_t s = 88172645463325252ull;
    for (size_t i = 0; i < nidx; ++i) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; h[i] = (uint32_t)(s >> 16); }
    CK(cudaMalloc(&rnd, nidx * sizeof(uint32_t))); CK(cudaMemcpy(rnd, h, nidx * sizeof(uint32_t), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&out, 256));
    run<0, 256, 16, 2>(buf, rnd, total, out, 2);      // the peel kernel's ring as committed
    run<0, 256, 16, 2, 1>(buf, rnd, total, out, 2);   // ... with its carry-save consumption
    run<0, 256, 16, 3, 1>(buf, rnd, total, out, 2);
    run<0, 256, 8, 4, 1>(buf, rnd, total, out, 2);
    run<0, 256, 16, 2, 1>(buf, rnd, total, out, 1);
    run<0, 256, 16, 4, 1>(buf, rnd, total, out, 1);
    run<0, 256, 16, 6, 1>(buf, rnd, total, out, 1);
    run<1, 256, 16, 2, 1>(buf, rnd, total, out, 2);
    run<1, 256, 16, 3, 1>(buf, rnd, total, out, 2);
    return 0;
}
