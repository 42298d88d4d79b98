// gather_probe.cu - how fast can B200 gather random 128-byte lines from HBM, as a function of the
// number of lines in flight per warp and of the warps per SM?  (Sizing experiment for the peel kernel.)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_probe gather_probe.cu && ./gather_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

template <int U>
__global__ void k_gather(const uint32_t* __restrict__ buf, const int* __restrict__ idx, int per_warp, uint32_t* out) {
    const int lane = threadIdx.x & 31;
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int* my = idx + (size_t)gw * per_warp;
    uint32_t acc = 0;
    for (int t = 0; t < per_warp; t += U) {
        uint32_t v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = __ldg(my + t + u);
            asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v[u]) : "l"(buf + (size_t)i * 32 + lane));
        }
#pragma unroll
        for (int u = 0; u < U; ++u) acc ^= v[u];
    }
    if (acc == 0x12345u) out[0] = acc;
}

int main() {
    const size_t rows = (size_t)1 << 26;            // 64M rows x 128 B = 8 GB
    uint32_t* buf; int* idx; uint32_t* out;
    CK(cudaMalloc(&buf, rows * 128));
    CK(cudaMemset(buf, 1, rows * 128));
    const int max_warps = 148 * 64, per_warp = 4096;
    const size_t nidx = (size_t)max_warps * per_warp;
    int* h = (int*)malloc(nidx * sizeof(int));
    uint64_t s = 88172645463325252ull;
    for (size_t i = 0; i < nidx; ++i) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; h[i] = (int)(s % rows); }
    CK(cudaMalloc(&idx, nidx * sizeof(int)));
    CK(cudaMemcpy(idx, h, nidx * sizeof(int), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&out, 256));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    printf("warps/SM  inflight/warp   GB/s    rows/us/SM   implied latency us (Little)\n");
    for (int wps : {4, 8, 16, 32, 64}) {
        for (int U : {4, 8, 16, 32}) {
            const int threads = 256, blocks = 148 * wps * 32 / threads;
            float best = 1e9f;
            for (int rep = 0; rep < 3; ++rep) {
                cudaEventRecord(e0);
                if (U == 4) k_gather<4><<<blocks, threads>>>(buf, idx, per_warp, out);
                if (U == 8) k_gather<8><<<blocks, threads>>>(buf, idx, per_warp, out);
                if (U == 16) k_gather<16><<<blocks, threads>>>(buf, idx, per_warp, out);
                if (U == 32) k_gather<32><<<blocks, threads>>>(buf, idx, per_warp, out);
                cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
                float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
            }
            const double nrows = (double)148 * wps * per_warp, gbs = nrows * 128 / (best * 1e-3) / 1e9;
            const double rps = nrows / 148 / (best * 1e3);
            printf("%6d %12d %10.0f %10.1f %12.2f\n", wps, U, gbs, rps, (double)wps * U / rps);
        }
    }
    return 0;
}
