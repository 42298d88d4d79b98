// gather_probe3.cu - random gather ceiling as a function of the FOOTPRINT (GB touched at random) and
// the row size.  32 warps per SM, 8 rows in flight per warp.  usage: gather_probe3 [GB ...]
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

template <typename T, int LANES, int U>
__global__ void k_gather(const T* __restrict__ buf, const uint32_t* __restrict__ rnd, int per_warp, uint64_t rows, T* out) {
    const int lane = threadIdx.x & 31;
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int sub = lane / LANES, l = lane % LANES;
    constexpr int RPI = 32 / LANES;
    const uint32_t* my = rnd + (size_t)gw * per_warp;
    unsigned acc = 0;
    for (int t = 0; t < per_warp; t += U * RPI) {
        T v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint64_t row = ((uint64_t)__ldg(my + t + u * RPI + sub) * rows) >> 32;
            v[u] = __ldg(buf + row * LANES + l);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) acc ^= *reinterpret_cast<unsigned*>(&v[u]);
    }
    if (acc == 0x12345u) *reinterpret_cast<unsigned*>(out) = acc;
}

template <typename T, int LANES>
int run(void* buf, uint32_t* rnd, size_t total_bytes, void* out) {
    const int row_bytes = LANES * sizeof(T);
    const uint64_t rows = total_bytes / row_bytes;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int wps = 32, per_warp = 4096, threads = 256, blocks = 148 * wps * 32 / threads;
    float best = 1e9f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        k_gather<T, LANES, 8><<<blocks, threads>>>((const T*)buf, rnd, per_warp, rows, (T*)out);
        cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    const double nrows = (double)148 * wps * per_warp;
    printf("footprint %5.0f GB  row %4d B : %7.0f GB/s  %8.1f rows/us/SM\n", total_bytes / 1073741824.0, row_bytes,
           nrows * row_bytes / (best * 1e-3) / 1e9, nrows / 148 / (best * 1e3));
    return 0;
}

int main(int argc, char** argv) {
    const size_t cap = (size_t)64 << 30;
    void* buf; uint32_t* rnd; void* out;
    CK(cudaMalloc(&buf, cap)); CK(cudaMemset(buf, 1, cap));
    const size_t nidx = (size_t)148 * 32 * 4096;
    uint32_t* h = (uint32_t*)malloc(nidx * sizeof(uint32_t));
    uint64_t s = 88172645463325252ull;
    for (size_t i = 0; i < nidx; ++i) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; h[i] = (uint32_t)(s >> 16); }
    CK(cudaMalloc(&rnd, nidx * sizeof(uint32_t))); CK(cudaMemcpy(rnd, h, nidx * sizeof(uint32_t), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&out, 256));
    const int gbs_default[] = {1, 4, 16, 32, 64};
    for (int i = 0; i < 5; ++i) {
        const size_t total = (size_t)gbs_default[i] << 30;
        run<uint32_t, 32>(buf, rnd, total, out);
        run<uint2, 32>(buf, rnd, total, out);
        run<uint4, 32>(buf, rnd, total, out);
    }
    return 0;
}
