// gather_probe2.cu - random gather ceiling as a function of the ROW SIZE (64 / 128 / 256 / 512 bytes).
// 32 warps per SM, 8 rows in flight per warp (the saturating point of gather_probe.cu).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

// LANES lanes fetch one row of LANES * sizeof(T) bytes; a warp fetches 32 / LANES rows per instruction.
template <typename T, int LANES, int U>
__global__ void k_gather(const T* __restrict__ buf, const int* __restrict__ idx, int per_warp, size_t row_elems, T* out) {
    const int lane = threadIdx.x & 31;
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int sub = lane / LANES, l = lane % LANES;
    constexpr int RPI = 32 / LANES;                  // rows per instruction
    const int* my = idx + (size_t)gw * per_warp;
    unsigned acc = 0;
    for (int t = 0; t < per_warp; t += U * RPI) {
        T v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = __ldg(buf + (size_t)__ldg(my + t + u * RPI + sub) * row_elems + l);
#pragma unroll
        for (int u = 0; u < U; ++u) acc ^= *reinterpret_cast<unsigned*>(&v[u]);
    }
    if (acc == 0x12345u) *reinterpret_cast<unsigned*>(out) = acc;
}

template <typename T, int LANES>
int run(const char* name, void* buf, int* idx, size_t total_bytes, void* out) {
    const int row_bytes = LANES * sizeof(T);
    const size_t rows = total_bytes / row_bytes;
    (void)rows;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int wps = 32, per_warp = 4096, threads = 256, blocks = 148 * wps * 32 / threads;
    float best = 1e9f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        k_gather<T, LANES, 8><<<blocks, threads>>>((const T*)buf, idx, per_warp, (size_t)LANES, (T*)out);
        cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    const double nrows = (double)148 * wps * per_warp;
    printf("%-28s row %4d B : %7.0f GB/s  %8.1f rows/us/SM\n", name, row_bytes, nrows * row_bytes / (best * 1e-3) / 1e9,
           nrows / 148 / (best * 1e3));
    return 0;
}

int main() {
    const size_t total = (size_t)8 << 30;
    void* buf; int* idx; void* out;
    CK(cudaMalloc(&buf, total)); CK(cudaMemset(buf, 1, total));
    const size_t nidx = (size_t)148 * 32 * 4096;
    int* h = (int*)malloc(nidx * sizeof(int));
    uint64_t s = 88172645463325252ull;
    for (size_t i = 0; i < nidx; ++i) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; h[i] = (int)(s % (total / 512)); }   // valid for every row size
    CK(cudaMalloc(&idx, nidx * sizeof(int))); CK(cudaMemcpy(idx, h, nidx * sizeof(int), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&out, 256));
    run<uint32_t, 16>("16 lanes x 4 B", buf, idx, total, out);
    run<uint32_t, 32>("32 lanes x 4 B", buf, idx, total, out);
    run<uint2, 32>("32 lanes x 8 B", buf, idx, total, out);
    run<uint4, 16>("16 lanes x 16 B", buf, idx, total, out);
    run<uint4, 32>("32 lanes x 16 B", buf, idx, total, out);
    run<uint4, 8>("8 lanes x 16 B", buf, idx, total, out);
    return 0;
}
