// relation.cu - general dominance relation between two point sets (SPEA2's entry point) and the
// device version of the reference's dense-matrix peel.
//
// Replaces utils/common.py:88-97 `dominate_relation(x, y)` (callers: algorithms/mo/spea2.py:24,
// operators/selection/non_dominate.py:62) and operators/selection/non_dominate.py:13-29
// `host_rank_from_domination_matrix`.
#include "common.cuh"

namespace evx {

constexpr int REL_TX = 128;      // threads per CTA
constexpr int REL_ROWS = 32;     // x rows per CTA
constexpr int REL_CPT = 16;      // y columns per thread (16 bytes of output per row)

// out[i, j] byte matrix.  Thread <-> 16 consecutive columns whose M objectives sit in registers
// (M <= 4; larger M re-reads them through L1), CTA <-> 32 rows staged in shared memory and broadcast.
// Each warp writes 512 contiguous bytes of one output row per step: the kernel is bound by the
// nx*ny-byte store stream the reference API demands, then by issue (2*M sign-bit compares per pair).
template <int M>
__global__ void __launch_bounds__(REL_TX)
k_relation_u8(const float* __restrict__ x, int64_t nx, const float* __restrict__ y, int64_t ny, int m_rt,
              uint8_t* __restrict__ out) {
    const int m = (M > 0) ? M : m_rt;
    extern __shared__ float xs[];                                    // [REL_ROWS][m]
    const int64_t row0 = (int64_t)blockIdx.y * REL_ROWS;
    const int rows = (int)min((int64_t)REL_ROWS, nx - row0);
    for (int t = threadIdx.x; t < rows * m; t += blockDim.x) xs[t] = x[row0 * m + t];
    __syncthreads();
    const int64_t j0 = ((int64_t)blockIdx.x * REL_TX + threadIdx.x) * REL_CPT;
    if (j0 >= ny) return;
    const int ncol = (int)min((int64_t)REL_CPT, ny - j0);
    const bool vec_ok = (ncol == REL_CPT) && ((ny & 15) == 0) && ((reinterpret_cast<size_t>(out) & 15) == 0);
    if constexpr (M > 0) {
        // Compares on the FMA pipe: for values without NaN and with -0 folded onto +0, b - a has its sign bit
        // set iff b < a (inf - inf is the canonical positive NaN: "not below"), so
        //   all_k(a_k <= b_k)  <=>  sign(OR_k (b_k - a_k)) == 0,     any_k(a_k < b_k)  <=>  sign(OR_k (a_k - b_k)) == 1.
        // Six FADD + three LOP3 per pair instead of six FSETP on the half-rate pipe.  A row or column that
        // holds a NaN dominates nothing and is dominated by nothing (IEEE compares): masked out at the end.
        float yv[REL_CPT][M];
        uint32_t cm[4] = {0, 0, 0, 0};                                 // 0x01 per column that exists and holds no NaN
#pragma unroll
        for (int c = 0; c < REL_CPT; ++c) {
            bool ok = c < ncol;
#pragma unroll
            for (int k = 0; k < M; ++k) {
                const float v = (c < ncol) ? __ldg(y + (j0 + c) * M + k) : 0.0f;
                ok = ok && (v == v);
                yv[c][k] = v + 0.0f;
            }
            cm[c >> 2] |= (uint32_t)ok << (8 * (c & 3));
        }
        for (int r = 0; r < rows; ++r) {
            float a[M];
            bool rok = true;
#pragma unroll
            for (int k = 0; k < M; ++k) { a[k] = xs[r * M + k] + 0.0f; rok = rok && (a[k] == a[k]); }
            uint32_t pk[4] = {0, 0, 0, 0};
#pragma unroll
            for (int c = 0; c < REL_CPT; ++c) {
                uint32_t below = 0u, above = 0u;
#pragma unroll
                for (int k = 0; k < M; ++k) {
                    below |= __float_as_uint(yv[c][k] - a[k]);         // sign: y < x in objective k
                    above |= __float_as_uint(a[k] - yv[c][k]);         // sign: x < y in objective k
                }
                pk[c >> 2] |= ((~below & above) >> 31) << (8 * (c & 3));
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) pk[q] = rok ? (pk[q] & cm[q]) : 0u;
            uint8_t* o = out + (row0 + r) * ny + j0;
            if (vec_ok) {
                *reinterpret_cast<uint4*>(o) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            } else {
                for (int c = 0; c < ncol; ++c) o[c] = (uint8_t)((pk[c >> 2] >> (8 * (c & 3))) & 0xFF);
            }
        }
    } else {
        for (int r = 0; r < rows; ++r) {
            uint32_t pk[4] = {0, 0, 0, 0};
            for (int c = 0; c < ncol; ++c) {
                const float* yj = y + (j0 + c) * m;
                bool le = true, lt = false;
                for (int k = 0; k < m; ++k) {
                    const float a = xs[r * m + k], b = __ldg(yj + k);
                    le = le && (a <= b);
                    lt = lt || (a < b);
                }
                pk[c >> 2] |= (uint32_t)(le && lt) << (8 * (c & 3));
            }
            uint8_t* o = out + (row0 + r) * ny + j0;
            if (vec_ok) {
                *reinterpret_cast<uint4*>(o) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            } else {
                for (int c = 0; c < ncol; ++c) o[c] = (uint8_t)((pk[c >> 2] >> (8 * (c & 3))) & 0xFF);
            }
        }
    }
}

// Bit-packed variant: lane <-> column, one ballot per (row, 32 columns).
__global__ void __launch_bounds__(256)
k_relation_bits(const float* __restrict__ x, int64_t nx, const float* __restrict__ y, int64_t ny, int m,
                uint32_t* __restrict__ out, int64_t words) {
    extern __shared__ float xs[];                                    // [REL_ROWS][m]
    const int64_t row0 = (int64_t)blockIdx.y * REL_ROWS;
    const int rows = (int)min((int64_t)REL_ROWS, nx - row0);
    for (int t = threadIdx.x; t < rows * m; t += blockDim.x) xs[t] = x[row0 * m + t];
    __syncthreads();
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t w = j >> 5;
    if (w >= words) return;                                          // whole warps leave together
    const bool valid = j < ny;
    for (int r = 0; r < rows; ++r) {
        bool le = valid, lt = false;
        for (int k = 0; k < m && valid; ++k) {
            const float a = xs[r * m + k], b = __ldg(y + j * m + k);
            le = le && (a <= b);
            lt = lt || (a < b);
        }
        const uint32_t word = __ballot_sync(0xFFFFFFFFu, le && lt);
        if ((threadIdx.x & 31) == 0) out[(row0 + r) * words + w] = word;
    }
}

// ---- dense-matrix peel (non_dominate.py:13-29) ---------------------------------------------------
__global__ void k_colsum_u8(const uint8_t* __restrict__ dom, int64_t n, int* __restrict__ count) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int64_t r0 = (int64_t)blockIdx.y * 256, r1 = min(n, r0 + 256);
    int s = 0;
    for (int64_t i = r0; i < r1; ++i) s += dom[i * n + j] != 0;
    if (s) atomicAdd(&count[j], s);
}

__global__ void k_dense_detect(int* __restrict__ count, int* __restrict__ rank, uint8_t* __restrict__ front,
                               int64_t n, int k, int* any) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const bool z = count[j] == 0;
    front[j] = z;
    if (z) { rank[j] = k; count[j] = -1; *any = 1; }
}

__global__ void k_dense_apply(const uint8_t* __restrict__ dom, const uint8_t* __restrict__ front, int64_t n,
                              int* __restrict__ count) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int64_t r0 = (int64_t)blockIdx.y * 256, r1 = min(n, r0 + 256);
    int s = 0;
    for (int64_t i = r0; i < r1; ++i)
        if (front[i]) s += dom[i * n + j] != 0;
    if (s) atomicSub(&count[j], s);
}

}  // namespace evx

using namespace evx;

extern "C" int evx_dominate_relation_u8(const float* x, int64_t nx, const float* y, int64_t ny, int32_t m,
                                        uint8_t* out, void* stream) {
    EVX_REQUIRE(nx >= 0 && ny >= 0 && m >= 1 && m <= EVX_MAX_OBJECTIVES, EVX_ERR_ARG,
                "evx_dominate_relation_u8: bad shape nx=%lld ny=%lld m=%d", (long long)nx, (long long)ny, m);
    if (nx == 0 || ny == 0) return EVX_OK;
    EVX_REQUIRE(x && y && out, EVX_ERR_ARG, "evx_dominate_relation_u8: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    dim3 grid((unsigned)ceil_div(ny, (int64_t)REL_TX * REL_CPT), (unsigned)ceil_div(nx, REL_ROWS));
    EVX_REQUIRE(grid.y <= 65535u, EVX_ERR_ARG, "evx_dominate_relation_u8: nx=%lld too large (max %d)", (long long)nx,
                65535 * REL_ROWS);
    const size_t smem = sizeof(float) * REL_ROWS * m;
    switch (m) {
        case 2: k_relation_u8<2><<<grid, REL_TX, smem, s>>>(x, nx, y, ny, m, out); break;
        case 3: k_relation_u8<3><<<grid, REL_TX, smem, s>>>(x, nx, y, ny, m, out); break;
        case 4: k_relation_u8<4><<<grid, REL_TX, smem, s>>>(x, nx, y, ny, m, out); break;
        default: k_relation_u8<0><<<grid, REL_TX, smem, s>>>(x, nx, y, ny, m, out); break;
    }
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}

extern "C" int evx_dominate_relation_bits(const float* x, int64_t nx, const float* y, int64_t ny, int32_t m,
                                          uint32_t* out, void* stream) {
    EVX_REQUIRE(nx >= 0 && ny >= 0 && m >= 1 && m <= EVX_MAX_OBJECTIVES, EVX_ERR_ARG,
                "evx_dominate_relation_bits: bad shape");
    if (nx == 0 || ny == 0) return EVX_OK;
    EVX_REQUIRE(x && y && out, EVX_ERR_ARG, "evx_dominate_relation_bits: null pointer");
    const int64_t words = ceil_div(ny, 32);
    dim3 grid((unsigned)ceil_div(words * 32, 256), (unsigned)ceil_div(nx, REL_ROWS));
    EVX_REQUIRE(grid.y <= 65535u, EVX_ERR_ARG, "evx_dominate_relation_bits: nx too large");
    k_relation_bits<<<grid, 256, sizeof(float) * REL_ROWS * m, (cudaStream_t)stream>>>(x, nx, y, ny, m, out, words);
    EVX_LAUNCH_CHECK();
    return EVX_OK;
}

extern "C" size_t evx_rank_from_matrix_workspace_bytes(int64_t n) {
    if (n < 0) return 0;
    return align_up(sizeof(int) * (size_t)n, 256) + align_up((size_t)n, 256) + 256 + 256;
}

extern "C" int evx_rank_from_matrix_u8(const uint8_t* dom, const int32_t* count_in, int64_t n, int32_t* rank,
                                       void* ws, size_t ws_bytes, void* stream) {
    EVX_REQUIRE(n >= 0, EVX_ERR_ARG, "evx_rank_from_matrix_u8: bad n");
    if (n == 0) return EVX_OK;
    EVX_REQUIRE(dom && rank, EVX_ERR_ARG, "evx_rank_from_matrix_u8: null pointer");
    EVX_REQUIRE(ws && ws_bytes >= evx_rank_from_matrix_workspace_bytes(n), EVX_ERR_WORKSPACE,
                "evx_rank_from_matrix_u8: workspace too small");
    EVX_REQUIRE(ceil_div(n, 256) <= 65535, EVX_ERR_ARG, "evx_rank_from_matrix_u8: n too large");
    cudaStream_t s = (cudaStream_t)stream;
    char* base = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(ws), 256));
    int* count = reinterpret_cast<int*>(base);
    uint8_t* front = reinterpret_cast<uint8_t*>(base + align_up(sizeof(int) * (size_t)n, 256));
    int* any = reinterpret_cast<int*>(base + align_up(sizeof(int) * (size_t)n, 256) + align_up((size_t)n, 256));
    dim3 g1((unsigned)ceil_div(n, 256)), g2((unsigned)ceil_div(n, 256), (unsigned)ceil_div(n, 256));
    if (count_in) {
        EVX_CUDA(cudaMemcpyAsync(count, count_in, sizeof(int) * (size_t)n, cudaMemcpyDeviceToDevice, s));
    } else {
        EVX_CUDA(cudaMemsetAsync(count, 0, sizeof(int) * (size_t)n, s));
        k_colsum_u8<<<g2, 256, 0, s>>>(dom, n, count);
        EVX_LAUNCH_CHECK();
    }
    EVX_CUDA(cudaMemsetAsync(rank, 0, sizeof(int) * (size_t)n, s));
    for (int k = 0;; ++k) {
        int h = 0;
        EVX_CUDA(cudaMemsetAsync(any, 0, sizeof(int), s));
        k_dense_detect<<<g1, 256, 0, s>>>(count, rank, front, n, k, any);
        EVX_LAUNCH_CHECK();
        EVX_CUDA(cudaMemcpyAsync(&h, any, sizeof(int), cudaMemcpyDeviceToHost, s));
        EVX_CUDA(cudaStreamSynchronize(s));
        if (!h) break;
        k_dense_apply<<<g2, 256, 0, s>>>(dom, front, n, count);
        EVX_LAUNCH_CHECK();
    }
    return EVX_OK;
}
