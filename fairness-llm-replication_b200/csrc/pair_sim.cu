// Dense per-user similarity matrices S[u] = X[u] X[u]^T (config 4; no reference code --
// definition: oracle/np_oracle.py:pair_similarity / normalize_rows_bf16).
#include <algorithm>
#include <cuda_bf16.h>

#include "common.cuh"

using namespace flr;

namespace {

// one warp per row: fp64 sum of squares (exact fp32 products), fp32 divide, RNE to bf16
__global__ void __launch_bounds__(256) normalize_rows_bf16_kernel(const float *__restrict__ x, int64_t rows, int D,
                                                                  __nv_bfloat16 *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);
    const int64_t nwarps = static_cast<int64_t>(gridDim.x) * 8;
    for (int64_t r = warp0; r < rows; r += nwarps) {
        const float *row = x + r * D;
        double ss = 0.0;
        for (int d = lane; d < D; d += 32) {
            const double v = row[d];
            ss = fma(v, v, ss);
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
        const double n = sqrt(ss);
        for (int d = lane; d < D; d += 32) {
            const float y = n > 0.0 ? static_cast<float>(static_cast<double>(row[d]) / n) : 0.0f;
            out[r * D + d] = __float2bfloat16_rn(y);
        }
    }
}

}  // namespace

extern "C" int flr_normalize_rows_bf16(const float *x, int64_t rows, int D, uint16_t *out_bf16, void *stream_) {
    FLR_REQUIRE(rows >= 0 && D >= 1, FLR_E_INVALID, "normalize_rows: bad shape");
    if (rows == 0) return FLR_OK;
    FLR_REQUIRE(x && out_bf16, FLR_E_INVALID, "normalize_rows: null pointer");
    const int grid = static_cast<int>(std::min<int64_t>((rows + 7) / 8, static_cast<int64_t>(device_sm_count()) * 8));
    normalize_rows_bf16_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream_)>>>(
        x, rows, D, reinterpret_cast<__nv_bfloat16 *>(out_bf16));
    FLR_LAUNCH_CHECK("normalize_rows_bf16_kernel");
    return FLR_OK;
}

extern "C" size_t flr_pair_similarity_scratch_bytes(int64_t U, int M, int D) {
    (void)U; (void)M; (void)D;
    return 256;
}

extern "C" int flr_pair_similarity_gemm(const uint16_t *x_bf16, int64_t U, int M, int D, float *out, void *scratch,
                                        size_t scratch_bytes, void *stream_) {
    (void)x_bf16; (void)U; (void)M; (void)D; (void)out; (void)scratch; (void)scratch_bytes; (void)stream_;
    set_error("pair_similarity_gemm: tcgen05 kernel not built in this revision");
    return FLR_E_UNSUPPORTED;
}
