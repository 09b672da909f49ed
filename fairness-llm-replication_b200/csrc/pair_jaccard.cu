// All-pairs list intersections between the M lists of each user (config 4, "pairwise
// list-similarity"; no reference code -- definition: oracle/np_oracle.py:pair_jaccard).
//
// One CTA per user: the M ranked lists become M `vocab`-bit item sets in shared memory
// (row stride padded to an odd word count so a warp's 32 rows hit 32 banks), every pair's
// intersection is a __popc sum over AND-ed words, and the M x M matrix is assembled in
// shared memory so the global store is one coalesced sweep.  Each thread owns a 2 x 2 block
// of pairs so every word loaded from shared memory feeds two ANDs.
#include <algorithm>

#include "common.cuh"

using namespace flr;

namespace {

constexpr int PJ_THREADS = 256;

__global__ void __launch_bounds__(PJ_THREADS) pair_jaccard_kernel(const int32_t *__restrict__ lists,
                                                                  const int32_t *__restrict__ lens, int64_t U, int M,
                                                                  int K, int vocab, int W, int Wp,
                                                                  int32_t *__restrict__ out_inter,
                                                                  int32_t *__restrict__ out_sizes) {
    extern __shared__ __align__(16) unsigned int s_mem[];
    unsigned int *s_bits = s_mem;                               // [M][Wp]
    int *s_inter = reinterpret_cast<int *>(s_mem + ((M * Wp + 3) & ~3));   // [M][M]
    const int t = threadIdx.x;
    const int Mh = (M + 1) >> 1;                                // 2x2 blocks per side

    for (int64_t u = blockIdx.x; u < U; u += gridDim.x) {
        for (int i = t; i < M * Wp; i += PJ_THREADS) s_bits[i] = 0u;
        __syncthreads();
        const int32_t *L = lists + u * M * K;
        for (int i = t; i < M * K; i += PJ_THREADS) {
            const int m = i / K, j = i - m * K;
            if (j < min(max(lens[u * M + m], 0), K)) {
                const unsigned id = static_cast<unsigned>(L[i]);
                if (id < static_cast<unsigned>(vocab)) atomicOr(&s_bits[m * Wp + (id >> 5)], 1u << (id & 31u));
            }
        }
        __syncthreads();
        // blocks (bm, bn) with bn >= bm; consecutive threads take consecutive bn
        for (int blk = t; blk < Mh * Mh; blk += PJ_THREADS) {
            const int bm = blk / Mh, bn = blk - bm * Mh;
            if (bn < bm) continue;
            // a block pairs rows {bm, bm+Mh} with {bn, bn+Mh}: lanes read consecutive rows (stride Wp, odd)
            const int m0 = bm, m1 = min(bm + Mh, M - 1), n0 = bn, n1 = min(bn + Mh, M - 1);
            const unsigned int *a0 = s_bits + m0 * Wp, *a1 = s_bits + m1 * Wp;
            const unsigned int *b0 = s_bits + n0 * Wp, *b1 = s_bits + n1 * Wp;
            int c00 = 0, c01 = 0, c10 = 0, c11 = 0;
            for (int w = 0; w < W; ++w) {
                const unsigned int x0 = a0[w], x1 = a1[w], y0 = b0[w], y1 = b1[w];
                c00 += __popc(x0 & y0); c01 += __popc(x0 & y1);
                c10 += __popc(x1 & y0); c11 += __popc(x1 & y1);
            }
            s_inter[m0 * M + n0] = c00; s_inter[n0 * M + m0] = c00;
            s_inter[m0 * M + n1] = c01; s_inter[n1 * M + m0] = c01;
            s_inter[m1 * M + n0] = c10; s_inter[n0 * M + m1] = c10;
            s_inter[m1 * M + n1] = c11; s_inter[n1 * M + m1] = c11;
        }
        __syncthreads();
        int32_t *dst = out_inter + u * M * M;
        for (int i = t; i < M * M; i += PJ_THREADS) dst[i] = s_inter[i];
        if (out_sizes)
            for (int m = t; m < M; m += PJ_THREADS) out_sizes[u * M + m] = s_inter[m * M + m];
        __syncthreads();
    }
}

}  // namespace

extern "C" int flr_pair_jaccard(const int32_t *lists, const int32_t *lens, int64_t U, int M, int K, int vocab,
                                int32_t *out_inter, int32_t *out_sizes, void *stream_) {
    FLR_REQUIRE(U >= 0 && M >= 1 && K >= 1 && vocab >= 1, FLR_E_INVALID, "pair_jaccard: bad shape");
    if (U == 0) return FLR_OK;
    FLR_REQUIRE(lists && lens && out_inter, FLR_E_INVALID, "pair_jaccard: null pointer");
    const int W = (vocab + 31) / 32, Wp = W | 1;
    const size_t smem = (static_cast<size_t>((M * Wp + 3) & ~3) + static_cast<size_t>(M) * M) * 4;
    FLR_REQUIRE(smem <= 227 * 1024, FLR_E_UNSUPPORTED,
                "pair_jaccard: M=%d lists of %d-bit sets need %zu B of shared memory (max 232448)", M, vocab, smem);
    FLR_CUDA(cudaFuncSetAttribute(pair_jaccard_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  static_cast<int>(smem)));
    const int per_sm = std::max<int>(1, static_cast<int>((227 * 1024) / (smem + 1024)));
    const int grid = static_cast<int>(std::min<int64_t>(U, static_cast<int64_t>(device_sm_count()) * per_sm));
    pair_jaccard_kernel<<<grid, PJ_THREADS, smem, static_cast<cudaStream_t>(stream_)>>>(lists, lens, U, M, K, vocab, W,
                                                                                        Wp, out_inter, out_sizes);
    FLR_LAUNCH_CHECK("pair_jaccard_kernel");
    return FLR_OK;
}
