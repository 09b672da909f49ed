// All-pairs list intersections between the M lists of each user (config 4, "pairwise
// list-similarity"; no reference code -- definition: oracle/np_oracle.py:pair_jaccard).
//
// One CTA per user: the M ranked lists become M item bit sets in shared memory (row stride padded to
// an odd word count so a warp's 32 rows hit 32 banks), every pair's intersection is a __popc sum over
// AND-ed words, and the M x M matrix is assembled in shared memory so the global store is one
// coalesced sweep.  Each thread owns a 4 x 4 block of pairs so every word loaded from shared memory
// feeds four ANDs.
#include <algorithm>

#include "common.cuh"

using namespace flr;

namespace {

constexpr int PJ_THREADS = 256;

// Per user the lists only touch a few hundred distinct items out of `vocab`, so the items are first
// renumbered into a dense per-user universe (claim a slot in a vocab-sized table with one CAS per
// element, number the winners with a shared counter); the bit sets are then n_local bits wide
// instead of vocab bits, which divides the popcount work by 116 / ceil(n_local / 32).
// StageT: element type of the M x M staging matrix -- an intersection is at most K, so uint8 when K <= 255
// (a quarter of the shared memory of int32: three CTAs per SM instead of one at M = 124, K = 20).
template <typename StageT>
__global__ void __launch_bounds__(PJ_THREADS) pair_jaccard_kernel(const int32_t *__restrict__ lists,
                                                                  const int32_t *__restrict__ lens, int64_t U, int M,
                                                                  int K, int vocab, int Wmax, bool ids_staged,
                                                                  int32_t *__restrict__ out_inter,
                                                                  int32_t *__restrict__ out_sizes) {
    extern __shared__ __align__(16) unsigned int s_mem[];
    constexpr unsigned EMPTY = 0xFFFFFFFFu, PENDING = 0xFFFFFFFEu;
    unsigned int *s_tab = s_mem;                                           // [vocab] item -> local index
    unsigned int *s_bits = s_tab + ((vocab + 3) & ~3);                      // [M][Wp], Wp odd
    const int Wp_max = Wmax | 1;
    StageT *s_inter = reinterpret_cast<StageT *>(s_bits + ((M * Wp_max + 3) & ~3));   // [M][M]
    // [M * K] the user's ids, read from global memory once per user; 0xFFFF = nothing here (past the row's
    // length, or outside the vocabulary).  Present only when vocab < 0xFFFF (ids_staged).
    unsigned short *s_ids = reinterpret_cast<unsigned short *>(
        reinterpret_cast<unsigned char *>(s_inter) + ((static_cast<size_t>(M) * M * sizeof(StageT) + 15) & ~static_cast<size_t>(15)));
    constexpr unsigned NONE = 0xFFFFu;
    __shared__ unsigned int s_count;
    const int t = threadIdx.x;
    constexpr int PB = 4;                                       // a thread owns a PB x PB block of pairs
    const int Mh = (M + PB - 1) / PB;                           // blocks per side
    const int n_blocks = Mh * (Mh + 1) / 2;                     // blocks (bm, bn) with bn >= bm

    for (int i = t; i < vocab; i += PJ_THREADS) s_tab[i] = EMPTY;
    for (int i = t; i < M * Wp_max; i += PJ_THREADS) s_bits[i] = 0u;
    if (t == 0) s_count = 0u;
    __syncthreads();

    for (int64_t u = blockIdx.x; u < U; u += gridDim.x) {
        const int32_t *L = lists + u * M * K;
        const int32_t *ln = lens + u * M;
        if (ids_staged) {
            // the three passes below walk the user's ids; from global memory each was a chain of dependent
            // loads (ncu: 45 % of the kernel waiting on them) -- one coalesced sweep into shared memory instead
            for (int i = t; i < M * K; i += PJ_THREADS) {
                const int m = i / K, j = i - m * K;
                const unsigned id = static_cast<unsigned>(L[i]);
                s_ids[i] = static_cast<unsigned short>((j < min(max(ln[m], 0), K) && id < static_cast<unsigned>(vocab)) ? id : NONE);
            }
            __syncthreads();
        }
        auto item = [&](int i) -> unsigned {          // id at flat position i of this user, NONE if there is none
            if (ids_staged) return s_ids[i];
            const int m = i / K, j = i - m * K;
            const unsigned id = static_cast<unsigned>(L[i]);
            return (j < min(max(ln[m], 0), K) && id < static_cast<unsigned>(vocab)) ? id : ~0u;
        };
        // pass 1: the first thread to touch an item claims it and takes the next local index
        for (int i = t; i < M * K; i += PJ_THREADS) {
            const unsigned id = item(i);
            if (id < static_cast<unsigned>(vocab) && s_tab[id] == EMPTY && atomicCAS(&s_tab[id], EMPTY, PENDING) == EMPTY)
                s_tab[id] = atomicAdd(&s_count, 1u);
        }
        __syncthreads();
        const int W = static_cast<int>((s_count + 31u) >> 5);             // words actually used (<= Wmax)
        const int Wp = W | 1;                                              // odd stride: a warp's rows hit 32 banks
        // pass 2: set bits in the dense universe
        for (int i = t; i < M * K; i += PJ_THREADS) {
            const unsigned id = item(i);
            if (id < static_cast<unsigned>(vocab)) {
                const unsigned l = s_tab[id];
                atomicOr(&s_bits[(i / K) * Wp + (l >> 5)], 1u << (l & 31u));
            }
        }
        __syncthreads();
        // blocks (bm, bn) with bn >= bm; consecutive threads take consecutive bn
        for (int blk = t; blk < n_blocks; blk += PJ_THREADS) {
            // blk-th block of the upper triangle, rows first: row bm starts at bm * Mh - bm (bm - 1) / 2
            int bm = static_cast<int>((2.0f * Mh + 1.0f - sqrtf((2.0f * Mh + 1.0f) * (2.0f * Mh + 1.0f) - 8.0f * blk)) * 0.5f);
            bm = min(max(bm, 0), Mh - 1);
            while (bm > 0 && bm * Mh - bm * (bm - 1) / 2 > blk) --bm;
            while ((bm + 1) * Mh - (bm + 1) * bm / 2 <= blk) ++bm;
            const int bn = bm + (blk - (bm * Mh - bm * (bm - 1) / 2));
            // a block pairs rows {bm + i Mh} with rows {bn + j Mh}, i, j < PB (clipped to M - 1: a clipped row
            // repeats a value that is written anyway); lanes read consecutive rows (stride Wp, odd).  Every
            // word loaded from shared memory feeds PB ANDs: the loop is bound by those loads.
            const unsigned int *ar[PB], *br[PB];
            int mi[PB], nj[PB], c[PB][PB];
#pragma unroll
            for (int i = 0; i < PB; ++i) {
                mi[i] = min(bm + i * Mh, M - 1);
                nj[i] = min(bn + i * Mh, M - 1);
                ar[i] = s_bits + mi[i] * Wp;
                br[i] = s_bits + nj[i] * Wp;
#pragma unroll
                for (int j = 0; j < PB; ++j) c[i][j] = 0;
            }
            for (int w = 0; w < W; ++w) {
                unsigned int x[PB], y[PB];
#pragma unroll
                for (int i = 0; i < PB; ++i) { x[i] = ar[i][w]; y[i] = br[i][w]; }
#pragma unroll
                for (int i = 0; i < PB; ++i)
#pragma unroll
                    for (int j = 0; j < PB; ++j) c[i][j] += __popc(x[i] & y[j]);
            }
#pragma unroll
            for (int i = 0; i < PB; ++i)
#pragma unroll
                for (int j = 0; j < PB; ++j) {
                    const StageT v = static_cast<StageT>(c[i][j]);
                    s_inter[mi[i] * M + nj[j]] = v;
                    s_inter[nj[j] * M + mi[i]] = v;
                }
        }
        __syncthreads();
        int32_t *dst = out_inter + u * M * M;
        for (int i = t; i < M * M; i += PJ_THREADS) dst[i] = static_cast<int32_t>(s_inter[i]);
        if (out_sizes)
            for (int m = t; m < M; m += PJ_THREADS) out_sizes[u * M + m] = static_cast<int32_t>(s_inter[m * M + m]);
        // restore: forget this user's items, wipe the words that were used
        for (int i = t; i < M * K; i += PJ_THREADS) {
            const unsigned id = item(i);
            if (id < static_cast<unsigned>(vocab)) s_tab[id] = EMPTY;
        }
        for (int i = t; i < M * Wp; i += PJ_THREADS) s_bits[i] = 0u;
        if (t == 0) s_count = 0u;
        __syncthreads();
    }
}

}  // namespace

extern "C" int flr_pair_jaccard(const int32_t *lists, const int32_t *lens, int64_t U, int M, int K, int vocab,
                                int32_t *out_inter, int32_t *out_sizes, void *stream_) {
    FLR_REQUIRE(U >= 0 && M >= 1 && K >= 1 && vocab >= 1, FLR_E_INVALID, "pair_jaccard: bad shape");
    if (U == 0) return FLR_OK;
    FLR_REQUIRE(lists && lens && out_inter, FLR_E_INVALID, "pair_jaccard: null pointer");
    // a user's M lists hold at most M*K distinct items: that bounds the dense universe
    const int Wmax = static_cast<int>((std::min<int64_t>(static_cast<int64_t>(M) * K, vocab) + 31) / 32);
    const size_t stage_elem = K <= 255 ? 1 : K <= 65535 ? 2 : 4;
    const bool ids_staged = vocab < 0xFFFF;
    const size_t smem = (static_cast<size_t>((vocab + 3) & ~3) + static_cast<size_t>((M * (Wmax | 1) + 3) & ~3)) * 4 +
                        ((static_cast<size_t>(M) * M * stage_elem + 15) & ~static_cast<size_t>(15)) +
                        (ids_staged ? ((static_cast<size_t>(M) * K * 2 + 15) & ~static_cast<size_t>(15)) : 0);
    FLR_REQUIRE(smem <= 226 * 1024, FLR_E_UNSUPPORTED,
                "pair_jaccard: M=%d lists of up to %d items over %d ids need %zu B of shared memory (max 231424)", M, K,
                vocab, smem);
    auto kern = stage_elem == 1 ? pair_jaccard_kernel<uint8_t>
              : stage_elem == 2 ? pair_jaccard_kernel<uint16_t> : pair_jaccard_kernel<int32_t>;
    FLR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    const int per_sm = std::max<int>(1, static_cast<int>((226 * 1024) / (smem + 1024)));
    const int grid = static_cast<int>(std::min<int64_t>(U, static_cast<int64_t>(device_sm_count()) * per_sm));
    kern<<<grid, PJ_THREADS, smem, static_cast<cudaStream_t>(stream_)>>>(lists, lens, U, M, K, vocab, Wmax, ids_staged,
                                                                         out_inter, out_sizes);
    FLR_LAUNCH_CHECK("pair_jaccard_kernel");
    return FLR_OK;
}
