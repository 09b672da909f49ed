// Host and device drivers of the synthetic generator (synth.cuh).  Generation is set-up
// work, never inside a timed region; the device form exists because the 10M-user sweep
// (476 GB of embeddings over 8 GPUs) cannot be staged through the host.
#include <vector>

#include "common.cuh"
#include "synth.cuh"

using namespace flr;
using namespace flr::synth;

namespace {

__global__ void synth_lists_kernel(uint64_t seed, int64_t user0, int64_t U, int V, int K, int vocab,
                                   const float *__restrict__ cdf, int32_t *__restrict__ neutral,
                                   int32_t *__restrict__ neutral_len, int32_t *__restrict__ lists,
                                   int32_t *__restrict__ lens) {
    int64_t idx = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    int64_t total = U * (V + 1);
    if (idx >= total) return;
    int64_t u = idx / (V + 1);
    int v = static_cast<int>(idx % (V + 1)) - 1;
    int32_t clean[kMaxK], row[kMaxK];
    int len = gen_row(seed, user0 + u, v, V, K, vocab, cdf, clean, row);
    int32_t *dst = (v < 0) ? neutral + u * K : lists + (u * V + v) * K;
    for (int j = 0; j < K; ++j) dst[j] = row[j];
    if (v < 0) neutral_len[u] = len; else lens[u * V + v] = len;
}

// one thread per (row, element pair); row 0 of a user is the neutral (anchor) embedding
__global__ void synth_emb_kernel(uint64_t seed, int64_t user0, int64_t U, int V, int D,
                                 float *__restrict__ emb_neutral, float *__restrict__ emb) {
    int half = D >> 1;
    int64_t idx = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    int64_t total = U * (V + 1) * half;
    if (idx >= total) return;
    int d = static_cast<int>(idx % half) * 2;
    int64_t rowi = idx / half;
    int64_t u = rowi / (V + 1);
    int v = static_cast<int>(rowi % (V + 1)) - 1;
    float b0, b1;
    gauss_pair(seed, user0 + u, 0x20000u, d, b0, b1);
    if (v < 0) {
        *reinterpret_cast<float2 *>(emb_neutral + u * D + d) = make_float2(b0, b1);
    } else {
        float n0, n1, sg = sigma_of(v, V);
        gauss_pair(seed, user0 + u, 0x20001u + static_cast<uint32_t>(v), d, n0, n1);
        *reinterpret_cast<float2 *>(emb + (u * V + v) * D + d) =
            make_float2(mul_add_no_fma(sg, n0, b0), mul_add_no_fma(sg, n1, b1));
    }
}

int check_list_args(int64_t U, int V, int K, int vocab) {
    FLR_REQUIRE(U >= 0 && V >= 1 && K >= 1 && K <= kMaxK, FLR_E_INVALID,
                "synth lists: need U>=0, V>=1, 1<=K<=%d (got U=%lld V=%d K=%d)", kMaxK, (long long)U, V, K);
    FLR_REQUIRE(vocab >= 2 * K + 2, FLR_E_INVALID, "synth lists: vocab must be >= 2K+2");
    return FLR_OK;
}

}  // namespace

extern "C" int flr_synth_zipf_cdf_host(float *cdf_host, int vocab) {
    FLR_REQUIRE(cdf_host && vocab >= 1, FLR_E_INVALID, "zipf cdf: null pointer or vocab < 1");
    double h = 0.0;
    for (int i = 0; i < vocab; ++i) h += 1.0 / (i + 1);
    double c = 0.0;
    for (int i = 0; i < vocab; ++i) {
        c += 1.0 / (i + 1);
        cdf_host[i] = static_cast<float>(c / h);
    }
    cdf_host[vocab - 1] = 1.0f;   // u < 1 always lands inside the table
    return FLR_OK;
}

extern "C" int flr_synth_lists_host(uint64_t seed, int64_t user0, int64_t U, int V, int K, int vocab,
                                    const float *cdf, int32_t *neutral, int32_t *neutral_len,
                                    int32_t *lists, int32_t *lens) {
    int rc = check_list_args(U, V, K, vocab);
    if (rc) return rc;
    FLR_REQUIRE(cdf && neutral && neutral_len && lists && lens, FLR_E_INVALID, "synth lists: null pointer");
#pragma omp parallel for schedule(static)
    for (int64_t u = 0; u < U; ++u) {
        int32_t clean[kMaxK], row[kMaxK];
        for (int v = -1; v < V; ++v) {
            int len = gen_row(seed, user0 + u, v, V, K, vocab, cdf, clean, row);
            int32_t *dst = (v < 0) ? neutral + u * K : lists + (u * V + v) * K;
            for (int j = 0; j < K; ++j) dst[j] = row[j];
            if (v < 0) neutral_len[u] = len; else lens[u * V + v] = len;
        }
    }
    return FLR_OK;
}

extern "C" int flr_synth_lists(uint64_t seed, int64_t user0, int64_t U, int V, int K, int vocab,
                               const float *cdf, int32_t *neutral, int32_t *neutral_len,
                               int32_t *lists, int32_t *lens, void *stream) {
    int rc = check_list_args(U, V, K, vocab);
    if (rc) return rc;
    FLR_REQUIRE(cdf && neutral && neutral_len && lists && lens, FLR_E_INVALID, "synth lists: null pointer");
    if (U == 0) return FLR_OK;
    int64_t total = U * (V + 1);
    int threads = 128;
    int64_t blocks = (total + threads - 1) / threads;
    FLR_REQUIRE(blocks <= 0x7FFFFFFF, FLR_E_UNSUPPORTED, "synth lists: too many rows for one launch");
    synth_lists_kernel<<<static_cast<unsigned>(blocks), threads, 0, static_cast<cudaStream_t>(stream)>>>(
        seed, user0, U, V, K, vocab, cdf, neutral, neutral_len, lists, lens);
    FLR_LAUNCH_CHECK("synth_lists_kernel");
    return FLR_OK;
}

extern "C" int flr_synth_embeddings_host(uint64_t seed, int64_t user0, int64_t U, int V, int D,
                                         float *emb_neutral, float *emb) {
    FLR_REQUIRE(U >= 0 && V >= 1 && D >= 2 && (D % 2) == 0, FLR_E_INVALID, "synth embeddings: bad shape");
    FLR_REQUIRE(emb_neutral && emb, FLR_E_INVALID, "synth embeddings: null pointer");
#pragma omp parallel for schedule(static)
    for (int64_t u = 0; u < U; ++u) {
        for (int d = 0; d < D; d += 2) {
            float b0, b1;
            gauss_pair(seed, user0 + u, 0x20000u, d, b0, b1);
            emb_neutral[u * D + d] = b0;
            emb_neutral[u * D + d + 1] = b1;
            for (int v = 0; v < V; ++v) {
                float n0, n1, sg = sigma_of(v, V);
                gauss_pair(seed, user0 + u, 0x20001u + static_cast<uint32_t>(v), d, n0, n1);
                emb[(u * V + v) * D + d] = mul_add_no_fma(sg, n0, b0);
                emb[(u * V + v) * D + d + 1] = mul_add_no_fma(sg, n1, b1);
            }
        }
    }
    return FLR_OK;
}

extern "C" int flr_synth_embeddings(uint64_t seed, int64_t user0, int64_t U, int V, int D,
                                    float *emb_neutral, float *emb, void *stream) {
    FLR_REQUIRE(U >= 0 && V >= 1 && D >= 2 && (D % 2) == 0, FLR_E_INVALID, "synth embeddings: bad shape");
    FLR_REQUIRE(emb_neutral && emb, FLR_E_INVALID, "synth embeddings: null pointer");
    if (U == 0) return FLR_OK;
    int64_t total = U * (V + 1) * (D / 2);
    int threads = 256;
    int64_t blocks = (total + threads - 1) / threads;
    FLR_REQUIRE(blocks <= 0x7FFFFFFF, FLR_E_UNSUPPORTED,
                "synth embeddings: too many elements for one launch; generate in user chunks");
    synth_emb_kernel<<<static_cast<unsigned>(blocks), threads, 0, static_cast<cudaStream_t>(stream)>>>(
        seed, user0, U, V, D, emb_neutral, emb);
    FLR_LAUNCH_CHECK("synth_emb_kernel");
    return FLR_OK;
}
