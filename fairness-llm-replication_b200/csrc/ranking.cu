// Ranking-fairness metrics of phase 2, batched over rankings:
//   ndcg_kernel               utils.calculate_ndcg            src/utils.py:113-132
//   exposure_by_group_kernel  calculate_exposure_ratio        src/phase2_cross_model_eval.py:216-254,
//                                                             src/utils.py:277-305
// The reference evaluates these on <= 20 items per call; one thread per ranking is plenty.  What
// matters here is bit-for-bit agreement, so the kernels add in exactly the order the reference does:
// NDCG accumulates `rel / log2(i + 2)` left to right (a Python loop / sum()), the exposure mean of a
// group is numpy's pairwise sum of the gathered exposures (eight strided partial sums for n >= 8)
// divided by their count.  log2(i + 2) and 1 / log2(i + 2) tables come from the caller (numpy), so the
// transcendental is the reference's own.
#include "common.cuh"

using namespace flr;

namespace {

__global__ void ndcg_kernel(const double *__restrict__ rel, const int32_t *__restrict__ rel_len, int n,
                            const double *__restrict__ ideal, const int32_t *__restrict__ ideal_len, int m, int k,
                            const double *__restrict__ log2tab, int64_t B, double *__restrict__ out) {
    const int64_t b = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const int nr = min(min(rel_len[b], n), k), ni = min(min(ideal_len[b], m), k);
    double dcg = 0.0, idcg = 0.0;
    for (int i = 0; i < nr; ++i) dcg = __dadd_rn(dcg, __ddiv_rn(rel[b * n + i], log2tab[i]));
    for (int i = 0; i < ni; ++i) idcg = __dadd_rn(idcg, __ddiv_rn(ideal[b * m + i], log2tab[i]));
    out[b] = idcg == 0.0 ? 0.0 : __ddiv_rn(dcg, idcg);
}

constexpr int EXP_MAX_N = 128;     // numpy sums blocks of up to 128 values with one 8-way unrolled loop

// numpy's pairwise_sum for n <= 128 (numpy/core/src/umath/loops_utils.h.src), in double
__device__ double numpy_sum(const double *a, int n) {
    if (n < 8) {
        double res = 0.0;
        for (int i = 0; i < n; ++i) res = __dadd_rn(res, a[i]);
        return res;
    }
    double r[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = a[j];
    int i = 8;
    for (; i < n - (n % 8); i += 8)
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = __dadd_rn(r[j], a[i + j]);
    double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                           __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
    for (; i < n; ++i) res = __dadd_rn(res, a[i]);
    return res;
}

__global__ void exposure_by_group_kernel(const uint8_t *__restrict__ group, const int32_t *__restrict__ len, int n,
                                         int G, const double *__restrict__ weight, int64_t B,
                                         double *__restrict__ out_mean, int32_t *__restrict__ out_cnt) {
    const int64_t b = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const int L = min(len[b], n);
    double vals[EXP_MAX_N];
    for (int g = 0; g < G; ++g) {
        int c = 0;
        for (int pos = 0; pos < L; ++pos)
            if (group[b * n + pos] == g) vals[c++] = weight[pos];
        out_cnt[b * G + g] = c;
        out_mean[b * G + g] = c ? __ddiv_rn(numpy_sum(vals, c), static_cast<double>(c)) : 0.0;
    }
}

}  // namespace

extern "C" int flr_ndcg(const double *rel, const int32_t *rel_len, int n, const double *ideal, const int32_t *ideal_len,
                        int m, int k, const double *log2tab, int64_t B, double *out, void *stream_) {
    FLR_REQUIRE(B >= 0 && n >= 1 && m >= 1 && k >= 0, FLR_E_INVALID, "ndcg: bad shape");
    if (B == 0) return FLR_OK;
    FLR_REQUIRE(rel && rel_len && ideal && ideal_len && log2tab && out, FLR_E_INVALID, "ndcg: null pointer");
    const int64_t blocks = (B + 127) / 128;
    FLR_REQUIRE(blocks <= 0x7FFFFFFF, FLR_E_UNSUPPORTED, "ndcg: too many rankings for one launch");
    ndcg_kernel<<<static_cast<unsigned>(blocks), 128, 0, static_cast<cudaStream_t>(stream_)>>>(
        rel, rel_len, n, ideal, ideal_len, m, k, log2tab, B, out);
    FLR_LAUNCH_CHECK("ndcg_kernel");
    return FLR_OK;
}

extern "C" int flr_exposure_by_group(const uint8_t *group, const int32_t *len, int n, int n_groups,
                                     const double *weight, int64_t B, double *out_mean, int32_t *out_cnt,
                                     void *stream_) {
    FLR_REQUIRE(B >= 0 && n >= 1 && n <= EXP_MAX_N && n_groups >= 1 && n_groups <= 255, FLR_E_UNSUPPORTED,
                "exposure_by_group: need 1 <= n <= %d ranked items and 1 <= groups <= 255", EXP_MAX_N);
    if (B == 0) return FLR_OK;
    FLR_REQUIRE(group && len && weight && out_mean && out_cnt, FLR_E_INVALID, "exposure_by_group: null pointer");
    const int64_t blocks = (B + 63) / 64;
    FLR_REQUIRE(blocks <= 0x7FFFFFFF, FLR_E_UNSUPPORTED, "exposure_by_group: too many rankings for one launch");
    exposure_by_group_kernel<<<static_cast<unsigned>(blocks), 64, 0, static_cast<cudaStream_t>(stream_)>>>(
        group, len, n, n_groups, weight, B, out_mean, out_cnt);
    FLR_LAUNCH_CHECK("exposure_by_group_kernel");
    return FLR_OK;
}
