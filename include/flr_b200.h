/*
 * flr_b200.h -- C ABI of the B200 (sm_100a) scoring path for
 * saakshipatel/fairness-llm-replication.
 *
 * The reference is pure Python and has no FFI of its own; each
 * entry point below names the reference function (path:line under the reference tree)
 * whose inner arithmetic it replaces.  The Python package in
 * fairness-llm-replication_b200/ keeps the reference's function signatures and calls
 * these through ctypes; INTEGRATION.md shows the stub a reference maintainer would add.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - functions only enqueue work; they never synchronise and never allocate
 *     (scratch is caller-owned, sized by the *_scratch_bytes query);
 *   - return value: 0 = ok, negative = error (FLR_E_*); flr_last_error() returns a
 *     thread-local message for the last failing call.
 *   - the library is sm_100a only; there is no CPU fallback behind any entry point.
 */
#ifndef FLR_B200_H
#define FLR_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FLR_OK            0
#define FLR_E_INVALID    -1   /* bad argument (null pointer, size out of range, ...) */
#define FLR_E_UNSUPPORTED -2  /* shape outside what the kernels were built for */
#define FLR_E_CUDA       -3   /* a CUDA runtime call failed; see flr_last_error() */
#define FLR_E_SCRATCH    -4   /* scratch buffer too small */

#define FLR_EMPTY_ANY_IS_ZERO 0   /* phase3_final.jaccard, src/phase3_final.py:112-117 */
#define FLR_EMPTY_BOTH_IS_ONE 1   /* utils.calculate_individual_fairness, src/utils.py:232-237 */

#define FLR_MAX_RANK_K     64     /* SERP / PRAG need list ranks to fit a 64-bit mask */
#define FLR_SELECT_BINS    2048   /* 11-bit radix digits */
#define FLR_MAX_ROUNDS     8      /* violation rounds packed into one flag byte */
#define FLR_DTYPE_F32      0
#define FLR_DTYPE_F64      1

int         flr_version(void);
const char *flr_last_error(void);
/* Number of SMs / compute capability of the current device. */
int         flr_device_info(int *sm_count, int *cc_major, int *cc_minor);

/* ---------------------------------------------------------------------------------------
 * List similarity: Jaccard / SERP* / PRAG* of each (user, variant) list against the user's
 * neutral list.
 * Replaces the per-pair body of utils.calculate_individual_fairness (src/utils.py:224-239),
 * measure_quality_preservation (src/phase3_facter_mitigation.py:341-365, k_cap = 10) and
 * phase3_final.jaccard (src/phase3_final.py:112-117, empty_policy = FLR_EMPTY_ANY_IS_ZERO).
 * SERP* / PRAG* have no reference code (definitions: oracle/np_oracle.py:list_metrics).
 *
 *   neutral      int32 [U, K]     neutral_len int32 [U]
 *   lists        int32 [U, V, K]  lens        int32 [U, V]
 *   ids must lie in [0, vocab); positions >= len are never read as ids.
 *   k_cap        rows are cut to min(len, k_cap); also the K of the SERP/PRAG normalisers.
 *   rank_metrics 0 = Jaccard only (any K), 1 = also SERP* and PRAG* (k_cap <= 64).
 * Outputs (any may be NULL): out_inter/out_union int32 [U,V] (exact), out_jac/out_serp/
 * out_prag float [U,V], out_variant_sums double [V,4] = per variant {sum jaccard, sum
 * serp_num, sum prag_num, user count} (summed in a fixed order: bit-reproducible).
 * out_status int32 [1]: number of ids found outside [0, vocab) (such ids are skipped): counted per
 * list row, a neutral row once per user.
 * ------------------------------------------------------------------------------------- */
size_t flr_list_metrics_scratch_bytes(int64_t U, int V, int K, int vocab);
int flr_list_metrics(const int32_t *neutral, const int32_t *neutral_len,
                     const int32_t *lists, const int32_t *lens,
                     int64_t U, int V, int K, int k_cap, int vocab,
                     int empty_policy, int rank_metrics,
                     int32_t *out_inter, int32_t *out_union,
                     float *out_jac, float *out_serp, float *out_prag,
                     double *out_variant_sums, int32_t *out_status,
                     void *scratch, size_t scratch_bytes, void *stream);

/* flr_list_metrics for callers that cannot hand over a scratch buffer.  (No counterpart in the
 * reference; it exists only because of how this library is embedded.)
 *
 * Some hosts run inside a memory plan they do not control: all of HBM already belongs to one
 * pool, or the user shard was sized to fill the device, or the step loop must never meet the
 * device-wide synchronisation a cudaMalloc brings.  They can pass inputs and outputs they
 * already own, but not one more buffer.  This variant therefore takes no scratch argument and
 * calls no allocating routine: the per-thread partial sums go to a fixed array compiled into
 * the library (FLR_LIST_METRICS_STATIC_SCRATCH_BYTES; every GPU that loads the library holds
 * that much until the process exits).
 *
 * Arguments and results are those of flr_list_metrics.  Limits and rules:
 *   - a shape whose partial sums do not fit the array is refused with FLR_E_UNSUPPORTED and a
 *     message naming both sizes.  Shapes served by the table kernel (2 <= V <= 128,
 *     vocab <= 16384, k_cap <= 64) always fit on a B200, whatever U is; compare-kernel shapes
 *     (V = 1 pair lists, larger vocabularies) are run with as many thread blocks as the array has
 *     room for (each thread then walks more pairs), so they fit for any U as long as V blocks do
 *     (V <= 512);
 *   - the first call on a device looks up the array's device address
 *     (cudaGetSymbolAddress, with that device current); make that first call before any stream
 *     capture starts;
 *   - all calls on a device share the one array and it keeps whatever the last call wrote.
 *     Nothing in it needs to be zero at entry (each launch overwrites the partials it later
 *     reads), but two calls whose kernels overlap on the GPU -- different streams, or
 *     different host threads -- race on it.  Keep calls on one device in one stream order. */
#define FLR_LIST_METRICS_STATIC_SCRATCH_BYTES 2097152
int flr_list_metrics_static(const int32_t *neutral, const int32_t *neutral_len,
                            const int32_t *lists, const int32_t *lens,
                            int64_t U, int V, int K, int k_cap, int vocab,
                            int empty_policy, int rank_metrics,
                            int32_t *out_inter, int32_t *out_union,
                            float *out_jac, float *out_serp, float *out_prag,
                            double *out_variant_sums, int32_t *out_status, void *stream);

/* All-pairs intersections between the M lists of each user (config 4 "pairwise
 * list-similarity"): 3706-bit item bitsets in shared memory, __popc over AND-ed words.
 *   lists int32 [U, M, K], lens int32 [U, M] -> out_inter int32 [U, M, M], out_sizes int32 [U, M]
 * No reference code (phase 2 of the reference compares no lists, SURVEY.md Fact 2). */
int flr_pair_jaccard(const int32_t *lists, const int32_t *lens, int64_t U, int M, int K,
                     int vocab, int32_t *out_inter, int32_t *out_sizes, void *stream);

/* ---------------------------------------------------------------------------------------
 * Cosine nonconformity  s = max(0, 1 - <a,b> / (|a||b|)),  zero-norm row -> 1.
 * Replaces the score column of generate_calibration_data
 * (src/phase3_facter_mitigation.py:125-139, a simulated |conf - actual|) with the
 * embedding form north_star prescribes.
 *   a float [U, D] (anchor / neutral row), b float [U, V, D]; V = 1 is the calibration form.
 *   D % 4 == 0, D <= 1024, rows 16-byte aligned.  fp32 products accumulated in fp64.
 *   out_score float [U, V].
 *   hist0 (nullable) uint64 [FLR_SELECT_BINS]: pass-0 radix histogram of the scores, added
 *   to (caller zeroes it) -- saves the select's first read of the scores.
 * ------------------------------------------------------------------------------------- */
int flr_cosine_nonconformity(const float *a, const float *b, int64_t U, int V, int D,
                             float *out_score, uint64_t *hist0, void *stream);

/* ---------------------------------------------------------------------------------------
 * Exact order statistic (radix select over the float bit patterns), per group.
 * Replaces np.sort + index in calculate_conformal_thresholds
 * (src/phase3_facter_mitigation.py:165-178).  Passes are split so that a multi-GPU caller
 * can all-reduce `hist` between flr_select_hist and flr_select_scan.
 *
 *   state: uint64 [G, FLR_SELECT_STATE_WORDS], zeroed by the caller before pass 0.
 *   hist : uint64 [G, FLR_SELECT_BINS],        zeroed by the caller before every pass.
 *   n_passes = flr_select_num_passes(dtype) (3 for f32, 6 for f64).
 *   flr_select_hist(pass p): histogram of digit p over the scores whose higher digits
 *     match the prefix chosen so far (pass 0: all scores).  group: uint8 [N] or NULL (G=1).
 *   flr_select_scan(pass p): picks the digit holding rank k and narrows the prefix.  On
 *     pass 0 it derives, per group, n = sum(hist) and
 *       k = clamp(int(ceil((n + 1) * one_minus_alpha)) - 1, 0, n - 1)
 *     (the reference's rule, src/phase3_facter_mitigation.py:174-175) unless k_override
 *     (uint64 [G], nullable) is given.  After the last pass out_thr[g] holds the selected
 *     element (0.5 for an empty group, :169-171), typed like the scores, and the state holds
 *     n, k, #less, #equal (layout below).
 * NaNs order last (as in np.sort); -0.0 is treated as +0.0.
 * ------------------------------------------------------------------------------------- */
#define FLR_SELECT_STATE_WORDS 8
#define FLR_ST_PREFIX 0   /* orderable key bits fixed so far            */
#define FLR_ST_KREM   1   /* rank still to resolve inside the prefix    */
#define FLR_ST_N      2   /* group size                                 */
#define FLR_ST_K      3   /* selected 0-based rank                      */
#define FLR_ST_LESS   4   /* number of scores strictly below the result */
#define FLR_ST_EQUAL  5   /* number of scores equal to the result       */
#define FLR_ST_EMPTY  6   /* 1 when the group is empty                  */
int flr_select_num_passes(int dtype);
int flr_select_hist(const void *scores, int dtype, const uint8_t *group, int64_t N, int G,
                    int pass, const uint64_t *state, uint64_t *hist, void *stream);
int flr_select_scan(int dtype, int G, int pass, double one_minus_alpha,
                    const uint64_t *k_override, const uint64_t *hist, uint64_t *state,
                    void *out_thr, void *stream);

/* ---------------------------------------------------------------------------------------
 * Counterfactual violation detection with round-synchronous threshold updates
 * (no reference code; definition: oracle/np_oracle.py:violation_rounds).
 *
 * flr_threshold_schedule: thr[0] = q0[0]; thr[r+1] = max(fl32(gamma*thr[r]), q_min) when
 *   the select state says some score lies above q0 (n - less - equal > 0), else thr[r].
 *   Valid closed form of the round recurrence for 0 < gamma <= 1 and q0 >= 0.  When some score lies
 *   above q0 and q0 < 0, the recurrence depends on per-round counts this call does not have: every
 *   out_thr entry is written as NaN (a NaN threshold flags nothing), so the unsupported case cannot
 *   pass for a result.
 *   q0: float [1] (device), state: the select state of group 0, out_thr: float [rounds+1].
 * flr_violation_detect: flags[i] bit r = scores[i] > thr[r]; counts[r] += #flags (uint64,
 *   added to; caller zeroes), variant_counts[v*rounds + r] likewise for v = i % V (nullable).
 * ------------------------------------------------------------------------------------- */
int flr_threshold_schedule(const float *q0, const uint64_t *state, float gamma, float q_min,
                           int rounds, float *out_thr, void *stream);
int flr_violation_detect(const float *scores, int64_t N, int V, const float *thr, int rounds,
                         uint8_t *flags, uint64_t *counts, uint64_t *variant_counts,
                         void *stream);

/* Conformal filtering as a prefix length (src/phase3_facter_mitigation.py:205-216):
 * keep = #{rank < len : 1.0 - rank*0.05 >= thr[group]} (float64, as the reference), raised
 * to 3 when fewer than 3 survive and len >= 3.  group uint8 [U] or NULL (thr[0]). */
int flr_conformal_keep_len(const int32_t *lens, const uint8_t *group, const double *thr,
                           int64_t U, int32_t *out_keep, void *stream);

/* ---------------------------------------------------------------------------------------
 * Demographic parity.
 * flr_group_item_histogram replaces the counting loop of utils.calculate_demographic_parity
 * (src/utils.py:180-187): hist[g, x] += occurrences of item x in the rows of group g, every
 * occurrence counted.  lists int32 [R, K], lens int32 [R]; group uint8 [R], or NULL for the dense
 * [U, V, K] layout where row r belongs to group r % n_groups.  hist uint64 [n_groups, vocab] is
 * added to (caller zeroes it; multi-GPU callers all-reduce it afterwards).
 * flr_js_distance_matrix replaces calculate_js_divergence (src/utils.py:89-102) for every group
 * pair i < j in row-major order: out_dist double [G(G-1)/2], out_valid uint8 (0 where a group is
 * empty and the reference skips the pair, src/utils.py:200).
 * ------------------------------------------------------------------------------------- */
int flr_group_item_histogram(const int32_t *lists, const int32_t *lens, const uint8_t *group,
                             int64_t R, int K, int k_cap, int n_groups, int vocab,
                             uint64_t *hist, void *stream);
int flr_js_distance_matrix(const uint64_t *hist, int n_groups, int vocab, double *out_dist,
                           uint8_t *out_valid, void *stream);
/* Equal opportunity from the same histogram -- replaces the set
 * intersection and the length count of utils.calculate_equal_opportunity (src/utils.py:254-266):
 * out_relevant[g] = number of distinct items x with relevant[x] != 0 that group g was recommended,
 * out_total[g] = number of recommendations made to group g.  relevant uint8 [vocab]. */
int flr_equal_opportunity_counts(const uint64_t *hist, const uint8_t *relevant, int n_groups, int vocab,
                                 uint64_t *out_relevant, uint64_t *out_total, void *stream);

/* ---------------------------------------------------------------------------------------
 * Ranking-fairness metrics of phase 2, batched over B rankings.
 * flr_ndcg replaces the loops of utils.calculate_ndcg (src/utils.py:118-132): rel double [B, n] =
 * relevance of the ranked items in rank order, ideal double [B, m] = the ground-truth relevances
 * sorted descending, log2tab[i] = log2(i + 2) (from the caller, so it is the reference's own
 * transcendental); out[b] = DCG@k / IDCG@k, 0 when IDCG is 0.  Sums run left to right as in the
 * reference: the result is bit-identical.
 * flr_exposure_by_group replaces the per-group exposure means of calculate_exposure_ratio
 * (src/phase2_cross_model_eval.py:228-243, src/utils.py:286-296): group uint8 [B, n] = protected
 * group of the item at each rank (255 = item unknown, skipped), weight[pos] = 1 / log2(pos + 2);
 * out_mean double [B, G] = numpy-order mean of the group's exposures (0 when it has none), out_cnt
 * int32 [B, G].  n <= 128.
 * ------------------------------------------------------------------------------------- */
int flr_ndcg(const double *rel, const int32_t *rel_len, int n, const double *ideal,
             const int32_t *ideal_len, int m, int k, const double *log2tab, int64_t B,
             double *out, void *stream);
int flr_exposure_by_group(const uint8_t *group, const int32_t *len, int n, int n_groups,
                          const double *weight, int64_t B, double *out_mean, int32_t *out_cnt,
                          void *stream);

/* ---------------------------------------------------------------------------------------
 * Dense per-user similarity matrices S[u] = X[u] X[u]^T (config 4; no reference code).
 *   x bf16 [U, M, D] (rows L2-normalised by flr_normalize_rows_bf16), out float [U, M, M].
 *   tcgen05.mma (kind::f16, fp32 accumulate in TMEM), operands staged by TMA.
 *   M <= 128, D % 64 == 0, D <= 512.
 * ------------------------------------------------------------------------------------- */
int flr_normalize_rows_bf16(const float *x, int64_t rows, int D, uint16_t *out_bf16, void *stream);
size_t flr_pair_similarity_scratch_bytes(int64_t U, int M, int D);
int flr_pair_similarity_gemm(const uint16_t *x_bf16, int64_t U, int M, int D, float *out,
                             void *scratch, size_t scratch_bytes, void *stream);

/* ---------------------------------------------------------------------------------------
 * Synthetic ML-1M-schema inputs; Philox4x32-10, counter =
 * (user, variant, draw).  The *_host and device forms produce identical bytes, so the CPU
 * oracle sees exactly what the kernels see.  `user0` offsets the user index (shards).
 * zipf_cdf: float [vocab] cumulative popularity (flr_synth_zipf_cdf_host fills it; copy it
 * to the device for the device form).
 * ------------------------------------------------------------------------------------- */
int flr_synth_zipf_cdf_host(float *cdf_host, int vocab);
int flr_synth_lists_host(uint64_t seed, int64_t user0, int64_t U, int V, int K, int vocab,
                         const float *zipf_cdf_host, int32_t *neutral_host,
                         int32_t *neutral_len_host, int32_t *lists_host, int32_t *lens_host);
int flr_synth_lists(uint64_t seed, int64_t user0, int64_t U, int V, int K, int vocab,
                    const float *zipf_cdf, int32_t *neutral, int32_t *neutral_len,
                    int32_t *lists, int32_t *lens, void *stream);
int flr_synth_embeddings_host(uint64_t seed, int64_t user0, int64_t U, int V, int D,
                              float *emb_neutral_host, float *emb_host);
int flr_synth_embeddings(uint64_t seed, int64_t user0, int64_t U, int V, int D,
                         float *emb_neutral, float *emb, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* FLR_B200_H */
