// Counterfactual violation detection with round-synchronous threshold updates, and the
// conformal filter as a prefix length.  Definitions: oracle/np_oracle.py:violation_rounds,
// threshold_schedule, conformal_keep_len (the latter follows
// src/phase3_facter_mitigation.py:205-216).
//
// For 0 < gamma <= 1 and q0 >= 0 the thresholds never rise, so whether round r+1 shrinks
// the threshold depends only on whether ANY score lies above q0 -- which the radix select
// already knows (n - less - equal).  All R rounds therefore collapse into one pass over the
// scores: R compares per score, one flag byte out, per-round counters.
#include <algorithm>

#include "common.cuh"

using namespace flr;

namespace {

__global__ void threshold_schedule_kernel(const float *__restrict__ q0, const unsigned long long *__restrict__ state,
                                          float gamma, float q_min, int rounds, float *__restrict__ out_thr) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const unsigned long long n = state[FLR_ST_N], less = state[FLR_ST_LESS], eq = state[FLR_ST_EQUAL];
    const bool any_above = state[FLR_ST_EMPTY] == 0ull && n > less + eq;
    float q = q0[0];
    // The closed form needs q0 >= 0: then gamma * q never exceeds q, and a floor above q0 is reached in one
    // step and kept whatever the later counts are.  For q0 < 0 (scores may be negative in general; the
    // cosine scores never are) gamma * q0 rises above q0 and the recurrence depends on counts this kernel
    // does not have: every threshold is written as NaN, which flags nothing and cannot pass for a result.
    if (any_above && q < 0.f) {
        for (int r = 0; r <= rounds; ++r) out_thr[r] = __int_as_float(0x7fc00000);
        return;
    }
    for (int r = 0; r < rounds; ++r) {
        out_thr[r] = q;
        if (any_above) q = fmaxf(__fmul_rn(gamma, q), q_min);
    }
    out_thr[rounds] = q;
}

constexpr int VIO_THREADS = 256;

// grid * VIO_THREADS is a multiple of V, so thread `tid` only ever sees variant tid % V.
// R = number of rounds, a template parameter: with the loops sized exactly the kernel is memory-bound
// (a fixed 8-round loop made it instruction-bound at 2 TB/s).
template <int R>
__global__ void __launch_bounds__(VIO_THREADS) violation_kernel(const float *__restrict__ scores, int64_t N, int V,
                                                                const float *__restrict__ thr, int rounds,
                                                                uint8_t *__restrict__ flags,
                                                                unsigned long long *__restrict__ counts,
                                                                unsigned long long *__restrict__ variant_counts,
                                                                bool vec_ok) {
    extern __shared__ unsigned int s_vc[];                 // [V][rounds] when variant_counts, else [rounds]
    const int n_slots = (variant_counts ? V : 1) * rounds;
    for (int i = threadIdx.x; i < n_slots; i += VIO_THREADS) s_vc[i] = 0u;
    __syncthreads();
    float q[R];
#pragma unroll
    for (int r = 0; r < R; ++r) q[r] = thr[r];
    const int64_t stride = static_cast<int64_t>(gridDim.x) * VIO_THREADS;
    const int64_t tid = static_cast<int64_t>(blockIdx.x) * VIO_THREADS + threadIdx.x;
    auto flag_of = [&](float s, unsigned int (&cnt)[R]) {
        unsigned int f = 0u;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const unsigned int hit = s > q[r] ? 1u : 0u;   // NaN never flags
            f |= hit << r;
            cnt[r] += hit;
        }
        return f;
    };
    auto publish = [&](int64_t first_elem, const unsigned int (&cnt)[R]) {
        const int v = variant_counts ? static_cast<int>(first_elem % V) : 0;
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (cnt[r]) atomicAdd(&s_vc[v * rounds + r], cnt[r]);
    };
    // Vector path: a thread owns the elements 4 * i + k (i = tid, tid + stride, ...); 4 * stride is a
    // multiple of V, so for each k the variant is fixed and the counts stay in registers.  One 16-byte
    // load and one 4-byte store per four scores, four of them in flight.
    const bool vec = vec_ok;
    const int64_t n4 = vec ? N / 4 : 0;
    if (vec) {
        unsigned int cnt4[4][R];
#pragma unroll
        for (int k = 0; k < 4; ++k)
#pragma unroll
            for (int r = 0; r < R; ++r) cnt4[k][r] = 0u;
        const float4 *s4 = reinterpret_cast<const float4 *>(scores);
        uchar4 *f4 = reinterpret_cast<uchar4 *>(flags);
        for (int64_t i = tid; i < n4; i += 4 * stride) {
            float4 x[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) x[j] = i + j * stride < n4 ? __ldcs(s4 + i + j * stride) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (i + j * stride < n4) {
                    uchar4 o;
                    o.x = static_cast<unsigned char>(flag_of(x[j].x, cnt4[0]));
                    o.y = static_cast<unsigned char>(flag_of(x[j].y, cnt4[1]));
                    o.z = static_cast<unsigned char>(flag_of(x[j].z, cnt4[2]));
                    o.w = static_cast<unsigned char>(flag_of(x[j].w, cnt4[3]));
                    f4[i + j * stride] = o;
                }
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) publish(4 * tid + k, cnt4[k]);
    }
    {   // scalar path: everything when not vectorised, else the last N % 4 scores
        unsigned int cnt[R];
#pragma unroll
        for (int r = 0; r < R; ++r) cnt[r] = 0u;
        for (int64_t i = 4 * n4 + tid; i < N; i += stride) flags[i] = static_cast<uint8_t>(flag_of(__ldg(scores + i), cnt));
        publish(4 * n4 + tid, cnt);
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < n_slots; idx += VIO_THREADS) {
        const unsigned int c = s_vc[idx];
        if (!c) continue;
        if (variant_counts) atomicAdd(variant_counts + idx, static_cast<unsigned long long>(c));
        atomicAdd(counts + (idx % rounds), static_cast<unsigned long long>(c));
    }
}

__global__ void keep_len_kernel(const int32_t *__restrict__ lens, const uint8_t *__restrict__ group,
                                const double *__restrict__ thr, int64_t U, int32_t *__restrict__ out) {
    const int64_t u = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (u >= U) return;
    const int len = max(lens[u], 0);
    const double q = thr[group ? group[u] : 0];
    int kept = 0;
    for (int rank = 0; rank < len; ++rank) {
        // 1.0 - (rank * 0.05) with the product rounded first, as CPython evaluates it
        const double conf = __dsub_rn(1.0, __dmul_rn(static_cast<double>(rank), 0.05));
        kept += conf >= q ? 1 : 0;
    }
    if (kept < 3 && len >= 3) kept = 3;
    out[u] = kept;
}

}  // namespace

extern "C" int flr_threshold_schedule(const float *q0, const uint64_t *state, float gamma, float q_min, int rounds,
                                      float *out_thr, void *stream_) {
    FLR_REQUIRE(q0 && state && out_thr, FLR_E_INVALID, "threshold_schedule: null pointer");
    FLR_REQUIRE(rounds >= 1 && rounds <= FLR_MAX_ROUNDS, FLR_E_INVALID, "threshold_schedule: rounds must be 1..%d",
                FLR_MAX_ROUNDS);
    FLR_REQUIRE(gamma > 0.f && gamma <= 1.f, FLR_E_UNSUPPORTED,
                "threshold_schedule: closed form needs 0 < gamma <= 1 (got %g)", static_cast<double>(gamma));
    threshold_schedule_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream_)>>>(
        q0, reinterpret_cast<const unsigned long long *>(state), gamma, q_min, rounds, out_thr);
    FLR_LAUNCH_CHECK("threshold_schedule_kernel");
    return FLR_OK;
}

extern "C" int flr_violation_detect(const float *scores, int64_t N, int V, const float *thr, int rounds,
                                    uint8_t *flags, uint64_t *counts, uint64_t *variant_counts, void *stream_) {
    FLR_REQUIRE(N >= 0 && V >= 1 && V <= 4096, FLR_E_INVALID, "violation_detect: need N>=0, 1<=V<=4096");
    FLR_REQUIRE(rounds >= 1 && rounds <= FLR_MAX_ROUNDS, FLR_E_INVALID, "violation_detect: rounds must be 1..%d",
                FLR_MAX_ROUNDS);
    if (N == 0) return FLR_OK;
    FLR_REQUIRE(scores && thr && flags && counts, FLR_E_INVALID, "violation_detect: null pointer");
    const int sms = device_sm_count();
    // at least 64 scores per thread when there are enough of them: the per-thread counters are published once
    int64_t blocks = std::min<int64_t>((N + VIO_THREADS * 64 - 1) / (VIO_THREADS * 64), static_cast<int64_t>(sms) * 8);
    blocks = std::max<int64_t>(blocks, 1);
    blocks = ((blocks + V - 1) / V) * V;                   // stride % V == 0
    const bool vec_ok = aligned16(scores) && (reinterpret_cast<uintptr_t>(flags) & 3u) == 0;
    const size_t smem = static_cast<size_t>(variant_counts ? V : 1) * rounds * sizeof(unsigned int);
    using Kern = void (*)(const float *, int64_t, int, const float *, int, uint8_t *, unsigned long long *,
                          unsigned long long *, bool);
    static const Kern kernels[FLR_MAX_ROUNDS] = {violation_kernel<1>, violation_kernel<2>, violation_kernel<3>,
                                                 violation_kernel<4>, violation_kernel<5>, violation_kernel<6>,
                                                 violation_kernel<7>, violation_kernel<8>};
    const Kern kern = kernels[rounds - 1];
    if (smem > 48 * 1024)
        FLR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    kern<<<static_cast<unsigned>(blocks), VIO_THREADS, smem, static_cast<cudaStream_t>(stream_)>>>(
        scores, N, V, thr, rounds, flags, reinterpret_cast<unsigned long long *>(counts),
        reinterpret_cast<unsigned long long *>(variant_counts), vec_ok);
    FLR_LAUNCH_CHECK("violation_kernel");
    return FLR_OK;
}

extern "C" int flr_conformal_keep_len(const int32_t *lens, const uint8_t *group, const double *thr, int64_t U,
                                      int32_t *out_keep, void *stream_) {
    FLR_REQUIRE(U >= 0, FLR_E_INVALID, "keep_len: U < 0");
    if (U == 0) return FLR_OK;
    FLR_REQUIRE(lens && thr && out_keep, FLR_E_INVALID, "keep_len: null pointer");
    const int64_t blocks = (U + 255) / 256;
    FLR_REQUIRE(blocks <= 0x7FFFFFFF, FLR_E_UNSUPPORTED, "keep_len: too many users for one launch");
    keep_len_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream_)>>>(lens, group, thr, U,
                                                                                                    out_keep);
    FLR_LAUNCH_CHECK("keep_len_kernel");
    return FLR_OK;
}
