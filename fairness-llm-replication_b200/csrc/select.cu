// Exact order statistic by radix select over the float bit patterns -- replaces
// np.sort(scores)[ceil((n+1)(1-alpha))-1] of src/phase3_facter_mitigation.py:165-178.
//
// 11-bit digits, most significant first (3 passes for f32 keys, 6 for f64).  A pass is two
// launches so that a multi-GPU caller can all-reduce the 2048-bin histogram in between:
//   select_hist : histogram of the current digit over the scores whose higher digits match
//                 the prefix fixed so far (warp-aggregated shared-memory atomics, one global
//                 atomic per non-empty bin per CTA);
//   select_scan : one CTA per group prefix-sums the bins, finds the one holding rank k and
//                 narrows the prefix.  Ties are exact: the last pass leaves #less and #equal.
#include <algorithm>

#include "common.cuh"
#include "select_key.cuh"

using namespace flr;
using namespace flr::select_key;

namespace {

constexpr int HIST_THREADS = 256;
constexpr int kSmemGroups = 8;     // groups whose histograms fit shared memory together

template <typename T> struct KeyOf;
template <> struct KeyOf<float> {
    using key_t = uint32_t;
    static constexpr int bits = 32;
    __device__ static key_t key(float x) { return key_f32(x); }
};
template <> struct KeyOf<double> {
    using key_t = uint64_t;
    static constexpr int bits = 64;
    __device__ static key_t key(double x) { return key_f64(x); }
};

// warp-aggregated add: lanes that hit the same bin elect one leader (bin < 0 = no vote)
__device__ __forceinline__ void aggregated_add(unsigned int *hist, int bin) {
    const unsigned peers = __match_any_sync(0xffffffffu, bin);
    if (bin >= 0 && (threadIdx.x & 31) == (__ffs(peers) - 1)) atomicAdd(hist + bin, __popc(peers));
}

template <typename T, bool SMEM>
__global__ void __launch_bounds__(HIST_THREADS) select_hist_kernel(const T *__restrict__ scores,
                                                                   const uint8_t *__restrict__ group, int64_t N,
                                                                   int G, int pass,
                                                                   const unsigned long long *__restrict__ state,
                                                                   unsigned long long *__restrict__ hist) {
    using K = KeyOf<T>;
    using key_t = typename K::key_t;
    extern __shared__ unsigned int s_hist[];          // [G][BINS] when SMEM
    const int shift = pass_shift(K::bits, pass);
    const int nbits = pass_nbits(K::bits, pass);
    const int hi_shift = shift + nbits;               // bits above the current digit
    const unsigned mask = (1u << nbits) - 1u;
    if (SMEM) {
        for (int i = threadIdx.x; i < G * FLR_SELECT_BINS; i += HIST_THREADS) s_hist[i] = 0u;
        __syncthreads();
    }
    const int64_t stride = static_cast<int64_t>(gridDim.x) * HIST_THREADS;
    const int64_t start = static_cast<int64_t>(blockIdx.x) * HIST_THREADS + threadIdx.x;
    // whole warps iterate together so the aggregated add always sees 32 lanes
    const int64_t n_iter = (N + stride - 1) / stride;
    for (int64_t it = 0; it < n_iter; ++it) {
        const int64_t i = start + it * stride;
        int bin = -1;
        if (i < N) {
            const int g = group ? group[i] : 0;
            if (g < G) {
                const key_t key = K::key(scores[i]);
                bool match = true;
                if (pass > 0) {
                    const key_t prefix = static_cast<key_t>(state[g * FLR_SELECT_STATE_WORDS + FLR_ST_PREFIX]);
                    const bool empty = state[g * FLR_SELECT_STATE_WORDS + FLR_ST_EMPTY] != 0ull;
                    match = !empty && hi_shift < K::bits && ((key >> hi_shift) == (prefix >> hi_shift));
                }
                if (match) bin = g * FLR_SELECT_BINS + static_cast<int>((key >> shift) & mask);
            }
        }
        if (SMEM) {
            aggregated_add(s_hist, bin);
        } else if (bin >= 0) {
            atomicAdd(hist + bin, 1ull);
        }
    }
    if (SMEM) {
        __syncthreads();
        for (int i = threadIdx.x; i < G * FLR_SELECT_BINS; i += HIST_THREADS) {
            const unsigned int c = s_hist[i];
            if (c) atomicAdd(hist + i, static_cast<unsigned long long>(c));
        }
    }
}

// Single group, no group ids (the pipeline's global quantile): a pure stream.  Each thread reads 16-byte
// vectors, several in flight, and feeds its matching digits through a one-entry run-length buffer --
// neighbouring scores often share a digit in the early passes, and in the late passes almost nothing
// matches the prefix -- so shared-memory atomics are rare and no warp-wide MATCH is needed.
constexpr int HIST_UNR = 4;

template <typename T>
__global__ void __launch_bounds__(HIST_THREADS) select_hist_stream_kernel(const T *__restrict__ scores, int64_t N,
                                                                          int pass,
                                                                          const unsigned long long *__restrict__ state,
                                                                          unsigned long long *__restrict__ hist) {
    using K = KeyOf<T>;
    using key_t = typename K::key_t;
    constexpr int VEC = 16 / sizeof(T);
    __shared__ unsigned int s_hist[FLR_SELECT_BINS];
    const int shift = pass_shift(K::bits, pass);
    const int nbits = pass_nbits(K::bits, pass);
    const int hi_shift = shift + nbits;
    const unsigned mask = (1u << nbits) - 1u;
    for (int i = threadIdx.x; i < FLR_SELECT_BINS; i += HIST_THREADS) s_hist[i] = 0u;
    __syncthreads();
    bool live = true;
    key_t want_hi = 0;
    if (pass > 0) {
        live = state[FLR_ST_EMPTY] == 0ull && hi_shift < K::bits;
        want_hi = live ? static_cast<key_t>(state[FLR_ST_PREFIX]) >> hi_shift : 0;
    }
    int run_bin = -1;
    unsigned run_cnt = 0;
    auto feed = [&](T x) {
        const key_t key = K::key(x);
        if (pass > 0 && (key >> hi_shift) != want_hi) return;
        const int bin = static_cast<int>((key >> shift) & mask);
        if (bin == run_bin) { ++run_cnt; return; }
        if (run_cnt) atomicAdd(s_hist + run_bin, run_cnt);
        run_bin = bin; run_cnt = 1;
    };
    if (live) {
        const int64_t n_vec = N / VEC;
        const int4 *v4 = reinterpret_cast<const int4 *>(scores);
        const int64_t stride = static_cast<int64_t>(gridDim.x) * HIST_THREADS;
        for (int64_t i0 = static_cast<int64_t>(blockIdx.x) * HIST_THREADS + threadIdx.x; i0 < n_vec; i0 += stride * HIST_UNR) {
            int4 q[HIST_UNR];
#pragma unroll
            for (int j = 0; j < HIST_UNR; ++j) {
                const int64_t i = i0 + j * stride;
                q[j] = i < n_vec ? __ldcs(v4 + i) : make_int4(0, 0, 0, 0);
            }
#pragma unroll
            for (int j = 0; j < HIST_UNR; ++j) {
                if (i0 + j * stride < n_vec) {
                    const T *x = reinterpret_cast<const T *>(&q[j]);
#pragma unroll
                    for (int e = 0; e < VEC; ++e) feed(x[e]);
                }
            }
        }
        // the last N % VEC scores
        const int64_t tail = n_vec * VEC + static_cast<int64_t>(blockIdx.x) * HIST_THREADS + threadIdx.x;
        if (tail < N) feed(scores[tail]);
        if (run_cnt) atomicAdd(s_hist + run_bin, run_cnt);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < FLR_SELECT_BINS; i += HIST_THREADS) {
        const unsigned int c = s_hist[i];
        if (c) atomicAdd(hist + i, static_cast<unsigned long long>(c));
    }
}

template <typename T>
__global__ void __launch_bounds__(256) select_scan_kernel(int pass, int n_passes, double one_minus_alpha,
                                                          const unsigned long long *__restrict__ k_override,
                                                          const unsigned long long *__restrict__ hist,
                                                          unsigned long long *__restrict__ state,
                                                          T *__restrict__ out_thr) {
    using K = KeyOf<T>;
    __shared__ unsigned long long s_scan[256];
    __shared__ unsigned long long s_k;
    __shared__ int s_skip;
    const int g = blockIdx.x, t = threadIdx.x;
    const unsigned long long *h = hist + static_cast<size_t>(g) * FLR_SELECT_BINS;
    unsigned long long *st = state + static_cast<size_t>(g) * FLR_SELECT_STATE_WORDS;
    constexpr int PER = FLR_SELECT_BINS / 256;
    unsigned long long loc[PER], sum = 0;
#pragma unroll
    for (int i = 0; i < PER; ++i) { loc[i] = h[t * PER + i]; sum += loc[i]; }
    s_scan[t] = sum;
    __syncthreads();
    for (int off = 1; off < 256; off <<= 1) {
        const unsigned long long v = t >= off ? s_scan[t - off] : 0ull;
        __syncthreads();
        s_scan[t] += v;
        __syncthreads();
    }
    const unsigned long long total = s_scan[255];
    if (t == 0) {
        s_skip = 0;
        if (pass == 0) {
            st[FLR_ST_N] = total;
            st[FLR_ST_PREFIX] = 0ull;
            st[FLR_ST_LESS] = 0ull;
            st[FLR_ST_EQUAL] = 0ull;
            if (total == 0ull) {
                st[FLR_ST_EMPTY] = 1ull; st[FLR_ST_K] = 0ull; st[FLR_ST_KREM] = 0ull;
                out_thr[g] = static_cast<T>(0.5);      // src/phase3_facter_mitigation.py:169-171
                s_skip = 1;
            } else {
                st[FLR_ST_EMPTY] = 0ull;
                unsigned long long k;
                if (k_override) {
                    k = k_override[g];
                } else {   // int(np.ceil((n + 1) * (1 - alpha))) - 1, src/phase3_facter_mitigation.py:174
                    const double x = ceil(__dmul_rn(static_cast<double>(total + 1ull), one_minus_alpha));
                    const long long idx = static_cast<long long>(x) - 1;
                    k = idx < 0 ? 0ull : static_cast<unsigned long long>(idx);
                }
                if (k > total - 1ull) k = total - 1ull;
                st[FLR_ST_K] = k;
                s_k = k;
            }
        } else {
            if (st[FLR_ST_EMPTY] != 0ull) s_skip = 1;
            s_k = st[FLR_ST_KREM];
        }
    }
    __syncthreads();
    if (s_skip) return;
    const unsigned long long k = s_k;
    unsigned long long before = s_scan[t] - sum;           // exclusive prefix of this thread's bins
    if (k >= before && k < before + sum) {                 // exactly one thread owns rank k
        int b = 0;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            if (k >= before + loc[i]) { before += loc[i]; b = i + 1; } else break;
        }
        const int bin = t * PER + b;
        const int shift = pass_shift(K::bits, pass);
        const unsigned long long prefix = st[FLR_ST_PREFIX] | (static_cast<unsigned long long>(bin) << shift);
        st[FLR_ST_PREFIX] = prefix;
        st[FLR_ST_KREM] = k - before;
        st[FLR_ST_LESS] += before;
        if (pass == n_passes - 1) {
            st[FLR_ST_EQUAL] = loc[b];
            if constexpr (K::bits == 32) out_thr[g] = unkey_f32(static_cast<uint32_t>(prefix));
            else out_thr[g] = unkey_f64(prefix);
        }
    }
}

template <typename T>
int launch_hist(const void *scores, const uint8_t *group, int64_t N, int G, int pass, const uint64_t *state,
                uint64_t *hist, cudaStream_t stream) {
    const int sms = device_sm_count();
    const int64_t need = (N + HIST_THREADS - 1) / HIST_THREADS;
    const int grid = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(need, static_cast<int64_t>(sms) * 8)));
    auto *st = reinterpret_cast<const unsigned long long *>(state);
    auto *hh = reinterpret_cast<unsigned long long *>(hist);
    if (G == 1 && !group && (reinterpret_cast<uintptr_t>(scores) & 15u) == 0) {
        const int64_t n_vec = N / (16 / sizeof(T));
        const int64_t want = (n_vec + HIST_THREADS * HIST_UNR - 1) / (HIST_THREADS * HIST_UNR);
        const int g2 = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(want, static_cast<int64_t>(sms) * 8)));
        select_hist_stream_kernel<T><<<g2, HIST_THREADS, 0, stream>>>(static_cast<const T *>(scores), N, pass, st, hh);
        FLR_LAUNCH_CHECK("select_hist_stream_kernel");
        return FLR_OK;
    }
    if (G <= kSmemGroups) {
        const size_t smem = static_cast<size_t>(G) * FLR_SELECT_BINS * sizeof(unsigned int);
        if (smem > 48 * 1024)
            FLR_CUDA(cudaFuncSetAttribute(select_hist_kernel<T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          static_cast<int>(smem)));
        select_hist_kernel<T, true><<<grid, HIST_THREADS, smem, stream>>>(static_cast<const T *>(scores), group, N, G,
                                                                          pass, st, hh);
    } else {
        select_hist_kernel<T, false><<<grid, HIST_THREADS, 0, stream>>>(static_cast<const T *>(scores), group, N, G,
                                                                        pass, st, hh);
    }
    FLR_LAUNCH_CHECK("select_hist_kernel");
    return FLR_OK;
}

}  // namespace

extern "C" int flr_select_num_passes(int dtype) {
    return dtype == FLR_DTYPE_F32 ? 3 : dtype == FLR_DTYPE_F64 ? 6 : -1;
}

extern "C" int flr_select_hist(const void *scores, int dtype, const uint8_t *group, int64_t N, int G, int pass,
                               const uint64_t *state, uint64_t *hist, void *stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int np = flr_select_num_passes(dtype);
    FLR_REQUIRE(np > 0, FLR_E_INVALID, "select: unknown dtype %d", dtype);
    FLR_REQUIRE(N >= 0 && G >= 1 && G <= 256, FLR_E_INVALID, "select: need N>=0 and 1<=G<=256");
    FLR_REQUIRE(pass >= 0 && pass < np, FLR_E_INVALID, "select: pass %d out of range", pass);
    FLR_REQUIRE(state && hist, FLR_E_INVALID, "select: null state/hist");
    FLR_REQUIRE(G == 1 || group, FLR_E_INVALID, "select: G > 1 needs group ids");
    if (N == 0) return FLR_OK;
    FLR_REQUIRE(scores, FLR_E_INVALID, "select: null scores");
    return dtype == FLR_DTYPE_F32 ? launch_hist<float>(scores, group, N, G, pass, state, hist, stream)
                                  : launch_hist<double>(scores, group, N, G, pass, state, hist, stream);
}

extern "C" int flr_select_scan(int dtype, int G, int pass, double one_minus_alpha, const uint64_t *k_override,
                               const uint64_t *hist, uint64_t *state, void *out_thr, void *stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int np = flr_select_num_passes(dtype);
    FLR_REQUIRE(np > 0, FLR_E_INVALID, "select: unknown dtype %d", dtype);
    FLR_REQUIRE(G >= 1 && G <= 256 && pass >= 0 && pass < np, FLR_E_INVALID, "select: bad G or pass");
    FLR_REQUIRE(hist && state && out_thr, FLR_E_INVALID, "select: null pointer");
    auto *hh = reinterpret_cast<const unsigned long long *>(hist);
    auto *st = reinterpret_cast<unsigned long long *>(state);
    auto *ko = reinterpret_cast<const unsigned long long *>(k_override);
    if (dtype == FLR_DTYPE_F32)
        select_scan_kernel<float><<<G, 256, 0, stream>>>(pass, np, one_minus_alpha, ko, hh, st, static_cast<float *>(out_thr));
    else
        select_scan_kernel<double><<<G, 256, 0, stream>>>(pass, np, one_minus_alpha, ko, hh, st, static_cast<double *>(out_thr));
    FLR_LAUNCH_CHECK("select_scan_kernel");
    return FLR_OK;
}
