// Demographic parity: per-group item histograms and the pairwise
// Jensen-Shannon distances between them -- utils.calculate_demographic_parity /
// calculate_js_divergence, src/utils.py:172-215 and :89-102.
//
// dense_group_histogram_kernel (dense [U, V, K] layout): see the comment above it.
// group_item_histogram_kernel (explicit group ids): one CTA per (group, slice of rows).  The CTA keeps the group's
// `vocab` counters in shared memory.  Recommendation lists are Zipf-skewed -- most of a warp often
// holds the same few titles -- so the 32 most popular ids are counted in registers from warp ballots
// and only the long tail uses shared-memory atomics; non-zero counters then go to the global uint64
// histogram.  Every occurrence counts (no de-duplication, as in the reference).
//
// js_distance_kernel: one CTA per group pair (i < j).  Keys = items either group holds; a missing
// key counts 1e-10; both sides are renormalised, then normalised once more exactly as
// scipy.spatial.distance.jensenshannon does, m = (p + q) / 2, sqrt((KL(p||m) + KL(q||m)) / 2),
// all in fp64 with fixed-order block reductions.
#include <algorithm>

#include "common.cuh"

using namespace flr;

namespace {

constexpr int GH_THREADS = 256;

__global__ void __launch_bounds__(GH_THREADS) group_item_histogram_kernel(
    const int32_t *__restrict__ lists, const int32_t *__restrict__ lens, const uint8_t *__restrict__ group,
    int64_t R, int K, int k_cap, int G, int V, int vocab, unsigned long long *__restrict__ hist) {
    extern __shared__ unsigned int s_cnt[];                 // [vocab]
    const int g = blockIdx.y;
    for (int i = threadIdx.x; i < vocab; i += GH_THREADS) s_cnt[i] = 0u;
    __syncthreads();
    // rows of group g: explicit ids (group[r] == g) or, when group == NULL, the rows r % V == g
    const int64_t stride = static_cast<int64_t>(gridDim.x) * GH_THREADS;
    const int64_t n_mine = group ? R : (R - g + V - 1) / V;
    const int64_t n_iter = (n_mine + stride - 1) / stride;
    unsigned inv[5];                                        // bit b of my lane clear -> flip ballot b
#pragma unroll
    for (int b = 0; b < 5; ++b) inv[b] = ((threadIdx.x >> b) & 1) ? 0u : 0xffffffffu;
    unsigned hot_cnt = 0;                                   // occurrences of id == my lane index
    for (int64_t it = 0; it < n_iter; ++it) {               // whole warps stay together for the vote
        const int64_t q = static_cast<int64_t>(blockIdx.x) * GH_THREADS + threadIdx.x + it * stride;
        int64_t r = -1;
        if (q < n_mine) {
            r = group ? q : q * V + g;
            if (group && group[r] != g) r = -1;
        }
        const int len = r >= 0 ? max(0, min(min(lens[r], K), k_cap)) : 0;
        const int max_len = __reduce_max_sync(0xffffffffu, len);
        const int32_t *row = lists + (r >= 0 ? r : 0) * K;
        // a lane's row is contiguous: fetch it in chunks of CH ids with independent loads (one exposed
        // memory latency per chunk instead of one per id), then vote on the chunk out of registers
        constexpr int CH = 16;
        const bool vec2 = (K % 2 == 0) && ((reinterpret_cast<uintptr_t>(lists) & 7u) == 0);
        for (int j0 = 0; j0 < max_len; j0 += CH) {
            int buf[CH];
            if (vec2 && j0 + CH <= K) {
#pragma unroll
                for (int e = 0; e < CH; e += 2) {
                    int2 q = make_int2(-1, -1);
                    if (j0 + e < len) q = *reinterpret_cast<const int2 *>(row + j0 + e);
                    buf[e] = q.x; buf[e + 1] = q.y;
                }
            } else {
#pragma unroll
                for (int e = 0; e < CH; ++e) buf[e] = (j0 + e < len) ? row[j0 + e] : -1;
            }
#pragma unroll
            for (int e = 0; e < CH; ++e) {
                if (j0 + e >= max_len) break;                      // warp-uniform
                int id = (j0 + e < len) ? buf[e] : -1;
                if (static_cast<unsigned>(id) >= static_cast<unsigned>(vocab)) id = -1;
                // Popular titles dominate (Zipf): ids 0..31 are counted with six ballots -- lane L keeps the
                // running count of id L in a register -- and only the long tail goes to shared-memory
                // atomics, where two lanes rarely meet.  (MATCH.ANY-based merging measured 3x slower.)
                const bool hot = static_cast<unsigned>(id) < 32u;
                unsigned m = __ballot_sync(0xffffffffu, hot);
#pragma unroll
                for (int b = 0; b < 5; ++b) m &= __ballot_sync(0xffffffffu, (id >> b) & 1) ^ inv[b];
                hot_cnt += __popc(m);
                if (id >= 32) atomicAdd(&s_cnt[id], 1u);
            }
        }
    }
    if (hot_cnt && (threadIdx.x & 31) < vocab) atomicAdd(&s_cnt[threadIdx.x & 31], hot_cnt);
    __syncthreads();
    unsigned long long *h = hist + static_cast<size_t>(g) * vocab;
    for (int i = threadIdx.x; i < vocab; i += GH_THREADS) {
        const unsigned int c = s_cnt[i];
        if (c) atomicAdd(h + i, static_cast<unsigned long long>(c));
    }
}

// Dense layout ([U, V, K], row r in group r % V), all V groups at once.  A persistent CTA walks a
// contiguous range of users and keeps ALL groups' counters in shared memory as 16-bit halves of 32-bit
// words (id x of group g lives in word g * Wd + x % Wd, half x / Wd, Wd = ceil(vocab / 2): 31 x 3706
// counters = 224 KB).  A warp reads whole rows -- RPW = 32 / (K / VEC) rows side by side, each lane a
// fixed (row slot, position) -- so row number, group and length cost a few instructions per row instead
// of per item, the loads are contiguous, and the ids of a row being (almost always) distinct, the lanes
// of a warp hit distinct counters: popular titles cost no more than rare ones.  An item that does not
// count (past the row's length, outside the vocabulary) is added to a spare word (one per lane, never
// read), so the loop has no branch and the idle adds do not pile onto one address.
// A counter half holds 65535: the CTA flushes to the global uint64 histogram after every chunk of
// 65535 / min(K, k_cap) users, which cannot add more than that to any counter.
constexpr int DH_THREADS = 1024;
constexpr int DH_WARPS = DH_THREADS / 32;
constexpr size_t kDenseHistSmemBudget = 226 * 1024;

__device__ __forceinline__ void dense_hist_flush(unsigned int *cnt, int V, int Wd, int vocab,
                                                 unsigned long long *__restrict__ hist) {
    for (int g = 0; g < V; ++g) {
        unsigned int *c = cnt + g * Wd;
        unsigned long long *h = hist + static_cast<size_t>(g) * vocab;
        for (int x = threadIdx.x; x < Wd; x += DH_THREADS) {
            const unsigned int w = c[x];
            if (w) {
                if (w & 0xffffu) atomicAdd(h + x, static_cast<unsigned long long>(w & 0xffffu));
                if (w >> 16) atomicAdd(h + x + Wd, static_cast<unsigned long long>(w >> 16));
                c[x] = 0u;
            }
        }
    }
}

template <int VEC>
__global__ void __launch_bounds__(DH_THREADS, 1) dense_group_histogram_kernel(
    const int32_t *__restrict__ lists, const int32_t *__restrict__ lens, int64_t U, int K, int k_cap, int V,
    int vocab, int chunk_users, unsigned long long *__restrict__ hist) {
    extern __shared__ unsigned int cnt[];                   // [V][Wd]
    const int Wd = (vocab + 1) >> 1;
    const int n_words = V * Wd;
    for (int i = threadIdx.x; i < n_words; i += DH_THREADS) cnt[i] = 0u;
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int Kv = K / VEC;                                 // vectors per row (K % VEC == 0, Kv <= 32)
    const int RPW = 32 / Kv;                                // rows a warp reads side by side
    const int sub = lane / Kv, pos = (lane - sub * Kv) * VEC;
    const bool lane_used = sub < RPW;
    const int Kc = min(K, k_cap);
    const int pass_rows = DH_WARPS * RPW;                   // rows the CTA covers per pass
    const int g_step = pass_rows % V;
    const int64_t per = (U + gridDim.x - 1) / gridDim.x;
    const int64_t ub = per * blockIdx.x, ue = min(U, ub + per);
    constexpr int UNR = 8;                                  // passes in flight: 8 x 32 warps x 200 B per SM
    for (int64_t uc = ub; uc < ue; uc += chunk_users) {
        const int64_t row0 = uc * V;
        const int n_rows = static_cast<int>(min(static_cast<int64_t>(chunk_users), ue - uc)) * V;
        const int32_t *base = lists + row0 * K;
        const int32_t *lbase = lens + row0;
        int row = warp * RPW + sub;                         // chunk-relative; group of row0 is 0 (row0 = uc * V)
        int g = row % V;
        // running pointers (one 64-bit add per pass each) instead of an address rebuilt from offsets per load
        const int32_t *ip = base + (static_cast<int64_t>(row) * K + pos);
        const int32_t *lp = lbase + row;
        const int pass_off = pass_rows * K;
        for (; __any_sync(0xffffffffu, row < n_rows); ) {
            int ids[UNR][VEC], wbase[UNR], len[UNR];
#pragma unroll
            for (int j = 0; j < UNR; ++j) {
                const bool ok = lane_used && row < n_rows;
                wbase[j] = g * Wd;
                len[j] = ok ? min(__ldg(lp), Kc) : 0;
                if constexpr (VEC == 2) {
                    const int2 v = ok ? __ldcs(reinterpret_cast<const int2 *>(ip)) : make_int2(-1, -1);
                    ids[j][0] = v.x; ids[j][1] = v.y;
                } else {
                    ids[j][0] = ok ? __ldcs(ip) : -1;
                }
                row += pass_rows; ip += pass_off; lp += pass_rows; g += g_step;
                if (g >= V) g -= V;
            }
#pragma unroll
            for (int j = 0; j < UNR; ++j) {
#pragma unroll
                for (int e = 0; e < VEC; ++e) {
                    const unsigned id = static_cast<unsigned>(ids[j][e]);
                    const bool counts = pos + e < len[j] && id < static_cast<unsigned>(vocab);
                    const bool hi = id >= static_cast<unsigned>(Wd);
                    // an item that does not count lands in a spare word past the tables (one per lane), never read
                    const unsigned x = counts ? wbase[j] + (hi ? id - Wd : id) : static_cast<unsigned>(n_words + lane);
                    atomicAdd(cnt + x, hi ? 0x10000u : 1u);
                }
            }
        }
        __syncthreads();
        if (uc + chunk_users < ue) {                        // a full chunk may have filled a counter half
            dense_hist_flush(cnt, V, Wd, vocab, hist);
            __syncthreads();
        }
    }
    dense_hist_flush(cnt, V, Wd, vocab, hist);
}

constexpr int JS_THREADS = 256;

// fixed-order block sum of two running values (shared-memory tree); every thread gets the result
__device__ __forceinline__ void block_sum2(double &a, double &b, double (*s)[2]) {
    const int t = threadIdx.x;
    s[t][0] = a; s[t][1] = b;
    __syncthreads();
    for (int half = JS_THREADS / 2; half >= 1; half >>= 1) {
        if (t < half) { s[t][0] += s[t + half][0]; s[t][1] += s[t + half][1]; }
        __syncthreads();
    }
    a = s[0][0]; b = s[0][1];
    __syncthreads();
}

__global__ void __launch_bounds__(JS_THREADS) js_distance_kernel(const unsigned long long *__restrict__ hist, int G,
                                                                 int vocab, double *__restrict__ out,
                                                                 unsigned char *__restrict__ valid) {
    __shared__ double s_red[JS_THREADS][2];
    // pair index -> (i, j), i < j, row-major like the reference's double loop
    int i = 0, rem = blockIdx.x;
    while (rem >= G - 1 - i) { rem -= G - 1 - i; ++i; }
    const int j = i + 1 + rem;
    const unsigned long long *ci = hist + static_cast<size_t>(i) * vocab, *cj = hist + static_cast<size_t>(j) * vocab;
    const int t = threadIdx.x;

    double ti = 0.0, tj = 0.0;                                  // group totals (exact integers)
    for (int x = t; x < vocab; x += JS_THREADS) { ti += static_cast<double>(ci[x]); tj += static_cast<double>(cj[x]); }
    block_sum2(ti, tj, s_red);
    if (!(ti > 0.0) || !(tj > 0.0)) {                           // an empty group is skipped (src/utils.py:200)
        if (t == 0) { out[blockIdx.x] = 0.0; valid[blockIdx.x] = 0; }
        return;
    }
    // p_x = count / total, or 1e-10 for a key only the other group holds (src/utils.py:95-97)
    auto pq = [&](int x, double &p, double &q) -> bool {
        const unsigned long long a = ci[x], b = cj[x];
        if (a == 0ull && b == 0ull) return false;
        p = a ? static_cast<double>(a) / ti : 1e-10;
        q = b ? static_cast<double>(b) / tj : 1e-10;
        return true;
    };
    double sp = 0.0, sq = 0.0;                                  // first renormalisation (:99-100)
    for (int x = t; x < vocab; x += JS_THREADS) { double p, q; if (pq(x, p, q)) { sp += p; sq += q; } }
    block_sum2(sp, sq, s_red);
    double sp2 = 0.0, sq2 = 0.0;                                // scipy normalises once more
    for (int x = t; x < vocab; x += JS_THREADS) { double p, q; if (pq(x, p, q)) { sp2 += p / sp; sq2 += q / sq; } }
    block_sum2(sp2, sq2, s_red);
    double left = 0.0, right = 0.0;
    for (int x = t; x < vocab; x += JS_THREADS) {
        double p, q;
        if (pq(x, p, q)) {
            p = (p / sp) / sp2;
            q = (q / sq) / sq2;
            const double m = (p + q) / 2.0;
            left += p * log(p / m);
            right += q * log(q / m);
        }
    }
    block_sum2(left, right, s_red);
    if (t == 0) {
        out[blockIdx.x] = sqrt((left + right) / 2.0);
        valid[blockIdx.x] = 1;
    }
}

// Equal opportunity (src/utils.py:254-266) from the same histogram: per group the number of DISTINCT
// relevant items that were recommended (counter > 0 on a relevant id) and the number of
// recommendations made (sum of the counters).  One CTA per group, fixed-order reduction.
__global__ void __launch_bounds__(JS_THREADS) equal_opportunity_kernel(const unsigned long long *__restrict__ hist,
                                                                       const uint8_t *__restrict__ relevant, int vocab,
                                                                       unsigned long long *__restrict__ out_relevant,
                                                                       unsigned long long *__restrict__ out_total) {
    __shared__ unsigned long long s_r[JS_THREADS], s_t[JS_THREADS];
    const int g = blockIdx.x, t = threadIdx.x;
    const unsigned long long *h = hist + static_cast<size_t>(g) * vocab;
    unsigned long long rel = 0, tot = 0;
    for (int x = t; x < vocab; x += JS_THREADS) {
        const unsigned long long c = h[x];
        tot += c;
        rel += (c != 0ull && relevant[x] != 0) ? 1ull : 0ull;
    }
    s_r[t] = rel; s_t[t] = tot;
    __syncthreads();
    for (int half = JS_THREADS / 2; half >= 1; half >>= 1) {
        if (t < half) { s_r[t] += s_r[t + half]; s_t[t] += s_t[t + half]; }
        __syncthreads();
    }
    if (t == 0) { out_relevant[g] = s_r[0]; out_total[g] = s_t[0]; }
}

}  // namespace

extern "C" int flr_group_item_histogram(const int32_t *lists, const int32_t *lens, const uint8_t *group, int64_t R,
                                        int K, int k_cap, int n_groups, int vocab, uint64_t *hist, void *stream_) {
    FLR_REQUIRE(R >= 0 && K >= 1 && k_cap >= 1 && n_groups >= 1 && n_groups <= 256 && vocab >= 1, FLR_E_INVALID,
                "group_item_histogram: bad shape");
    FLR_REQUIRE(static_cast<size_t>(vocab) * 4 <= 200 * 1024, FLR_E_UNSUPPORTED,
                "group_item_histogram: vocab %d does not fit the shared-memory counters (max 51200)", vocab);
    if (R == 0) return FLR_OK;
    FLR_REQUIRE(lists && lens && hist, FLR_E_INVALID, "group_item_histogram: null pointer");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    {   // dense layout with every group's counters resident: one pass of coalesced reads
        const size_t dense_smem = (static_cast<size_t>(n_groups) * ((vocab + 1) / 2) + 32) * sizeof(unsigned int);   // + a spare word per lane
        const int Kc = std::min(K, k_cap);
        const int chunk_users = 65535 / Kc;                 // a chunk adds at most 2^16 - 1 to any counter
        const int vec = (K % 2 == 0 && (reinterpret_cast<uintptr_t>(lists) & 7u) == 0) ? 2 : 1;
        if (!group && R % n_groups == 0 && dense_smem <= kDenseHistSmemBudget && chunk_users >= 1 && K / vec <= 32) {
            const int64_t U = R / n_groups;
            const int grid = static_cast<int>(std::min<int64_t>(device_sm_count(), (U + 63) / 64));
            auto kern = vec == 2 ? dense_group_histogram_kernel<2> : dense_group_histogram_kernel<1>;
            FLR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          static_cast<int>(dense_smem)));
            kern<<<grid, DH_THREADS, dense_smem, stream>>>(lists, lens, U, K, k_cap, n_groups, vocab, chunk_users,
                                                           reinterpret_cast<unsigned long long *>(hist));
            FLR_LAUNCH_CHECK("dense_group_histogram_kernel");
            return FLR_OK;
        }
    }
    const size_t smem = static_cast<size_t>(vocab) * sizeof(unsigned int);
    if (smem > 48 * 1024)
        FLR_CUDA(cudaFuncSetAttribute(group_item_histogram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      static_cast<int>(smem)));
    const int64_t rows_per_group = group ? R : (R + n_groups - 1) / n_groups;
    const int sms = device_sm_count();
    int bx = static_cast<int>(std::min<int64_t>((rows_per_group + GH_THREADS - 1) / GH_THREADS,
                                                std::max(1, sms * 8 / n_groups)));
    bx = std::max(bx, 1);
    dim3 grid(bx, n_groups);
    group_item_histogram_kernel<<<grid, GH_THREADS, smem, static_cast<cudaStream_t>(stream_)>>>(
        lists, lens, group, R, K, k_cap, n_groups, n_groups, vocab, reinterpret_cast<unsigned long long *>(hist));
    FLR_LAUNCH_CHECK("group_item_histogram_kernel");
    return FLR_OK;
}

extern "C" int flr_js_distance_matrix(const uint64_t *hist, int n_groups, int vocab, double *out_dist,
                                      uint8_t *out_valid, void *stream_) {
    FLR_REQUIRE(n_groups >= 1 && vocab >= 1, FLR_E_INVALID, "js_distance_matrix: bad shape");
    const int n_pairs = n_groups * (n_groups - 1) / 2;
    if (n_pairs == 0) return FLR_OK;
    FLR_REQUIRE(hist && out_dist && out_valid, FLR_E_INVALID, "js_distance_matrix: null pointer");
    js_distance_kernel<<<n_pairs, JS_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(
        reinterpret_cast<const unsigned long long *>(hist), n_groups, vocab, out_dist, out_valid);
    FLR_LAUNCH_CHECK("js_distance_kernel");
    return FLR_OK;
}

extern "C" int flr_equal_opportunity_counts(const uint64_t *hist, const uint8_t *relevant, int n_groups, int vocab,
                                            uint64_t *out_relevant, uint64_t *out_total, void *stream_) {
    FLR_REQUIRE(n_groups >= 0 && vocab >= 1, FLR_E_INVALID, "equal_opportunity_counts: bad shape");
    if (n_groups == 0) return FLR_OK;
    FLR_REQUIRE(hist && relevant && out_relevant && out_total, FLR_E_INVALID, "equal_opportunity_counts: null pointer");
    equal_opportunity_kernel<<<n_groups, JS_THREADS, 0, static_cast<cudaStream_t>(stream_)>>>(
        reinterpret_cast<const unsigned long long *>(hist), relevant, vocab,
        reinterpret_cast<unsigned long long *>(out_relevant), reinterpret_cast<unsigned long long *>(out_total));
    FLR_LAUNCH_CHECK("equal_opportunity_kernel");
    return FLR_OK;
}
