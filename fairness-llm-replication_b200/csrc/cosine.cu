// Cosine nonconformity  s = max(0, 1 - <a,b>/(|a||b|))  -- the embedding form of the score the
// reference simulates at src/phase3_facter_mitigation.py:125-139.
//
// HBM-bound row-dot kernel: one warp per user.  The anchor (neutral) row is read once and
// kept in registers as fp64; each of the user's V counterfactual rows is streamed with
// 128-bit loads, NC per lane, two rows in flight.  fp32 products are exact in fp64, so
// dot / |a|^2 / |b|^2 are accumulated in fp64 (lane-strided chunks, then an xor butterfly):
// the score stays accurate where cos -> 1, and oracle/np_oracle.py:cosine_nonconformity_mirror
// reproduces it bit for bit.  The pass-0 radix histogram of the scores (11 top key bits) is
// built in shared memory on the way out, so the select never re-reads them for pass 0.
#include <algorithm>

#include "common.cuh"
#include "select_key.cuh"

using namespace flr;

namespace {

constexpr int COS_THREADS = 256;
constexpr int COS_WARPS = COS_THREADS / 32;

__device__ __forceinline__ double warp_sum(double x) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
    return x;
}

__device__ __forceinline__ void fma4(const double (&a)[4], const float4 &b, double &dot, double &nb) {
    const double b0 = b.x, b1 = b.y, b2 = b.z, b3 = b.w;
    dot = fma(a[0], b0, dot); nb = fma(b0, b0, nb);
    dot = fma(a[1], b1, dot); nb = fma(b1, b1, nb);
    dot = fma(a[2], b2, dot); nb = fma(b2, b2, nb);
    dot = fma(a[3], b3, dot); nb = fma(b3, b3, nb);
}

__device__ __forceinline__ float score_from(double dot, double na, double nb) {
    const double den = sqrt(na * nb);
    if (!(den > 0.0)) return 1.0f;
    const double s = 1.0 - dot / den;
    return static_cast<float>(s > 0.0 ? s : 0.0);
}

// NC = float4 chunks per lane = ceil(D / 128)
template <int NC>
__global__ void __launch_bounds__(COS_THREADS) cosine_kernel(const float *__restrict__ a, const float *__restrict__ b,
                                                             int64_t U, int V, int D, float *__restrict__ out,
                                                             unsigned long long *__restrict__ hist0) {
    __shared__ unsigned int s_hist[FLR_SELECT_BINS];
    const bool do_hist = hist0 != nullptr;
    if (do_hist) {
        for (int i = threadIdx.x; i < FLR_SELECT_BINS; i += COS_THREADS) s_hist[i] = 0u;
        __syncthreads();
    }
    const int lane = threadIdx.x & 31;
    const int nchunk = D >> 2;
    const int64_t warp0 = static_cast<int64_t>(blockIdx.x) * COS_WARPS + (threadIdx.x >> 5);
    const int64_t nwarps = static_cast<int64_t>(gridDim.x) * COS_WARPS;

    for (int64_t u = warp0; u < U; u += nwarps) {
        double av[NC][4];
        double na = 0.0;
        const float4 *arow = reinterpret_cast<const float4 *>(a + u * D);
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const int ch = lane + 32 * c;
            float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ch < nchunk) q = ldg_stream_f4(arow + ch);
            av[c][0] = q.x; av[c][1] = q.y; av[c][2] = q.z; av[c][3] = q.w;
#pragma unroll
            for (int e = 0; e < 4; ++e) na = fma(av[c][e], av[c][e], na);
        }
        na = warp_sum(na);

        const float4 *brow = reinterpret_cast<const float4 *>(b + u * V * D);
        float mine = 0.f;          // lane (v & 31) keeps the score of variant v
        int v = 0;
        for (; v + 2 <= V; v += 2) {
            float4 q0[NC], q1[NC];
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                const int ch = lane + 32 * c;
                q0[c] = make_float4(0.f, 0.f, 0.f, 0.f);
                q1[c] = q0[c];
                if (ch < nchunk) {
                    q0[c] = ldg_stream_f4(brow + static_cast<size_t>(v) * nchunk + ch);
                    q1[c] = ldg_stream_f4(brow + static_cast<size_t>(v + 1) * nchunk + ch);
                }
            }
            double d0 = 0.0, n0 = 0.0, d1 = 0.0, n1 = 0.0;
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                fma4(av[c], q0[c], d0, n0);
                fma4(av[c], q1[c], d1, n1);
            }
            d0 = warp_sum(d0); n0 = warp_sum(n0);
            d1 = warp_sum(d1); n1 = warp_sum(n1);
            const float s0 = score_from(d0, na, n0), s1 = score_from(d1, na, n1);
            if (lane == (v & 31)) mine = s0;
            if (lane == ((v + 1) & 31)) mine = s1;
            if (((v + 2) & 31) == 0) {         // 32 scores gathered: one coalesced store
                const int vv = v + 2 - 32 + lane;
                out[u * V + vv] = mine;
                if (do_hist) atomicAdd(&s_hist[select_key::key_f32(mine) >> 21], 1u);
            }
        }
        if (v < V) {
            double d0 = 0.0, n0 = 0.0;
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                const int ch = lane + 32 * c;
                float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
                if (ch < nchunk) q = ldg_stream_f4(brow + static_cast<size_t>(v) * nchunk + ch);
                fma4(av[c], q, d0, n0);
            }
            d0 = warp_sum(d0); n0 = warp_sum(n0);
            const float s0 = score_from(d0, na, n0);
            if (lane == (v & 31)) mine = s0;
            ++v;
        }
        const int rem = V & 31;                 // scores not yet written
        if (rem && lane < rem) {
            const int vv = V - rem + lane;
            out[u * V + vv] = mine;
            if (do_hist) atomicAdd(&s_hist[select_key::key_f32(mine) >> 21], 1u);
        }
    }
    if (do_hist) {
        __syncthreads();
        for (int i = threadIdx.x; i < FLR_SELECT_BINS; i += COS_THREADS) {
            const unsigned int c = s_hist[i];
            if (c) atomicAdd(hist0 + i, static_cast<unsigned long long>(c));
        }
    }
}

// ---------------------------------------------------------------------------------------
// TMA-pipelined form (the fast path).  One persistent CTA per SM: a producer thread streams
// groups of G users -- the anchor rows and the G*V counterfactual rows are each one contiguous
// block of HBM -- into a ring of shared-memory stages with bulk TMA copies (cp.async.bulk +
// mbarrier tx-count); sixteen consumer warps take batches of four rows out of the stage that
// just landed (sixteen consumer warps: ncu showed the consumers, not HBM, on the critical path with eight).  HBM requests are therefore always RING bytes deep per SM (~190 KB),
// independent of what the consumers are doing.  The arithmetic is the one described at the
// top of this file, bit for bit: lane-strided fp64 accumulation, xor butterfly 16,8,4,2,1 --
// the butterfly is "transposed" (eight running sums are halved at each step instead of all
// being shuffled), which pairs exactly the same operands.
// ---------------------------------------------------------------------------------------
constexpr int TMA_CONSUMER_WARPS = 16;
constexpr int TMA_THREADS = (TMA_CONSUMER_WARPS + 1) * 32;
constexpr int TMA_MAX_STAGES = 4;

struct CosTmaParams {
    const float *a, *b;
    int64_t U;
    int V, D, G, stages;
    uint32_t stage_bytes, a_bytes;       // per stage: [a chunk | b chunk], a_bytes = G*D*4 rounded to 128
    float *out;
    unsigned long long *hist0;
};

template <int NC>
__device__ __forceinline__ void load_row_f64(const float *row, int lane, int nchunk, double (&v)[NC][4]) {
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        const int ch = lane + 32 * c;
        float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ch < nchunk) q = *reinterpret_cast<const float4 *>(row + 4 * ch);
        v[c][0] = q.x; v[c][1] = q.y; v[c][2] = q.z; v[c][3] = q.w;
    }
}

// Reduce the eight running sums of a four-row batch (x[2i] = dot_i, x[2i+1] = |b_i|^2) over the
// warp and return this lane's row score: lane l ends up with row (l >> 3).  Transposed xor
// butterfly: at offsets 16, 8, 4 each lane keeps half of its sums and ships the other half, then
// two plain steps -- every addition pairs the same two operands as the plain butterfly would.
__device__ __forceinline__ float reduce_batch_scores(const double (&x)[8], const double (&na_row)[4], int lane) {
    double y4[4], y2[2], y1;
    {
        const bool hi = lane & 16;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const double keep = hi ? x[k + 4] : x[k], send = hi ? x[k] : x[k + 4];
            y4[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
    }
    {
        const bool hi = lane & 8;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const double keep = hi ? y4[k + 2] : y4[k], send = hi ? y4[k] : y4[k + 2];
            y2[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
    }
    {
        const bool hi = lane & 4;
        const double keep = hi ? y2[1] : y2[0], send = hi ? y2[0] : y2[1];
        y1 = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    y1 += __shfl_xor_sync(0xffffffffu, y1, 2);
    y1 += __shfl_xor_sync(0xffffffffu, y1, 1);
    // lane l now holds sum number (l >> 2) & 7 = 2 * row + kind, kind = (l >> 2) & 1 (0 dot, 1 |b|^2)
    const double other = __shfl_xor_sync(0xffffffffu, y1, 4);
    const bool is_nb = (lane >> 2) & 1;
    const double dot = is_nb ? other : y1, nb = is_nb ? y1 : other;
    const int my_i = lane >> 3;
    const double my_na = my_i == 0 ? na_row[0] : my_i == 1 ? na_row[1] : my_i == 2 ? na_row[2] : na_row[3];
    return score_from(dot, my_na, nb);
}

// The same, with the anchor row's running |a|^2 reduced alongside: its plain butterfly (16, 8, 4, 2, 1 -- the
// bits warp_sum() gives) runs step by step beside the batch's, so the two dependency chains hide each other's
// shuffle latency instead of running one after the other.
__device__ __forceinline__ float reduce_batch_scores_na(const double (&x)[8], double acc_a, int lane) {
    double y4[4], y2[2], y1;
    {
        const bool hi = lane & 16;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const double keep = hi ? x[k + 4] : x[k], send = hi ? x[k] : x[k + 4];
            y4[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
        acc_a += __shfl_xor_sync(0xffffffffu, acc_a, 16);
    }
    {
        const bool hi = lane & 8;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const double keep = hi ? y4[k + 2] : y4[k], send = hi ? y4[k] : y4[k + 2];
            y2[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
        acc_a += __shfl_xor_sync(0xffffffffu, acc_a, 8);
    }
    {
        const bool hi = lane & 4;
        const double keep = hi ? y2[1] : y2[0], send = hi ? y2[0] : y2[1];
        y1 = keep + __shfl_xor_sync(0xffffffffu, send, 4);
        acc_a += __shfl_xor_sync(0xffffffffu, acc_a, 4);
    }
    y1 += __shfl_xor_sync(0xffffffffu, y1, 2);
    acc_a += __shfl_xor_sync(0xffffffffu, acc_a, 2);
    y1 += __shfl_xor_sync(0xffffffffu, y1, 1);
    acc_a += __shfl_xor_sync(0xffffffffu, acc_a, 1);
    const double other = __shfl_xor_sync(0xffffffffu, y1, 4);
    const bool is_nb = (lane >> 2) & 1;
    const double dot = is_nb ? other : y1, nb = is_nb ? y1 : other;
    return score_from(dot, acc_a, nb);
}

// Reduce four running sums over the warp with the same transposed xor butterfly (offsets 16, 8 halve the
// set each lane carries, then 4, 2, 1): lane l ends up with sum number (l >> 3) & 3.  Every addition pairs
// the operands the plain butterfly would, so each sum has the bits warp_sum() gives it.
__device__ __forceinline__ double reduce4_transposed(const double (&v)[4], int lane) {
    double y2[2], y1;
    {
        const bool hi = lane & 16;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const double keep = hi ? v[k + 2] : v[k], send = hi ? v[k] : v[k + 2];
            y2[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
    }
    {
        const bool hi = lane & 8;
        const double keep = hi ? y2[1] : y2[0], send = hi ? y2[0] : y2[1];
        y1 = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    y1 += __shfl_xor_sync(0xffffffffu, y1, 4);
    y1 += __shfl_xor_sync(0xffffffffu, y1, 2);
    y1 += __shfl_xor_sync(0xffffffffu, y1, 1);
    return y1;
}

// FAST: D == NC * 128 (no partial chunk) and batches of four rows that never straddle two users --
// the benchmark shape (two users per stage, sixteen batches, one per consumer warp).
template <int NC, bool FAST>
__global__ void __launch_bounds__(TMA_THREADS, 1) cosine_tma_kernel(const CosTmaParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem);                 // [stages]
    uint64_t *empty = full + TMA_MAX_STAGES;                             // [stages]
    unsigned int *s_hist = reinterpret_cast<unsigned int *>(smem + 128); // [2048]
    unsigned char *ring = smem + 128 + FLR_SELECT_BINS * 4;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool do_hist = p.hist0 != nullptr;
    if (do_hist)
        for (int i = tid; i < FLR_SELECT_BINS; i += TMA_THREADS) s_hist[i] = 0u;
    if (tid == 0) {
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(full + s, 1);
            mbar_init(empty + s, TMA_CONSUMER_WARPS);
        }
        fence_mbar_init();
    }
    __syncthreads();

    const int V = p.V, D = p.D, G = p.G;
    const int nchunk = D >> 2;
    const int64_t n_groups = (p.U + G - 1) / G;

    if (warp == TMA_CONSUMER_WARPS) {
        // ===== producer: one thread keeps the ring full =====
        if (lane == 0) {
            int s = 0;
            uint32_t ph = 1;                                  // first pass over the ring never waits
            for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
                mbar_wait(empty + s, ph);
                const int64_t u0 = g * G;
                const uint32_t nu = static_cast<uint32_t>(min(static_cast<int64_t>(G), p.U - u0));
                const uint32_t ab = nu * D * 4u, bb = nu * V * D * 4u;
                unsigned char *dst = ring + static_cast<size_t>(s) * p.stage_bytes;
                mbar_expect_tx(full + s, ab + bb);
                bulk_g2s_stream(dst, p.a + u0 * D, ab, full + s);
                bulk_g2s_stream(dst + p.a_bytes, p.b + u0 * V * D, bb, full + s);
                if (++s == p.stages) { s = 0; ph ^= 1u; }
            }
        }
    } else {
        // ===== consumers: batches of four rows =====
        int s = 0;
        uint32_t ph = 0;
        int64_t it = 0;
        for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x, ++it) {
            mbar_wait(full + s, ph);
            const int64_t u0 = g * G;
            const float *sa = reinterpret_cast<const float *>(ring + static_cast<size_t>(s) * p.stage_bytes);
            const float *sb = reinterpret_cast<const float *>(ring + static_cast<size_t>(s) * p.stage_bytes + p.a_bytes);
            // rotate the batch -> warp map from stage to stage so short stages still use every warp
            const int first = static_cast<int>((warp + TMA_CONSUMER_WARPS - (it % TMA_CONSUMER_WARPS)) % TMA_CONSUMER_WARPS);
            double av[NC][4];
            double na = 0.0;
            if (V == 1) {
                // calibration form (one counterfactual row per anchor): a warp takes one (a, b) pair at a time, so
                // a 16-user stage keeps all sixteen warps busy; its three sums go through one transposed butterfly
                const int nu = static_cast<int>(min(static_cast<int64_t>(G), p.U - u0));
                for (int row = first; row < nu; row += TMA_CONSUMER_WARPS) {
                    const float *arow = sa + static_cast<size_t>(row) * D, *brow = sb + static_cast<size_t>(row) * D;
                    double acc = 0.0, d = 0.0, n = 0.0;
#pragma unroll
                    for (int c = 0; c < NC; ++c) {
                        const int ch = lane + 32 * c;
                        float4 qa = make_float4(0.f, 0.f, 0.f, 0.f), qb = qa;
                        if (FAST || ch < nchunk) {
                            qa = *reinterpret_cast<const float4 *>(arow + 4 * ch);
                            qb = *reinterpret_cast<const float4 *>(brow + 4 * ch);
                        }
                        av[c][0] = qa.x; av[c][1] = qa.y; av[c][2] = qa.z; av[c][3] = qa.w;
#pragma unroll
                        for (int e = 0; e < 4; ++e) acc = fma(av[c][e], av[c][e], acc);
                        fma4(av[c], qb, d, n);
                    }
                    const double v4[4] = {d, n, acc, 0.0};
                    const double y = reduce4_transposed(v4, lane);          // lanes 0-7: dot, 8-15: |b|^2, 16-23: |a|^2
                    const double nb_all = __shfl_sync(0xffffffffu, y, 8), na_all = __shfl_sync(0xffffffffu, y, 16);
                    if (lane == 0) {
                        const float sc = score_from(y, na_all, nb_all);
                        p.out[u0 + row] = sc;
                        if (do_hist) atomicAdd(&s_hist[select_key::key_f32(sc) >> 21], 1u);
                    }
                }
            } else if constexpr (FAST) {
                // batches never straddle users: user slot = batch / pub, rows (batch % pub) * 4 .. + 3
                const int nu = static_cast<int>(min(static_cast<int64_t>(G), p.U - u0));
                const int pub = (V + 3) >> 2;                  // batches per user
                const int n_batches = nu * pub;
                int cur_slot = -1;
                for (int batch = first; batch < n_batches; batch += TMA_CONSUMER_WARPS) {
                    const int slot = batch / pub;
                    const int r0 = (batch - slot * pub) * 4;
                    if (slot != cur_slot) {                    // warp-uniform: this user's anchor row as fp64
                        const float *arow = sa + static_cast<size_t>(slot) * D;
                        double acc = 0.0;
#pragma unroll
                        for (int c = 0; c < NC; ++c) {
                            const float4 q = *reinterpret_cast<const float4 *>(arow + 4 * (lane + 32 * c));
                            av[c][0] = q.x; av[c][1] = q.y; av[c][2] = q.z; av[c][3] = q.w;
#pragma unroll
                            for (int e = 0; e < 4; ++e) acc = fma(av[c][e], av[c][e], acc);
                        }
                        na = acc;                              // this lane's share; summed over the warp with the batch
                        cur_slot = slot;
                    }
                    const float *srow = sb + static_cast<size_t>(slot) * V * D;
                    double x[8];
                    if (r0 + 4 <= V) {                         // full batch: twelve independent 128-bit loads
                        float4 q[4][NC];
#pragma unroll
                        for (int i = 0; i < 4; ++i)
#pragma unroll
                            for (int c = 0; c < NC; ++c)
                                q[i][c] = *reinterpret_cast<const float4 *>(srow + static_cast<size_t>(r0 + i) * D + 4 * (lane + 32 * c));
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            double d = 0.0, n = 0.0;
#pragma unroll
                            for (int c = 0; c < NC; ++c) fma4(av[c], q[i][c], d, n);
                            x[2 * i] = d; x[2 * i + 1] = n;
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            double d = 0.0, n = 0.0;
                            if (r0 + i < V) {
#pragma unroll
                                for (int c = 0; c < NC; ++c)
                                    fma4(av[c], *reinterpret_cast<const float4 *>(srow + static_cast<size_t>(r0 + i) * D + 4 * (lane + 32 * c)), d, n);
                            }
                            x[2 * i] = d; x[2 * i + 1] = n;
                        }
                    }
                    const float sc = reduce_batch_scores_na(x, na, lane);
                    const int r = r0 + (lane >> 3);
                    if ((lane & 7) == 0 && r < V) {
                        p.out[(u0 + slot) * V + r] = sc;
                        if (do_hist) atomicAdd(&s_hist[select_key::key_f32(sc) >> 21], 1u);
                    }
                }
            } else {
                const int nu = static_cast<int>(min(static_cast<int64_t>(G), p.U - u0));
                const int rows = nu * V;
                const int n_batches = (rows + 3) >> 2;
                int cur_g = -1;
                for (int batch = first; batch < n_batches; batch += TMA_CONSUMER_WARPS) {
                    double x[8];
                    double na_row[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int r = batch * 4 + i;
                        x[2 * i] = 0.0; x[2 * i + 1] = 0.0;
                        na_row[i] = 0.0;
                        if (r < rows) {
                            const int gi = r / V;
                            if (gi != cur_g) {                     // warp-uniform
                                load_row_f64<NC>(sa + static_cast<size_t>(gi) * D, lane, nchunk, av);
                                double acc = 0.0;
#pragma unroll
                                for (int c = 0; c < NC; ++c)
#pragma unroll
                                    for (int e = 0; e < 4; ++e) acc = fma(av[c][e], av[c][e], acc);
                                na = warp_sum(acc);
                                cur_g = gi;
                            }
                            na_row[i] = na;
                            const float *brow = sb + static_cast<size_t>(r) * D;
                            double d = 0.0, n = 0.0;
#pragma unroll
                            for (int c = 0; c < NC; ++c) {
                                const int ch = lane + 32 * c;
                                float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
                                if (ch < nchunk) q = *reinterpret_cast<const float4 *>(brow + 4 * ch);
                                fma4(av[c], q, d, n);
                            }
                            x[2 * i] = d; x[2 * i + 1] = n;
                        }
                    }
                    const float sc = reduce_batch_scores(x, na_row, lane);
                    const int r = batch * 4 + (lane >> 3);
                    if ((lane & 7) == 0 && r < rows) {
                        p.out[u0 * V + r] = sc;
                        if (do_hist) atomicAdd(&s_hist[select_key::key_f32(sc) >> 21], 1u);
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty + s);
            if (++s == p.stages) { s = 0; ph ^= 1u; }
        }
    }
    __syncthreads();
    if (do_hist)
        for (int i = tid; i < FLR_SELECT_BINS; i += TMA_THREADS) {
            const unsigned int c = s_hist[i];
            if (c) atomicAdd(p.hist0 + i, static_cast<unsigned long long>(c));
        }
}

// Returns false when the shape does not suit the TMA kernel (caller falls back to the LDG kernel).
template <int NC>
bool launch_cos_tma(const float *a, const float *b, int64_t U, int V, int D, float *out, uint64_t *hist0,
                    cudaStream_t stream, int *rc) {
    const size_t user_bytes = static_cast<size_t>(V + 1) * D * 4;
    const size_t budget = 200 * 1024;                        // ring bytes per SM
    if (user_bytes * 2 > budget) return false;               // fewer than two stages would fit
    // users per stage: half the ring.  Two large stages beat more, smaller ones at the same ring size (at
    // V = 1, D = 384: 4 x 48 KB -> 5.07 TB/s, 2 x 96 KB -> 5.37 TB/s; 4 x 12 KB -> 2.4 TB/s)
    int G = static_cast<int>(std::max<size_t>(1, (budget / 2) / user_bytes));
    auto stage_bytes_for = [&](int g) {
        const size_t ab = (static_cast<size_t>(g) * D * 4 + 127) & ~static_cast<size_t>(127);
        return (ab + static_cast<size_t>(g) * V * D * 4 + 127) & ~static_cast<size_t>(127);
    };
    while (G > 1 && stage_bytes_for(G) * 2 > budget) --G;    // the 128-byte rounding must not cost the second stage
    CosTmaParams p{};
    p.a = a; p.b = b; p.U = U; p.V = V; p.D = D; p.G = G; p.out = out;
    p.hist0 = reinterpret_cast<unsigned long long *>(hist0);
    p.a_bytes = static_cast<uint32_t>((static_cast<size_t>(G) * D * 4 + 127) & ~static_cast<size_t>(127));
    p.stage_bytes = static_cast<uint32_t>((p.a_bytes + static_cast<size_t>(G) * V * D * 4 + 127) & ~static_cast<size_t>(127));
    p.stages = static_cast<int>(std::min<size_t>(TMA_MAX_STAGES, budget / p.stage_bytes));
    if (p.stages < 2) return false;
    const size_t smem = 128 + FLR_SELECT_BINS * 4 + static_cast<size_t>(p.stages) * p.stage_bytes;
    const int64_t n_groups = (U + G - 1) / G;
    const int grid = static_cast<int>(std::min<int64_t>(n_groups, device_sm_count()));
    const bool fast = (D == NC * 128);
    auto kern = fast ? cosine_tma_kernel<NC, true> : cosine_tma_kernel<NC, false>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) {
        set_error("cudaFuncSetAttribute(cosine_tma_kernel) failed: %s", cudaGetErrorString(e));
        *rc = FLR_E_CUDA;
        return true;
    }
    kern<<<grid, TMA_THREADS, smem, stream>>>(p);
    e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("launch of cosine_tma_kernel failed: %s", cudaGetErrorString(e));
        *rc = FLR_E_CUDA;
        return true;
    }
    *rc = FLR_OK;
    return true;
}

template <int NC>
int launch_cos(const float *a, const float *b, int64_t U, int V, int D, float *out, uint64_t *hist0,
               cudaStream_t stream) {
    int rc = FLR_OK;
    if (launch_cos_tma<NC>(a, b, U, V, D, out, hist0, stream, &rc)) return rc;
    const int sms = device_sm_count();
    const int64_t need = (U + COS_WARPS - 1) / COS_WARPS;
    const int grid = static_cast<int>(std::min<int64_t>(need, static_cast<int64_t>(sms) * 8));
    cosine_kernel<NC><<<grid, COS_THREADS, 0, stream>>>(a, b, U, V, D, out,
                                                         reinterpret_cast<unsigned long long *>(hist0));
    FLR_LAUNCH_CHECK("cosine_kernel");
    return FLR_OK;
}

}  // namespace

extern "C" int flr_cosine_nonconformity(const float *a, const float *b, int64_t U, int V, int D, float *out_score,
                                        uint64_t *hist0, void *stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    FLR_REQUIRE(U >= 0 && V >= 1, FLR_E_INVALID, "cosine: need U>=0, V>=1");
    FLR_REQUIRE(D >= 4 && D % 4 == 0 && D <= 1024, FLR_E_UNSUPPORTED, "cosine: D must be a multiple of 4 in [4,1024] (got %d)", D);
    if (U == 0) return FLR_OK;
    FLR_REQUIRE(a && b && out_score, FLR_E_INVALID, "cosine: null pointer");
    FLR_REQUIRE(aligned16(a) && aligned16(b), FLR_E_INVALID, "cosine: embedding rows must be 16-byte aligned");
    switch ((D + 127) / 128) {
        case 1: return launch_cos<1>(a, b, U, V, D, out_score, hist0, stream);
        case 2: return launch_cos<2>(a, b, U, V, D, out_score, hist0, stream);
        case 3: return launch_cos<3>(a, b, U, V, D, out_score, hist0, stream);
        case 4: return launch_cos<4>(a, b, U, V, D, out_score, hist0, stream);
        case 5: return launch_cos<5>(a, b, U, V, D, out_score, hist0, stream);
        case 6: return launch_cos<6>(a, b, U, V, D, out_score, hist0, stream);
        case 7: return launch_cos<7>(a, b, U, V, D, out_score, hist0, stream);
        default: return launch_cos<8>(a, b, U, V, D, out_score, hist0, stream);
    }
}
