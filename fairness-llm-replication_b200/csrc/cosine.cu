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

template <int NC>
int launch_cos(const float *a, const float *b, int64_t U, int V, int D, float *out, uint64_t *hist0,
               cudaStream_t stream) {
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
