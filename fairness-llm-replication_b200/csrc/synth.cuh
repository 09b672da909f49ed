// Synthetic ML-1M-schema generator shared by host and device code.
//
// Everything is integer arithmetic or single correctly-rounded float operations, so the host
// build (g++) and the device build (sm_100a) produce identical bytes.  RNG: Philox4x32-10,
// key = seed, counter = (user lo, user hi, stream, draw index).
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define FLR_HD __host__ __device__ __forceinline__
#else
#define FLR_HD inline
#endif

namespace flr {
namespace synth {

constexpr int kMaxK = 64;

struct U4 {
    uint32_t x, y, z, w;
};

FLR_HD void mulhilo(uint32_t a, uint32_t b, uint32_t &hi, uint32_t &lo) {
    uint64_t p = static_cast<uint64_t>(a) * b;
    hi = static_cast<uint32_t>(p >> 32);
    lo = static_cast<uint32_t>(p);
}

FLR_HD U4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0, lo0, hi1, lo1;
        mulhilo(0xD2511F53u, c0, hi0, lo0);
        mulhilo(0xCD9E8D57u, c2, hi1, lo1);
        uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return U4{c0, c1, c2, c3};
}

// One 24-bit uniform per draw (word 0 of the block; the rest is discarded on purpose: draws
// inside rejection loops stay independent of how many were needed before).
struct Stream {
    uint32_t u_lo, u_hi, stream, k0, k1, n;
    FLR_HD uint32_t next24() {
        U4 r = philox4x32_10(u_lo, u_hi, stream, n++, k0, k1);
        return r.x >> 8;
    }
};

FLR_HD Stream make_stream(uint64_t seed, int64_t user, uint32_t stream) {
    Stream s;
    s.u_lo = static_cast<uint32_t>(static_cast<uint64_t>(user));
    s.u_hi = static_cast<uint32_t>(static_cast<uint64_t>(user) >> 32);
    s.stream = stream;
    s.k0 = static_cast<uint32_t>(seed);
    s.k1 = static_cast<uint32_t>(seed >> 32);
    s.n = 0;
    return s;
}

// Zipf(1.0)-weighted item: inverse CDF by binary search (first index with cdf[i] > u).
FLR_HD int zipf_item(const float *cdf, int vocab, uint32_t r24) {
    float u = static_cast<float>(r24) * 5.9604644775390625e-08f;   // r24 / 2^24, exact
    int lo = 0, hi = vocab - 1;
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (cdf[mid] > u) hi = mid; else lo = mid + 1;
    }
    return lo;
}

FLR_HD bool contains(const int32_t *lst, int n, int x) {
    for (int i = 0; i < n; ++i)
        if (lst[i] == x) return true;
    return false;
}

// Neutral list of user `user`: K distinct Zipf-popular items (rejection on repeats).
FLR_HD void neutral_items(uint64_t seed, int64_t user, int K, int vocab, const float *cdf, int32_t *out) {
    Stream s = make_stream(seed, user, 0u);
    int n = 0, guard = 0;
    while (n < K) {
        int x = zipf_item(cdf, vocab, s.next24());
        if (++guard > 64 * K) {                 // cannot happen for vocab >= 2K; keeps the loop bounded
            x = 0;
            while (contains(out, n, x)) ++x;
        }
        if (!contains(out, n, x)) out[n++] = x;
    }
}

// Shorten 1% of rows (length uniform in [0, K)) and plant a duplicate id in 0.1% of rows.
FLR_HD int mutilate(Stream &s, int32_t *row, int K) {
    int len = K;
    uint32_t a = s.next24(), b = s.next24();
    if (a < 167772u) len = static_cast<int>(b % static_cast<uint32_t>(K));
    uint32_t c = s.next24(), d = s.next24(), e = s.next24();
    if (c < 16777u && len >= 2) {
        int j1 = static_cast<int>(d % static_cast<uint32_t>(len));
        int j2 = static_cast<int>(e % static_cast<uint32_t>(len));
        if (j1 != j2) row[j2] = row[j1];
    }
    for (int j = len; j < K; ++j) row[j] = -1;   // padding must never be read as an id
    return len;
}

// Row of variant v (0-based) or the neutral row (v = -1) of one user.  `clean` receives the
// un-mutilated neutral list (scratch, K ints).
FLR_HD int gen_row(uint64_t seed, int64_t user, int v, int V, int K, int vocab, const float *cdf,
                   int32_t *clean, int32_t *row) {
    neutral_items(seed, user, K, vocab, cdf, clean);
    for (int j = 0; j < K; ++j) row[j] = clean[j];
    Stream s = make_stream(seed, user, static_cast<uint32_t>(v + 1) | 0x10000u);
    if (v >= 0) {
        // replacement probability p_v in [0.1, 0.6], fixed per variant, as a 24-bit threshold
        uint32_t thr = 1677722u + (V > 1 ? static_cast<uint32_t>((8388608ull * static_cast<uint32_t>(v)) /
                                                                  static_cast<uint32_t>(V - 1))
                                          : 0u);
        for (int j = 0; j < K; ++j) {
            if (s.next24() < thr) {
                int x, guard = 0;
                do {
                    x = zipf_item(cdf, vocab, s.next24());
                    if (++guard > 4096) {
                        x = 0;
                        while (contains(row, K, x)) ++x;
                    }
                } while (contains(row, K, x));
                row[j] = x;
            }
        }
        for (int j = 0; j + 1 < K; ++j) {       // adjacent swaps so rank metrics vary
            if (s.next24() < 3355443u) {
                int32_t t = row[j]; row[j] = row[j + 1]; row[j + 1] = t;
            }
        }
    }
    return mutilate(s, row, K);
}

// Approximately N(0,1): centred sum of four 16-bit uniforms (Irwin-Hall), one exact multiply.
FLR_HD float gauss_from(uint32_t w0, uint32_t w1) {
    int s = static_cast<int>((w0 & 0xFFFFu) + (w0 >> 16) + (w1 & 0xFFFFu) + (w1 >> 16)) - 131070;
    return static_cast<float>(s) * 2.6428906e-05f;
}

FLR_HD float sigma_of(int v, int V) {
    int milli = 50 + (V > 1 ? (450 * v) / (V - 1) : 0);
    return static_cast<float>(milli) * 0.001f;
}

// Elements d, d+1 (d even) of the base (stream 0x20000) or noise (0x20001 + v) vector.
FLR_HD void gauss_pair(uint64_t seed, int64_t user, uint32_t stream, int d, float &g0, float &g1) {
    U4 r = philox4x32_10(static_cast<uint32_t>(static_cast<uint64_t>(user)),
                         static_cast<uint32_t>(static_cast<uint64_t>(user) >> 32), stream,
                         static_cast<uint32_t>(d >> 1), static_cast<uint32_t>(seed),
                         static_cast<uint32_t>(seed >> 32));
    g0 = gauss_from(r.x, r.y);
    g1 = gauss_from(r.z, r.w);
}

FLR_HD float mul_add_no_fma(float a, float b, float c) {
#ifdef __CUDA_ARCH__
    return __fadd_rn(__fmul_rn(a, b), c);
#else
    volatile float t = a * b;
    return t + c;
#endif
}

}  // namespace synth
}  // namespace flr
