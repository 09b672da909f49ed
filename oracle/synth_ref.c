/* Synthetic ML-1M-schema inputs for the CPU arms -- TEST INFRASTRUCTURE, not product code.
 *
 * A plain-C restatement of the benchmark's input generator (Philox4x32-10
 * keyed by the seed, counter = (user, stream, draw); Zipf(1.0) item popularity; per-variant
 * replacement probability 0.1 .. 0.6; adjacent swaps; 1 % short rows; 0.1 % rows with a repeated id;
 * embeddings = base + sigma_v * noise from four-term Irwin-Hall normals).  It exists so that the
 * reference arm of bench.py can build its inputs without loading the CUDA library;
 * tests/test_oracle_golden.py checks that it produces the same bytes as the product's host generator.
 *
 * Built by __graft_entry__.build() / oracle/build_ref.py:  gcc -O2 -shared -fPIC -ffp-contract=off
 */
#include <stdint.h>

typedef struct { uint32_t w[4]; } block_t;

static block_t philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
    for (int round = 0; round < 10; ++round) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    block_t b = {{c0, c1, c2, c3}};
    return b;
}

typedef struct { uint64_t seed; int64_t user; uint32_t stream, n; } draws_t;

/* one 24-bit uniform per Philox block (word 0) */
static uint32_t draw24(draws_t *d) {
    block_t b = philox((uint32_t)(uint64_t)d->user, (uint32_t)((uint64_t)d->user >> 32), d->stream, d->n++,
                       (uint32_t)d->seed, (uint32_t)(d->seed >> 32));
    return b.w[0] >> 8;
}

static int zipf_pick(const float *cdf, int vocab, uint32_t r24) {
    float u = (float)r24 * 5.9604644775390625e-08f;      /* r24 / 2^24 */
    int lo = 0, hi = vocab - 1;
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (cdf[mid] > u) hi = mid; else lo = mid + 1;
    }
    return lo;
}

static int holds(const int32_t *a, int n, int x) {
    for (int i = 0; i < n; ++i) if (a[i] == x) return 1;
    return 0;
}

static void neutral_list(uint64_t seed, int64_t user, int K, int vocab, const float *cdf, int32_t *out) {
    draws_t d = {seed, user, 0u, 0u};
    int n = 0, tries = 0;
    while (n < K) {
        int x = zipf_pick(cdf, vocab, draw24(&d));
        if (++tries > 64 * K) { x = 0; while (holds(out, n, x)) ++x; }
        if (!holds(out, n, x)) out[n++] = x;
    }
}

/* row of variant v (v = -1: the neutral row); returns its length */
static int make_row(uint64_t seed, int64_t user, int v, int V, int K, int vocab, const float *cdf, int32_t *row) {
    int32_t base[64];
    neutral_list(seed, user, K, vocab, cdf, base);
    for (int j = 0; j < K; ++j) row[j] = base[j];
    draws_t d = {seed, user, (uint32_t)(v + 1) | 0x10000u, 0u};
    if (v >= 0) {
        uint32_t thr = 1677722u + (V > 1 ? (uint32_t)((8388608ull * (uint32_t)v) / (uint32_t)(V - 1)) : 0u);
        for (int j = 0; j < K; ++j) {
            if (draw24(&d) < thr) {
                int x, tries = 0;
                do {
                    x = zipf_pick(cdf, vocab, draw24(&d));
                    if (++tries > 4096) { x = 0; while (holds(row, K, x)) ++x; }
                } while (holds(row, K, x));
                row[j] = x;
            }
        }
        for (int j = 0; j + 1 < K; ++j)
            if (draw24(&d) < 3355443u) { int32_t t = row[j]; row[j] = row[j + 1]; row[j + 1] = t; }
    }
    int len = K;
    uint32_t a = draw24(&d), b = draw24(&d);
    if (a < 167772u) len = (int)(b % (uint32_t)K);
    uint32_t c = draw24(&d), e = draw24(&d), f = draw24(&d);
    if (c < 16777u && len >= 2) {
        int j1 = (int)(e % (uint32_t)len), j2 = (int)(f % (uint32_t)len);
        if (j1 != j2) row[j2] = row[j1];
    }
    for (int j = len; j < K; ++j) row[j] = -1;
    return len;
}

void synth_ref_zipf_cdf(float *cdf, int vocab) {
    double h = 0.0, c = 0.0;
    for (int i = 0; i < vocab; ++i) h += 1.0 / (i + 1);
    for (int i = 0; i < vocab; ++i) { c += 1.0 / (i + 1); cdf[i] = (float)(c / h); }
    cdf[vocab - 1] = 1.0f;
}

void synth_ref_lists(uint64_t seed, int64_t user0, int64_t U, int V, int K, int vocab, const float *cdf,
                     int32_t *neutral, int32_t *neutral_len, int32_t *lists, int32_t *lens) {
    for (int64_t u = 0; u < U; ++u) {
        neutral_len[u] = make_row(seed, user0 + u, -1, V, K, vocab, cdf, neutral + u * K);
        for (int v = 0; v < V; ++v)
            lens[u * V + v] = make_row(seed, user0 + u, v, V, K, vocab, cdf, lists + (u * V + v) * K);
    }
}

static float normal_from(uint32_t w0, uint32_t w1) {
    int s = (int)((w0 & 0xFFFFu) + (w0 >> 16) + (w1 & 0xFFFFu) + (w1 >> 16)) - 131070;
    return (float)s * 2.6428906e-05f;
}

void synth_ref_embeddings(uint64_t seed, int64_t user0, int64_t U, int V, int D, float *emb_neutral, float *emb) {
    for (int64_t u = 0; u < U; ++u) {
        uint32_t lo = (uint32_t)(uint64_t)(user0 + u), hi = (uint32_t)((uint64_t)(user0 + u) >> 32);
        for (int d = 0; d < D; d += 2) {
            block_t b = philox(lo, hi, 0x20000u, (uint32_t)(d >> 1), (uint32_t)seed, (uint32_t)(seed >> 32));
            float b0 = normal_from(b.w[0], b.w[1]), b1 = normal_from(b.w[2], b.w[3]);
            emb_neutral[u * D + d] = b0;
            emb_neutral[u * D + d + 1] = b1;
            for (int v = 0; v < V; ++v) {
                block_t n = philox(lo, hi, 0x20001u + (uint32_t)v, (uint32_t)(d >> 1), (uint32_t)seed, (uint32_t)(seed >> 32));
                float sigma = (float)(50 + (V > 1 ? (450 * v) / (V - 1) : 0)) * 0.001f;
                volatile float p0 = sigma * normal_from(n.w[0], n.w[1]), p1 = sigma * normal_from(n.w[2], n.w[3]);
                emb[(u * V + v) * D + d] = p0 + b0;
                emb[(u * V + v) * D + d + 1] = p1 + b1;
            }
        }
    }
}
