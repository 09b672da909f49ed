// List similarity of every (user, variant) list against the user's neutral list:
// Jaccard (src/utils.py:224-239), SERP* and PRAG* (oracle/np_oracle.py:list_metrics).
//
// Four kernels share two data structures -- per user a `vocab`-entry byte table that maps item id ->
// rank in the neutral list, per (user, variant) row a private `vocab`-bit set that collapses repeated
// ids exactly as the reference's set() does; PRAG*'s pair count is a running 64-bit mask of the neutral
// ranks already seen:
//   warp kernel     4 <= V <= 128 (V * K >= 80), K <= 64 (the benchmark shapes): one warp per user, one lane per variant,
//                   no block barrier, an optimistic test-and-set main loop with an exact warp-cooperative
//                   redo of the rare rows that hold a repeat or a foreign id (described at its definition);
//                   more than 32 variants run as one launch per chunk of 32;
//   register kernel V = 2, 3 and short rows of few variants, K <= 62: one thread per pair, one CTA per tile of users, rows in
//                   registers, tiles staged by bulk TMA copies one tile ahead;
//   table kernel    K = 63, 64: the same tile structure with the row read from shared memory;
//   compare kernel  everything else (V = 1 pair lists of the string API, vocabularies beyond 16384 ids,
//                   K > 64 Jaccard-only): vocabulary-independent O(K^2) compares straight from global memory.
#include <algorithm>
#include <cstdlib>
#include <mutex>

#include "common.cuh"

using namespace flr;

namespace {

constexpr int NT = 128;                 // threads per CTA (table kernel)
constexpr int kTableSmemBudget = 110 * 1024;   // keeps two CTAs resident per SM

struct LmParams {
    const int32_t *neutral, *neutral_len, *lists, *lens;
    int64_t U;
    int V, K, k_cap, vocab, empty_policy;
    int T, use_bulk, vocab_pad, W;
    int32_t *out_inter, *out_union;
    float *out_jac, *out_serp, *out_prag;
    double *partial;
    int32_t *status;
    int v0, Vt;        // chunked warp kernel only: first variant of this launch's chunk, variants per user in memory
};

struct LmSmem {
    size_t off_lists, off_neutral, off_lens, off_nlen, off_na, off_tab, off_bits, total;
};

__host__ __device__ inline size_t align16(size_t x) { return (x + 15) & ~static_cast<size_t>(15); }

__host__ __device__ inline LmSmem lm_layout(int T, int V, int K, int vocab_pad, int W) {
    LmSmem s;
    size_t o = 16;                                   // mbarrier
    s.off_lists = o;   o = align16(o + static_cast<size_t>(T) * V * K * 4);
    s.off_neutral = o; o = align16(o + static_cast<size_t>(T) * K * 4);
    s.off_lens = o;    o = align16(o + static_cast<size_t>(T) * V * 4);
    s.off_nlen = o;    o = align16(o + static_cast<size_t>(T) * 4);
    s.off_na = o;      o = align16(o + static_cast<size_t>(T) * 4);
    s.off_tab = o;     o = align16(o + static_cast<size_t>(T) * vocab_pad);
    s.off_bits = o;    o = align16(o + static_cast<size_t>(W + 1) * NT * 4);   // +1: per-thread spare word
    s.total = o;
    return s;
}

template <int VEC>
__device__ __forceinline__ void load_ids(const int32_t *p, int (&ids)[VEC]) {
    if constexpr (VEC == 4) {
        int4 q = *reinterpret_cast<const int4 *>(p);
        ids[0] = q.x; ids[1] = q.y; ids[2] = q.z; ids[3] = q.w;
    } else if constexpr (VEC == 2) {
        int2 q = *reinterpret_cast<const int2 *>(p);
        ids[0] = q.x; ids[1] = q.y;
    } else {
        ids[0] = *p;
    }
}

template <int VEC, bool RANK>
__global__ void __launch_bounds__(NT, 2) list_metrics_table_kernel(const LmParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const LmSmem L = lm_layout(p.T, p.V, p.K, p.vocab_pad, p.W);
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem);
    int32_t *s_lists = reinterpret_cast<int32_t *>(smem + L.off_lists);
    int32_t *s_neutral = reinterpret_cast<int32_t *>(smem + L.off_neutral);
    int32_t *s_lens = reinterpret_cast<int32_t *>(smem + L.off_lens);
    int32_t *s_nlen = reinterpret_cast<int32_t *>(smem + L.off_nlen);
    int32_t *s_na = reinterpret_cast<int32_t *>(smem + L.off_na);
    uint8_t *s_tab = smem + L.off_tab;
    uint32_t *s_bits = reinterpret_cast<uint32_t *>(smem + L.off_bits);

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int V = p.V, K = p.K, T = p.T, Kn = p.k_cap;
    const unsigned vocab = static_cast<unsigned>(p.vocab);
    const int pairs_per_tile = T * V;
    const int slot = t / V;
    const bool pair_thread = t < pairs_per_tile;

    {   // one-time init: tables to "absent", private bit sets to empty
        int4 *tab4 = reinterpret_cast<int4 *>(s_tab);
        const int n4 = T * p.vocab_pad / 16;
        for (int i = t; i < n4; i += NT) tab4[i] = make_int4(-1, -1, -1, -1);
        for (int w = 0; w < p.W; ++w) s_bits[w * NT + t] = 0u;
        if (t == 0) {
            mbar_init(bar, 1);
            fence_mbar_init();
        }
    }
    __syncthreads();

    uint32_t phase = 0;
    double acc_j = 0.0;
    long long acc_s = 0, acc_p = 0;
    int acc_n = 0, oov = 0;
    const int64_t n_tiles = (p.U + T - 1) / T;
    const uint32_t bytes_lists = static_cast<uint32_t>(T) * V * K * 4, bytes_neutral = static_cast<uint32_t>(T) * K * 4;
    const uint32_t bytes_lens = static_cast<uint32_t>(T) * V * 4, bytes_nlen = static_cast<uint32_t>(T) * 4;

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t u0 = tile * T;
        const int tu = static_cast<int>(min(static_cast<int64_t>(T), p.U - u0));

        // ---- stage the tile ----
        if (p.use_bulk && tu == T) {
            if (t == 0) {
                fence_proxy_async();
                mbar_expect_tx(bar, bytes_lists + bytes_neutral + bytes_lens + bytes_nlen);
                bulk_g2s(s_lists, p.lists + u0 * V * K, bytes_lists, bar);
                bulk_g2s(s_neutral, p.neutral + u0 * K, bytes_neutral, bar);
                bulk_g2s(s_lens, p.lens + u0 * V, bytes_lens, bar);
                bulk_g2s(s_nlen, p.neutral_len + u0, bytes_nlen, bar);
            }
            mbar_wait(bar, phase);
            phase ^= 1u;
        } else {
            const int32_t *gl = p.lists + u0 * V * K;
            for (int i = t; i < tu * V * K; i += NT) s_lists[i] = gl[i];
            const int32_t *gn = p.neutral + u0 * K;
            for (int i = t; i < tu * K; i += NT) s_neutral[i] = gn[i];
            for (int i = t; i < tu * V; i += NT) s_lens[i] = p.lens[u0 * V + i];
            for (int i = t; i < tu; i += NT) s_nlen[i] = p.neutral_len[u0 + i];
            __syncthreads();
        }

        // ---- rank tables: tab[id] = first position of id in the neutral list ----
        for (int s = warp; s < tu; s += NT / 32) {
            const int Ka = max(0, min(s_nlen[s], Kn));
            const int32_t *A = s_neutral + s * K;
            uint8_t *tab = s_tab + s * p.vocab_pad;
            for (int base = ((Ka - 1) >> 5) << 5; base >= 0; base -= 32) {   // later halves first
                const int j = base + lane;
                if (j < Ka) {
                    const unsigned id = static_cast<unsigned>(A[j]);
                    if (id < vocab) tab[id] = static_cast<uint8_t>(j); else ++oov;
                }
                __syncwarp();
            }
            bool again;
            do {    // duplicates inside one 32-wide step raced: let the smallest position win
                again = false;
                for (int base = 0; base < Ka; base += 32) {
                    const int j = base + lane;
                    if (j < Ka) {
                        const unsigned id = static_cast<unsigned>(A[j]);
                        if (id < vocab && tab[id] > j) { tab[id] = static_cast<uint8_t>(j); again = true; }
                    }
                }
                __syncwarp();
                again = __any_sync(0xffffffffu, again);
            } while (again);
            int cnt = 0;
            for (int base = 0; base < Ka; base += 32) {
                const int j = base + lane;
                bool first = false;
                if (j < Ka) {
                    const unsigned id = static_cast<unsigned>(A[j]);
                    first = id < vocab && tab[id] == j;
                }
                cnt += __popc(__ballot_sync(0xffffffffu, first));
            }
            if (lane == 0) s_na[s] = cnt;
        }
        __syncthreads();

        // ---- one thread per pair ----
        if (pair_thread && slot < tu) {
            const int Kb = max(0, min(s_lens[t], Kn));
            const int32_t *row = s_lists + t * K;
            const uint8_t *tab = s_tab + slot * p.vocab_pad;
            uint32_t *mybits = s_bits + t;
            uint64_t seen = 0;
            int n_b = 0, inter = 0, serp = 0, prag = 0;
            for (int j0 = 0; j0 < Kb; j0 += VEC) {
                int ids[VEC];
                load_ids<VEC>(row + j0, ids);
#pragma unroll
                for (int e = 0; e < VEC; ++e) {
                    const int j = j0 + e;
                    const unsigned id = static_cast<unsigned>(ids[e]);
                    if (j < Kb) {
                        if (id < vocab) {
                            uint32_t *wp = mybits + (id >> 5) * NT;
                            const uint32_t w = *wp, b = 1u << (id & 31u);
                            if (!(w & b)) {
                                *wp = w | b;
                                ++n_b;
                                const unsigned r = tab[id];
                                if (r != 0xFFu) {
                                    ++inter;
                                    if constexpr (RANK) {
                                        serp += Kn - j;
                                        prag += __popcll(seen & ((1ull << r) - 1ull));
                                        seen |= 1ull << r;
                                    }
                                } else if constexpr (RANK) {
                                    prag += __popcll(seen);   // r = +inf: every earlier hit outranks it
                                }
                            }
                        } else {
                            ++oov;
                        }
                    }
                }
            }
            const int n_a = s_na[slot];
            const int uni = n_a + n_b - inter;
            double jac;
            if (p.empty_policy == FLR_EMPTY_BOTH_IS_ONE)
                jac = uni > 0 ? static_cast<double>(inter) / static_cast<double>(uni) : 1.0;
            else
                jac = (n_a == 0 || n_b == 0) ? 0.0 : static_cast<double>(inter) / static_cast<double>(uni);
            const int64_t pair = u0 * V + t;
            if (p.out_inter) p.out_inter[pair] = inter;
            if (p.out_union) p.out_union[pair] = uni;
            if (p.out_jac) p.out_jac[pair] = static_cast<float>(jac);
            if constexpr (RANK) {
                if (p.out_serp) p.out_serp[pair] = static_cast<float>(serp) / static_cast<float>(Kn * (Kn + 1) / 2);
                if (p.out_prag) p.out_prag[pair] = static_cast<float>(prag) / static_cast<float>(Kn * (Kn + 1));
            }
            acc_j += jac;
            acc_s += serp;
            acc_p += prag;
            acc_n += 1;
            // wipe the private bit set for the next tile
            for (int j0 = 0; j0 < Kb; j0 += VEC) {
                int ids[VEC];
                load_ids<VEC>(row + j0, ids);
#pragma unroll
                for (int e = 0; e < VEC; ++e) {
                    const unsigned id = static_cast<unsigned>(ids[e]);
                    if (j0 + e < Kb && id < vocab) mybits[(id >> 5) * NT] = 0u;
                }
            }
        }
        __syncthreads();

        // ---- un-mark the rank tables ----
        for (int s = warp; s < tu; s += NT / 32) {
            const int Ka = max(0, min(s_nlen[s], Kn));
            const int32_t *A = s_neutral + s * K;
            uint8_t *tab = s_tab + s * p.vocab_pad;
            for (int j = lane; j < Ka; j += 32) {
                const unsigned id = static_cast<unsigned>(A[j]);
                if (id < vocab) tab[id] = 0xFFu;
            }
        }
        __syncthreads();
    }

    double *dst = p.partial + (static_cast<size_t>(blockIdx.x) * NT + t) * 4;
    dst[0] = acc_j;
    dst[1] = static_cast<double>(acc_s);
    dst[2] = static_cast<double>(acc_p);
    dst[3] = static_cast<double>(acc_n);
    if (oov && p.status) atomicAdd(p.status, oov);
}

// ---------------------------------------------------------------------------------------
// Register kernel: the table kernel with a compile-time row capacity KT in {10, 20, 32, 50, 62}; a
// row of K <= KT ids is padded with the spare id (K = 10, 20, 50 -- the benchmark shapes -- are exact).  Same data structures and results; what changes is the
// schedule:
//   * each thread copies its row from the staged tile into registers and the CTA immediately
//     re-arms the bulk TMA copy of its NEXT tile into the same buffer, so the HBM read of tile
//     i+1 overlaps all the work on tile i;
//   * all KT rank look-ups are issued back to back (independent loads) and packed beside the
//     ids; the only serial chain left per element is the private bit set's load -> store;
//   * every loop over the row is fully unrolled.
// ---------------------------------------------------------------------------------------
template <int KT, bool RANK>
__global__ void __launch_bounds__(NT, 2) list_metrics_reg_kernel(const LmParams p) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int K = p.K;                                   // row length in memory, K <= KT (the unroll capacity)
    const LmSmem L = lm_layout(p.T, p.V, K, p.vocab_pad, p.W);
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem);
    int32_t *s_lists = reinterpret_cast<int32_t *>(smem + L.off_lists);
    int32_t *s_neutral = reinterpret_cast<int32_t *>(smem + L.off_neutral);
    int32_t *s_lens = reinterpret_cast<int32_t *>(smem + L.off_lens);
    int32_t *s_nlen = reinterpret_cast<int32_t *>(smem + L.off_nlen);
    int32_t *s_na = reinterpret_cast<int32_t *>(smem + L.off_na);
    uint8_t *s_tab = smem + L.off_tab;
    uint32_t *s_bits = reinterpret_cast<uint32_t *>(smem + L.off_bits);
    int32_t *s_keep = reinterpret_cast<int32_t *>(smem + L.total);          // [T][KT] neutral ids of the live tile (vocab = none)
    double *s_recip = reinterpret_cast<double *>(smem + L.total + align16(static_cast<size_t>(p.T) * (KT * 4 + 4)));  // [2 KT + 1]

    constexpr int VEC = (KT % 4 == 0) ? 4 : 2;
    static_assert(KT % 2 == 0 && KT <= 62, "row length must be even; rank 63 is the absent marker");

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int V = p.V, T = p.T, Kn = p.k_cap;
    const unsigned vocab = static_cast<unsigned>(p.vocab);
    const int pairs_per_tile = T * V;
    const int slot = t / V;
    const bool pair_thread = t < pairs_per_tile;

    {
        int4 *tab4 = reinterpret_cast<int4 *>(s_tab);
        const int n4 = T * p.vocab_pad / 16;
        for (int i = t; i < n4; i += NT) tab4[i] = make_int4(0x3F3F3F3F, 0x3F3F3F3F, 0x3F3F3F3F, 0x3F3F3F3F);   // 63 = absent
        for (int w = 0; w < p.W; ++w) s_bits[w * NT + t] = 0u;
        s_bits[(vocab >> 5) * NT + t] = 1u << (vocab & 31u);  // the spare element `vocab` is always "seen before"
        for (int i = t; i <= 2 * KT; i += NT) s_recip[i] = i > 0 ? 1.0 / static_cast<double>(i) : 0.0;
        if (t == 0) {
            mbar_init(bar, 1);
            fence_mbar_init();
        }
    }
    __syncthreads();

    const int64_t n_tiles = (p.U + T - 1) / T;
    const uint32_t bytes_lists = static_cast<uint32_t>(T) * V * K * 4, bytes_neutral = static_cast<uint32_t>(T) * K * 4;
    const uint32_t bytes_lens = static_cast<uint32_t>(T) * V * 4, bytes_nlen = static_cast<uint32_t>(T) * 4;
    auto issue_tma = [&](int64_t tile) {       // thread 0 only; the stage buffer must be free
        const int64_t u0 = tile * T;
        fence_proxy_async();
        mbar_expect_tx(bar, bytes_lists + bytes_neutral + bytes_lens + bytes_nlen);
        bulk_g2s(s_lists, p.lists + u0 * V * K, bytes_lists, bar);
        bulk_g2s(s_neutral, p.neutral + u0 * K, bytes_neutral, bar);
        bulk_g2s(s_lens, p.lens + u0 * V, bytes_lens, bar);
        bulk_g2s(s_nlen, p.neutral_len + u0, bytes_nlen, bar);
    };
    auto tile_is_bulk = [&](int64_t tile) { return p.use_bulk && tile < n_tiles && (tile + 1) * T <= p.U; };

    const float inv_serp = 1.0f / static_cast<float>(Kn * (Kn + 1) / 2), inv_prag = 1.0f / static_cast<float>(Kn * (Kn + 1));
    uint32_t phase = 0;
    double acc_j = 0.0;
    long long acc_s = 0, acc_p = 0;
    int acc_n = 0, oov = 0;
    if (t == 0 && tile_is_bulk(blockIdx.x)) issue_tma(blockIdx.x);

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t u0 = tile * T;
        const int tu = static_cast<int>(min(static_cast<int64_t>(T), p.U - u0));
        if (tile_is_bulk(tile)) {
            mbar_wait(bar, phase);
            phase ^= 1u;
        } else {
            const int32_t *gl = p.lists + u0 * V * K;
            for (int i = t; i < tu * V * K; i += NT) s_lists[i] = gl[i];
            const int32_t *gn = p.neutral + u0 * K;
            for (int i = t; i < tu * K; i += NT) s_neutral[i] = gn[i];
            for (int i = t; i < tu * V; i += NT) s_lens[i] = p.lens[u0 * V + i];
            for (int i = t; i < tu; i += NT) s_nlen[i] = p.neutral_len[u0 + i];
            __syncthreads();
        }

        // ---- move the tile out of the stage buffer: own row -> registers, neutral lists -> s_keep ----
        unsigned ids[KT];
        int Kb = K;
        const bool live = pair_thread && slot < tu;
        if (live) {
            Kb = max(0, min(s_lens[t], Kn));
            const int32_t *row = s_lists + t * K;
            if (K == KT) {                                   // warp-uniform: vector loads of the whole row
#pragma unroll
                for (int j0 = 0; j0 < KT; j0 += VEC) {
                    int tmp[VEC];
                    load_ids<VEC>(row + j0, tmp);
#pragma unroll
                    for (int e = 0; e < VEC; ++e) ids[j0 + e] = static_cast<unsigned>(tmp[e]);
                }
            } else {                                         // shorter rows in memory: positions >= K are spare
#pragma unroll
                for (int j = 0; j < KT; ++j) ids[j] = j < K ? static_cast<unsigned>(row[j]) : vocab;
            }
        }
        // Rows shorter than KT are rare: only a warp that holds one pays for masking the tail.  A
        // masked position holds the id `vocab`, the spare element (see the main loop).
        if (__any_sync(0xffffffffu, Kb < K)) {
#pragma unroll
            for (int j = 0; j < KT; ++j)
                if (j >= Kb) ids[j] = vocab;
        }
        // ---- rank tables: tab[id] = first position of id in the neutral list (63 = absent).  Slot s is
        //      always handled by warp s % 4 -- build here, un-mark at the end of the tile -- so the two
        //      need no barrier between them; the ids are parked in s_keep for the un-mark. ----
        for (int s = warp; s < tu; s += NT / 32) {
            const int Ka = max(0, min(s_nlen[s], Kn));
            const int32_t *A = s_neutral + s * K;
            uint8_t *tab = s_tab + s * p.vocab_pad;
            const unsigned id0 = lane < Ka ? static_cast<unsigned>(A[lane]) : vocab;
            const unsigned id1 = lane + 32 < Ka ? static_cast<unsigned>(A[lane + 32]) : vocab;
            if (lane < KT) s_keep[s * KT + lane] = static_cast<int32_t>(id0);
            if (lane + 32 < KT) s_keep[s * KT + lane + 32] = static_cast<int32_t>(id1);
            oov += (lane < Ka && id0 >= vocab ? 1 : 0) + (lane + 32 < Ka && id1 >= vocab ? 1 : 0);
            if (id1 < vocab) tab[id1] = static_cast<uint8_t>(lane + 32);         // later half first,
            __syncwarp();
            if (id0 < vocab) tab[id0] = static_cast<uint8_t>(lane);              // so the earlier half wins
            __syncwarp();
            int cnt;
            bool again;
            do {    // repeats inside one 32-wide store raced: let the smallest position win, then count
                const unsigned v0 = id0 < vocab ? tab[id0] : 0u, v1 = id1 < vocab ? tab[id1] : 0u;
                again = false;
                if (id0 < vocab && v0 > static_cast<unsigned>(lane)) { tab[id0] = static_cast<uint8_t>(lane); again = true; }
                if (id1 < vocab && v1 > static_cast<unsigned>(lane + 32)) { tab[id1] = static_cast<uint8_t>(lane + 32); again = true; }
                cnt = __popc(__ballot_sync(0xffffffffu, id0 < vocab && v0 == static_cast<unsigned>(lane))) +
                      __popc(__ballot_sync(0xffffffffu, id1 < vocab && v1 == static_cast<unsigned>(lane + 32)));
                again = __any_sync(0xffffffffu, again);
                __syncwarp();
            } while (again);
            if (lane == 0) s_na[s] = cnt;
        }
        __syncthreads();       // stage buffer fully consumed, every table built
        if (t == 0 && tile_is_bulk(tile + gridDim.x)) issue_tma(tile + gridDim.x);   // overlaps everything below

        if (live) {
            const uint8_t *tab = s_tab + slot * p.vocab_pad;
            // One fused, branch-free pass over the row (fully unrolled, so the compiler hoists the
            // independent rank look-ups ahead of the bit set's load -> store chain).
            //   * The id `vocab` is a spare element: its bit in this thread's set is always 1 ("seen
            //     before") and its table entry always reads "absent".  min(id, vocab) sends padding
            //     (-1 or anything >= vocab as unsigned), masked tail positions and out-of-vocabulary ids
            //     there, so they contribute nothing and need no test.
            //   * An absent item has rank 63: "hits ranked below it" then counts every hit so far, and
            //     2 << 63 adds no bit, so the rank mask update needs no condition (a repeated present
            //     item sets a bit that is already set).
            // seen2 = (mask of neutral ranks already hit) << 1; hitpos = positions of the hits.
            char *bits_base = reinterpret_cast<char *>(s_bits + t);
            uint64_t seen2 = 0, hitpos = 0;
            int n_b = 0, prag = 0;
#pragma unroll
            for (int j = 0; j < KT; ++j) {
                const unsigned idx = min(ids[j], vocab);
                uint32_t *wp = reinterpret_cast<uint32_t *>(bits_base + (idx >> 5) * (NT * 4));
                const uint32_t w = *wp;
                const uint32_t bit = 1u << (idx & 31u);
                *wp = w | bit;
                const bool first = (w & bit) == 0u;
                if (first) ++n_b;
                const unsigned rr = tab[idx];                  // 63 when absent
                __builtin_assume(rr <= 63u);                   // lets the compiler drop a mask before the 64-bit shifts
                if constexpr (RANK) {
                    const int below = __popcll(seen2 << (rr ^ 63u));
                    if (first) prag += below;
                    if (first && rr != 63u) hitpos |= 1ull << j;
                }
                seen2 |= 2ull << rr;
            }
            const int inter = __popcll(seen2);
            int serp = 0;
            if constexpr (RANK) {       // sum over hits of (Kn - j), from the position mask
                const uint32_t lo = static_cast<uint32_t>(hitpos), hi = static_cast<uint32_t>(hitpos >> 32);
                int pos_sum = 32 * __popc(hi);
#pragma unroll
                for (int bshift = 0; bshift < 5; ++bshift) {
                    const uint32_t m = bshift == 0 ? 0xAAAAAAAAu : bshift == 1 ? 0xCCCCCCCCu : bshift == 2 ? 0xF0F0F0F0u
                                     : bshift == 3 ? 0xFF00FF00u : 0xFFFF0000u;
                    pos_sum += (__popc(lo & m) + __popc(hi & m)) << bshift;
                }
                serp = inter * Kn - pos_sum;
            }
            if (n_b != Kb) {            // repeats or foreign ids in this row (rare): count the foreign ids
                int spare_hits = 0;
#pragma unroll
                for (int j = 0; j < KT; ++j) spare_hits += ids[j] >= vocab ? 1 : 0;
                oov += spare_hits - (KT - Kb);                      // positions >= Kb all hold the spare id
            }
            const int n_a = s_na[slot];
            const int uni = n_a + n_b - inter;
            // inter * (1/uni) from a table instead of an fp64 division per pair: within one ulp of the
            // quotient, which the fp32 output and the 1e-12 tolerance on the sums never see
            double jac = static_cast<double>(inter) * s_recip[uni];
            if (p.empty_policy == FLR_EMPTY_BOTH_IS_ONE) jac = uni > 0 ? jac : 1.0;
            const int64_t pair = u0 * V + t;
            if (p.out_inter) p.out_inter[pair] = inter;
            if (p.out_union) p.out_union[pair] = uni;
            if (p.out_jac) p.out_jac[pair] = static_cast<float>(jac);
            if constexpr (RANK) {
                if (p.out_serp) p.out_serp[pair] = static_cast<float>(serp) * inv_serp;
                if (p.out_prag) p.out_prag[pair] = static_cast<float>(prag) * inv_prag;
            }
            acc_j += jac;
            acc_s += serp;
            acc_p += prag;
            acc_n += 1;
        }
        __syncthreads();

        // ---- un-set the bits this thread set (every one lives in the word of one of its ids: K word stores
        //      instead of a sweep over all the bit sets), re-arm the spare element, un-mark the rank tables ----
        if (live) {
            char *bits_base = reinterpret_cast<char *>(s_bits + t);
#pragma unroll
            for (int j = 0; j < KT; ++j)
                *reinterpret_cast<uint32_t *>(bits_base + (min(ids[j], vocab) >> 5) * (NT * 4)) = 0u;
            s_bits[(vocab >> 5) * NT + t] = 1u << (vocab & 31u);
        }
        for (int s = warp; s < tu; s += NT / 32) {
            const int32_t *A = s_keep + s * KT;
            uint8_t *tab = s_tab + s * p.vocab_pad;
            for (int j = lane; j < KT; j += 32) {
                const unsigned id = static_cast<unsigned>(A[j]);
                if (id < vocab) tab[id] = 63u;
            }
            __syncwarp();      // every lane's un-mark lands before any lane of this warp re-marks the table
        }
        // no barrier here: a thread's bit set is its own, and each table is rebuilt by the warp that just
        // un-marked it
    }

    double *dst = p.partial + (static_cast<size_t>(blockIdx.x) * NT + t) * 4;
    dst[0] = acc_j;
    dst[1] = static_cast<double>(acc_s);
    dst[2] = static_cast<double>(acc_p);
    dst[3] = static_cast<double>(acc_n);
    if (oov && p.status) atomicAdd(p.status, oov);
}

// ---------------------------------------------------------------------------------------
// Warp kernel (the fast path for 17 <= V <= 128, K <= 64; V > 32 in chunks of 32): one warp per user, one lane per variant,
// and no block-wide barrier anywhere -- every warp of a CTA runs its own pipeline over its own
// shared-memory region (stage buffer, rank table, bit sets, mbarrier), so while one warp builds a
// table or writes results another is in its main loop.
//
// Per user the warp
//   1. waits for the bulk TMA copy of the user's V x K id block (issued one user ahead), and every
//      lane copies its row into registers, clamping foreign ids to `vocab`; lane 0 then re-arms the
//      copy for the warp's next user;
//   2. builds the rank table of the neutral list (two positions per lane), t = 63 - rank, 0 = absent;
//   3. runs the optimistic main loop.  Nearly every row is "clean": no id repeats inside it and every
//      id lies in [0, vocab).  For a clean row every element is a first occurrence, so the exact
//      loop's per-element conditions disappear and what is left per element is
//        * one shared-memory atomicOr on the lane's private bit set -- it returns the old word, so
//          "seen before" is (old & bit), OR-ed into one flag per row; unlike load -> or -> store,
//          consecutive atomics have no ordering hazard, so a row's elements are independent
//          (profiles/micro/smem_rmw_bench.cu measures both forms);
//        * one byte look-up of t: the number of earlier hits ranked below is popc(seen << t)
//          (absent: every hit), the hit's own bit is 1 << (64 - t), which is 0 for an absent item;
//        * a multiply-add for the SERP* position sum;
//   4. redoes flagged rows (a repeat, or a foreign id: `vocab` stands for those and its bit is always
//      set) with the exact sequential loop -- rare, one lane at a time, from a copy of the row in
//      shared memory;
//   5. writes the pair's results, un-sets the bits it set (one word store per id) and un-marks the table.
// Positions past a row's length hold spare ids (vocab + 1 + position): private bits, table entries
// that read "absent"; each one counts as a miss that follows every hit, which is subtracted at the end.
// ---------------------------------------------------------------------------------------
// Shared-memory address of the word of a lane's bit set that holds bit `idx` (sets are laid out [word][lane],
// base = the lane's word 0): written as a shift and a multiply-add so that it stays two instructions (the
// compiler's own form, (idx * 4 & ~127) + base, is three).
__device__ __forceinline__ uint32_t word_addr32(uint32_t base, unsigned idx) {
    uint32_t a;
    asm("{\n\t.reg .u32 w;\n\tshr.u32 w, %2, 5;\n\tmad.lo.u32 %0, w, 128, %1;\n\t}" : "=r"(a) : "r"(base), "r"(idx));
    return a;
}

__device__ __forceinline__ uint64_t shl64_clamped(uint64_t v, unsigned amount) {   // amount may be 64 (-> 0)
    uint64_t r;
    asm("shl.b64 %0, %1, %2;" : "=l"(r) : "l"(v), "r"(amount));
    return r;
}

__device__ __forceinline__ uint64_t shr64_clamped(uint64_t v, unsigned amount) {   // amount may be >= 64 (-> 0)
    uint64_t r;
    asm("shr.b64 %0, %1, %2;" : "=l"(r) : "l"(v), "r"(amount));
    return r;
}

__device__ __forceinline__ uint32_t shl32_clamped(uint32_t v, unsigned amount) {   // amount may be >= 32 (-> 0)
    uint32_t r;
    asm("shl.b32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(amount));
    return r;
}
__device__ __forceinline__ uint32_t shr32_clamped(uint32_t v, unsigned amount) {
    uint32_t r;
    asm("shr.b32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(amount));
    return r;
}

constexpr int WK_WARPS = NT / 32;

struct WkSmem {
    size_t off_recip, warp0, off_stage, off_tab, off_bits, off_row, warp_bytes, total;
};

// Words between two rows of a staged user block.  Lane l reads row l, so the stride decides the bank
// conflicts of the row copy: 64-bit loads are conflict-free for K = 2 (mod 4) (half a warp covers 32 banks),
// 128-bit loads for a stride of 4 (mod 8) (a quarter warp does); K = 0 (mod 8) -- K = 32, 64: every lane on
// the same banks -- gets four words of padding per row and is copied row by row (from K = 32 up: below that a
// row is under 128 bytes, too small a bulk copy to be worth it, and the conflicts are 2- to 4-way only).
__host__ __device__ inline int wk_row_stride(int K) { return (K % 8 == 0 && K >= 32) ? K + 4 : K; }

__host__ __device__ inline WkSmem wk_layout(int V, int K, int KT, int vocab) {
    WkSmem s;
    s.off_recip = 0;
    s.warp0 = align16((2 * KT + 1) * sizeof(double));
    size_t o = 16;                                                    // mbarrier
    s.off_stage = o; o = align16(o + static_cast<size_t>(V) * wk_row_stride(K) * 4 + 16);   // + 16: the copy starts at a 16-byte boundary
    s.off_tab = o;   o = align16(o + static_cast<size_t>(vocab) + 65);          // ids up to vocab + 64 (the spare of position 63) are looked up
    s.off_bits = o;  o = align16(o + (static_cast<size_t>(vocab >> 5) + 3) * 128);
    s.off_row = o;   o = align16(o + static_cast<size_t>(KT) * 4);
    s.warp_bytes = (o + 127) & ~static_cast<size_t>(127);
    s.total = s.warp0 + s.warp_bytes * WK_WARPS;
    return s;
}

// CHUNKED: users with more than 32 variants are scored in launches of up to 32 variants each -- this launch
// takes variants [p.v0, p.v0 + p.V) of the p.Vt every user holds.  A compile-time flag, so that the common
// one-chunk kernels carry none of the extra address arithmetic.
template <int KT, bool RANK, bool CHUNKED>
__global__ void __launch_bounds__(NT, 2) list_metrics_warp_kernel(const LmParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr unsigned FULL = 0xffffffffu;
    // Table entry of a neutral item of rank r: TOP - r; of an id that is not in the neutral list: TOP + 1.
    // Rows of at most 32 items keep their rank masks in one 32-bit word.
    constexpr bool NARROW = KT <= 32;
    constexpr unsigned TOP = NARROW ? 31u : 63u, ABSENT = TOP + 1u, ABSENT_SHIFT = NARROW ? 5u : 6u;
    constexpr int ABSENT4 = static_cast<int>(ABSENT * 0x01010101u);
    static_assert(KT % 2 == 0 && KT <= 64, "row length must be even; ranks must fit the 64-bit mask");
    const int K = p.K, V = p.V, Kn = p.k_cap;
    const int Vt = CHUNKED ? p.Vt : V, v0 = CHUNKED ? p.v0 : 0;
    const unsigned vocab = static_cast<unsigned>(p.vocab);
    const WkSmem L = wk_layout(V, K, KT, p.vocab);
    const int t = threadIdx.x, lane = t & 31;
    const int warp = __shfl_sync(FULL, t >> 5, 0);       // tells the compiler the value is warp-uniform
    double *s_recip = reinterpret_cast<double *>(smem + L.off_recip);
    unsigned char *wbase = smem + L.warp0 + L.warp_bytes * warp;
    uint64_t *bar = reinterpret_cast<uint64_t *>(wbase);
    unsigned char *s_stage = wbase + L.off_stage;
    uint8_t *tab = wbase + L.off_tab;
    uint32_t *s_bits = reinterpret_cast<uint32_t *>(wbase + L.off_bits);      // [word][lane]
    uint32_t *s_row = reinterpret_cast<uint32_t *>(wbase + L.off_row);
    const int tab_bytes16 = static_cast<int>((L.off_bits - L.off_tab) / 16);

    // Id space of the private bit sets: [0, vocab) real ids; `vocab` stands for every foreign id and its
    // bit is always set; vocab + 1 + j is the spare of row position j.  Words own_row .. own_row + 2 hold
    // the foreign bit and the spares.
    const int own_row = static_cast<int>(vocab >> 5);
    const uint32_t foreign_bit = 1u << (vocab & 31u);

    for (int i = t; i <= 2 * KT; i += NT) s_recip[i] = i > 0 ? 1.0 / static_cast<double>(i) : 0.0;
    for (int i = lane; i < tab_bytes16; i += 32)
        reinterpret_cast<int4 *>(tab)[i] = make_int4(ABSENT4, ABSENT4, ABSENT4, ABSENT4);
    for (int w = 0; w < own_row + 3; ++w) s_bits[w * 32 + lane] = w == own_row ? foreign_bit : 0u;
    if (lane == 0) {
        mbar_init(bar, 1);
        fence_mbar_init();
    }
    __syncthreads();          // the only block-wide barrier: the reciprocal table is shared

    const int64_t n_warps = static_cast<int64_t>(gridDim.x) * WK_WARPS;
    const bool live = lane < V;
    const int idle_row = live ? lane : (lane & 15) < V ? (lane & 15) : 0;     // the staged row a lane copies out
    const int rs = wk_row_stride(K);                     // words between staged rows
    const bool padded = rs != K, vec4 = (K & 3) == 0;    // padded rows are copied one by one, each by its lane
    // user_bytes: what one launch copies per user; a user's rows are user_stride apart in memory
    const int64_t user_bytes = static_cast<int64_t>(K) * 4 * V, user_stride = static_cast<int64_t>(K) * 4 * Vt;
    const int64_t total_bytes = user_stride * p.U;
    const char *lists_b = reinterpret_cast<const char *>(p.lists) + static_cast<int64_t>(v0) * K * 4;
    // bulk copy of user u's block: from the 16-byte boundary at or below it, a multiple of 16 bytes
    auto bulk_ok = [&](int64_t uu) {
        if (!p.use_bulk || uu >= p.U) return false;
        if (padded) return true;                 // K % 8 == 0: every row starts and ends on a 16-byte boundary
        const int64_t off = uu * user_stride, a = off & ~15LL;       // relative to lists_b (itself 16-byte aligned)
        return a + ((off - a + user_bytes + 15) & ~15LL) <= total_bytes - static_cast<int64_t>(v0) * K * 4;
    };
    auto issue_tma = [&](int64_t uu) {          // whole warp; the stage buffer must be free
        const int64_t off = uu * user_stride;
        if (padded) {
            if (lane == 0) {
                fence_proxy_async();
                mbar_expect_tx(bar, static_cast<uint32_t>(user_bytes));
            }
            __syncwarp();
            if (live) bulk_g2s(s_stage + static_cast<size_t>(lane) * rs * 4, lists_b + off + static_cast<int64_t>(lane) * K * 4,
                               static_cast<uint32_t>(K) * 4u, bar);
        } else if (lane == 0) {
            const int64_t a = off & ~15LL;
            const uint32_t bytes = static_cast<uint32_t>((off - a + user_bytes + 15) & ~15LL);
            fence_proxy_async();
            mbar_expect_tx(bar, bytes);
            bulk_g2s(s_stage, lists_b + a, bytes, bar);
        }
    };

    const float inv_serp = 1.0f / static_cast<float>(Kn * (Kn + 1) / 2), inv_prag = 1.0f / static_cast<float>(Kn * (Kn + 1));
    uint32_t phase = 0;
    double acc_j = 0.0;
    long long acc_s = 0, acc_p = 0;
    int acc_n = 0, oov = 0;

    int64_t u = static_cast<int64_t>(blockIdx.x) * WK_WARPS + warp;
    // The small per-user inputs travel in registers, fetched one user ahead -- after the last use of the
    // current ones, so that waiting for the current values never waits for the loads just issued.
    unsigned nid0 = vocab, nid1 = vocab;
    int nlen = 0, blen = 0;
    auto fetch_small = [&](int64_t uu, unsigned &a0, unsigned &a1, int &nl, int &bl) {
        if (uu < p.U) {
            const int32_t *A = p.neutral + uu * K;
            a0 = lane < K ? static_cast<unsigned>(A[lane]) : vocab;
            a1 = lane + 32 < K ? static_cast<unsigned>(A[lane + 32]) : vocab;
            nl = p.neutral_len[uu];
            bl = live ? p.lens[uu * Vt + v0 + lane] : K;
        }
    };
    fetch_small(u, nid0, nid1, nlen, blen);
    bool bulk_cur = bulk_ok(u);
    if (bulk_cur) issue_tma(u);

    for (; u < p.U; u += n_warps) {
        const int64_t un = u + n_warps;
        const bool bulk_next = bulk_ok(un);

        // ---- stage -> registers ----
        int delta = 0;
        if (bulk_cur) {
            delta = static_cast<int>((u * user_stride) & 15);
            mbar_wait(bar, phase);
            phase ^= 1u;
        } else {
            const int32_t *gl = p.lists + (u * Vt + v0) * K;
            for (int i = lane; i < V * K; i += 32) reinterpret_cast<int32_t *>(s_stage)[(i / K) * rs + i % K] = gl[i];
            __syncwarp();
        }
        unsigned ids[KT];
        const int Kb = live ? max(0, min(blen, Kn)) : K;       // an idle lane redoes another lane's row; nothing of it is kept
        {
            // An idle lane reads the row sixteen lanes down: rows 16 apart start on the same banks (the stride is
            // even), and that row is read by the other half of the warp -- no bank conflict in the idle lane's own
            // half.  (Row 0, the obvious choice, sits on lane 16's banks: one more wavefront per load for V = 31.)
            const int32_t *row = reinterpret_cast<const int32_t *>(s_stage + delta) + idle_row * rs;
            if (vec4) {                              // K % 4 == 0: rows are 16-byte aligned, whole 128-bit chunks
#pragma unroll
                for (int j0 = 0; j0 < KT; j0 += 4) {
                    int4 q = make_int4(0, 0, 0, 0);
                    if (j0 < K) q = *reinterpret_cast<const int4 *>(row + j0);
                    const int qq[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                        if (j0 + e < KT) ids[j0 + e] = j0 < K ? min(static_cast<unsigned>(qq[e]), vocab) : vocab + 1u + j0 + e;
                }
            } else if (K == KT) {
#pragma unroll
                for (int j0 = 0; j0 < KT; j0 += 2) {
                    const int2 q = *reinterpret_cast<const int2 *>(row + j0);
                    ids[j0] = min(static_cast<unsigned>(q.x), vocab);
                    ids[j0 + 1] = min(static_cast<unsigned>(q.y), vocab);
                }
            } else {
#pragma unroll
                for (int j = 0; j < KT; ++j) ids[j] = j < K ? min(static_cast<unsigned>(row[j]), vocab) : vocab + 1u + j;
            }
        }
        if (__any_sync(FULL, Kb < K)) {      // short rows are rare: only their warps mask the tail
            unsigned spare0 = vocab + 1u;
            asm volatile("" : "+r"(spare0));         // keeps the KT spare ids out of registers outside this block
#pragma unroll
            for (int j = 0; j < KT; ++j)
                if (j >= Kb) ids[j] = spare0 + j;
        }
        __syncwarp();
        if (bulk_next) issue_tma(un);                   // overlaps everything below

        // ---- rank table: tab[id] = 63 - (first position of id in the neutral list), 64 = absent ----
        int n_a;
        unsigned un_pack;       // the two table entries this lane marks, 16 bits each, kept for the un-mark at the end
        {
            const int Ka = max(0, min(nlen, Kn));
            const unsigned id0 = lane < Ka ? nid0 : vocab, id1 = lane + 32 < Ka ? nid1 : vocab;
            un_pack = min(id0, vocab) | (min(id1, vocab) << 16);        // foreign -> entry `vocab`, which reads absent anyway
            if (v0 == 0) oov += (lane < Ka && id0 >= vocab ? 1 : 0) + (lane + 32 < Ka && id1 >= vocab ? 1 : 0);   // once per user
            const unsigned t0 = TOP - lane, t1 = TOP - 32u - lane;      // t of positions lane and lane + 32 (t1 unused when NARROW)
            if (id1 < vocab) tab[id1] = static_cast<uint8_t>(t1);       // later half first,
            __syncwarp();
            if (id0 < vocab) tab[id0] = static_cast<uint8_t>(t0);       // so the earlier half wins
            __syncwarp();
            bool again;
            do {    // repeats inside one 32-wide store raced: let the first position (largest t) win, then count
                const unsigned v0 = id0 < vocab ? tab[id0] : 255u, v1 = id1 < vocab ? tab[id1] : 255u;
                again = false;
                if (id0 < vocab && v0 < t0) { tab[id0] = static_cast<uint8_t>(t0); again = true; }
                if (id1 < vocab && v1 < t1) { tab[id1] = static_cast<uint8_t>(t1); again = true; }
                n_a = __popc(__ballot_sync(FULL, id0 < vocab && v0 == t0)) + __popc(__ballot_sync(FULL, id1 < vocab && v1 == t1));
                again = __any_sync(FULL, again);
                __syncwarp();
            } while (again);
        }
        // Next user's small inputs, used one iteration from now.  Their current values are dead here (the un-mark
        // keeps its own packed copy), so the loads land in the loop-carried registers themselves: with a copy of
        // the old ids still alive, the compiler loaded into temporaries and moved them over at the end of this
        // iteration -- a wait for the loads half an iteration after issuing them (12 % of the K = 20 kernel).
        fetch_small(un, nid0, nid1, nlen, blen);

        // ---- optimistic pass over the row ----
        const uint32_t bits32 = smem_u32(s_bits + lane);
        const uint32_t tab32 = smem_u32(tab);
        int below_hi = 0;                       // the two halves' popcounts go to two sums: adds the FMA pipe can take
        uint32_t seen_lo = 0, seen_hi = 0;      // bit TOP - r = neutral rank r was hit
        uint32_t repeat = 0, absent_pos = 0;
        uint32_t tpack = 0;
        int below = 0;
#pragma unroll
        for (int j = 0; j < KT; ++j) {
            const unsigned idx = ids[j];                   // in [0, vocab], or a spare
            const uint32_t bit = 1u << (idx & 31u);
            uint32_t old;                                  // the word's address: a shift and a multiply-add
            asm volatile("atom.shared.or.b32 %0, [%1], %2;" : "=r"(old) : "r"(word_addr32(bits32, idx)), "r"(bit) : "memory");
            repeat |= old & bit;
            unsigned tt;                                   // TOP - rank, TOP + 1 when absent
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(tt) : "r"(tab32 + idx));   // one 32-bit base: a two-input add
            if constexpr (RANK) {
                // four table bytes per register (a multiply-add each: the bytes are disjoint), then one masked
                // byte dot product with the positions as weights: ABSENT * sum of the absent positions
                tpack = (j & 3) == 0 ? tt : tt * (1u << (8 * (j & 3))) + tpack;
                if ((j & 3) == 3 || j == KT - 1) {
                    const unsigned j0 = static_cast<unsigned>(j & ~3);
                    const unsigned wts = j0 | ((j0 + 1) << 8) | ((j0 + 2) << 16) | ((j0 + 3) << 24);
                    absent_pos = __dp4a(tpack & (ABSENT * 0x01010101u), wts, absent_pos);
                }
            }
            if constexpr (NARROW) {
                if constexpr (RANK) below += __popc(shr32_clamped(seen_lo, tt));    // hits so far ranked before this one (absent: 0)
                seen_lo += shl32_clamped(1u, tt);                                    // absent: 0; bits of a clean row are distinct: + is |
            } else {
                if constexpr (RANK) {
                    const uint64_t seen = (static_cast<uint64_t>(seen_hi) << 32) | seen_lo;
                    const uint64_t sh = shr64_clamped(seen, tt);     // hits so far ranked before this one (absent: 0)
                    below += __popc(static_cast<uint32_t>(sh));
                    below_hi += __popc(static_cast<uint32_t>(sh >> 32));
                }
                const uint64_t own = shl64_clamped(1ull, tt);
                seen_lo += static_cast<uint32_t>(own);
                seen_hi += static_cast<uint32_t>(own >> 32);
            }
        }
        // A clean row of Kb elements with hits at positions p_1 < ... < p_h: a miss counts the hits before
        // it, which adds up to h (Kb - 1) - sum p_i - h (h - 1) / 2; spares past the end count nothing.
        int inter = __popc(seen_lo) + __popc(seen_hi);
        below += below_hi;
        absent_pos >>= ABSENT_SHIFT;
        const int pos_sum = KT * (KT - 1) / 2 - static_cast<int>(absent_pos);
        int serp = inter * Kn - pos_sum;
        int prag = below + inter * (Kb - 1) - pos_sum - inter * (inter - 1) / 2;
        int n_b = Kb;

        // ---- flagged rows, one at a time, redone exactly by the whole warp: the owner parks its row in
        //      shared memory, lane l takes positions l and l + 32.  First occurrences come from a shuffle
        //      sweep over the row; the hits ranked below an element are a prefix-OR of one-hot rank masks
        //      over the earlier positions. ----
        unsigned flagged = __ballot_sync(FULL, live && repeat != 0u);
        while (flagged) {
            const int owner = __ffs(flagged) - 1;
            flagged &= flagged - 1u;
            if (lane == owner) {
#pragma unroll
                for (int j = 0; j < KT; ++j) s_row[j] = ids[j];
            }
            const int Ko = __shfl_sync(FULL, Kb, owner);
            __syncwarp();
            const unsigned id_a = s_row[lane < KT ? lane : 0];
            const unsigned id_b = s_row[lane + 32 < KT ? lane + 32 : 0];
            const bool in_a = lane < Ko, in_b = KT > 32 && lane + 32 < Ko;
            bool dup_a = false, dup_b = false;
            for (int k = 0; k < 32; ++k) {           // every lane sees every id of the row
                const unsigned xa = __shfl_sync(FULL, id_a, k);
                dup_a |= k < lane && xa == id_a;
                if constexpr (KT > 32) {
                    const unsigned xb = __shfl_sync(FULL, id_b, k);
                    dup_b |= xa == id_b || (k < lane && xb == id_b);      // position k + 32 < Ko whenever lane + 32 is
                }
            }
            const bool first_a = in_a && id_a < vocab && !dup_a;
            const bool first_b = in_b && id_b < vocab && !dup_b;
            const int foreign_n = __popc(__ballot_sync(FULL, in_a && id_a >= vocab)) +
                                  __popc(__ballot_sync(FULL, in_b && id_b >= vocab));
            const unsigned t_a = first_a ? tab[id_a] : ABSENT, t_b = first_b ? tab[id_b] : ABSENT;
            const bool hit_a = t_a != ABSENT, hit_b = t_b != ABSENT;
            const uint64_t m_a = hit_a ? 1ull << (TOP - t_a) : 0ull;        // one-hot neutral ranks
            const uint64_t m_b = hit_b ? 1ull << (TOP - t_b) : 0ull;
            uint64_t inc_a = m_a, inc_b = m_b;                               // inclusive prefix-ORs over lanes
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const uint64_t ya = __shfl_up_sync(FULL, inc_a, off), yb = __shfl_up_sync(FULL, inc_b, off);
                if (lane >= off) { inc_a |= ya; inc_b |= yb; }
            }
            uint64_t before_a = __shfl_up_sync(FULL, inc_a, 1), before_b = __shfl_up_sync(FULL, inc_b, 1);
            if (lane == 0) { before_a = 0; before_b = 0; }
            before_b |= __shfl_sync(FULL, inc_a, 31);
            const int c_nb = __popc(__ballot_sync(FULL, first_a)) + __popc(__ballot_sync(FULL, first_b));
            const int c_in = __popc(__ballot_sync(FULL, hit_a)) + __popc(__ballot_sync(FULL, hit_b));
            int c_serp = (hit_a ? Kn - lane : 0) + (hit_b ? Kn - lane - 32 : 0);
            int c_prag = (first_a ? __popcll(hit_a ? before_a & (m_a - 1ull) : before_a) : 0) +
                         (first_b ? __popcll(hit_b ? before_b & (m_b - 1ull) : before_b) : 0);
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) {
                c_serp += __shfl_xor_sync(FULL, c_serp, off);
                c_prag += __shfl_xor_sync(FULL, c_prag, off);
            }
            if (lane == owner) {
                n_b = c_nb; inter = c_in; serp = c_serp; prag = c_prag;
                oov += foreign_n;
            }
            __syncwarp();
        }

        if (live) {
            const int uni = n_a + n_b - inter;
            // inter * (1/uni) from a table instead of an fp64 division per pair: within one ulp of the
            // quotient, which the fp32 output and the 1e-12 tolerance on the sums never see
            double jac = static_cast<double>(inter) * s_recip[uni];
            if (p.empty_policy == FLR_EMPTY_BOTH_IS_ONE) jac = uni > 0 ? jac : 1.0;
            const int64_t pair = u * Vt + v0 + lane;
            if (p.out_inter) p.out_inter[pair] = inter;
            if (p.out_union) p.out_union[pair] = uni;
            if (p.out_jac) p.out_jac[pair] = static_cast<float>(jac);
            if constexpr (RANK) {
                if (p.out_serp) p.out_serp[pair] = static_cast<float>(serp) * inv_serp;
                if (p.out_prag) p.out_prag[pair] = static_cast<float>(prag) * inv_prag;
            }
            acc_j += jac;
            acc_s += serp;
            acc_p += prag;
            acc_n += 1;
        }

        // ---- un-set the bit sets and un-mark the rank table.  Every bit a lane set lives in the word of one of
        //      its ids, so 50 conflict-free word stores per lane clear them all -- less than half the shared-
        //      memory traffic of sweeping all 118 rows, which is what bounded the kernel ----
        __syncwarp();
        {
#pragma unroll
            for (int j = 0; j < KT; ++j) asm volatile("st.shared.u32 [%0], %1;" :: "r"(word_addr32(bits32, ids[j])), "r"(0u) : "memory");
            s_bits[own_row * 32 + lane] = foreign_bit;
        }
        tab[un_pack & 0xffffu] = static_cast<uint8_t>(ABSENT);
        tab[un_pack >> 16] = static_cast<uint8_t>(ABSENT);
        __syncwarp();
        bulk_cur = bulk_next;
    }

    double *dst = p.partial + (static_cast<size_t>(blockIdx.x) * NT + t) * 4;
    dst[0] = acc_j;
    dst[1] = static_cast<double>(acc_s);
    dst[2] = static_cast<double>(acc_p);
    dst[3] = static_cast<double>(acc_n);
    if (oov && p.status) atomicAdd(p.status, oov);
}

// Fixed-order reduction of the warp kernel's per-thread partials: variant v lives in lane v of every
// warp; one warp per variant adds the (block, warp) entries l, l + 32, ... in ascending order, then an
// xor butterfly -- the order depends only on the launch geometry.
__global__ void __launch_bounds__(128) list_metrics_warp_finalize_kernel(const double *__restrict__ partial, int grid,
                                                                         int V, double *__restrict__ out) {
    const int v = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (v >= V) return;
    double a[4] = {0, 0, 0, 0};
    const int n = grid * WK_WARPS;
    for (int e = lane; e < n; e += 32) {
        const double4 q = *reinterpret_cast<const double4 *>(partial + (static_cast<size_t>(e) * 32 + v) * 4);
        a[0] += q.x; a[1] += q.y; a[2] += q.z; a[3] += q.w;
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1)
#pragma unroll
        for (int k = 0; k < 4; ++k) a[k] += __shfl_xor_sync(0xffffffffu, a[k], off);
    if (lane == 0) { out[v * 4 + 0] = a[0]; out[v * 4 + 1] = a[1]; out[v * 4 + 2] = a[2]; out[v * 4 + 3] = a[3]; }
}

// Fixed-order reduction of the per-thread partials, one warp per variant: lane l adds the
// (block, slot) entries l, l+32, ... in ascending order, then an xor butterfly -- the order
// depends only on the launch geometry, so the sums are bit-reproducible.
__global__ void __launch_bounds__(128) list_metrics_finalize_kernel(const double *__restrict__ partial, int grid,
                                                                    int threads, int T, int V,
                                                                    double *__restrict__ out) {
    const int v = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (v >= V) return;
    double a[4] = {0, 0, 0, 0};
    const int n = grid * T;
    for (int e = lane; e < n; e += 32) {
        const int b = e / T, sl = e - b * T;
        const double4 *q = reinterpret_cast<const double4 *>(partial + (static_cast<size_t>(b) * threads + sl * V + v) * 4);
        const double4 lo = *q;
        a[0] += lo.x; a[1] += lo.y; a[2] += lo.z; a[3] += lo.w;
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1)
#pragma unroll
        for (int k = 0; k < 4; ++k) a[k] += __shfl_xor_sync(0xffffffffu, a[k], off);
    if (lane == 0) { out[v * 4 + 0] = a[0]; out[v * 4 + 1] = a[1]; out[v * 4 + 2] = a[2]; out[v * 4 + 3] = a[3]; }
}

// ---------------------------------------------------------------------------------------
// Compare kernel: thread per pair, rows read straight from global memory.
// ---------------------------------------------------------------------------------------
constexpr int CMP_THREADS = 128;

template <bool RANK>
__global__ void __launch_bounds__(CMP_THREADS) list_metrics_cmp_kernel(const LmParams p) {
    const int V = p.V, K = p.K, Kn = p.k_cap;
    const int64_t total = p.U * V;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;   // multiple of V by construction
    const int64_t first = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    double acc_j = 0.0;
    long long acc_s = 0, acc_p = 0;
    int acc_n = 0, oov = 0;
    const bool check_vocab = p.vocab > 0;
    const unsigned vocab = static_cast<unsigned>(p.vocab);
    for (int64_t pair = first; pair < total; pair += stride) {
        const int64_t u = pair / V;
        const int32_t *A = p.neutral + u * K;
        const int32_t *B = p.lists + pair * K;
        const int Ka = max(0, min(p.neutral_len[u], Kn));
        const int Kb = max(0, min(p.lens[pair], Kn));
        int n_a = 0;
        for (int i = 0; i < Ka; ++i) {
            const int x = A[i];
            if (check_vocab && static_cast<unsigned>(x) >= vocab) { oov += (pair % V == 0); continue; }   // once per user, like the table kernels
            bool dup = false;
            for (int i2 = 0; i2 < i; ++i2) dup |= (A[i2] == x);
            n_a += !dup;
        }
        uint64_t seen = 0;
        int n_b = 0, inter = 0, serp = 0, prag = 0;
        for (int j = 0; j < Kb; ++j) {
            const int x = B[j];
            if (check_vocab && static_cast<unsigned>(x) >= vocab) { ++oov; continue; }
            bool dup = false;
            for (int j2 = 0; j2 < j; ++j2) dup |= (B[j2] == x);
            if (dup) continue;
            ++n_b;
            int r = -1;
            for (int i = Ka - 1; i >= 0; --i) r = (A[i] == x) ? i : r;      // first occurrence
            if (r >= 0) {
                ++inter;
                if constexpr (RANK) {
                    serp += Kn - j;
                    prag += __popcll(seen & ((1ull << r) - 1ull));
                    seen |= 1ull << r;
                }
            } else if constexpr (RANK) {
                prag += __popcll(seen);
            }
        }
        const int uni = n_a + n_b - inter;
        double jac;
        if (p.empty_policy == FLR_EMPTY_BOTH_IS_ONE)
            jac = uni > 0 ? static_cast<double>(inter) / static_cast<double>(uni) : 1.0;
        else
            jac = (n_a == 0 || n_b == 0) ? 0.0 : static_cast<double>(inter) / static_cast<double>(uni);
        if (p.out_inter) p.out_inter[pair] = inter;
        if (p.out_union) p.out_union[pair] = uni;
        if (p.out_jac) p.out_jac[pair] = static_cast<float>(jac);
        if constexpr (RANK) {
            if (p.out_serp) p.out_serp[pair] = static_cast<float>(serp) / static_cast<float>(Kn * (Kn + 1) / 2);
            if (p.out_prag) p.out_prag[pair] = static_cast<float>(prag) / static_cast<float>(Kn * (Kn + 1));
        }
        acc_j += jac; acc_s += serp; acc_p += prag; acc_n += 1;
    }
    double *dst = p.partial + static_cast<size_t>(first) * 4;
    dst[0] = acc_j;
    dst[1] = static_cast<double>(acc_s);
    dst[2] = static_cast<double>(acc_p);
    dst[3] = static_cast<double>(acc_n);
    if (oov && p.status) atomicAdd(p.status, oov);
}

// Block v sums the partials of all threads whose first pair has variant v (index = v mod V):
// thread t takes every 256th of them in ascending order, then a fixed shared-memory tree.
__global__ void __launch_bounds__(256) list_metrics_cmp_finalize_kernel(const double *__restrict__ partial,
                                                                        int64_t n_threads, int V,
                                                                        double *__restrict__ out) {
    __shared__ double s_acc[256][4];
    const int v = blockIdx.x, t = threadIdx.x;
    double a[4] = {0, 0, 0, 0};
    for (int64_t i = v + static_cast<int64_t>(V) * t; i < n_threads; i += static_cast<int64_t>(V) * 256) {
        const double *q = partial + i * 4;
        a[0] += q[0]; a[1] += q[1]; a[2] += q[2]; a[3] += q[3];
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) s_acc[t][k] = a[k];
    __syncthreads();
    for (int half = 128; half >= 1; half >>= 1) {
        if (t < half)
#pragma unroll
            for (int k = 0; k < 4; ++k) s_acc[t][k] += s_acc[t + half][k];
        __syncthreads();
    }
    if (t < 4) out[v * 4 + t] = s_acc[0][t];
}

struct Plan {
    bool table;        // table / register / test-and-set kernel (else compare kernel)
    int reg_k;         // unroll capacity of the register kernels (>= K), 0 = generic table kernel (K = 63, 64)
    bool warp;         // warp kernel (one warp per user) instead of the register kernel
    int chunks;        // warp kernel: launches per call, one per 32 variants
    int T, vocab_pad, W, grid;
    size_t smem;
    int64_t cmp_threads;
};

Plan make_plan(int64_t U, int V, int K, int k_cap, int vocab) {
    Plan pl{};
    const int sms = device_sm_count();
    pl.vocab_pad = (vocab + 16) & ~15;       // > vocab: entry [vocab] is a spare that always reads "absent"
    pl.W = (vocab + 31) / 32;
    pl.table = false;
    if (vocab > 0 && vocab <= 16384 && k_cap <= FLR_MAX_RANK_K && V >= 2 && V <= NT) {
        const int cap = K <= 10 ? 10 : K <= 20 ? 20 : K <= 32 ? 32 : K <= 50 ? 50 : K <= 62 ? 62 : 0;   // unroll capacity (register kernel)
        const int wcap = K > 64 ? 0 : K > 50 ? 64 : cap;                                                 // ... of the warp kernel: K <= 64
        const bool reg = cap != 0;
        const char *pick = getenv("FLR_LM_KERNEL");
        // the warp kernel's four per-warp regions must fit one SM (vocabularies up to ~9000 ids do; with more
        // than half of it only one CTA is resident per SM); larger vocabularies fall through to the tile
        // kernels and from there to the compare kernel
        const int Vw = std::min(V, 32);          // variants per launch of the warp kernel
        auto plan_warp = [&]() {
            pl.table = true;
            pl.warp = true;
            pl.chunks = (V + 31) / 32;
            pl.reg_k = wcap;
            pl.T = WK_WARPS;
            pl.smem = wk_layout(Vw, K, wcap, vocab).total;
            const int64_t tiles = (U + WK_WARPS - 1) / WK_WARPS;
            pl.grid = static_cast<int>(std::min<int64_t>(std::max<int64_t>(tiles, 1), 2LL * sms));
        };
        // more than 32 variants: one launch per chunk of 32 (a chunk starts 128 * K bytes into the user's block,
        // so the bulk copies' alignment arithmetic is that of the one-chunk case)
        const bool warp_fits = wcap != 0 && !(pick && pick[0] == 'r') &&
                               wk_layout(Vw, K, wcap, vocab).total <= static_cast<size_t>(226 * 1024);
        // Fewer than 17 variants leave lanes idle, but a warp's time per user hardly depends on V, while the tile
        // kernels' grows with the pairs: measured, the warp kernel is ahead from four variants of 20 items up
        // (V=4, K=20: 1.7x; V=16, K=50: 1.8x; V=8..16, K=20: 5-8 %), behind at V=3, K=10 (0.7x).
        // FLR_LM_KERNEL=w / =r force one or the other for A/B timing.
        if (warp_fits && (V >= 17 || (V >= 4 && V * K >= 80) || (pick && pick[0] == 'w'))) {
            plan_warp();
            return pl;
        }
        // register kernel extras: neutral ids + lengths of the live tile, and the reciprocal table
        auto bytes = [&](int T) {
            return lm_layout(T, V, K, pl.vocab_pad, pl.W).total +
                   (reg ? align16(static_cast<size_t>(T) * (cap * 4 + 4)) + align16((2 * cap + 1) * sizeof(double)) : 0);
        };
        int T = NT / V;
        if (T >= 4) T &= ~3;                 // multiples of 4 keep every bulk copy 16-byte sized
        while (T >= 1 && bytes(T) > static_cast<size_t>(kTableSmemBudget))
            T = (T > 4) ? ((T - 1) & ~3) : T - 1;
        if (T >= 1 && T * V * 2 >= NT) {      // at least half of the CTA's threads own a pair
            pl.table = true;
            pl.reg_k = cap;
            pl.T = T;
            pl.smem = bytes(T);
            const int64_t tiles = (U + T - 1) / T;
            pl.grid = static_cast<int>(std::min<int64_t>(std::max<int64_t>(tiles, 1), 2LL * sms));
        } else if (warp_fits) {
            // few variants with long rows: no tile of users fits the tile kernels' shared memory; a warp per
            // user with V of its 32 lanes busy is still far ahead of the O(K^2) compare kernel
            plan_warp();
        }
    }
    if (!pl.table) {
        // grid*threads is a multiple of V so a thread keeps one variant over its whole stride
        int64_t want = std::min<int64_t>(std::max<int64_t>(U * V, 1), 16LL * sms * CMP_THREADS);
        int64_t blocks = (want + CMP_THREADS - 1) / CMP_THREADS;
        blocks = ((blocks + V - 1) / V) * V;
        pl.grid = static_cast<int>(blocks);
        pl.cmp_threads = blocks * CMP_THREADS;
    }
    return pl;
}

// Bytes of per-thread partial sums the planned launch writes (4 doubles per thread).
size_t partial_bytes(const Plan &pl) {
    return (pl.table ? static_cast<size_t>(pl.grid) * NT : static_cast<size_t>(pl.cmp_threads)) * 4 * sizeof(double);
}

using TableKernel = void (*)(const LmParams);

TableKernel pick_table_kernel(const Plan &pl, int K, bool rank) {
    if (pl.warp) {
#define FLR_WARP_CASE(KT_) case KT_: return pl.chunks > 1 ? (rank ? list_metrics_warp_kernel<KT_, true, true> : list_metrics_warp_kernel<KT_, false, true>) \
                                                          : (rank ? list_metrics_warp_kernel<KT_, true, false> : list_metrics_warp_kernel<KT_, false, false>);
        switch (pl.reg_k) {
            FLR_WARP_CASE(10) FLR_WARP_CASE(20) FLR_WARP_CASE(32) FLR_WARP_CASE(50) FLR_WARP_CASE(64)
            default: break;
        }
#undef FLR_WARP_CASE
    }
    switch (pl.reg_k) {
        case 10: return rank ? list_metrics_reg_kernel<10, true> : list_metrics_reg_kernel<10, false>;
        case 20: return rank ? list_metrics_reg_kernel<20, true> : list_metrics_reg_kernel<20, false>;
        case 32: return rank ? list_metrics_reg_kernel<32, true> : list_metrics_reg_kernel<32, false>;
        case 50: return rank ? list_metrics_reg_kernel<50, true> : list_metrics_reg_kernel<50, false>;
        case 62: return rank ? list_metrics_reg_kernel<62, true> : list_metrics_reg_kernel<62, false>;
        default: break;
    }
    const int vec = (K % 4 == 0) ? 4 : (K % 2 == 0) ? 2 : 1;
    if (rank) return vec == 4 ? list_metrics_table_kernel<4, true> : vec == 2 ? list_metrics_table_kernel<2, true> : list_metrics_table_kernel<1, true>;
    return vec == 4 ? list_metrics_table_kernel<4, false> : vec == 2 ? list_metrics_table_kernel<2, false> : list_metrics_table_kernel<1, false>;
}

int check_args(const char *who, const int32_t *neutral, const int32_t *neutral_len, const int32_t *lists,
               const int32_t *lens, int64_t U, int V, int K, int &k_cap, int empty_policy, int rank_metrics) {
    FLR_REQUIRE(U >= 0 && V >= 1 && K >= 1, FLR_E_INVALID, "%s: need U>=0, V>=1, K>=1", who);
    FLR_REQUIRE(k_cap >= 1, FLR_E_INVALID, "%s: k_cap must be >= 1", who);
    FLR_REQUIRE(empty_policy == FLR_EMPTY_ANY_IS_ZERO || empty_policy == FLR_EMPTY_BOTH_IS_ONE, FLR_E_INVALID,
                "%s: unknown empty_policy %d", who, empty_policy);
    if (k_cap > K) k_cap = K;
    FLR_REQUIRE(!rank_metrics || k_cap <= FLR_MAX_RANK_K, FLR_E_UNSUPPORTED,
                "%s: SERP*/PRAG* need k_cap <= %d (got %d)", who, FLR_MAX_RANK_K, k_cap);
    FLR_REQUIRE(U * V < (1LL << 40), FLR_E_UNSUPPORTED, "%s: too many pairs", who);
    FLR_REQUIRE(U == 0 || (neutral && neutral_len && lists && lens), FLR_E_INVALID, "%s: null input pointer", who);
    return FLR_OK;
}

// Launches the planned kernels; `partial` must hold partial_bytes(pl).
int launch_planned(const Plan &pl, LmParams p, int rank_metrics, double *out_variant_sums, cudaStream_t stream) {
    p.vocab_pad = pl.vocab_pad;
    p.W = pl.W;
    if (pl.table) {
        p.T = pl.T;
        p.use_bulk = pl.warp ? aligned16(p.lists)
                             : (pl.T % 4 == 0) && aligned16(p.neutral) && aligned16(p.neutral_len) && aligned16(p.lists) &&
                                   aligned16(p.lens);
        TableKernel kern = pick_table_kernel(pl, p.K, rank_metrics != 0);
        FLR_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(pl.smem)));
        if (pl.warp) {
            // one launch per chunk of 32 variants (one launch in all for V <= 32); the chunks reuse the partials
            const int Vt = p.V;
            for (int c = 0; c < pl.chunks; ++c) {
                p.Vt = Vt;
                p.v0 = 32 * c;
                p.V = std::min(32, Vt - p.v0);
                kern<<<pl.grid, NT, pl.smem, stream>>>(p);
                FLR_LAUNCH_CHECK("list_metrics_warp_kernel");
                if (out_variant_sums) {
                    list_metrics_warp_finalize_kernel<<<(p.V + 3) / 4, 128, 0, stream>>>(p.partial, pl.grid, p.V,
                                                                                         out_variant_sums + 4 * p.v0);
                    FLR_LAUNCH_CHECK("list_metrics_warp_finalize_kernel");
                }
            }
            return FLR_OK;
        }
        kern<<<pl.grid, NT, pl.smem, stream>>>(p);
        FLR_LAUNCH_CHECK("list_metrics table/register kernel");
        if (out_variant_sums) {
            list_metrics_finalize_kernel<<<(p.V + 3) / 4, 128, 0, stream>>>(p.partial, pl.grid, NT, pl.T, p.V,
                                                                             out_variant_sums);
            FLR_LAUNCH_CHECK("list_metrics_finalize_kernel");
        }
    } else {
        if (rank_metrics) list_metrics_cmp_kernel<true><<<pl.grid, CMP_THREADS, 0, stream>>>(p);
        else              list_metrics_cmp_kernel<false><<<pl.grid, CMP_THREADS, 0, stream>>>(p);
        FLR_LAUNCH_CHECK("list_metrics_cmp_kernel");
        if (out_variant_sums) {
            list_metrics_cmp_finalize_kernel<<<p.V, 256, 0, stream>>>(p.partial, pl.cmp_threads, p.V, out_variant_sums);
            FLR_LAUNCH_CHECK("list_metrics_cmp_finalize_kernel");
        }
    }
    return FLR_OK;
}

LmParams make_params(const int32_t *neutral, const int32_t *neutral_len, const int32_t *lists, const int32_t *lens,
                     int64_t U, int V, int K, int k_cap, int vocab, int empty_policy, int32_t *out_inter,
                     int32_t *out_union, float *out_jac, float *out_serp, float *out_prag, int32_t *out_status,
                     double *partial) {
    LmParams p{};
    p.neutral = neutral; p.neutral_len = neutral_len; p.lists = lists; p.lens = lens;
    p.U = U; p.V = V; p.K = K; p.k_cap = k_cap; p.vocab = vocab; p.empty_policy = empty_policy;
    p.out_inter = out_inter; p.out_union = out_union; p.out_jac = out_jac;
    p.out_serp = out_serp; p.out_prag = out_prag;
    p.partial = partial;
    p.status = out_status;
    return p;
}

}  // namespace

extern "C" size_t flr_list_metrics_scratch_bytes(int64_t U, int V, int K, int vocab) {
    if (U < 0 || V < 1 || K < 1) return 0;
    // covers both kernels for any k_cap: table partials (<= 2*SMs CTAs) or compare partials
    const int sms = device_sm_count();
    size_t table = static_cast<size_t>(2) * sms * NT * 4 * sizeof(double);
    int64_t want = std::min<int64_t>(std::max<int64_t>(U * V, 1), 16LL * sms * CMP_THREADS);
    int64_t blocks = (want + CMP_THREADS - 1) / CMP_THREADS;
    blocks = ((blocks + V - 1) / V) * V;
    size_t cmp = static_cast<size_t>(blocks) * CMP_THREADS * 4 * sizeof(double);
    (void)K; (void)vocab;
    return std::max(table, cmp);
}

extern "C" int flr_list_metrics(const int32_t *neutral, const int32_t *neutral_len, const int32_t *lists,
                                const int32_t *lens, int64_t U, int V, int K, int k_cap, int vocab,
                                int empty_policy, int rank_metrics, int32_t *out_inter, int32_t *out_union,
                                float *out_jac, float *out_serp, float *out_prag, double *out_variant_sums,
                                int32_t *out_status, void *scratch, size_t scratch_bytes, void *stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int rc = check_args("list_metrics", neutral, neutral_len, lists, lens, U, V, K, k_cap, empty_policy,
                              rank_metrics);
    if (rc != FLR_OK) return rc;
    if (U == 0) {
        if (out_variant_sums) FLR_CUDA(cudaMemsetAsync(out_variant_sums, 0, sizeof(double) * 4 * V, stream));
        return FLR_OK;
    }
    FLR_REQUIRE(scratch && scratch_bytes >= flr_list_metrics_scratch_bytes(U, V, K, vocab), FLR_E_SCRATCH,
                "list_metrics: scratch too small (%zu < %zu)", scratch_bytes,
                flr_list_metrics_scratch_bytes(U, V, K, vocab));
    const Plan pl = make_plan(U, V, K, k_cap, vocab);
    return launch_planned(pl, make_params(neutral, neutral_len, lists, lens, U, V, K, k_cap, vocab, empty_policy,
                                          out_inter, out_union, out_jac, out_serp, out_prag, out_status,
                                          static_cast<double *>(scratch)),
                          rank_metrics, out_variant_sums, stream);
}

// ---------------------------------------------------------------------------------------
// flr_list_metrics_static: the same computation for callers that cannot supply scratch.
//
// Who that is: a process whose GPU memory was planned by someone else -- a serving worker
// whose HBM was handed to one pool up front, a job that sized its user shard to fill the
// device, a step loop that must never run into cudaMalloc's device-wide synchronisation.
// Such a caller owns its input and output buffers but has no way to come up with one more
// buffer for the per-thread partial sums.  Here those partials live in the fixed array
// below, which is part of the library image: every GPU that loads the library gives up
// FLR_LIST_METRICS_STATIC_SCRATCH_BYTES for as long as the process lives.
//
// In host code `g_lm_static_scratch` is only a handle; the address the kernels need comes
// from cudaGetSymbolAddress, looked up once per device, with that device current, by the
// first call made on it.
// ---------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_lm_static_scratch[FLR_LIST_METRICS_STATIC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
double *g_static_ptr[kMaxDevices];     // device address of g_lm_static_scratch, per device ordinal
bool g_static_ready[kMaxDevices];
std::mutex g_static_mutex;

// Runs the one-time lookup for the current device if it has not been done yet.
int static_scratch_for_current_device(double **out) {
    int dev = -1;
    FLR_CUDA(cudaGetDevice(&dev));
    FLR_REQUIRE(dev >= 0 && dev < kMaxDevices, FLR_E_UNSUPPORTED,
                "list_metrics_static: device ordinal %d outside [0, %d)", dev, kMaxDevices);
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_ready[dev]) {
        void *ptr = nullptr;
        FLR_CUDA(cudaGetSymbolAddress(&ptr, g_lm_static_scratch));
        FLR_REQUIRE(ptr != nullptr, FLR_E_CUDA, "list_metrics_static: cudaGetSymbolAddress returned a null address");
        g_static_ptr[dev] = static_cast<double *>(ptr);
        g_static_ready[dev] = true;
    }
    *out = g_static_ptr[dev];
    return FLR_OK;
}

}  // namespace

extern "C" int flr_list_metrics_static(const int32_t *neutral, const int32_t *neutral_len, const int32_t *lists,
                                       const int32_t *lens, int64_t U, int V, int K, int k_cap, int vocab,
                                       int empty_policy, int rank_metrics, int32_t *out_inter, int32_t *out_union,
                                       float *out_jac, float *out_serp, float *out_prag, double *out_variant_sums,
                                       int32_t *out_status, void *stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    int rc = check_args("list_metrics_static", neutral, neutral_len, lists, lens, U, V, K, k_cap, empty_policy,
                        rank_metrics);
    if (rc != FLR_OK) return rc;

    // How many partials this shape writes: 4 doubles per thread of the kernel that will run.  The compare
    // kernel walks its pairs with a grid stride, so its grid can be cut down to what the array holds (a
    // multiple of V blocks, so that a thread still sees one variant only).
    Plan pl = make_plan(U, V, K, k_cap, vocab);
    if (!pl.table) {
        const int64_t room = static_cast<int64_t>(FLR_LIST_METRICS_STATIC_SCRATCH_BYTES) / (CMP_THREADS * 4 * sizeof(double));
        const int64_t blocks = std::min<int64_t>(pl.grid, (room / V) * V);
        if (blocks >= V) {
            pl.grid = static_cast<int>(blocks);
            pl.cmp_threads = blocks * CMP_THREADS;
        }
    }
    const size_t need = partial_bytes(pl);
    FLR_REQUIRE(need <= static_cast<size_t>(FLR_LIST_METRICS_STATIC_SCRATCH_BYTES), FLR_E_UNSUPPORTED,
                "list_metrics_static: this shape needs %zu bytes of partial sums but the built-in scratch holds "
                "%d; use flr_list_metrics with caller-provided scratch (U=%lld V=%d K=%d vocab=%d)",
                need, FLR_LIST_METRICS_STATIC_SCRATCH_BYTES, static_cast<long long>(U), V, K, vocab);

    double *partial = nullptr;
    rc = static_scratch_for_current_device(&partial);
    if (rc != FLR_OK) return rc;
    if (U == 0) {
        if (out_variant_sums) FLR_CUDA(cudaMemsetAsync(out_variant_sums, 0, sizeof(double) * 4 * V, stream));
        return FLR_OK;
    }
    // Nothing in the array has to start at zero: every thread of the launch overwrites its own four
    // partials before the reduction kernel reads them, and the reduction reads only those.  What the
    // previous call left behind is never looked at, so nothing is cleared here.
    return launch_planned(pl, make_params(neutral, neutral_len, lists, lens, U, V, K, k_cap, vocab, empty_policy,
                                          out_inter, out_union, out_jac, out_serp, out_prag, out_status, partial),
                          rank_metrics, out_variant_sums, stream);
}
