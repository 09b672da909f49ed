// Dense per-user similarity matrices S[u] = X[u] X[u]^T (config 4; no reference code --
// definition: oracle/np_oracle.py:pair_similarity / normalize_rows_bf16).
//
// The one GEMM-shaped piece of the path, so it runs on the 5th-generation tensor cores:
//   * X[u] (M <= 128 rows, D bf16 columns, K-major) is staged by TMA (cp.async.bulk.tensor.2d,
//     128-byte swizzle) as D/64 boxes of [128 rows x 64 columns]; the SAME shared-memory tile is
//     both the A and the B operand (S = X X^T);
//   * one thread issues tcgen05.mma (kind::f16, bf16 x bf16 -> fp32, M = N = 128, K = 16 per
//     instruction) into a 128-column TMEM accumulator, double-buffered so the epilogue of user i
//     overlaps the MMAs of user i+1;
//   * four epilogue warps read their 32-lane TMEM quarter with tcgen05.ld and store the M x M block
//     through a small shared-memory stage, as full 128-byte row segments.
// Warp roles: 0 = TMA producer, 1 = TMEM allocator + MMA issuer, 2..5 = epilogue.  Persistent
// CTAs, one per SM; smem ring of 2 stages (96 KB each at D = 384).
// The kernel moves 95 KB in and 61.5 KB out per user for 11.8 MFLOP: HBM-bound (75 flop/B).
#include <algorithm>
#include <cuda.h>
#include <cuda_bf16.h>
#include <mutex>

#include "common.cuh"

using namespace flr;

namespace {

// one warp per row: fp64 sum of squares (exact fp32 products), fp32 divide, RNE to bf16.
// NC > 0: the row is read once with 128-bit loads and kept in registers (D % 4 == 0, D <= NC * 128);
// a warp keeps RPW rows in flight per trip so that enough bytes are outstanding per SM.  NC == 0: any D.
template <int NC>
__global__ void __launch_bounds__(256) normalize_rows_bf16_kernel(const float *__restrict__ x, int64_t rows, int D,
                                                                  __nv_bfloat16 *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);
    const int64_t nwarps = static_cast<int64_t>(gridDim.x) * 8;
    if constexpr (NC > 0) {
        constexpr int RPW = NC <= 2 ? 4 : NC <= 4 ? 2 : 1;
        const int nchunk = D >> 2;
        for (int64_t r0 = warp0 * RPW; r0 < rows; r0 += nwarps * RPW) {
            float4 q[RPW][NC];
#pragma unroll
            for (int i = 0; i < RPW; ++i) {
                const float4 *row = reinterpret_cast<const float4 *>(x + (r0 + i) * D);
#pragma unroll
                for (int c = 0; c < NC; ++c) {
                    const int ch = lane + 32 * c;
                    q[i][c] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (r0 + i < rows && ch < nchunk) q[i][c] = ldg_stream_f4(row + ch);
                }
            }
#pragma unroll
            for (int i = 0; i < RPW; ++i) {
                if (r0 + i >= rows) break;           // warp-uniform
                // lane l sums chunks l, l+32, ... (the scalar form below strides by element instead; the two
                // differ only in the last bits of |x|, far below bf16 resolution)
                double ss = 0.0;
#pragma unroll
                for (int c = 0; c < NC; ++c) {
                    const double a = q[i][c].x, b2 = q[i][c].y, c2 = q[i][c].z, d2 = q[i][c].w;
                    ss = fma(a, a, ss); ss = fma(b2, b2, ss); ss = fma(c2, c2, ss); ss = fma(d2, d2, ss);
                }
#pragma unroll
                for (int off = 16; off >= 1; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
                const double n = sqrt(ss);
                const float inv = n > 0.0 ? static_cast<float>(1.0 / n) : 0.0f;
#pragma unroll
                for (int c = 0; c < NC; ++c) {
                    const int ch = lane + 32 * c;
                    if (ch < nchunk) {
                        // one fp64 division per row, then fp32 products: x * fp32(1/n) is within 1.2e-7 relative
                        // of x / n, 2^-15 of a bf16 step -- the rounding to bf16 differs from the exact quotient's
                        // for about 3 elements in 10^5, by one bf16 step (the fp64 product and its conversion
                        // back cost two more conversion-pipe instructions per element, which bounded the kernel)
                        const float y0 = q[i][c].x * inv, y1 = q[i][c].y * inv, y2 = q[i][c].z * inv, y3 = q[i][c].w * inv;
                        __nv_bfloat162 lo = __floats2bfloat162_rn(y0, y1), hi = __floats2bfloat162_rn(y2, y3);
                        uint2 pk;
                        pk.x = *reinterpret_cast<uint32_t *>(&lo);
                        pk.y = *reinterpret_cast<uint32_t *>(&hi);
                        *reinterpret_cast<uint2 *>(out + (r0 + i) * D + 4 * ch) = pk;
                    }
                }
            }
        }
    } else {
        for (int64_t r = warp0; r < rows; r += nwarps) {
            const float *row = x + r * D;
            double ss = 0.0;
            for (int d = lane; d < D; d += 32) {
                const double v = row[d];
                ss = fma(v, v, ss);
            }
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
            const double n = sqrt(ss);
            for (int d = lane; d < D; d += 32) {
                const float y = n > 0.0 ? static_cast<float>(static_cast<double>(row[d]) / n) : 0.0f;
                out[r * D + d] = __float2bfloat16_rn(y);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// tcgen05 / TMEM / TMA wrappers (PTX names; SASS: UTCHMMA, LDTM, UTMALDG)
// ------------------------------------------------------------------------------------------
constexpr int PS_THREADS = 192;
constexpr int PS_TILE = 128;                 // UMMA M = N
constexpr int PS_KBOX = 64;                  // bf16 columns per TMA box = one 128-byte swizzle row
constexpr uint32_t PS_BOX_BYTES = PS_TILE * PS_KBOX * 2;     // 16 KB
constexpr int PS_TMEM_COLS = 256;            // two 128-column fp32 accumulators
constexpr int PS_STAGE_LD = 36;              // floats per row of an epilogue warp's [32][36] staging chunk
constexpr size_t PS_EPI_BYTES = 4 * 32 * PS_STAGE_LD * sizeof(float);

__device__ __forceinline__ void tma_load_2d(void *dst_smem, const CUtensorMap *tmap, int c0, int c1, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(smem_u32(dst_smem)),
        "l"(reinterpret_cast<uint64_t>(tmap)), "r"(c0), "r"(c1), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, bf16 inputs, fp32 accumulate
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}"
        ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// K-major operand tile, 128-byte swizzle: 8-row groups are 1024 bytes apart (SBO), LBO unused.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);             // start address
    d |= 1ull << 16;                                                     // leading byte offset: unused for one swizzle atom of K
    d |= static_cast<uint64_t>(1024u >> 4) << 32;                        // stride byte offset
    d |= 1ull << 46;                                                     // descriptor version (sm_100)
    d |= 2ull << 61;                                                     // SWIZZLE_128B
    return d;
}
// kind::f16 instruction descriptor: D fp32, A/B bf16, both K-major, N = 128, M = 128
constexpr uint32_t kIdescBf16_128x128 = (1u << 4) | (1u << 7) | (1u << 10) | ((128u >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct PsParams {
    int64_t U;
    int M, D, kblocks, stages;
    uint32_t stage_bytes;
    float *out;
};

__global__ void __launch_bounds__(PS_THREADS, 1) pair_similarity_kernel(const __grid_constant__ CUtensorMap tmap,
                                                                        const PsParams p) {
    extern __shared__ unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full[2], empty[2], tfull[2], tempty[2];
    __shared__ uint32_t tmem_base_slot;
    // 128-byte swizzle atoms need 1024-byte aligned tiles
    unsigned char *tiles = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(full + i, 1);
            mbar_init(empty + i, 1);
            mbar_init(tfull + i, 1);
            mbar_init(tempty + i, 4);
        }
        fence_mbar_init();
    }
    if (warp == 1) {     // one warp allocates TMEM and owns the deallocation
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)),
                     "r"(PS_TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = tmem_base_slot;

    const int64_t U = p.U;
    const int M = p.M;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int s = 0;
            uint32_t ph = 1;
            for (int64_t u = blockIdx.x; u < U; u += gridDim.x) {
                mbar_wait(empty + s, ph);
                mbar_expect_tx(full + s, p.stage_bytes);
                unsigned char *dst = tiles + static_cast<size_t>(s) * p.stage_bytes;
                const int row0 = static_cast<int>(u * M);
                for (int kb = 0; kb < p.kblocks; ++kb)
                    tma_load_2d(dst + static_cast<size_t>(kb) * PS_BOX_BYTES, &tmap, kb * PS_KBOX, row0, full + s);
                if (++s == p.stages) { s = 0; ph ^= 1u; }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one thread) =====
        if (lane == 0) {
            int s = 0, a = 0;
            uint32_t ph = 0, aph = 1;
            for (int64_t u = blockIdx.x; u < U; u += gridDim.x) {
                mbar_wait(tempty + a, aph);                  // epilogue has drained this accumulator
                mbar_wait(full + s, ph);                     // operands have landed
                tcgen05_fence_after();
                const uint32_t tile = smem_u32(tiles + static_cast<size_t>(s) * p.stage_bytes);
                const uint32_t acc = tmem_base + static_cast<uint32_t>(a) * PS_TILE;
                for (int kb = 0; kb < p.kblocks; ++kb) {
#pragma unroll
                    for (int k = 0; k < PS_KBOX / 16; ++k) {
                        const uint64_t desc = umma_desc_sw128(tile + kb * PS_BOX_BYTES + k * 32);
                        umma_bf16(acc, desc, desc, kIdescBf16_128x128, (kb | k) != 0 ? 1u : 0u);
                    }
                }
                tcgen05_commit(empty + s);                   // smem stage reusable once the MMAs retire
                tcgen05_commit(tfull + a);                   // accumulator ready for the epilogue
                if (++s == p.stages) { s = 0; ph ^= 1u; }
                a ^= 1;
                if (a == 0) aph ^= 1u;
            }
        }
    } else {
        // ===== epilogue: warp w owns TMEM lanes [32 * (w % 4), +32) =====
        // tcgen05.ld hands lane l row l of a 32 x 32 chunk.  Stored straight from there, one warp store
        // would touch 32 different rows, 16 bytes each; so the chunk goes through a small shared-memory
        // stage ([32][36] floats per warp: 128-bit accesses are conflict-free both ways) and leaves as
        // full 128-byte row segments, four rows per store instruction.
        const int q = warp & 3;
        float *stage = reinterpret_cast<float *>(tiles + static_cast<size_t>(p.stages) * p.stage_bytes) + q * (32 * PS_STAGE_LD);
        int a = 0;
        uint32_t aph = 0;
        const bool vec4 = (M % 4) == 0;
        for (int64_t u = blockIdx.x; u < U; u += gridDim.x) {
            mbar_wait(tfull + a, aph);
            tcgen05_fence_after();
            const int row = q * 32 + lane;
            float *orow = p.out + (static_cast<size_t>(u) * M + row) * M;
            float *oblock = p.out + (static_cast<size_t>(u) * M + q * 32) * M;
#pragma unroll 1
            for (int c = 0; c < PS_TILE / 32; ++c) {
                if (c * 32 >= M) break;
                uint32_t r[32];
                tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(a * PS_TILE + c * 32), r);
                if (vec4) {
                    float4 *srow = reinterpret_cast<float4 *>(stage + lane * PS_STAGE_LD);
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        srow[j] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                              __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
                    __syncwarp();
                    const int col = c * 32 + (lane & 7) * 4;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int rl = i * 4 + (lane >> 3);
                        if (q * 32 + rl < M && col < M)
                            *reinterpret_cast<float4 *>(oblock + static_cast<size_t>(rl) * M + col) =
                                *reinterpret_cast<const float4 *>(stage + rl * PS_STAGE_LD + (lane & 7) * 4);
                    }
                    __syncwarp();
                } else if (row < M) {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (c * 32 + j < M) orow[c * 32 + j] = __uint_as_float(r[j]);
                }
            }
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty + a);
            a ^= 1;
            if (a == 0) aph ^= 1u;
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(PS_TMEM_COLS) : "memory");
    }
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time libcuda dependency,
// so the library still loads on a machine without a driver).
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void *ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(ptr);
    });
    return fn;
}

}  // namespace

extern "C" int flr_normalize_rows_bf16(const float *x, int64_t rows, int D, uint16_t *out_bf16, void *stream_) {
    FLR_REQUIRE(rows >= 0 && D >= 1, FLR_E_INVALID, "normalize_rows: bad shape");
    if (rows == 0) return FLR_OK;
    FLR_REQUIRE(x && out_bf16, FLR_E_INVALID, "normalize_rows: null pointer");
    const int grid = static_cast<int>(std::min<int64_t>((rows + 7) / 8, static_cast<int64_t>(device_sm_count()) * 8));
    const bool vec = (D % 4 == 0) && D <= 1024 && aligned16(x) && ((reinterpret_cast<uintptr_t>(out_bf16) & 7u) == 0);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    __nv_bfloat16 *out = reinterpret_cast<__nv_bfloat16 *>(out_bf16);
    switch (vec ? (D + 127) / 128 : 0) {
        case 1: normalize_rows_bf16_kernel<1><<<grid, 256, 0, stream>>>(x, rows, D, out); break;
        case 2: normalize_rows_bf16_kernel<2><<<grid, 256, 0, stream>>>(x, rows, D, out); break;
        case 3: normalize_rows_bf16_kernel<3><<<grid, 256, 0, stream>>>(x, rows, D, out); break;
        case 4: normalize_rows_bf16_kernel<4><<<grid, 256, 0, stream>>>(x, rows, D, out); break;
        case 5: case 6: normalize_rows_bf16_kernel<6><<<grid, 256, 0, stream>>>(x, rows, D, out); break;
        case 7: case 8: normalize_rows_bf16_kernel<8><<<grid, 256, 0, stream>>>(x, rows, D, out); break;
        default: normalize_rows_bf16_kernel<0><<<grid, 256, 0, stream>>>(x, rows, D, out); break;
    }
    FLR_LAUNCH_CHECK("normalize_rows_bf16_kernel");
    return FLR_OK;
}

extern "C" size_t flr_pair_similarity_scratch_bytes(int64_t U, int M, int D) {
    (void)U; (void)M; (void)D;
    return 0;       // operands are staged in shared memory, accumulators in TMEM: no global scratch
}

extern "C" int flr_pair_similarity_gemm(const uint16_t *x_bf16, int64_t U, int M, int D, float *out, void *scratch,
                                        size_t scratch_bytes, void *stream_) {
    (void)scratch; (void)scratch_bytes;
    FLR_REQUIRE(U >= 0 && M >= 1 && M <= PS_TILE, FLR_E_UNSUPPORTED, "pair_similarity: need 1 <= M <= %d (got %d)", PS_TILE, M);
    FLR_REQUIRE(D >= PS_KBOX && D % PS_KBOX == 0 && D <= 512, FLR_E_UNSUPPORTED,
                "pair_similarity: D must be a multiple of %d in [%d, 512] (got %d)", PS_KBOX, PS_KBOX, D);
    if (U == 0) return FLR_OK;
    FLR_REQUIRE(x_bf16 && out, FLR_E_INVALID, "pair_similarity: null pointer");
    FLR_REQUIRE(aligned16(x_bf16) && aligned16(out), FLR_E_INVALID, "pair_similarity: x and out must be 16-byte aligned");
    FLR_REQUIRE(U * M < (1LL << 31), FLR_E_UNSUPPORTED, "pair_similarity: U*M must fit 31 bits; split the batch");
    EncodeTiledFn encode = encode_tiled_fn();
    FLR_REQUIRE(encode != nullptr, FLR_E_CUDA, "pair_similarity: cuTensorMapEncodeTiled is not available from this driver");

    CUtensorMap tmap;
    const cuuint64_t gdim[2] = {static_cast<cuuint64_t>(D), static_cast<cuuint64_t>(U * M)};
    const cuuint64_t gstride[1] = {static_cast<cuuint64_t>(D) * 2};
    const cuuint32_t box[2] = {PS_KBOX, PS_TILE};
    const cuuint32_t estride[2] = {1, 1};
    const CUresult cr = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<uint16_t *>(x_bf16), gdim, gstride,
                               box, estride, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    FLR_REQUIRE(cr == CUDA_SUCCESS, FLR_E_CUDA, "pair_similarity: cuTensorMapEncodeTiled failed with CUresult %d",
                static_cast<int>(cr));

    PsParams p{};
    p.U = U; p.M = M; p.D = D; p.kblocks = D / PS_KBOX; p.out = out;
    p.stage_bytes = static_cast<uint32_t>(p.kblocks) * PS_BOX_BYTES;
    p.stages = std::max(1, std::min<int>(2, static_cast<int>((200u * 1024u) / p.stage_bytes)));   // + 18 KB epilogue stage <= 227 KB
    const size_t smem = static_cast<size_t>(p.stages) * p.stage_bytes + PS_EPI_BYTES + 1024;
    FLR_CUDA(cudaFuncSetAttribute(pair_similarity_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  static_cast<int>(smem)));
    const int grid = static_cast<int>(std::min<int64_t>(U, device_sm_count()));
    pair_similarity_kernel<<<grid, PS_THREADS, smem, static_cast<cudaStream_t>(stream_)>>>(tmap, p);
    FLR_LAUNCH_CHECK("pair_similarity_kernel");
    return FLR_OK;
}
