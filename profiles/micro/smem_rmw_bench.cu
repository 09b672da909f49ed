// Micro-benchmark: what one "test-and-set on a private shared-memory bit set" costs on sm_100a,
// in the layout the list-metric kernels use (word-interleaved across the 128 threads of a CTA, so
// every warp access is bank-conflict free).  Variants of the per-element step:
//   0  address + bit arithmetic only (the floor every variant pays)
//   1  atomicOr returning the old word        (ATOMS.OR)
//   2  atomicOr, result unused                (ATOMS.OR without a consumer / RED)
//   3  load -> or -> store, program order     (LDS, LOP3, STS: each load waits for the previous store)
//   4  load issued before the previous store, patched when both hit the same word
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o smem_rmw_bench smem_rmw_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int NT = 128, WORDS = 118, ELEMS = 50;

__device__ __forceinline__ uint32_t lds(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts(uint32_t addr, uint32_t v) {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v));
}

template <int MODE>
__global__ void __launch_bounds__(NT, 2) rmw_kernel(int rounds, uint32_t seed, uint32_t *out) {
    extern __shared__ uint32_t bits[];
    const int t = threadIdx.x;
    for (int w = 0; w < WORDS; ++w) bits[w * NT + t] = 0;
    __syncthreads();
    uint32_t base = static_cast<uint32_t>(__cvta_generic_to_shared(bits + t));
    uint32_t x = seed ^ (blockIdx.x * 131u + t * 2654435761u);
    uint32_t acc = 0;
    for (int r = 0; r < rounds; ++r) {
        uint32_t ids[ELEMS];
#pragma unroll
        for (int j = 0; j < ELEMS; ++j) {
            x = x * 1664525u + 1013904223u;
            ids[j] = (x >> 8) % 3706u;
        }
        uint32_t prev_addr = 0xffffffffu, prev_new = 0, pending_addr = 0, pending_val = 0;
        bool have_pending = false;
#pragma unroll
        for (int j = 0; j < ELEMS; ++j) {
            const uint32_t idx = ids[j];
            const uint32_t addr = base + (idx >> 5) * (NT * 4);
            const uint32_t bit = 1u << (idx & 31u);
            if constexpr (MODE == 0) {
                acc += addr ^ bit;
            } else if constexpr (MODE == 1) {
                const uint32_t old = atomicOr(reinterpret_cast<unsigned int *>(__cvta_shared_to_generic(addr)), bit);
                acc |= old & bit;
            } else if constexpr (MODE == 2) {
                atomicOr(reinterpret_cast<unsigned int *>(__cvta_shared_to_generic(addr)), bit);
                acc += bit;
            } else if constexpr (MODE == 3) {
                const uint32_t w = lds(addr);
                sts(addr, w | bit);
                acc |= w & bit;
            } else {
                // the load of element j is issued before the store of element j - 1
                uint32_t w = lds(addr);
                if (have_pending) sts(pending_addr, pending_val);
                if (addr == prev_addr) w = prev_new;
                acc |= w & bit;
                prev_new = w | bit;
                prev_addr = addr;
                pending_addr = addr; pending_val = prev_new; have_pending = true;
            }
        }
        if constexpr (MODE == 4) sts(pending_addr, pending_val);
        __syncthreads();
        // wipe (as the real kernel does) so later rounds start clean
        for (int w = 0; w < WORDS; ++w) bits[w * NT + t] = 0;
        __syncthreads();
    }
    out[blockIdx.x * NT + t] = acc;
}

template <int MODE>
float run(int rounds, uint32_t *out) {
    const int smem = WORDS * NT * 4;
    cudaFuncSetAttribute(rmw_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    rmw_kernel<MODE><<<296, NT, smem>>>(rounds / 10 + 1, 1u, out);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    rmw_kernel<MODE><<<296, NT, smem>>>(rounds, 7u, out);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms;
}

int main() {
    uint32_t *out;
    cudaMalloc(&out, 296 * NT * 4);
    const int rounds = 2000;
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const float ms[5] = {run<0>(rounds, out), run<1>(rounds, out), run<2>(rounds, out), run<3>(rounds, out), run<4>(rounds, out)};
    const char *name[5] = {"arith only", "atomicOr (old used)", "atomicOr (old unused)", "ld/or/st in order", "ld early + patch"};
    // 2 CTAs x 4 warps per SM; per SM: 8 warps x rounds x ELEMS element-steps
    for (int m = 0; m < 5; ++m) {
        const double steps_per_sm = 8.0 * rounds * ELEMS;
        const double ns_per_step = ms[m] * 1e6 / steps_per_sm;
        printf("{\"variant\": \"%s\", \"ms\": %.3f, \"ns_per_warp_element_per_sm\": %.3f, \"cycles_at_max_clock\": %.2f}\n",
               name[m], ms[m], ns_per_step, ns_per_step * clk_khz * 1e-6);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
