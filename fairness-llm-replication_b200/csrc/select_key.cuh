// Order-preserving integer keys for the radix select: unsigned order of the key == the order
// np.sort gives the floats (NaN last); -0.0 is folded onto +0.0.
#pragma once
#include <stdint.h>

namespace flr {
namespace select_key {

__host__ __device__ __forceinline__ uint32_t key_f32(float x) {
    if (x != x) return 0xFFFFFFFFu;
    if (x == 0.0f) return 0x80000000u;
#ifdef __CUDA_ARCH__
    uint32_t u = __float_as_uint(x);
#else
    union { float f; uint32_t u; } c; c.f = x; uint32_t u = c.u;
#endif
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float unkey_f32(uint32_t k) {
    uint32_t u = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}
__host__ __device__ __forceinline__ uint64_t key_f64(double x) {
    if (x != x) return 0xFFFFFFFFFFFFFFFFull;
    if (x == 0.0) return 0x8000000000000000ull;
#ifdef __CUDA_ARCH__
    uint64_t u = static_cast<uint64_t>(__double_as_longlong(x));
#else
    union { double f; uint64_t u; } c; c.f = x; uint64_t u = c.u;
#endif
    return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
__host__ __device__ __forceinline__ double unkey_f64(uint64_t k) {
    uint64_t u = (k >> 63) ? (k & 0x7FFFFFFFFFFFFFFFull) : ~k;
#ifdef __CUDA_ARCH__
    return __longlong_as_double(static_cast<long long>(u));
#else
    union { double f; uint64_t u; } c; c.u = u; return c.f;
#endif
}

// digit layout: pass p covers key bits [shift, shift + nbits)
__host__ __device__ __forceinline__ int pass_shift(int key_bits, int pass) {
    int s = key_bits - 11 * (pass + 1);
    return s < 0 ? 0 : s;
}
__host__ __device__ __forceinline__ int pass_nbits(int key_bits, int pass) {
    int rem = key_bits - 11 * pass;
    return rem < 11 ? rem : 11;
}

}  // namespace select_key
}  // namespace flr
