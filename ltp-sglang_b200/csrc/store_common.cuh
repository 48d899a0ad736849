// Bit-exact conversions shared by the KV-store kernels (kv_store.cu, rope_kv_store.cu).
#pragma once

#include <string.h>

#include "common.cuh"

namespace b200attn {

// fp32 -> float8_e4m3fn bits, round-to-nearest-even, NON-saturating: every
// |x| >= 480 (and NaN) becomes 0x7F|sign, 464 < |x| < 480 rounds up into 0x7F,
// subnormals are kept.  This is the arithmetic torch's `.to(float8_e4m3fn)`
// performs on both CPU and CUDA (hardware cvt.satfinite would clamp to 448).
__host__ __device__ inline uint8_t f32_to_e4m3fn(float f) {
  uint32_t bits;
#ifdef __CUDA_ARCH__
  bits = __float_as_uint(f);
#else
  memcpy(&bits, &f, 4);
#endif
  const uint32_t sign = bits & 0x80000000u;
  bits ^= sign;
  uint8_t r;
  if (bits >= (1087u << 20)) {  // >= 480.0 or NaN
    r = 0x7f;
  } else if (bits < (121u << 23)) {  // < 2^-6: lands in the subnormal range
    const uint32_t magic = 141u << 23;
    float a, m;
#ifdef __CUDA_ARCH__
    a = __uint_as_float(bits);
    m = __uint_as_float(magic);
    bits = __float_as_uint(__fadd_rn(a, m));
#else
    memcpy(&a, &bits, 4);
    memcpy(&m, &magic, 4);
    a = a + m;
    memcpy(&bits, &a, 4);
#endif
    r = static_cast<uint8_t>(bits - magic);
  } else {
    const uint32_t odd = (bits >> 20) & 1u;
    bits += (static_cast<uint32_t>(7 - 127) << 23) + 0x7FFFFu;
    bits += odd;
    r = static_cast<uint8_t>(bits >> 20);
  }
  return r | static_cast<uint8_t>(sign >> 24);
}

// fp32 -> float8_e5m2 bits the way torch casts (c10/util/Float8_e5m2.h): round-to-nearest-even, NON-saturating --
// |x| >= 61440 rounds into infinity (0x7C), NaN becomes 0x7F, subnormals are kept.
__host__ __device__ inline uint8_t f32_to_e5m2(float f) {
  uint32_t bits;
#ifdef __CUDA_ARCH__
  bits = __float_as_uint(f);
#else
  memcpy(&bits, &f, 4);
#endif
  const uint32_t sign = bits & 0x80000000u;
  bits ^= sign;
  uint8_t r;
  if (bits >= (143u << 23)) {  // >= 2^16, infinity or NaN
    r = bits > (255u << 23) ? 0x7f : 0x7c;
  } else if (bits < (113u << 23)) {  // < 2^-14: lands in the subnormal range
    const uint32_t magic = 134u << 23;
    float a, m;
#ifdef __CUDA_ARCH__
    a = __uint_as_float(bits);
    m = __uint_as_float(magic);
    bits = __float_as_uint(__fadd_rn(a, m));
#else
    memcpy(&a, &bits, 4);
    memcpy(&m, &magic, 4);
    a = a + m;
    memcpy(&bits, &a, 4);
#endif
    r = static_cast<uint8_t>(bits - magic);
  } else {
    const uint32_t odd = (bits >> 21) & 1u;
    bits += (static_cast<uint32_t>(15 - 127) << 23) + 0xFFFFFu;
    bits += odd;
    r = static_cast<uint8_t>(bits >> 21);
  }
  return r | static_cast<uint8_t>(sign >> 24);
}
__host__ __device__ inline uint8_t f32_to_fp8(float f, bool e5m2) { return e5m2 ? f32_to_e5m2(f) : f32_to_e4m3fn(f); }

template <typename T>
__device__ __forceinline__ float to_f32(T x);
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 x) {
  return __bfloat162float(x);
}
template <>
__device__ __forceinline__ float to_f32<__half>(__half x) {
  return __half2float(x);
}
template <typename T>
__device__ __forceinline__ T from_f32(float x);
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float x) {
  return __float2bfloat16_rn(x);
}
template <>
__device__ __forceinline__ __half from_f32<__half>(float x) {
  return __float2half_rn(x);
}

// x / s the way the reference's in-place div_ does it (memory_pool.py:385-389): fp32 arithmetic,
// rounded back to the source dtype.  `recip`: torch's CUDA kernel for a Python-float divisor
// multiplies by the fp32 reciprocal.
template <typename T>
__device__ __forceinline__ T div_scale(T x, float s, float inv, bool recip) {
  return recip ? from_f32<T>(__fmul_rn(to_f32<T>(x), inv)) : from_f32<T>(__fdiv_rn(to_f32<T>(x), s));
}

}  // namespace b200attn
