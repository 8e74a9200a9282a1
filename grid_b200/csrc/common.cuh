// Shared host/device helpers for libgridb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>

#include "../../include/gridb200.h"

namespace g2 {

void set_error(const char* fmt, ...);
extern std::atomic<uint64_t> g_launches;

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

#define G2_REQUIRE(cond, code, ...)            \
  do {                                         \
    if (!(cond)) {                             \
      g2::set_error(__VA_ARGS__);              \
      return (code);                           \
    }                                          \
  } while (0)

#define G2_CUDA(expr)                                                                  \
  do {                                                                                 \
    cudaError_t e__ = (expr);                                                          \
    if (e__ != cudaSuccess) {                                                          \
      g2::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, \
                    __LINE__);                                                         \
      return GRIDB200_ECUDA;                                                           \
    }                                                                                  \
  } while (0)

// Count the launch and surface launch-configuration errors immediately.
#define G2_LAUNCHED()                                                                        \
  do {                                                                                       \
    g2::g_launches.fetch_add(1, std::memory_order_relaxed);                                  \
    cudaError_t e__ = cudaGetLastError();                                                    \
    if (e__ != cudaSuccess) {                                                                \
      g2::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e__), __FILE__,   \
                    __LINE__);                                                               \
      return GRIDB200_ECUDA;                                                                 \
    }                                                                                        \
  } while (0)

constexpr int kSMs = 148;  // B200

// ---- device helpers ------------------------------------------------------------------------------------

// depth (fp32, a multiple of 0.01 below 41943.04) -> exact hundredths.  One FFMA rounds the EXACT product
// x*100 to the nearest integer (ulp at 1.5*2^23 is 1); the low 22 mantissa bits are the integer.
__device__ __forceinline__ int depth_hundredths(float x) {
  return __float_as_int(__fmaf_rn(x, 100.0f, 12582912.0f)) & 0x3FFFFF;
}

// exact (double)h for 0 <= h < 2^32 without the conversion pipe
__device__ __forceinline__ double u32_to_double(unsigned h) {
  return __hiloint2double(0x43300000, (int)h) - 4503599627370496.0;
}

// Round-half-even to an integer WITHOUT the conversion pipe: (t + 1.5*2^52) - 1.5*2^52, |t| < 2^51.
constexpr double kRoundMagic = 6755399441055744.0;
__device__ __forceinline__ double round_even(double t) { return (t + kRoundMagic) - kRoundMagic; }
// exact int of an integer-valued double with |q| < 2^31 (low word of the biased sum is two's complement)
__device__ __forceinline__ int double_to_int_exact(double q) { return __double2loint(q + kRoundMagic); }

// Integer n with "%.*f" % v == n / 10^d  (mult = 10^d): printf rounds the exact binary value half-to-even.
// t = v*mult may land exactly on k+0.5 only through rounding; the exact residual r = fma(v,mult,-t) decides.
__device__ __forceinline__ double printf_round(double v, double mult) {
  const double t = v * mult;
  double q = round_even(t);
  if (fabs(t - q) == 0.5) {
    const double r = fma(v, mult, -t);
    if (r > 0.0) q = t + 0.5;
    else if (r < 0.0) q = t - 0.5;
  }
  return q;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ long long warp_sum(long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// streaming 128-bit load that does not allocate in L1 (the matrices are read once per pass)
__device__ __forceinline__ float4 ld_stream(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}

}  // namespace g2
