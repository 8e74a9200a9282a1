// Device twin of grid_b200/synth.py: counter-based hashes + IEEE add/mul only (explicit _rn intrinsics so the
// compiler cannot contract into FMA), therefore bit-identical to the numpy generator.
#include "common.cuh"

namespace g2 {

__device__ __host__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
  x += 0x9E3779B97F4A7C15ULL;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
  return x ^ (x >> 31);
}
__device__ __forceinline__ double znorm(unsigned long long h) {
  const long long s = (long long)((h & 0xFFFF) + ((h >> 16) & 0xFFFF) + ((h >> 32) & 0xFFFF) + (h >> 48)) - 131070;
  return __dmul_rn((double)s, 1.0 / 37837.22);
}

__global__ void __launch_bounds__(256) k_synth_depth(unsigned long long key3, unsigned long long key4, int64_t n,
                                                     int64_t b, int64_t row0, int64_t col0,
                                                     const double* __restrict__ c, const double* __restrict__ rc,
                                                     const double* __restrict__ bcol, const double* __restrict__ rb,
                                                     float* __restrict__ depth, int64_t ld) {
  const int64_t j0 = ((int64_t)blockIdx.x * 256 + threadIdx.x) * 4;
  if (j0 >= b) return;
  double bj[4], rbj[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const bool in = j0 + q < b;
    bj[q] = in ? bcol[j0 + q] : 0.0;
    rbj[q] = in ? rb[j0 + q] : 0.0;
  }
  for (int64_t i = blockIdx.y; i < n; i += gridDim.y) {
    const unsigned long long gi = (unsigned long long)(row0 + i);
    const double ci = c[i], rci = rc[i];
    float o[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const unsigned long long gj = (unsigned long long)(col0 + j0 + q);
      const double e = __dadd_rn(1.0, __dmul_rn(0.05, znorm(splitmix64(key3 ^ (((gi & 15ULL) << 32) | gj)))));
      const double lam = __dmul_rn(__dmul_rn(ci, bj[q]), e);
      const unsigned long long h4 = splitmix64(key4 ^ ((gi << 32) | gj));
      const double d = __dmul_rn(lam, __dadd_rn(1.0, __dmul_rn(znorm(h4), __dmul_rn(rci, rbj[q]))));
      double h = d > 0.0 ? floor(__dadd_rn(__dmul_rn(d, 100.0), 0.5)) : 0.0;
      if (splitmix64(h4) % 1000ULL == 0) h = 0.0;
      o[q] = (float)__ddiv_rn(h, 100.0);
    }
    float* dst = depth + i * ld + j0;
    if (j0 + 4 <= b) {
      *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
    } else {
      for (int q = 0; q < 4 && j0 + q < b; ++q) dst[q] = o[q];
    }
  }
}

}  // namespace g2

using namespace g2;

extern "C" int gridb200_synth_depth(uint64_t seed, int64_t n, int64_t b, int64_t row0, int64_t col0, const double* c,
                                    const double* rc, const double* bcol, const double* rb, float* depth, int64_t ld,
                                    void* stream) {
  G2_REQUIRE(c && rc && bcol && rb && depth, GRIDB200_EINVAL, "null pointer");
  G2_REQUIRE(n > 0 && b > 0 && ld >= b && ld % 4 == 0, GRIDB200_EINVAL, "bad shape");
  G2_REQUIRE((reinterpret_cast<uintptr_t>(depth) & 15) == 0, GRIDB200_EINVAL, "depth must be 16-byte aligned");
  G2_REQUIRE(row0 + n < (1LL << 32) && col0 + b < (1LL << 32), GRIDB200_EINVAL, "indices exceed 32 bits");
  const unsigned long long key3 = splitmix64((unsigned long long)seed * 0x100ULL + 3ULL);
  const unsigned long long key4 = splitmix64((unsigned long long)seed * 0x100ULL + 4ULL);
  const int64_t cb = (b + 1023) / 1024;
  const unsigned gy = (unsigned)std::min<int64_t>(n, std::max<int64_t>(1, (int64_t)kSMs * 16 / cb));
  k_synth_depth<<<dim3((unsigned)cb, gy), 256, 0, as_stream(stream)>>>(key3, key4, n, b, row0, col0, c, rc, bcol, rb,
                                                                      depth, ld);
  G2_LAUNCHED();
  return 0;
}
