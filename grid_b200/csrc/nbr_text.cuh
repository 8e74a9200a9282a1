// Number formatting of the step-5 neighbours file (find_neighbors.py:259-266): the printed distance
// "%.2f" % (dist**2 / (2 * R_use)) from the exact integer squared distance, and decimal output.
// Plain C++ (no CUDA dependency besides the rounding intrinsics under __CUDA_ARCH__): nbrio.cu inlines it as device
// code and tests/csrc/step5_host.cpp compiles the same functions with g++ (-ffp-contract=off) to compare them with
// Python's own float formatting.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define G2N_HD __host__ __device__ __forceinline__
#else
#define G2N_HD inline
#endif

namespace g2 {

// IEEE double operations that the compiler must not fuse (device code contracts a*b+c into an FMA by default)
#if defined(__CUDA_ARCH__)
G2N_HD double n_mul(double a, double b) { return __dmul_rn(a, b); }
G2N_HD double n_add(double a, double b) { return __dadd_rn(a, b); }
G2N_HD double n_div(double a, double b) { return __ddiv_rn(a, b); }
G2N_HD double n_sqrt(double a) { return __dsqrt_rn(a); }
G2N_HD double n_fma(double a, double b, double c) { return __fma_rn(a, b, c); }
#else
G2N_HD double n_mul(double a, double b) { return a * b; }
G2N_HD double n_add(double a, double b) { return a + b; }
G2N_HD double n_div(double a, double b) { return a / b; }
G2N_HD double n_sqrt(double a) { return sqrt(a); }
G2N_HD double n_fma(double a, double b, double c) { return fma(a, b, c); }
#endif

// Integer q with "%.2f" % v == q / 100 for 0 <= v < 2^44: printf rounds the EXACT binary value half-to-even.
// t = fl(v * 100) cannot cross a half-integer (those are representable), it can only land on one; then the exact
// residual fma(v, 100, -t) says on which side the true product lies.
G2N_HD uint64_t fixed2_hundredths(double v) {
  const double magic = 6755399441055744.0;  // 1.5 * 2^52
  const double t = n_mul(v, 100.0);
  double q = n_add(n_add(t, magic), -magic);
  if (fabs(n_add(t, -q)) == 0.5) {
    const double r = n_fma(v, 100.0, -t);
    if (r > 0.0) q = n_add(t, 0.5);
    else if (r < 0.0) q = n_add(t, -0.5);
  }
  return (uint64_t)q;
}

// d2: exact squared distance in units of 1e-4 (integer, < 2^53).  The reference takes sklearn's distance
// sqrt(d^2) and squares it again (find_neighbors.py:222) before dividing by 2 * R_use (:263): same operations here,
// each correctly rounded.
G2N_HD uint64_t nbr_dist_hundredths(long long d2, double denom) {
  const double s = n_sqrt(n_div((double)d2, 1e4));
  return fixed2_hundredths(n_div(n_mul(s, s), denom));
}

G2N_HD int dec_digits(uint64_t v) {
  int n = 1;
  while (v >= 10) v /= 10, ++n;
  return n;
}

// bytes of "<ip>.<dd>" for q hundredths
G2N_HD int fixed2_len(uint64_t q) { return dec_digits(q / 100) + 3; }

G2N_HD int put_fixed2(char* dst, uint64_t q) {
  uint64_t ip = q / 100;
  const int fr = (int)(q % 100), nd = dec_digits(ip);
  for (int t = nd - 1; t >= 0; --t) dst[t] = (char)('0' + ip % 10), ip /= 10;
  dst[nd] = '.', dst[nd + 1] = (char)('0' + fr / 10), dst[nd + 2] = (char)('0' + fr % 10);
  return nd + 3;
}

}  // namespace g2
