// Test harness (never shipped): the number formatting of grid_b200/csrc/nbr_text.cuh -- the functions nbrio.cu inlines --
// compiled for the host so that tests/test_step5_host.py can compare it with Python's float formatting.
#include "../../grid_b200/csrc/nbr_text.cuh"

extern "C" unsigned long long nt_dist_hundredths(long long d2, double denom) { return g2::nbr_dist_hundredths(d2, denom); }
extern "C" unsigned long long nt_fixed2_hundredths(double v) { return g2::fixed2_hundredths(v); }
extern "C" int nt_put_fixed2(unsigned long long q, char* buf) {
  const int n = g2::put_fixed2(buf, q);
  return n == g2::fixed2_len(q) ? n : -1;
}
