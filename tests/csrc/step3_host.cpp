// Test harness (never shipped): drives the line grammar and read predicate of grid_b200/csrc/region_line.cuh -- the very
// functions the CUDA kernels inline -- on the CPU, with the same line-start rule, key packing and prefix-sum lookup as
// k_region_coverage / k_read_hist / k_region_counts.  Built by tests/test_step3_host.py with g++.
#include <vector>

#include "../../grid_b200/csrc/region_line.cuh"

extern "C" void rl_region_coverage(const char* text, long long n, const char* chrom, int chrom_len, long long start,
                                   long long end, unsigned long long* out4) {
  g2::RegionParams P = {};
  for (int t = 0; t < chrom_len && t < 64; ++t) P.chrom[t] = chrom[t];
  P.chrom_len = chrom_len, P.start = start, P.end = end;
  out4[0] = out4[1] = out4[2] = out4[3] = 0;
  for (long long p = 0; p < n; ++p) {
    if (!(p == 0 || text[p - 1] == '\n')) continue;
    const g2::LineResult r = g2::rc_parse_line(text, n, p, P);
    if (r.status == g2::RC_LINE_IRREGULAR) ++out4[g2::RC_IRREGULAR];
    else if (r.status == g2::RC_LINE_COUNTS)
      out4[g2::RC_WEIGHTED] += (unsigned long long)r.h * (unsigned long long)r.ov, out4[g2::RC_COVERED] += r.ov,
          ++out4[g2::RC_LINES];
  }
}

extern "C" void rl_count_reads(const uint16_t* flag, const uint8_t* mapq, const int32_t* tid, const int32_t* ntid,
                               const int32_t* pos, long long n, const uint32_t* flag_bits, int min_mapq,
                               const long long* bounds, int nb, const int32_t* lo, const int32_t* hi, int g,
                               long long* counts) {
  std::vector<uint32_t> hist(nb + 1, 0);
  long long r = 0;
  for (; r + 8 <= n; r += 8) {  // the kernel's vector path: eight keys searched in lockstep, rejected reads as key -1
    long long key[8];
    int e[8];
    for (int t = 0; t < 8; ++t) {
      const bool ok = g2::read_passes(flag[r + t], mapq[r + t], tid[r + t], ntid[r + t], flag_bits, min_mapq) &&
                      tid[r + t] >= 0 && pos[r + t] >= 0;
      key[t] = ok ? (((long long)tid[r + t] << 32) | (long long)(unsigned)pos[r + t]) : -1;
    }
    g2::interval_of_n<8>(key, bounds, nb, e);
    for (int t = 0; t < 8; ++t)
      if (e[t] != 0 && e[t] != nb) ++hist[e[t]];
  }
  for (; r < n; ++r) {  // the kernel's tail path
    if (!g2::read_passes(flag[r], mapq[r], tid[r], ntid[r], flag_bits, min_mapq)) continue;
    if (tid[r] < 0 || pos[r] < 0) continue;
    ++hist[g2::interval_of(((long long)tid[r] << 32) | (long long)(unsigned)pos[r], bounds, nb)];
  }
  for (int e = 1; e <= nb; ++e) hist[e] += hist[e - 1];
  for (int k = 0; k < g; ++k) counts[k] = hi[k] > lo[k] ? (long long)(hist[hi[k]] - hist[lo[k]]) : 0;
}

// interval_of_n against interval_of on arbitrary keys
extern "C" int rl_interval_check(const long long* keys, long long n, const long long* bounds, int nb) {
  for (long long r = 0; r + 8 <= n; r += 8) {
    long long key[8];
    int e[8];
    for (int t = 0; t < 8; ++t) key[t] = keys[r + t];
    g2::interval_of_n<8>(key, bounds, nb, e);
    for (int t = 0; t < 8; ++t)
      if (e[t] != g2::interval_of(key[t], bounds, nb)) return 0;
  }
  return 1;
}
