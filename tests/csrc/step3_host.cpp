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
                               const int32_t* first, int shift, int per_tid, int n_tid, long long* counts) {
  std::vector<uint32_t> hist(nb + 1, 0);
  const g2::BucketDir D{first, shift, per_tid, n_tid};
  for (long long r = 0; r < n; ++r) {  // the body of k_read_hist's take()
    if (!g2::read_passes(flag[r], mapq[r], tid[r], ntid[r], flag_bits, min_mapq)) continue;
    if (tid[r] < 0 || pos[r] < 0) continue;
    const int e = D.first ? g2::interval_by_bucket(tid[r], pos[r], bounds, D)
                          : g2::interval_of(((long long)tid[r] << 32) | (long long)(unsigned)pos[r], bounds, nb);
    if (e <= 0 || e >= nb) continue;
    ++hist[e];
  }
  for (int e = 1; e <= nb; ++e) hist[e] += hist[e - 1];
  for (int k = 0; k < g; ++k) counts[k] = hi[k] > lo[k] ? (long long)(hist[hi[k]] - hist[lo[k]]) : 0;
}

// interval_by_bucket against interval_of on arbitrary positions: equal wherever the directory answers, and it may
// decline (-1) only beyond every boundary of that reference
extern "C" int rl_bucket_check(const int32_t* tid, const int32_t* pos, long long n, const long long* bounds, int nb,
                               const int32_t* first, int shift, int per_tid, int n_tid) {
  const g2::BucketDir D{first, shift, per_tid, n_tid};
  for (long long r = 0; r < n; ++r) {
    const long long key = ((long long)tid[r] << 32) | (long long)(unsigned)pos[r];
    const int want = g2::interval_of(key, bounds, nb), got = g2::interval_by_bucket(tid[r], pos[r], bounds, D);
    if (got >= 0 ? got != want : !(want == nb || (want < nb && (bounds[want] >> 32) > tid[r]))) return 0;
  }
  return 1;
}
