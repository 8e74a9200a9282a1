// Per-line grammar of gridb200_region_coverage and the per-read predicate / interval lookup of gridb200_count_reads.
// Plain C++ with no CUDA dependency: the kernels in region.cu include it as device code, and tests/test_step3_host.py
// compiles the SAME functions with g++ to fuzz them against Python's str.strip()/split()/int()/float() rules on the CPU.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define G2_HD __host__ __device__ __forceinline__
#else
#define G2_HD inline
#endif

namespace g2 {

struct RegionParams {
  char chrom[64];
  int chrom_len;
  long long start, end;
};

enum { RC_WEIGHTED = 0, RC_COVERED = 1, RC_LINES = 2, RC_IRREGULAR = 3 };
enum { RC_LINE_SKIP = 0, RC_LINE_COUNTS = 1, RC_LINE_IRREGULAR = 2 };

struct LineResult {
  int status;
  long long h, ov;  // depth in hundredths, overlap in bp (status == RC_LINE_COUNTS)
};

G2_HD int rc_peek(const char* text, long long n, long long i) { return i < n ? (int)(unsigned char)text[i] : -1; }
G2_HD bool rc_digit(int c) { return c >= '0' && c <= '9'; }
G2_HD bool rc_blank(int c) { return c == ' ' || c == '\v' || c == '\f'; }  // ASCII blanks strip() removes (no \r: host)

// digits followed by a tab; at most 10 digits, so coordinates stay below 2^34
G2_HD bool rc_uint_tab(const char* text, long long n, long long& i, long long& out) {
  long long v = 0;
  int nd = 0, c;
  while (rc_digit(c = rc_peek(text, n, i))) {
    v = v * 10 + (c - '0');
    if (++nd > 10) return false;
    ++i;
  }
  if (nd == 0 || c != '\t') return false;
  ++i;
  out = v;
  return true;
}

// One line `chrom\tstart\tend\tdepth[\t...]` starting at text[i]; the text ends at n.  The reference converts start,
// end and depth of EVERY line before it looks at the chromosome (mosdepth.py:283-288), so a line outside this grammar
// is reported whatever its chromosome; the host replays such a file with Python semantics.
//   chrom : one or more bytes, the first a printable ASCII character (strip() eats leading blanks, also exotic ones)
//   start, end : 1-10 decimal digits
//   depth : 1-9 digits, optionally '.' and 0-2 digits, below 41943.04; then a tab, or blanks up to the newline
G2_HD LineResult rc_parse_line(const char* text, long long n, long long i, const RegionParams& P) {
  LineResult res = {RC_LINE_IRREGULAR, 0, 0};
  const int first = rc_peek(text, n, i);
  if (first < 0x21 || first > 0x7e) return res;  // blank line, leading blank, control or non-ASCII byte
  bool same = true;
  int len0 = 0, c;
  while ((c = rc_peek(text, n, i)) != '\t') {
    if (c == '\n' || c < 0) return res;  // fewer than four fields
    same = same && len0 < P.chrom_len && c == (int)(unsigned char)P.chrom[len0];
    ++len0, ++i;
  }
  ++i;
  same = same && len0 == P.chrom_len;
  long long s = 0, e = 0;
  if (!rc_uint_tab(text, n, i, s) || !rc_uint_tab(text, n, i, e)) return res;
  long long ip = 0;
  int nd = 0, nf = 0, frac = 0;
  while (rc_digit(c = rc_peek(text, n, i))) {
    ip = ip * 10 + (c - '0');
    if (++nd > 9) return res;
    ++i;
  }
  if (c == '.') {
    ++i;
    while (rc_digit(c = rc_peek(text, n, i))) {
      if (nf < 2) frac = frac * 10 + (c - '0');
      ++nf, ++i;
    }
  }
  if (nd == 0 || nf > 2) return res;
  if (c != '\t') {
    while (rc_blank(c = rc_peek(text, n, i))) ++i;
    if (c != '\n' && c >= 0) return res;
  }
  if (nf == 1) frac *= 10;
  const long long h = ip * 100 + frac;
  if (h >= (1LL << 22)) return res;
  res.status = RC_LINE_SKIP;
  if (!same) return res;                                                       // r_chr != chrom (:287)
  const long long ov = (P.end < e ? P.end : e) - (P.start > s ? P.start : s);  // :290-292
  if (ov <= 0) return res;
  res.status = RC_LINE_COUNTS, res.h = h, res.ov = ov;
  return res;
}

// count_reads.py:96-100: flag in proper_flags (an exact-value test: bit `flag` of a 65536-bit table -- 8 KB that stay in
// L1 once the read columns are streamed past it; comparing a handful of flags in registers instead cost more
// instructions and was 6 % slower), mapq >= min_mapq, mate on the same reference, not a duplicate (0x400), not
// secondary (0x100); the position test is the interval lookup below.
G2_HD bool read_passes(unsigned flag, int mapq, int tid, int ntid, const uint32_t* flag_bits, int min_mapq) {
  return ((flag_bits[flag >> 5] >> (flag & 31)) & 1u) && mapq >= min_mapq && tid == ntid && !(flag & 0x400u) &&
         !(flag & 0x100u);
}

// number of boundaries <= key: the elementary interval a position falls into (0 = before every region)
G2_HD int interval_of(long long key, const long long* bounds, int nb) {
  int lo = 0, hi = nb;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (bounds[mid] <= key) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

// Coarse directory over the boundary table: bucket b = tid * per_tid + (pos >> shift) holds first[b] = number of
// boundaries below the bucket's first key, so a position's interval is first[b] plus a search among the (usually zero
// or one) boundaries that fall inside its bucket: one 8-byte load instead of log2(nb) dependent ones.
// Returns -1 for a position no region can contain (reference beyond the last region's, or beyond the directory).
struct BucketDir {
  const int32_t* first;  // n_tid * per_tid + 1 entries
  int shift, per_tid, n_tid;
};

G2_HD int interval_by_bucket(int tid, int pos, const long long* bounds, const BucketDir& D) {
  const int col = pos >> D.shift;
  if (tid >= D.n_tid || col >= D.per_tid) return -1;
  const long long b = (long long)tid * D.per_tid + col;
  const int lo = D.first[b], hi = D.first[b + 1];
  const long long key = ((long long)tid << 32) | (long long)(unsigned)pos;
  return lo + interval_of(key, bounds + lo, hi - lo);
}

}  // namespace g2
