// Device ingest of one sample's decompressed mosdepth text (`chrom\tstart\tend\tdepth\n` per line) into one row
// of the packed fp32 depth matrix, with the line filters of normalize_mosdepth.py:262-285 / :321-352.
//
//   k_ingest_scan   : every thread owns 16 text bytes; a thread parses each line that STARTS inside its chunk
//                     (lines are ~30 bytes, so usually none or one).  A kept line claims its bin with
//                     atomicMax(winner[col], (line offset + 1) << 24 | hundredths): when a file repeats a bin the
//                     LAST line wins, as the reference's dict assignment does (:285), independent of scheduling.
//   k_ingest_finish : winner -> depth_row[col] = hundredths / 100 (correctly rounded fp32; 0.0 = absent).
//
// Anything the strict grammar below does not cover (signs, exponents, more than two decimals, irregular or
// unaligned bins, |depth| >= 41943.04) is COUNTED in stats[1] and skipped; the host then re-parses that file
// itself, so exotic inputs keep the reference's exact behaviour (Python int()/float() semantics).
#include "common.cuh"

namespace g2 {

struct IngestParams {
  char prefix[32];
  int prefix_len;
  int has_win;
  long long win_lo, win_hi;
  const uint32_t* mask;
  long long mask_kb;
  long long bin0, b;
};

enum { ST_KEPT = 0, ST_IRREGULAR = 1, ST_LINES = 2, ST_MASKED = 3, ST_FOREIGN = 4 };

__device__ __forceinline__ bool is_digit(int c) { return c >= '0' && c <= '9'; }
__device__ __forceinline__ bool is_eol(int c) { return c == '\n' || c < 0; }  // c < 0: end of text
// what Python's str.strip() removes at the end of a line (besides the newline itself)
__device__ __forceinline__ bool is_trail_ws(int c) { return c == '\r' || c == ' ' || c == '\v' || c == '\f'; }

struct Cursor {
  const char* text;
  long long n, i;
  __device__ __forceinline__ int peek() const { return i < n ? (int)(unsigned char)text[i] : -1; }
  __device__ __forceinline__ void next() { ++i; }
};

// unsigned decimal integer followed by a tab (the caller has already established that the tab exists)
__device__ bool parse_uint_tab(Cursor& c, long long& out) {
  long long v = 0;
  int nd = 0;
  while (is_digit(c.peek())) {
    v = v * 10 + (c.peek() - '0');
    if (++nd > 13) return false;
    c.next();
  }
  if (nd == 0 || c.peek() != '\t') return false;
  c.next();
  out = v;
  return true;
}

__device__ void parse_line(const char* text, long long n, long long start, const IngestParams& P,
                           unsigned long long* winner, unsigned long long* stats) {
  Cursor c{text, n, start};
  for (int t = 0; t < P.prefix_len; ++t) {  // line.startswith(chrom) -- a PREFIX test on the raw line (:264)
    if (c.peek() != (int)(unsigned char)P.prefix[t]) return;
    c.next();
  }
  // fields = line.strip().split("\t"); fewer than 4 -> `continue`, whatever they contain (:266-268).
  // `stop` = one past the last character strip() keeps; tabs at or after it do not separate fields.
  long long stop = start;
  int tabs_total = 0, tabs = 0;
  {
    Cursor q{text, n, start};
    int ch;
    while (!is_eol(ch = q.peek())) {
      if (ch == '\t') ++tabs_total;
      else if (!is_trail_ws(ch)) stop = q.i + 1, tabs = tabs_total;
      q.next();
    }
  }
  if (stop == start) return;  // blank line
  {
    const int first = (int)(unsigned char)text[start];
    if (first == '\t' || is_trail_ws(first)) {  // leading whitespace is stripped by the reference: leave it to the host
      atomicAdd(&stats[ST_IRREGULAR], 1ULL);
      return;
    }
  }
  atomicAdd(&stats[ST_LINES], 1ULL);
  if (tabs < 3) return;
  // rest of field 0
  bool foreign = false;
  while (c.peek() != '\t') {
    foreign = true;  // the chromosome field is longer than the prefix (chr1 also admits chr10)
    c.next();
  }
  c.next();
  if (foreign && P.prefix_len > 0) atomicAdd(&stats[ST_FOREIGN], 1ULL);
  long long s = 0, e = 0;
  if (!parse_uint_tab(c, s) || !parse_uint_tab(c, e)) {
    atomicAdd(&stats[ST_IRREGULAR], 1ULL);
    return;
  }
  // depth: digits [ '.' digits{0..2} ], ending at a tab (further mosdepth columns) or at the stripped end
  long long ip = 0;
  int nd = 0, nf = 0, frac = 0;
  while (is_digit(c.peek())) {
    ip = ip * 10 + (c.peek() - '0');
    if (++nd > 9) break;
    c.next();
  }
  if (c.peek() == '.') {
    c.next();
    while (is_digit(c.peek())) {
      if (nf < 2) frac = frac * 10 + (c.peek() - '0');
      ++nf;
      c.next();
    }
  }
  const bool ok = nd > 0 && nd <= 9 && nf <= 2 && (c.peek() == '\t' || c.i >= stop);
  if (nf == 1) frac *= 10;
  const long long h = ip * 100 + frac;
  if (!ok || h >= (1LL << 22)) {
    atomicAdd(&stats[ST_IRREGULAR], 1ULL);
    return;
  }
  if (h == 0) return;                                          // depth > 0 (:274, :278)
  if (P.has_win && !(e >= P.win_lo && s <= P.win_hi)) return;  // closed-interval window test (:274-276)
  const long long len = e - s, rel = s - P.bin0;
  if (len <= 0 || len > 1000 || rel < 0 || rel % 1000 != 0 || rel / 1000 >= P.b) {
    atomicAdd(&stats[ST_IRREGULAR], 1ULL);
    return;
  }
  if (P.mask) {  // any masked kb in [s//1000, e//1000] (:281-283)
    for (long long kb = s / 1000; kb <= e / 1000; ++kb) {
      if (kb < P.mask_kb && ((P.mask[kb >> 5] >> (kb & 31)) & 1u)) {
        atomicAdd(&stats[ST_MASKED], 1ULL);
        return;
      }
    }
  }
  atomicMax(&winner[rel / 1000], ((unsigned long long)(start + 1) << 24) | (unsigned long long)h);
  atomicAdd(&stats[ST_KEPT], 1ULL);
}

__global__ void __launch_bounds__(256) k_ingest_scan(const char* __restrict__ text, long long n, IngestParams P,
                                                     unsigned long long* __restrict__ winner,
                                                     unsigned long long* __restrict__ stats) {
  const long long c0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 16;
  if (c0 >= n) return;
  // bit t of `starts` = a line starts at byte c0 + t  (the byte before it is a newline, or it is byte 0)
  unsigned starts = 0;
  if ((c0 + 16 <= n) && ((reinterpret_cast<uintptr_t>(text) & 15) == 0)) {
    const uint4 v = *reinterpret_cast<const uint4*>(text + c0);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int t = 0; t < 15; ++t)
      if (((w[t >> 2] >> (8 * (t & 3))) & 0xff) == '\n') starts |= 1u << (t + 1);
  } else {
    for (int t = 0; t < 15 && c0 + t < n; ++t)
      if (text[c0 + t] == '\n') starts |= 1u << (t + 1);
  }
  if (c0 == 0 || text[c0 - 1] == '\n') starts |= 1u;
  while (starts) {
    const int t = __ffs(starts) - 1;
    starts &= starts - 1;
    if (c0 + t < n) parse_line(text, n, c0 + t, P, winner, stats);
  }
}

__global__ void k_ingest_finish(const unsigned long long* __restrict__ winner, long long b, float* __restrict__ depth_row) {
  const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= b) return;
  const unsigned long long w = winner[j];
  depth_row[j] = w ? __fdiv_rn((float)(unsigned)(w & 0xFFFFFFull), 100.0f) : 0.0f;
}

}  // namespace g2

extern "C" size_t gridb200_ingest_ws_bytes(int64_t b) { return (size_t)(b > 0 ? b : 0) * 8 + 256; }

extern "C" int gridb200_ingest_text(const char* text, int64_t nbytes, const char* chrom_prefix, int32_t chrom_len,
                                    int64_t win_lo, int64_t win_hi, const uint32_t* mask_bits, int64_t mask_kb,
                                    int64_t bin0, int64_t b, float* depth_row, int64_t* stats, void* ws,
                                    size_t ws_bytes, void* stream) {
  using namespace g2;
  G2_REQUIRE(depth_row && stats && b > 0 && nbytes >= 0, GRIDB200_EINVAL, "bad argument");
  G2_REQUIRE(text || nbytes == 0, GRIDB200_EINVAL, "null text");
  G2_REQUIRE(chrom_len >= 0 && chrom_len <= 32 && (chrom_len == 0 || chrom_prefix), GRIDB200_EINVAL,
             "chromosome prefix must be at most 32 bytes");
  G2_REQUIRE(ws && ws_bytes >= gridb200_ingest_ws_bytes(b), GRIDB200_EINVAL, "workspace too small: need %zu bytes",
             gridb200_ingest_ws_bytes(b));
  G2_REQUIRE(bin0 >= 0 && b < (1LL << 31), GRIDB200_EINVAL, "bad bin grid");
  cudaStream_t st = as_stream(stream);
  IngestParams P;
  memset(&P, 0, sizeof(P));
  for (int t = 0; t < chrom_len; ++t) P.prefix[t] = chrom_prefix[t];
  P.prefix_len = chrom_len;
  P.has_win = win_lo >= 0;
  P.win_lo = win_lo, P.win_hi = win_hi;
  P.mask = mask_bits, P.mask_kb = mask_bits ? mask_kb : 0;
  P.bin0 = bin0, P.b = b;
  unsigned long long* winner = static_cast<unsigned long long*>(ws);
  G2_CUDA(cudaMemsetAsync(winner, 0, (size_t)b * 8, st));
  G2_CUDA(cudaMemsetAsync(stats, 0, 8 * sizeof(int64_t), st));
  if (nbytes > 0) {
    const long long chunks = (nbytes + 15) / 16;
    k_ingest_scan<<<(unsigned)((chunks + 255) / 256), 256, 0, st>>>(text, nbytes, P, winner,
                                                                   reinterpret_cast<unsigned long long*>(stats));
    G2_LAUNCHED();
  }
  k_ingest_finish<<<(unsigned)((b + 255) / 256), 256, 0, st>>>(winner, b, depth_row);
  G2_LAUNCHED();
  return 0;
}

// ---- packed host format: uint16 hundredths --------------------------------------------------------------
// mosdepth depths are two-decimal numbers, so a cell is an integer number of hundredths; below 655.35 it fits
// 16 bits.  The host keeps (and PCIe carries) 2 bytes per cell instead of 4; 0xFFFF marks a cell whose value is in
// the escape list (cell index, fp32 value).  Unpacking reproduces the fp32 matrix bit for bit:
// h / 100 correctly rounded, the same conversion k_ingest_finish applies.
namespace g2 {

__global__ void __launch_bounds__(256) k_pack_u16(const float* __restrict__ depth, int64_t n, int64_t b, int64_t ld,
                                                  uint16_t* __restrict__ out, int64_t ldo, int64_t row0_global,
                                                  int64_t* __restrict__ esc_cell, float* __restrict__ esc_val,
                                                  int64_t esc_cap, unsigned long long* __restrict__ esc_count) {
  const int64_t j = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const int64_t i = blockIdx.y;
  if (j >= b) return;
  const float* p = depth + i * ld + j;
  uint16_t* o = out + i * ldo + j;
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    if (j + t >= b) break;
    const float x = p[t];
    const int h = depth_hundredths(x);
    uint16_t code = (uint16_t)h;
    if (h >= 65535 || !(x < 41943.04f)) {  // does not fit (or is beyond what depth_hundredths recovers exactly)
      code = 0xFFFF;
      const unsigned long long slot = atomicAdd(esc_count, 1ULL);
      if ((int64_t)slot < esc_cap) esc_cell[slot] = (row0_global + i) * b + (j + t), esc_val[slot] = x;
    }
    o[t] = code;
  }
}

__global__ void __launch_bounds__(256) k_unpack_u16(const uint16_t* __restrict__ src, int64_t n, int64_t b, int64_t lds,
                                                    float* __restrict__ depth, int64_t ld) {
  const int64_t j = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  const int64_t i = blockIdx.y;
  if (j >= b) return;
  const uint16_t* s = src + i * lds + j;
  float* d = depth + i * ld + j;
  if (j + 8 <= b && ((reinterpret_cast<uintptr_t>(s) & 15) == 0) && ((reinterpret_cast<uintptr_t>(d) & 15) == 0)) {
    const uint4 v = *reinterpret_cast<const uint4*>(s);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    float f[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) f[t] = __fdiv_rn((float)((w[t >> 1] >> (16 * (t & 1))) & 0xFFFFu), 100.0f);
    reinterpret_cast<float4*>(d)[0] = make_float4(f[0], f[1], f[2], f[3]);
    reinterpret_cast<float4*>(d)[1] = make_float4(f[4], f[5], f[6], f[7]);
  } else {
    for (int t = 0; t < 8 && j + t < b; ++t) d[t] = __fdiv_rn((float)s[t], 100.0f);
  }
}

// depth[cell / b - row0][cell % b] = value for the escape cells that fall into rows [row0, row0 + n)
__global__ void k_patch_depth(const int64_t* __restrict__ cell, const float* __restrict__ val, int64_t count, int64_t b,
                              int64_t row0, int64_t n, float* __restrict__ depth, int64_t ld) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= count) return;
  const int64_t r = cell[e] / b - row0, c = cell[e] % b;
  if (r >= 0 && r < n) depth[r * ld + c] = val[e];
}

// ---- bit-packed variant: `bits` (9..15) per cell instead of 16 ---------------------------------------------------------
// A 30x cohort's hundredths sit below 2^13, so 13 bits per cell carry the matrix and PCIe -- what the end-to-end step is
// bound by -- moves 19 % fewer bytes.  Layout: a row is a sequence of groups of 32 cells; a group is `bits` 32-bit words
// holding a little-endian bit stream (cell c of the group at bit c * bits).  The all-ones code is the escape.
//
// Relative codes (row_q / col_q given).  Coverage is, to first order, (sample's overall depth) x (bin's bias): with
// pred(i, j) = (row_q[i] * col_q[j] + 2048) >> 12  (col_q = the bin's mean hundredths, row_q = 4096 x the sample's mean
// over the cohort mean -- integers, so packer and unpacker agree exactly) a present cell is stored as
// h - pred + 2^(bits-1); 0 stays "absent" and all-ones the escape.  What is left is the cell's own noise, a few hundred
// hundredths wide: 11 bits where the absolute value needs 13.
__device__ __forceinline__ int rel_pred(int rq, int cq) { return (int)(((unsigned)rq * (unsigned)cq + 2048u) >> 12); }

__global__ void __launch_bounds__(256) k_pack_bits(const float* __restrict__ depth, int64_t b, int64_t ld, int bits,
                                                   const int* __restrict__ row_q, const int* __restrict__ col_q,
                                                   uint32_t* __restrict__ out, int64_t ldw, int64_t row0_global,
                                                   int64_t* __restrict__ esc_cell, float* __restrict__ esc_val,
                                                   int64_t esc_cap, unsigned long long* __restrict__ esc_count) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // group of 32 cells
  const int64_t i = blockIdx.y;
  if (g * 32 >= b) return;
  const float* p = depth + i * ld + g * 32;
  uint32_t* o = out + i * ldw + g * bits;
  const uint32_t esc = (1u << bits) - 1u;
  unsigned long long acc = 0;
  int nb = 0, w = 0;
  for (int c = 0; c < 32; ++c) {
    uint32_t code = 0;
    if (g * 32 + c < b) {
      const float x = p[c];
      const int h = depth_hundredths(x);
      code = (uint32_t)h;
      bool out_of_range = (uint32_t)h >= esc || !(x < 41943.04f);
      if (row_q && h > 0) {  // relative: the residual against the rank-one prediction, biased to the middle of the range
        const int r = h - rel_pred(row_q[i], col_q[g * 32 + c]) + (1 << (bits - 1));
        out_of_range = r < 1 || r >= (int)esc || h >= 0xFFFF || !(x < 41943.04f);
        code = (uint32_t)r;
      }
      if (out_of_range) {
        code = esc;
        const unsigned long long slot = atomicAdd(esc_count, 1ULL);
        if ((int64_t)slot < esc_cap) esc_cell[slot] = (row0_global + i) * b + (g * 32 + c), esc_val[slot] = x;
      }
    }
    acc |= (unsigned long long)code << nb;
    nb += bits;
    if (nb >= 32) {
      o[w++] = (uint32_t)acc;
      acc >>= 32;
      nb -= 32;
    }
  }
}

// thread = 4 consecutive cells -> one float4 store; the codes are read through L1 (adjacent threads share words)
__global__ void __launch_bounds__(256) k_unpack_bits(const uint32_t* __restrict__ src, int64_t b, int64_t ldw, int bits,
                                                     const int* __restrict__ row_q, const int* __restrict__ col_q,
                                                     float* __restrict__ depth, int64_t ld) {
  const int64_t j = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const int64_t i = blockIdx.y;
  if (j >= b) return;
  const uint32_t* s = src + i * ldw + (j >> 5) * bits;
  const uint32_t mask = (1u << bits) - 1u;
  float f[4];
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const int bit = (int)((j + t) & 31) * bits;   // 4 | 32: the four cells never leave the group
    const int w = bit >> 5, sh = bit & 31;
    uint32_t v = __ldg(s + w) >> sh;
    if (sh + bits > 32) v |= __ldg(s + w + 1) << (32 - sh);
    v &= mask;
    if (row_q && v != 0 && v != mask && j + t < b) v = (uint32_t)(rel_pred(row_q[i], col_q[j + t]) + (int)v - (1 << (bits - 1)));
    f[t] = __fdiv_rn((float)v, 100.0f);   // (the escape code is overwritten by k_patch_depth)
  }
  float* d = depth + i * ld + j;
  if (j + 4 <= b && (reinterpret_cast<uintptr_t>(d) & 15) == 0) *reinterpret_cast<float4*>(d) = make_float4(f[0], f[1], f[2], f[3]);
  else
    for (int t = 0; t < 4 && j + t < b; ++t) d[t] = f[t];
}

// bit-packed codes -> the 2-byte DEVICE matrix (uint16 hundredths; the width's escape code becomes 0xFFFF): what the
// step-4 kernels read when the matrix is kept in 2 bytes per cell (gridb200_*_u16)
__global__ void __launch_bounds__(256) k_unpack_bits_u16(const uint32_t* __restrict__ src, int64_t b, int64_t ldw, int bits,
                                                         const int* __restrict__ row_q, const int* __restrict__ col_q,
                                                         uint16_t* __restrict__ out, int64_t ldo) {
  const int64_t j = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const int64_t i = blockIdx.y;
  if (j >= b) return;
  const uint32_t* s = src + i * ldw + (j >> 5) * bits;
  const uint32_t mask = (1u << bits) - 1u;
  uint32_t c[4];
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const int bit = (int)((j + t) & 31) * bits;
    const int w = bit >> 5, sh = bit & 31;
    uint32_t v = __ldg(s + w) >> sh;
    if (sh + bits > 32) v |= __ldg(s + w + 1) << (32 - sh);
    v &= mask;
    const bool is_esc = v == mask;
    if (row_q && v != 0 && !is_esc && j + t < b) v = (uint32_t)(rel_pred(row_q[i], col_q[j + t]) + (int)v - (1 << (bits - 1)));
    c[t] = (j + t < b) ? (is_esc ? 0xFFFFu : (v & 0xFFFFu)) : 0u;
  }
  *reinterpret_cast<uint2*>(out + i * ldo + j) = make_uint2(c[0] | (c[1] << 16), c[2] | (c[3] << 16));  // ldo % 4 == 0
}

}  // namespace g2

extern "C" int gridb200_depth_unpack_bits_u16(const uint32_t* src, int64_t n, int64_t b, int64_t ldw, int32_t bits,
                                              const int32_t* row_q, const int32_t* col_q, uint16_t* out, int64_t ldo,
                                              void* stream) {
  using namespace g2;
  G2_REQUIRE(src && out && n > 0 && b > 0 && n <= 65535, GRIDB200_EINVAL, "bad argument (rows per call <= 65535)");
  G2_REQUIRE((row_q == nullptr) == (col_q == nullptr), GRIDB200_EINVAL, "relative codes need both row_q and col_q");
  G2_REQUIRE(bits >= 9 && bits <= 15 && ldw >= ((b + 31) / 32 * bits), GRIDB200_EINVAL, "bits must be 9..15 and ldw cover whole groups");
  G2_REQUIRE(ldo >= (b + 3) / 4 * 4 && ldo % 4 == 0 && (reinterpret_cast<uintptr_t>(out) & 7) == 0, GRIDB200_EINVAL,
             "output rows must be 8-byte aligned with ldo a multiple of 4 and >= roundup4(b)");
  k_unpack_bits_u16<<<dim3((unsigned)((b + 1023) / 1024), (unsigned)n), 256, 0, as_stream(stream)>>>(src, b, ldw, bits, row_q,
                                                                                                     col_q, out, ldo);
  G2_LAUNCHED();
  return 0;
}

extern "C" int64_t gridb200_depth_packed_row_words(int64_t b, int32_t bits) {
  if (b <= 0 || bits < 9 || bits > 15) return -1;
  return ((b + 31) / 32 * bits + 3) / 4 * 4;  // whole groups, rows 16-byte aligned
}

extern "C" int gridb200_depth_pack_bits(const float* depth, int64_t n, int64_t b, int64_t ld, int32_t bits,
                                        const int32_t* row_q, const int32_t* col_q, uint32_t* out, int64_t ldw, int64_t row0,
                                        int64_t* esc_cell, float* esc_val, int64_t esc_cap, int64_t* esc_count, void* stream) {
  using namespace g2;
  G2_REQUIRE((row_q == nullptr) == (col_q == nullptr), GRIDB200_EINVAL, "relative codes need both row_q and col_q");
  G2_REQUIRE(depth && out && esc_count && n > 0 && b > 0 && ld >= b && n <= 65535, GRIDB200_EINVAL,
             "bad argument (rows per call <= 65535)");
  G2_REQUIRE(bits >= 9 && bits <= 15 && ldw >= gridb200_depth_packed_row_words(b, bits), GRIDB200_EINVAL,
             "bits must be 9..15 and ldw at least gridb200_depth_packed_row_words(b, bits)");
  G2_REQUIRE(esc_cap == 0 || (esc_cell && esc_val), GRIDB200_EINVAL, "escape arrays missing");
  const int64_t groups = (b + 31) / 32;
  k_pack_bits<<<dim3((unsigned)((groups + 255) / 256), (unsigned)n), 256, 0, as_stream(stream)>>>(
      depth, b, ld, bits, row_q, col_q, out, ldw, row0, esc_cell, esc_val, esc_cap,
      reinterpret_cast<unsigned long long*>(esc_count));
  G2_LAUNCHED();
  return 0;
}

extern "C" int gridb200_depth_unpack_bits(const uint32_t* src, int64_t n, int64_t b, int64_t ldw, int32_t bits,
                                          const int32_t* row_q, const int32_t* col_q, float* depth, int64_t ld,
                                          const int64_t* esc_cell, const float* esc_val, int64_t esc_count, int64_t row0,
                                          void* stream) {
  using namespace g2;
  G2_REQUIRE((row_q == nullptr) == (col_q == nullptr), GRIDB200_EINVAL, "relative codes need both row_q and col_q");
  G2_REQUIRE(src && depth && n > 0 && b > 0 && ld >= b && n <= 65535, GRIDB200_EINVAL, "bad argument (rows per call <= 65535)");
  G2_REQUIRE(bits >= 9 && bits <= 15 && ldw >= gridb200_depth_packed_row_words(b, bits), GRIDB200_EINVAL,
             "bits must be 9..15 and ldw at least gridb200_depth_packed_row_words(b, bits)");
  G2_REQUIRE(esc_count == 0 || (esc_cell && esc_val), GRIDB200_EINVAL, "escape arrays missing");
  cudaStream_t st = as_stream(stream);
  k_unpack_bits<<<dim3((unsigned)((b + 1023) / 1024), (unsigned)n), 256, 0, st>>>(src, b, ldw, bits, row_q, col_q, depth, ld);
  G2_LAUNCHED();
  if (esc_count > 0) {
    k_patch_depth<<<(unsigned)((esc_count + 255) / 256), 256, 0, st>>>(esc_cell, esc_val, esc_count, b, row0, n, depth, ld);
    G2_LAUNCHED();
  }
  return 0;
}

extern "C" int gridb200_depth_pack_u16(const float* depth, int64_t n, int64_t b, int64_t ld, uint16_t* out, int64_t ldo,
                                       int64_t row0, int64_t* esc_cell, float* esc_val, int64_t esc_cap,
                                       int64_t* esc_count, void* stream) {
  using namespace g2;
  G2_REQUIRE(depth && out && esc_count && n > 0 && b > 0 && ld >= b && ldo >= b && n <= 65535, GRIDB200_EINVAL,
             "bad argument (rows per call <= 65535)");
  G2_REQUIRE(esc_cap == 0 || (esc_cell && esc_val), GRIDB200_EINVAL, "escape arrays missing");
  k_pack_u16<<<dim3((unsigned)((b + 1023) / 1024), (unsigned)n), 256, 0, as_stream(stream)>>>(
      depth, n, b, ld, out, ldo, row0, esc_cell, esc_val, esc_cap, reinterpret_cast<unsigned long long*>(esc_count));
  G2_LAUNCHED();
  return 0;
}

extern "C" int gridb200_depth_unpack_u16(const uint16_t* src, int64_t n, int64_t b, int64_t lds, float* depth,
                                         int64_t ld, const int64_t* esc_cell, const float* esc_val, int64_t esc_count,
                                         int64_t row0, void* stream) {
  using namespace g2;
  G2_REQUIRE(src && depth && n > 0 && b > 0 && lds >= b && ld >= b && n <= 65535, GRIDB200_EINVAL,
             "bad argument (rows per call <= 65535)");
  G2_REQUIRE(esc_count == 0 || (esc_cell && esc_val), GRIDB200_EINVAL, "escape arrays missing");
  cudaStream_t st = as_stream(stream);
  k_unpack_u16<<<dim3((unsigned)((b + 2047) / 2048), (unsigned)n), 256, 0, st>>>(src, n, b, lds, depth, ld);
  G2_LAUNCHED();
  if (esc_count > 0) {
    k_patch_depth<<<(unsigned)((esc_count + 255) / 256), 256, 0, st>>>(esc_cell, esc_val, esc_count, b, row0, n, depth, ld);
    G2_LAUNCHED();
  }
  return 0;
}
