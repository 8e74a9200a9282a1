// Step-4 hand-off file on the device: the text body of write_normalized_output
// (normalize_mosdepth.py:502-554) from the int16 hundredths codes, and its inverse for read_normalized_data
// (find_neighbors.py:81-124).  The host keeps what is small (ids, scales, the two header lines, gzip); the
// N x Rw cells -- 8.4e9 of them at config 2 -- never pass through Python objects.
//
//   format : k_fmt_len   per 2048-cell chunk: bytes of text                         (grid: chunks x rows)
//            k_chunk_scan per row: exclusive scan over its chunks, row length        (grid: rows)
//            k_row_scan   exclusive scan over the rows of the batch -> row_off       (one block)
//            k_fmt_write  chunk text built in shared memory at the destination's 16-byte phase, copied out with
//                         128-bit stores                                             (grid: chunks x rows)
//   parse  : k_parse_count tabs per 4096-byte chunk of a line; k_chunk_scan; k_parse_cells: a thread owns 16 bytes
//            and decodes every field that STARTS inside them (field index = tabs before it).
//
// A cell is "NA", "-0.00" or [-]d{1,3}.dd -- exactly what "%.2f" produces for |z| <= 327.66.  The parser accepts
// only that grammar (plus "nan"); anything else is counted and the host re-reads the file with Python semantics.
#include "common.cuh"

namespace g2 {
namespace textio {

constexpr int ZQ_NA = -32768, ZQ_NEG_ZERO = -32767, ZQ_LIMIT = 32766;
constexpr int THREADS = 256;
constexpr int CELLS = 8, CHUNK_CELLS = THREADS * CELLS;     // formatter: cells per thread / per block
constexpr int BYTES = 16, CHUNK_BYTES = THREADS * BYTES;    // parser: text bytes per thread / per block

// bytes of one cell including its separator
__device__ __forceinline__ int cell_len(int code) {
  if (code == ZQ_NA) return 3;
  if (code == ZQ_NEG_ZERO) return 6;
  const int a = abs(code);
  return (code < 0 ? 1 : 0) + (a >= 10000 ? 3 : (a >= 1000 ? 2 : 1)) + 4;
}

__device__ __forceinline__ int cell_put(int code, char sep, char* d) {
  int n = 0;
  if (code == ZQ_NA) {
    d[0] = 'N', d[1] = 'A', n = 2;
  } else {
    int a = code == ZQ_NEG_ZERO ? 0 : abs(code);
    if (code < 0) d[n++] = '-';  // ZQ_NEG_ZERO is negative too
    const int ip = a / 100, fr = a - ip * 100;
    if (ip >= 100) d[n++] = (char)('0' + ip / 100);
    if (ip >= 10) d[n++] = (char)('0' + (ip / 10) % 10);
    d[n++] = (char)('0' + ip % 10);
    d[n++] = '.';
    d[n++] = (char)('0' + fr / 10);
    d[n++] = (char)('0' + fr % 10);
  }
  d[n++] = sep;
  return n;
}

// the CELLS codes of one thread (ZQ_NA beyond the row is never used: `cnt` says how many are real)
__device__ __forceinline__ int load_cells(const int16_t* __restrict__ row, int64_t c0, int64_t rw, bool vec, int* code) {
  const int cnt = (int)max((int64_t)0, min((int64_t)CELLS, rw - c0));
  if (vec && cnt == CELLS) {
    const uint4 v = *reinterpret_cast<const uint4*>(row + c0);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < CELLS; ++i) code[i] = (int)(int16_t)(w[i >> 1] >> (16 * (i & 1)));
  } else {
#pragma unroll
    for (int i = 0; i < CELLS; ++i) code[i] = i < cnt ? (int)row[c0 + i] : 0;
  }
  return cnt;
}

// exclusive scan of one int per thread over the block; *total = block sum
__device__ __forceinline__ int block_excl_scan(int v, int* total) {
  __shared__ int s_warp[THREADS / 32];
  __shared__ int s_total;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int w = lane < THREADS / 32 ? s_warp[lane] : 0;
    int winc = w;
#pragma unroll
    for (int o = 1; o < THREADS / 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += t;
    }
    if (lane < THREADS / 32) s_warp[lane] = winc - w;
    if (lane == THREADS / 32 - 1) s_total = winc;
  }
  __syncthreads();
  *total = s_total;
  return s_warp[warp] + inc - v;
}

__global__ void __launch_bounds__(THREADS) k_fmt_len(const int16_t* __restrict__ zq, int64_t rw, int64_t ldq, bool vec,
                                                     int nchunks, uint32_t* __restrict__ chunk_len) {
  const int64_t r = blockIdx.y;
  const int64_t c0 = ((int64_t)blockIdx.x * THREADS + threadIdx.x) * CELLS;
  int code[CELLS];
  const int cnt = load_cells(zq + r * ldq, c0, rw, vec, code);
  int len = 0;
#pragma unroll
  for (int i = 0; i < CELLS; ++i)
    if (i < cnt) len += cell_len(code[i]);
  int total;
  block_excl_scan(len, &total);
  if (threadIdx.x == 0) chunk_len[r * nchunks + blockIdx.x] = (uint32_t)total;
}

// in place: vals[row][0..nchunks) -> exclusive prefix sums; row_total[row] = sum
__global__ void __launch_bounds__(THREADS) k_chunk_scan(uint32_t* __restrict__ vals, int nchunks,
                                                        int64_t* __restrict__ row_total) {
  uint32_t* v = vals + (int64_t)blockIdx.x * nchunks;
  int carry = 0;
  for (int base = 0; base < nchunks; base += THREADS) {
    const int i = base + threadIdx.x;
    const int x = i < nchunks ? (int)v[i] : 0;
    int total;
    const int ex = block_excl_scan(x, &total);
    if (i < nchunks) v[i] = (uint32_t)(carry + ex);
    carry += total;
    __syncthreads();  // block_excl_scan's shared scratch is reused by the next round
  }
  if (threadIdx.x == 0) row_total[blockIdx.x] = carry;
}

// row_off[0..rows] = exclusive scan of row_total[0..rows)   (one block; rows per batch is modest)
__global__ void __launch_bounds__(THREADS) k_row_scan(const int64_t* __restrict__ row_total, int64_t rows,
                                                      int64_t* __restrict__ row_off) {
  __shared__ long long s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int64_t base = 0; base < rows; base += THREADS) {
    const int64_t i = base + threadIdx.x;
    const int x = i < rows ? (int)row_total[i] : 0;  // one row of text is < 2^31 bytes (checked by the host wrapper)
    int total;
    const int ex = block_excl_scan(x, &total);
    const long long carry = s_carry;
    if (i < rows) row_off[i] = carry + ex;
    __syncthreads();
    if (threadIdx.x == 0) s_carry = carry + total;
    __syncthreads();
  }
  if (threadIdx.x == 0) row_off[rows] = s_carry;
}

__global__ void __launch_bounds__(THREADS) k_fmt_write(const int16_t* __restrict__ zq, int64_t rw, int64_t ldq, bool vec,
                                                       int nchunks, const uint32_t* __restrict__ chunk_off,
                                                       const int64_t* __restrict__ row_off, char* __restrict__ text,
                                                       int64_t cap) {
  __shared__ __align__(16) char s_txt[16 + CHUNK_CELLS * 8 + 16];
  const int64_t r = blockIdx.y;
  const int64_t c0 = ((int64_t)blockIdx.x * THREADS + threadIdx.x) * CELLS;
  int code[CELLS];
  const int cnt = load_cells(zq + r * ldq, c0, rw, vec, code);
  int len = 0;
#pragma unroll
  for (int i = 0; i < CELLS; ++i)
    if (i < cnt) len += cell_len(code[i]);
  int total;
  int off = block_excl_scan(len, &total);
  const int64_t dst0 = row_off[r] + chunk_off[r * nchunks + blockIdx.x];
  if (dst0 + total > cap) return;  // the wrapper sized `cap` from row_off; never taken when it did
  char* dst = text + dst0;
  const int shift = (int)(reinterpret_cast<uintptr_t>(dst) & 15);  // build the text at the destination's phase
  char* p = s_txt + shift + off;
#pragma unroll
  for (int i = 0; i < CELLS; ++i)
    if (i < cnt) p += cell_put(code[i], (c0 + i == rw - 1) ? '\n' : '\t', p);
  __syncthreads();
  const int head = min(total, (16 - shift) & 15);
  if ((int)threadIdx.x < head) dst[threadIdx.x] = s_txt[shift + threadIdx.x];
  const int nvec = (total - head) >> 4;
  const uint4* sv = reinterpret_cast<const uint4*>(s_txt + shift + head);
  uint4* dv = reinterpret_cast<uint4*>(dst + head);
  for (int i = threadIdx.x; i < nvec; i += THREADS) dv[i] = sv[i];
  const int done = head + (nvec << 4);
  if ((int)threadIdx.x < total - done) dst[done + threadIdx.x] = s_txt[shift + done + threadIdx.x];
}

// ---- parser ---------------------------------------------------------------------------------------------
enum { PS_CELLS = 0, PS_BAD_TOKENS = 1, PS_BAD_ROWS = 2 };

__device__ __forceinline__ int byte_at(const char* __restrict__ text, int64_t p, int64_t end) {
  return p < end ? (int)(unsigned char)text[p] : '\n';
}

__global__ void __launch_bounds__(THREADS) k_parse_count(const char* __restrict__ text,
                                                         const int64_t* __restrict__ line_off, int nchunks,
                                                         uint32_t* __restrict__ chunk_tabs) {
  const int64_t r = blockIdx.y;
  const int64_t ls = line_off[r], le = line_off[r + 1];
  const int64_t p0 = ls + (int64_t)blockIdx.x * CHUNK_BYTES + threadIdx.x * BYTES;
  int tabs = 0;
  if (p0 < le) {
#pragma unroll
    for (int i = 0; i < BYTES; ++i) tabs += (p0 + i < le && text[p0 + i] == '\t') ? 1 : 0;
  }
  int total;
  block_excl_scan(tabs, &total);
  if (threadIdx.x == 0) chunk_tabs[r * nchunks + blockIdx.x] = (uint32_t)total;
}

// One token starting at p (ends at a tab, a newline or the end of the line) -> code; false if outside the grammar.
__device__ __forceinline__ bool parse_cell(const char* __restrict__ text, int64_t p, int64_t end, int* code) {
  int c = byte_at(text, p, end);
  if (c == 'N' || c == 'n') {  // "NA" | "nan"   (find_neighbors.py: `t in ("NA", "nan")`)
    const int c1 = byte_at(text, p + 1, end), c2 = byte_at(text, p + 2, end), c3 = byte_at(text, p + 3, end);
    const bool na = c == 'N' && c1 == 'A' && (c2 == '\t' || c2 == '\n');
    const bool nan3 = c == 'n' && c1 == 'a' && c2 == 'n' && (c3 == '\t' || c3 == '\n');
    *code = ZQ_NA;
    return na || nan3;
  }
  bool neg = false;
  if (c == '-') neg = true, c = byte_at(text, ++p, end);
  int ip = 0, nd = 0;
  while (c >= '0' && c <= '9' && nd < 4) ip = ip * 10 + (c - '0'), ++nd, c = byte_at(text, ++p, end);
  if (nd < 1 || nd > 3 || c != '.') return false;
  const int d1 = byte_at(text, p + 1, end), d2 = byte_at(text, p + 2, end), t = byte_at(text, p + 3, end);
  if (d1 < '0' || d1 > '9' || d2 < '0' || d2 > '9' || (t != '\t' && t != '\n')) return false;
  const int v = ip * 100 + (d1 - '0') * 10 + (d2 - '0');
  if (v > ZQ_LIMIT) return false;
  *code = neg ? -v : v;  // "-0.00" reads back as 0, as float() * 100 -> rint does
  return true;
}

__global__ void __launch_bounds__(THREADS) k_parse_cells(const char* __restrict__ text,
                                                         const int64_t* __restrict__ line_off, int nchunks,
                                                         const uint32_t* __restrict__ chunk_tabs, int skip, int64_t rw,
                                                         int16_t* __restrict__ zq, int64_t ldq,
                                                         unsigned long long* __restrict__ stats) {
  const int64_t r = blockIdx.y;
  const int64_t ls = line_off[r], le = line_off[r + 1];
  const int64_t p0 = ls + (int64_t)blockIdx.x * CHUNK_BYTES + threadIdx.x * BYTES;
  // bit i of `starts`: a field starts at p0 + i;  tabs in my bytes
  unsigned starts = 0;
  int tabs = 0;
  if (p0 < le) {
    int prev = p0 == ls ? '\t' : (int)(unsigned char)text[p0 - 1];
#pragma unroll
    for (int i = 0; i < BYTES; ++i) {
      const int c = p0 + i < le ? (int)(unsigned char)text[p0 + i] : '\n';
      if (prev == '\t' && p0 + i < le) starts |= 1u << i;
      tabs += (c == '\t' && p0 + i < le) ? 1 : 0;
      prev = c;
    }
  }
  int total;
  const int before = block_excl_scan(tabs, &total);
  if (!starts) return;
  int field = (int)chunk_tabs[r * nchunks + blockIdx.x] + before;  // index of the first field starting in my bytes...
  // ...provided no tab of mine precedes it; walk my bytes again, counting tabs as they pass
  unsigned long long good = 0, bad = 0;
  for (int i = 0; i < BYTES; ++i) {
    if (p0 + i >= le) break;
    if (starts & (1u << i)) {
      const int64_t col = (int64_t)field - skip;
      if (col >= 0 && col < rw) {
        int code;
        if (parse_cell(text, p0 + i, le, &code)) zq[r * ldq + col] = (int16_t)code, ++good;
        else ++bad;
      }
    }
    if (text[p0 + i] == '\t') ++field;
  }
  if (good) atomicAdd(&stats[PS_CELLS], good);
  if (bad) atomicAdd(&stats[PS_BAD_TOKENS], bad);
}

// a row must have exactly skip + rw fields
__global__ void k_parse_check_rows(const int64_t* __restrict__ row_tabs, int64_t rows, int64_t want_tabs,
                                   unsigned long long* __restrict__ stats) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < rows && row_tabs[r] != want_tabs) atomicAdd(&stats[PS_BAD_ROWS], 1ULL);
}


// ---- step-5 file, reading side (load_neighbors, compute_dipcn.py:105-152) ------------------------------------------------
// A line is  id \t scale ( \t nbr_id \t nbr_scale \t dist )*.  In a file step 5 wrote, every "nbr_id\tnbr_scale" is byte for
// byte the "id\tscale" head of some line, so neighbours are resolved to row indices by hashing that token and comparing
// the bytes; a token that is nobody's head, a line whose field count is not 2 + 3m, or more entries than the caller
// allowed are counted and the host re-reads the file with the reference's rules.
enum { PN_ENTRIES = 0, PN_UNRESOLVED = 1, PN_BAD_ROWS = 2 };

__device__ __forceinline__ unsigned long long tok_hash_step(unsigned long long h, int c) {
  return (h ^ (unsigned long long)c) * 0x100000001B3ULL;  // FNV-1a
}
__device__ __forceinline__ long long tok_hash_finish(unsigned long long h) {
  h ^= h >> 30, h *= 0xBF58476D1CE4E5B9ULL, h ^= h >> 27, h *= 0x94D049BB133111EBULL, h ^= h >> 31;
  return (long long)h;
}

__global__ void k_tok_hash(const char* __restrict__ tok, const int64_t* __restrict__ tok_off, int64_t n,
                           long long* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  unsigned long long h = 0xCBF29CE484222325ULL;
  for (int64_t p = tok_off[i]; p < tok_off[i + 1]; ++p) h = tok_hash_step(h, (unsigned char)tok[p]);
  out[i] = tok_hash_finish(h);
}

struct NbrTable {
  const char* tok;
  const int64_t* tok_off;
  const long long* hash_sorted;  // ascending (signed)
  const int32_t* hash_row;       // row whose head has hash_sorted[i]
  int64_t n;
};

// the entry token starting at p: "id \t scale" up to (not including) the second tab, the newline or the end of the line
__device__ __forceinline__ int resolve_entry(const char* __restrict__ text, int64_t p, int64_t end, const NbrTable& T) {
  unsigned long long h = 0xCBF29CE484222325ULL;
  int tabs = 0;
  int64_t q = p;
  for (; q < end; ++q) {
    const int c = (unsigned char)text[q];
    if (c == '\n' || (c == '\t' && ++tabs == 2)) break;
    h = tok_hash_step(h, c);
  }
  if (tabs < 1) return -1;  // an id without a scale
  const long long key = tok_hash_finish(h);
  int64_t lo = 0, hi = T.n;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (T.hash_sorted[mid] < key) lo = mid + 1;
    else hi = mid;
  }
  if (lo >= T.n || T.hash_sorted[lo] != key) return -1;
  const int row = T.hash_row[lo];
  const int64_t t0 = T.tok_off[row], tl = T.tok_off[row + 1] - t0;
  if (tl != q - p) return -1;
  for (int64_t t = 0; t < tl; ++t)
    if (T.tok[t0 + t] != text[p + t]) return -1;
  return row;
}

__global__ void __launch_bounds__(THREADS) k_parse_nbr_entries(const char* __restrict__ text,
                                                               const int64_t* __restrict__ line_off, int nchunks,
                                                               const uint32_t* __restrict__ chunk_tabs, NbrTable T,
                                                               int32_t k_cap, int32_t* __restrict__ idx, int64_t ld,
                                                               unsigned long long* __restrict__ stats) {
  const int64_t r = blockIdx.y;
  const int64_t ls = line_off[r], le = line_off[r + 1];
  const int64_t p0 = ls + (int64_t)blockIdx.x * CHUNK_BYTES + threadIdx.x * BYTES;
  unsigned starts = 0;
  int tabs = 0;
  if (p0 < le) {
    int prev = p0 == ls ? '\t' : (int)(unsigned char)text[p0 - 1];
#pragma unroll
    for (int i = 0; i < BYTES; ++i) {
      const int c = p0 + i < le ? (int)(unsigned char)text[p0 + i] : '\n';
      if (prev == '\t' && p0 + i < le) starts |= 1u << i;
      tabs += (c == '\t' && p0 + i < le) ? 1 : 0;
      prev = c;
    }
  }
  int total;
  const int before = block_excl_scan(tabs, &total);
  if (!starts) return;
  int field = (int)chunk_tabs[r * nchunks + blockIdx.x] + before;
  unsigned long long good = 0, bad = 0;
  for (int i = 0; i < BYTES; ++i) {
    if (p0 + i >= le) break;
    if ((starts & (1u << i)) && field >= 2 && (field - 2) % 3 == 0) {
      const int e = (field - 2) / 3;
      const int row = e < k_cap ? resolve_entry(text, p0 + i, le, T) : -1;
      if (row >= 0) idx[r * ld + e] = row, ++good;
      else ++bad;
    }
    if (text[p0 + i] == '\t') ++field;
  }
  if (good) atomicAdd(&stats[PN_ENTRIES], good);
  if (bad) atomicAdd(&stats[PN_UNRESOLVED], bad);
}

// a row must have 2 + 3m fields; its entry count goes to n_entries
__global__ void k_parse_nbr_check_rows(const int64_t* __restrict__ row_tabs, int64_t rows, int32_t k_cap,
                                       int32_t* __restrict__ n_entries, unsigned long long* __restrict__ stats) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const int64_t fields = row_tabs[r] + 1;
  const bool ok = fields >= 2 && (fields - 2) % 3 == 0 && (fields - 2) / 3 <= k_cap;
  n_entries[r] = ok ? (int32_t)((fields - 2) / 3) : -1;
  if (!ok) atomicAdd(&stats[PN_BAD_ROWS], 1ULL);
}

__global__ void k_fill_i32(int32_t* __restrict__ p, int64_t rows, int64_t cols, int64_t ld, int32_t v) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < rows * cols) p[(i / cols) * ld + i % cols] = v;
}

}  // namespace textio
}  // namespace g2

using namespace g2;
using namespace g2::textio;

static inline int fmt_chunks(int64_t rw) { return (int)((rw + CHUNK_CELLS - 1) / CHUNK_CELLS); }

extern "C" size_t gridb200_format_z_ws_bytes(int64_t rows, int64_t rw) {
  if (rows <= 0 || rw <= 0) return 256;
  return (size_t)rows * fmt_chunks(rw) * 4 + (size_t)rows * 8 + 256;
}

extern "C" int gridb200_format_z_measure(const int16_t* zq, int64_t rows, int64_t rw, int64_t ldq, int64_t* row_off,
                                         void* ws, size_t ws_bytes, void* stream) {
  G2_REQUIRE(zq && row_off && ws, GRIDB200_EINVAL, "null pointer");
  G2_REQUIRE(rows > 0 && rows <= 65535 && rw > 0 && rw < (1LL << 27) && ldq >= rw, GRIDB200_EINVAL,
             "bad shape (rows per call <= 65535, rw < 2^27)");
  G2_REQUIRE(ws_bytes >= gridb200_format_z_ws_bytes(rows, rw), GRIDB200_EINVAL, "workspace too small: need %zu bytes",
             gridb200_format_z_ws_bytes(rows, rw));
  cudaStream_t st = as_stream(stream);
  const int nch = fmt_chunks(rw);
  uint32_t* chunk = static_cast<uint32_t*>(ws);
  int64_t* row_total = reinterpret_cast<int64_t*>(static_cast<char*>(ws) + (((size_t)rows * nch * 4 + 7) & ~(size_t)7));
  const bool vec = (ldq % 8 == 0) && ((reinterpret_cast<uintptr_t>(zq) & 15) == 0);
  k_fmt_len<<<dim3(nch, (unsigned)rows), THREADS, 0, st>>>(zq, rw, ldq, vec, nch, chunk);
  G2_LAUNCHED();
  k_chunk_scan<<<(unsigned)rows, THREADS, 0, st>>>(chunk, nch, row_total);
  G2_LAUNCHED();
  k_row_scan<<<1, THREADS, 0, st>>>(row_total, rows, row_off);
  G2_LAUNCHED();
  return 0;
}

extern "C" int gridb200_format_z_write(const int16_t* zq, int64_t rows, int64_t rw, int64_t ldq, const int64_t* row_off,
                                       const void* ws, char* text, int64_t text_cap, void* stream) {
  G2_REQUIRE(zq && row_off && ws && text, GRIDB200_EINVAL, "null pointer");
  G2_REQUIRE(rows > 0 && rows <= 65535 && rw > 0 && rw < (1LL << 27) && ldq >= rw && text_cap > 0, GRIDB200_EINVAL,
             "bad shape");
  cudaStream_t st = as_stream(stream);
  const int nch = fmt_chunks(rw);
  const bool vec = (ldq % 8 == 0) && ((reinterpret_cast<uintptr_t>(zq) & 15) == 0);
  k_fmt_write<<<dim3(nch, (unsigned)rows), THREADS, 0, st>>>(zq, rw, ldq, vec, nch, static_cast<const uint32_t*>(ws),
                                                            row_off, text, text_cap);
  G2_LAUNCHED();
  return 0;
}

static inline int parse_chunks(int64_t max_line) { return (int)((max_line + CHUNK_BYTES - 1) / CHUNK_BYTES); }

extern "C" size_t gridb200_parse_z_ws_bytes(int64_t rows, int64_t max_line_bytes) {
  if (rows <= 0 || max_line_bytes <= 0) return 256;
  return (size_t)rows * parse_chunks(max_line_bytes) * 4 + (size_t)rows * 8 + 256;
}

extern "C" int gridb200_parse_z_rows(const char* text, const int64_t* line_off, int64_t rows, int64_t max_line_bytes,
                                     int32_t skip_fields, int64_t rw, int16_t* zq, int64_t ldq, int64_t* stats, void* ws,
                                     size_t ws_bytes, void* stream) {
  G2_REQUIRE(text && line_off && zq && stats && ws, GRIDB200_EINVAL, "null pointer");
  G2_REQUIRE(rows > 0 && rows <= 65535 && rw > 0 && ldq >= rw && skip_fields >= 0 && max_line_bytes > 0 &&
                 max_line_bytes < (1LL << 31),
             GRIDB200_EINVAL, "bad shape (rows per call <= 65535)");
  G2_REQUIRE(ws_bytes >= gridb200_parse_z_ws_bytes(rows, max_line_bytes), GRIDB200_EINVAL,
             "workspace too small: need %zu bytes", gridb200_parse_z_ws_bytes(rows, max_line_bytes));
  cudaStream_t st = as_stream(stream);
  const int nch = parse_chunks(max_line_bytes);
  uint32_t* chunk = static_cast<uint32_t*>(ws);
  int64_t* row_tabs = reinterpret_cast<int64_t*>(static_cast<char*>(ws) + (((size_t)rows * nch * 4 + 7) & ~(size_t)7));
  G2_CUDA(cudaMemsetAsync(stats, 0, 4 * sizeof(int64_t), st));
  k_parse_count<<<dim3(nch, (unsigned)rows), THREADS, 0, st>>>(text, line_off, nch, chunk);
  G2_LAUNCHED();
  k_chunk_scan<<<(unsigned)rows, THREADS, 0, st>>>(chunk, nch, row_tabs);
  G2_LAUNCHED();
  k_parse_check_rows<<<(unsigned)((rows + 255) / 256), 256, 0, st>>>(row_tabs, rows, (int64_t)skip_fields + rw - 1,
                                                                    reinterpret_cast<unsigned long long*>(stats));
  G2_LAUNCHED();
  k_parse_cells<<<dim3(nch, (unsigned)rows), THREADS, 0, st>>>(text, line_off, nch, chunk, skip_fields, rw, zq, ldq,
                                                              reinterpret_cast<unsigned long long*>(stats));
  G2_LAUNCHED();
  return 0;
}

extern "C" int gridb200_hash_tokens(const char* tok, const int64_t* tok_off, int64_t n, int64_t* hash, void* stream) {
  G2_REQUIRE(tok && tok_off && hash && n > 0, GRIDB200_EINVAL, "bad argument");
  k_tok_hash<<<(unsigned)((n + 255) / 256), 256, 0, as_stream(stream)>>>(tok, tok_off, n, reinterpret_cast<long long*>(hash));
  G2_LAUNCHED();
  return 0;
}

extern "C" int gridb200_parse_nbrs_rows(const char* text, const int64_t* line_off, int64_t rows, int64_t max_line_bytes,
                                        const char* tok, const int64_t* tok_off, const int64_t* hash_sorted,
                                        const int32_t* hash_row, int64_t n_tok, int32_t k_cap, int32_t* idx, int64_t ld,
                                        int32_t* n_entries, int64_t* stats, void* ws, size_t ws_bytes, void* stream) {
  G2_REQUIRE(text && line_off && tok && tok_off && hash_sorted && hash_row && idx && n_entries && stats && ws,
             GRIDB200_EINVAL, "null pointer");
  G2_REQUIRE(rows > 0 && rows <= 65535 && n_tok > 0 && n_tok < (1LL << 31) && k_cap > 0 && ld >= k_cap &&
                 max_line_bytes > 0 && max_line_bytes < (1LL << 31),
             GRIDB200_EINVAL, "bad shape (rows per call <= 65535)");
  G2_REQUIRE(ws_bytes >= gridb200_parse_z_ws_bytes(rows, max_line_bytes), GRIDB200_EINVAL,
             "workspace too small: need %zu bytes", gridb200_parse_z_ws_bytes(rows, max_line_bytes));
  cudaStream_t st = as_stream(stream);
  const int nch = parse_chunks(max_line_bytes);
  uint32_t* chunk = static_cast<uint32_t*>(ws);
  int64_t* row_tabs = reinterpret_cast<int64_t*>(static_cast<char*>(ws) + (((size_t)rows * nch * 4 + 7) & ~(size_t)7));
  G2_CUDA(cudaMemsetAsync(stats, 0, 4 * sizeof(int64_t), st));
  k_fill_i32<<<(unsigned)((rows * k_cap + 255) / 256), 256, 0, st>>>(idx, rows, k_cap, ld, -1);
  G2_LAUNCHED();
  k_parse_count<<<dim3(nch, (unsigned)rows), THREADS, 0, st>>>(text, line_off, nch, chunk);
  G2_LAUNCHED();
  k_chunk_scan<<<(unsigned)rows, THREADS, 0, st>>>(chunk, nch, row_tabs);
  G2_LAUNCHED();
  k_parse_nbr_check_rows<<<(unsigned)((rows + 255) / 256), 256, 0, st>>>(row_tabs, rows, k_cap, n_entries,
                                                                        reinterpret_cast<unsigned long long*>(stats));
  G2_LAUNCHED();
  NbrTable T{tok, tok_off, reinterpret_cast<const long long*>(hash_sorted), hash_row, n_tok};
  k_parse_nbr_entries<<<dim3(nch, (unsigned)rows), THREADS, 0, st>>>(text, line_off, nch, chunk, T, k_cap, idx, ld,
                                                                    reinterpret_cast<unsigned long long*>(stats));
  G2_LAUNCHED();
  return 0;
}
