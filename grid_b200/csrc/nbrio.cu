// Step-5 hand-off text on the device (find_neighbors.py:231-267): the body of the neighbours file
//
//   id \t scale ( \t nbr_id \t nbr_scale \t dist )* \n          dist = "%.2f" % (d2 / (2 * R_use))
//
// from the k-best lists as they sit in HBM.  The host supplies one token per sample, "id\tscale" (ids are
// arbitrary strings, the scale is the step-4 "%.2f" field), so an entry is '\t' + token[j] + '\t' + dist.
//
//   k_nbr_measure : one block per row -> bytes of the row (entries stop at the first negative index, as the list does)
//   k_nbr_scan    : exclusive scan of the row lengths of the batch -> row_off (one block)
//   k_nbr_write   : one block per row; entry lengths -> block scan -> every thread writes its entries
#include "common.cuh"
#include "nbr_text.cuh"

namespace g2 {

constexpr int NT = 128;

struct NbrArgs {
  const char* tok;
  const long long* tok_off;
  const int* idx;
  const long long* d2;
  long long ld;  // row stride of idx / d2 in elements
  int k;
  long long n_tok;  // number of tokens (= samples of the cohort)
  double denom;
};

__device__ __forceinline__ int block_min(int v, int* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  int r = sh[0];
  for (int w = 1; w < NT / 32; ++w) r = min(r, sh[w]);
  __syncthreads();
  return r;
}

// entries [0, kk) of the row are written: kk = first j with idx < 0 (or an index outside the token table), else k
__device__ __forceinline__ int row_entries(const NbrArgs& A, const int* idx_row, int* sh) {
  int first = A.k;
  for (int j = threadIdx.x; j < A.k; j += NT) {
    const int v = idx_row[j];
    if (v < 0 || v >= A.n_tok) {
      first = j;
      break;
    }
  }
  return block_min(first, sh);
}

__device__ __forceinline__ int entry_len(const NbrArgs& A, int nb, long long d2) {
  return 2 + (int)(A.tok_off[nb + 1] - A.tok_off[nb]) + fixed2_len(nbr_dist_hundredths(d2, A.denom));
}

__global__ void __launch_bounds__(NT) k_nbr_measure(NbrArgs A, long long row0, long long* __restrict__ row_len) {
  __shared__ int sh[NT / 32];
  __shared__ long long acc;
  const long long i = blockIdx.x;
  const int* idx_row = A.idx + i * A.ld;
  const long long* d2_row = A.d2 + i * A.ld;
  const int kk = row_entries(A, idx_row, sh);
  if (threadIdx.x == 0) acc = 0;
  __syncthreads();
  long long mine = 0;
  for (int j = threadIdx.x; j < kk; j += NT) mine += entry_len(A, idx_row[j], d2_row[j]);
  mine = warp_sum(mine);
  if ((threadIdx.x & 31) == 0) atomicAdd(reinterpret_cast<unsigned long long*>(&acc), (unsigned long long)mine);
  __syncthreads();
  if (threadIdx.x == 0) row_len[i] = acc + (A.tok_off[row0 + i + 1] - A.tok_off[row0 + i]) + 1;  // own token + '\n'
}

__global__ void __launch_bounds__(1024) k_nbr_scan(const long long* __restrict__ row_len, long long rows,
                                                   long long* __restrict__ row_off) {
  __shared__ long long part[1024];
  const long long per = (rows + 1023) / 1024;
  const long long a = threadIdx.x * per, b = min(rows, a + per);
  long long s = 0;
  for (long long r = a; r < b; ++r) s += row_len[r];
  part[threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    long long run = 0;
    for (int t = 0; t < 1024; ++t) {
      const long long v = part[t];
      part[t] = run;
      run += v;
    }
    row_off[rows] = run;
  }
  __syncthreads();
  long long run = part[threadIdx.x];
  for (long long r = a; r < b; ++r) {
    row_off[r] = run;
    run += row_len[r];
  }
}

__global__ void __launch_bounds__(NT) k_nbr_write(NbrArgs A, long long row0, const long long* __restrict__ row_off,
                                                  char* __restrict__ text) {
  __shared__ int sh[NT / 32];
  __shared__ int wsum[NT / 32];
  const long long i = blockIdx.x;
  const int* idx_row = A.idx + i * A.ld;
  const long long* d2_row = A.d2 + i * A.ld;
  const int kk = row_entries(A, idx_row, sh);
  char* out = text + row_off[i];
  const long long own0 = A.tok_off[row0 + i], own_len = A.tok_off[row0 + i + 1] - own0;
  for (long long t = threadIdx.x; t < own_len; t += NT) out[t] = A.tok[own0 + t];
  long long base = own_len;  // running offset inside the row, the same in every thread
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int j0 = 0; j0 < kk; j0 += NT) {
    const int j = j0 + threadIdx.x;
    int nb = 0, len = 0;
    uint64_t q = 0;
    if (j < kk) {
      nb = idx_row[j];
      q = nbr_dist_hundredths(d2_row[j], A.denom);
      len = 2 + (int)(A.tok_off[nb + 1] - A.tok_off[nb]) + fixed2_len(q);
    }
    int incl = len;  // inclusive scan inside the warp
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    int before = 0, total = 0;
    for (int w = 0; w < NT / 32; ++w) {
      if (w < warp) before += wsum[w];
      total += wsum[w];
    }
    if (j < kk) {
      char* p = out + base + before + (incl - len);
      *p++ = '\t';
      const long long t0 = A.tok_off[nb], tl = A.tok_off[nb + 1] - t0;
      for (long long t = 0; t < tl; ++t) p[t] = A.tok[t0 + t];
      p += tl;
      *p++ = '\t';
      put_fixed2(p, q);
    }
    base += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) out[base] = '\n';
}

}  // namespace g2

extern "C" size_t gridb200_format_nbrs_ws_bytes(int64_t rows) { return (size_t)(rows > 0 ? rows : 0) * 8 + 256; }

static int nbr_args(g2::NbrArgs& A, const char* tok, const int64_t* tok_off, int64_t n_tok, const int32_t* idx,
                    const int64_t* d2, int64_t rows, int32_t k, int64_t ld, int64_t row0, int64_t r_use) {
  using namespace g2;
  G2_REQUIRE(tok && tok_off && rows > 0 && rows <= 65535 && k >= 0 && ld >= k && n_tok > 0, GRIDB200_EINVAL,
             "bad argument (rows per call <= 65535)");
  G2_REQUIRE(k == 0 || (idx && d2), GRIDB200_EINVAL, "null neighbour lists");
  G2_REQUIRE(row0 >= 0 && row0 + rows <= n_tok, GRIDB200_EINVAL, "rows outside the token table");
  G2_REQUIRE(r_use >= 0 && r_use < (1LL << 51), GRIDB200_EDOMAIN, "r_use out of range");
  A.tok = tok, A.tok_off = reinterpret_cast<const long long*>(tok_off), A.idx = idx;
  A.d2 = reinterpret_cast<const long long*>(d2), A.ld = ld, A.k = k, A.n_tok = n_tok;
  A.denom = (double)(2 * (r_use ? r_use : 1));  // R_use == 0 -> 1 (find_neighbors.py:256-257)
  return 0;
}

extern "C" int gridb200_format_nbrs_measure(const char* tok, const int64_t* tok_off, int64_t n_tok, const int32_t* idx,
                                            const int64_t* d2, int64_t rows, int32_t k, int64_t ld, int64_t row0,
                                            int64_t r_use, int64_t* row_off, void* ws, size_t ws_bytes, void* stream) {
  using namespace g2;
  NbrArgs A;
  if (int rc = nbr_args(A, tok, tok_off, n_tok, idx, d2, rows, k, ld, row0, r_use)) return rc;
  G2_REQUIRE(row_off && ws && ws_bytes >= gridb200_format_nbrs_ws_bytes(rows), GRIDB200_EINVAL,
             "workspace too small: need %zu bytes", gridb200_format_nbrs_ws_bytes(rows));
  cudaStream_t st = as_stream(stream);
  long long* row_len = static_cast<long long*>(ws);
  k_nbr_measure<<<(unsigned)rows, NT, 0, st>>>(A, row0, row_len);
  G2_LAUNCHED();
  k_nbr_scan<<<1, 1024, 0, st>>>(row_len, rows, reinterpret_cast<long long*>(row_off));
  G2_LAUNCHED();
  return 0;
}

extern "C" int gridb200_format_nbrs_write(const char* tok, const int64_t* tok_off, int64_t n_tok, const int32_t* idx,
                                          const int64_t* d2, int64_t rows, int32_t k, int64_t ld, int64_t row0,
                                          int64_t r_use, const int64_t* row_off, char* text, int64_t text_cap,
                                          void* stream) {
  using namespace g2;
  NbrArgs A;
  if (int rc = nbr_args(A, tok, tok_off, n_tok, idx, d2, rows, k, ld, row0, r_use)) return rc;
  G2_REQUIRE(row_off && text && text_cap > 0, GRIDB200_EINVAL, "null output");
  k_nbr_write<<<(unsigned)rows, NT, 0, as_stream(stream)>>>(A, row0, reinterpret_cast<const long long*>(row_off), text);
  G2_LAUNCHED();
  return 0;
}
