// Step 4 kernels: per-bin raw sums, per-sample sums, per-bin normalised statistics, transform + quantise.
//
// All four matrix passes are HBM-bound streaming reads of depth[n, ld] (fp32).  Thread mapping for the
// column-oriented passes: one thread owns 4 consecutive bins (one 16-byte load per row), a block owns 1024
// bins x a slab of rows, so a warp reads 512 contiguous bytes of every row it visits and the per-bin state
// lives in registers.  Arithmetic is exact integer (hundredths) for sums and fp64 for everything the
// reference does in fp64; see DESIGN.md "Numerics" for why fp32 is not enough before the %.2f rounding.
#include "common.cuh"

namespace g2 {

constexpr int kThreads = 256;
constexpr int kColsPerBlock = kThreads * 4;

// ---------------------------------------------------------------------------------------------------------
// Two storage types of the depth matrix share every kernel below (template parameter T):
//   float     the packed fp32 matrix (0.0f = absent), 4 bytes per cell: hundredths recovered with one FFMA per cell
//   uint16_t  the hundredths themselves (0 = absent), 2 bytes per cell: half the HBM traffic of the three step-4
//             passes.  The code 0xFFFF is an escape: the cell's true hundredths (>= 65535, i.e. depth >= 655.35) are in a
//             table sorted by cell key (global row * b + bin), looked up by binary search -- rare by construction.
// Both yield the SAME integers h, so everything downstream is bit-identical.
// ---------------------------------------------------------------------------------------------------------
struct Esc {
  const long long* key;  // sorted cell keys (esc_row0 + row) * b + bin
  const int* h;          // true hundredths of those cells
  long long n;
  long long row0;        // global row of the matrix's row 0
  long long b;           // bins per row in the key
};

__device__ __noinline__ int esc_lookup(const Esc& e, int64_t row, int64_t col) {
  const long long want = (e.row0 + row) * e.b + col;
  long long lo = 0, hi = e.n;
  while (lo < hi) {
    const long long mid = (lo + hi) >> 1;
    if (e.key[mid] < want) lo = mid + 1;
    else hi = mid;
  }
  return (lo < e.n && e.key[lo] == want) ? e.h[lo] : 65535;
}

template <typename T>
struct Cells;

template <>
struct Cells<float> {
  using Raw = float4;
  static constexpr bool kEsc = false;
  static __device__ __forceinline__ Raw zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
  static __device__ __forceinline__ Raw load(const float* p) { return ld_stream(reinterpret_cast<const float4*>(p)); }
  static __device__ __forceinline__ Raw load_cached(const float* p) { return *reinterpret_cast<const float4*>(p); }
  static __device__ __forceinline__ void hund(const Raw& v, const Esc&, int64_t, int64_t, int (&h)[4]) {
    h[0] = depth_hundredths(v.x), h[1] = depth_hundredths(v.y), h[2] = depth_hundredths(v.z), h[3] = depth_hundredths(v.w);
  }
  static __device__ __forceinline__ void hund_exact(const Raw& v, const Esc& e, int64_t r, int64_t c, int (&h)[4]) { hund(v, e, r, c, h); }
  static __device__ __forceinline__ int hund1(const float* p, const Esc&, int64_t, int64_t) { return depth_hundredths(*p); }
};

template <>
struct Cells<uint16_t> {
  using Raw = uint2;
  static constexpr bool kEsc = true;
  static __device__ __forceinline__ Raw zero() { return make_uint2(0u, 0u); }
  static __device__ __forceinline__ Raw load(const uint16_t* p) {
    uint2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
  }
  static __device__ __forceinline__ Raw load_cached(const uint16_t* p) { return *reinterpret_cast<const uint2*>(p); }
  // the streaming passes take the CODES (an escape cell counts as 65535) and small patch kernels put the true values in
  // afterwards (k_esc_*): any escape test inside the hot loops cost two instructions per cell and forty registers
  static __device__ __forceinline__ void hund(const Raw& v, const Esc&, int64_t, int64_t, int (&h)[4]) {
    h[0] = (int)(v.x & 0xFFFFu), h[1] = (int)(v.x >> 16), h[2] = (int)(v.y & 0xFFFFu), h[3] = (int)(v.y >> 16);
  }
  static __device__ __forceinline__ void hund_exact(const Raw& v, const Esc& e, int64_t row, int64_t col, int (&h)[4]) {
    hund(v, e, row, col, h);
    if (__vcmpeq2(v.x, 0xFFFFFFFFu) | __vcmpeq2(v.y, 0xFFFFFFFFu)) {  // some cell of the four is an escape: rare
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (h[q] == 0xFFFF) h[q] = esc_lookup(e, row, col + q);
    }
  }
  static __device__ __forceinline__ int hund1(const uint16_t* p, const Esc& e, int64_t row, int64_t col) {
    const int h = *p;
    return h == 0xFFFF ? esc_lookup(e, row, col) : h;
  }
};

// rows per slab so that the grid has at least ~8 blocks per SM when the bin count alone cannot provide them
static int64_t pick_row_slab(int64_t n, int64_t col_blocks, int64_t min_rows, int per_sm = 8) {
  const int64_t want_blocks = (int64_t)kSMs * per_sm;
  int64_t slabs = (want_blocks + col_blocks - 1) / col_blocks;
  int64_t rows = (n + slabs - 1) / slabs;
  if (rows < min_rows) rows = min_rows;
  if (rows > n) rows = n;
  return rows < 1 ? 1 : rows;
}

// Cells per thread of the two streaming statistics passes (K1, K4): 16 bytes per load for either storage -- 4 fp32 cells
// or 8 uint16 cells.  (With 8-byte loads the 2-byte matrix had half the bytes in flight and the passes took as long as on
// fp32: latency-bound instead of HBM-bound.)
template <typename T>
struct Wide : Cells<T> {
  static constexpr int kN = 4;
};
template <>
struct Wide<uint16_t> {
  using Raw = uint4;
  static constexpr int kN = 8;
  static __device__ __forceinline__ Raw zero() { return make_uint4(0u, 0u, 0u, 0u); }
  static __device__ __forceinline__ Raw load(const uint16_t* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
  }
  static __device__ __forceinline__ void hund(const Raw& v, const Esc&, int64_t, int64_t, int (&h)[8]) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int t = 0; t < 4; ++t) h[2 * t] = (int)(w[t] & 0xFFFFu), h[2 * t + 1] = (int)(w[t] >> 16);  // codes (see Cells)
  }
};

// ---------------------------------------------------------------------------------------------------------
// K1: colsum[j] = sum_i h_ij, colcnt[j] = #{h_ij > 0}; with kRows also the per-sample totals over ALL bins
// (rowsum_all[i] = sum_j h_ij, rowcnt_all[i] = #{h_ij > 0}), which let step K3 skip its full pass over the matrix
// when few bins fail the depth filter (see k_row_fix).
// ---------------------------------------------------------------------------------------------------------
// One warp's partial of one row into the block's shared accumulators.  rs < 2^29 (128 bins x h < 2^22): its low 16
// bits and the rest go to separate words, so eight warps cannot overflow either and no carry has to be read back --
// three fire-and-forget shared-memory reductions (written as asm: the compiler turns an atomicAdd inside a divergent
// branch into a dozen instructions of warp aggregation).
__device__ __forceinline__ void row_partial(unsigned* lo, unsigned* hi, int* cnt, unsigned rs, int rc) {
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(lo)), "r"(rs & 0xFFFFu) : "memory");
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(hi)), "r"(rs >> 16) : "memory");
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(cnt)), "r"((unsigned)rc) : "memory");
}

template <bool kRows, typename T>
__global__ void __launch_bounds__(kThreads) k_col_raw(const T* __restrict__ depth, const Esc esc, int64_t n, int64_t b,
                                                      int64_t ld, int64_t rows_per_slab,
                                                      unsigned long long* __restrict__ colsum,
                                                      int* __restrict__ colcnt,
                                                      unsigned long long* __restrict__ rowsum_all,
                                                      int* __restrict__ rowcnt_all) {
  // per-sample partials of this block's 1024 bins for one chunk of 512 rows: the warps add into shared memory and the
  // block sends ONE global atomic per row and chunk (a global atomic per warp and row -- 1.5e8 of them on 3202
  // addresses at config 2 -- was what held this pass at 0.75 of HBM speed)
  __shared__ unsigned s_lo[kRows ? 512 : 1], s_hi[kRows ? 512 : 1];  // sums split at bit 16: plain 32-bit reductions
  __shared__ int s_rc[kRows ? 512 : 1];
  using C = Wide<T>;
  constexpr int kN = C::kN;
  const int64_t j0 = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * kN;
  const bool live = j0 < b;  // dead threads of the last block stay for the collectives and barriers (they add zeros)
  if (!kRows && !live) return;
  const bool lane0 = (threadIdx.x & 31) == 0;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_slab;
  const int64_t r1 = min(n, r0 + rows_per_slab);
  unsigned long long s[kN];
  int c[kN];
#pragma unroll
  for (int q = 0; q < kN; ++q) s[q] = 0, c[q] = 0;
  const T* p = depth + r0 * ld + (live ? j0 : 0);
  const typename C::Raw zero4 = C::zero();
  for (int64_t base = r0; base < r1; base += 512) {  // uint32 partials cannot overflow within 512 rows
    const int64_t lim = (r1 < base + 512 ? r1 : base + 512);
    if (kRows) {
      for (int t = threadIdx.x; t < 512; t += kThreads) s_lo[t] = 0, s_hi[t] = 0, s_rc[t] = 0;
      __syncthreads();
    }
    unsigned ps[kN];
#pragma unroll
    for (int q = 0; q < kN; ++q) ps[q] = 0;
    // one row's cells of this thread: per-bin partials, and (kRows) the warp's share of the row total into shared memory
    auto take = [&](const typename C::Raw& v, int64_t row) {
      int hh[kN];
      C::hund(v, esc, row, j0, hh);
      unsigned hs = 0;
      int ks = 0;
#pragma unroll
      for (int q = 0; q < kN; ++q) {
        ps[q] += hh[q];
        const int k = hh[q] > 0;
        c[q] += k;
        hs += (unsigned)hh[q];
        ks += k;
      }
      if (kRows) {  // 32 lanes x kN bins x h < 2^22 fit 32 bits
        const unsigned rs = __reduce_add_sync(0xffffffffu, hs);
        const int rc = __reduce_add_sync(0xffffffffu, ks);
        if (lane0 && rc) row_partial(&s_lo[row - base], &s_hi[row - base], &s_rc[row - base], rs, rc);
      }
    };
    int64_t i = base;
    for (; i + 8 <= lim; i += 8) {
      typename C::Raw v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = live ? C::load(p + u * ld) : zero4;
      p += 8 * ld;
#pragma unroll
      for (int u = 0; u < 8; ++u) take(v[u], i + u);
    }
    for (; i < lim; ++i) {
      const typename C::Raw v = live ? C::load(p) : zero4;
      p += ld;
      take(v, i);
    }
#pragma unroll
    for (int q = 0; q < kN; ++q) s[q] += ps[q];
    if (kRows) {
      __syncthreads();
      for (int t = threadIdx.x; t < (int)(lim - base); t += kThreads) {
        if (s_rc[t]) {
          atomicAdd(&rowsum_all[base + t], ((unsigned long long)s_hi[t] << 16) + s_lo[t]);
          atomicAdd(&rowcnt_all[base + t], s_rc[t]);
        }
      }
      __syncthreads();
    }
  }
  if (!live) return;
#pragma unroll
  for (int q = 0; q < kN; ++q) {
    if (j0 + q < b) {
      if (gridDim.y == 1) {
        colsum[j0 + q] = s[q];
        colcnt[j0 + q] = c[q];
      } else {
        atomicAdd(&colsum[j0 + q], s[q]);
        atomicAdd(&colcnt[j0 + q], c[q]);
      }
    }
  }
}

// K1 on the 2-byte matrix.  Two cells per 32-bit word, never unpacked: run through the generic kernel the pass was
// ALU-bound at 9 instructions per cell (ncu: ALU pipe 70 %, 96 registers, 4.1 ms against a 2.9 ms HBM floor).  Per word
// here: a packed 16-bit min against 1 (the two presence bits), one packed add for the counts (512 rows fit a half),
// one dot-product-accumulate for the low half's sum, one shift-add for the high half's and one dot-product-accumulate
// for the row total.  The row totals of a warp go to lane (row % 32) and are stored 32 rows at a time into the warp's
// own shared-memory column -- no shared atomics; the eight columns are summed once per 512-row chunk.  Escape cells
// count as 65535 here; k_esc_raw adds the rest afterwards.
__global__ void __launch_bounds__(kThreads, 3) k_col_raw_u16(const uint16_t* __restrict__ depth, int64_t n, int64_t b,
                                                             int64_t ld, int64_t rows_per_slab,
                                                             unsigned long long* __restrict__ colsum,
                                                             int* __restrict__ colcnt,
                                                             unsigned long long* __restrict__ rowsum_all,
                                                             int* __restrict__ rowcnt_all) {
  constexpr int kWarps = kThreads / 32;
  __shared__ uint2 s_wp[kWarps * 512];  // (row total, packed presence count) per warp and row of the chunk
  using C = Wide<uint16_t>;
  const int64_t j0 = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * 8;
  const bool live = j0 < b;  // dead threads of the last block stay for the collectives and barriers (they add zeros)
  const int lane = threadIdx.x & 31;
  uint2* const mycol = s_wp + (threadIdx.x >> 5) * 512;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_slab;
  const int64_t r1 = min(n, r0 + rows_per_slab);
  unsigned long long s[8];
  int c[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) s[q] = 0, c[q] = 0;
  const uint16_t* p = depth + r0 * ld + (live ? j0 : 0);
  const uint4 zero4 = C::zero();
  for (int64_t base = r0; base < r1; base += 512) {  // 32-bit sums and 16-bit counts cannot overflow within 512 rows
    const int rows = (int)(r1 - base < 512 ? r1 - base : 512);
    unsigned ps[8], cp[4];
#pragma unroll
    for (int q = 0; q < 8; ++q) ps[q] = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) cp[q] = 0;
    uint2 mine = make_uint2(0u, 0u);
    auto take = [&](const uint4& v, int slot) {
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
      unsigned hs = 0, km = 0;
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const unsigned m = __vminu2(w[t], 0x00010001u);
        cp[t] += m;
        km += m;
        ps[2 * t] = __dp2a_lo(w[t], 0x0001u, ps[2 * t]);
        ps[2 * t + 1] += w[t] >> 16;
        hs = __dp2a_lo(w[t], 0x0101u, hs);
      }
      // 32 lanes x 8 cells x 65535 < 2^24; per half at most 4 present cells per lane, so the packed counts survive the sum
      const unsigned rs = __reduce_add_sync(0xffffffffu, hs);
      const unsigned rk = __reduce_add_sync(0xffffffffu, km);
      if (lane == (slot & 31)) mine = make_uint2(rs, rk);
    };
    int i = 0;
    for (; i + 8 <= rows; i += 8) {
      uint4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = live ? C::load(p + u * ld) : zero4;
      p += 8 * ld;
#pragma unroll
      for (int u = 0; u < 8; ++u) take(v[u], i + u);
      if (((i + 8) & 31) == 0) mycol[i + 8 - 32 + lane] = mine;
    }
    for (; i < rows; ++i) {
      const uint4 v = live ? C::load(p) : zero4;
      p += ld;
      take(v, i);
    }
    if (lane < (rows & 31)) mycol[(rows & ~31) + lane] = mine;
#pragma unroll
    for (int q = 0; q < 8; ++q) s[q] += ps[q];
#pragma unroll
    for (int q = 0; q < 4; ++q) c[2 * q] += (int)(cp[q] & 0xFFFFu), c[2 * q + 1] += (int)(cp[q] >> 16);
    __syncthreads();
    for (int t = threadIdx.x; t < rows; t += kThreads) {
      unsigned rs = 0, rk = 0;  // 8 warps x 2^24, 8 x 128 per half
#pragma unroll
      for (int w = 0; w < kWarps; ++w) rs += s_wp[w * 512 + t].x, rk += s_wp[w * 512 + t].y;
      const int rc = (int)((rk & 0xFFFFu) + (rk >> 16));
      if (rc) {
        atomicAdd(&rowsum_all[base + t], (unsigned long long)rs);
        atomicAdd(&rowcnt_all[base + t], rc);
      }
    }
    __syncthreads();
  }
  if (!live) return;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    if (j0 + q < b) {
      if (gridDim.y == 1) {
        colsum[j0 + q] = s[q];
        colcnt[j0 + q] = c[q];
      } else {
        atomicAdd(&colsum[j0 + q], s[q]);
        atomicAdd(&colcnt[j0 + q], c[q]);
      }
    }
  }
}

__global__ void k_depth_filter(const long long* __restrict__ colsum, const int* __restrict__ colcnt, int64_t b,
                               double lo, double hi, uint8_t* __restrict__ valid) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= b) return;
  const int c = colcnt[j];
  bool ok = false;
  if (c > 0) {
    const double m = (double)colsum[j] / (100.0 * (double)c);
    ok = (lo <= m) && (m <= hi);
  }
  valid[j] = ok ? 1 : 0;
}

// ---------------------------------------------------------------------------------------------------------
// K3: rowsum[i] = sum over valid bins of h_ij, rowcnt[i] = #present.  grid = (bin chunks, rows).
// ---------------------------------------------------------------------------------------------------------
constexpr int kRowUnroll = 16;
constexpr int kRowChunk = kThreads * 4 * kRowUnroll;  // 16384 bins per block
constexpr int64_t kRowGridY = 32;                     // rows are strided over this many blocks per bin chunk

template <typename T>
__global__ void __launch_bounds__(kThreads) k_row_stats(const T* __restrict__ depth, const Esc esc, int64_t row_base,
                                                        int64_t b, int64_t ld, const uint8_t* __restrict__ valid,
                                                        unsigned long long* __restrict__ rowsum,
                                                        int* __restrict__ rowcnt, const int* __restrict__ skip_if,
                                                        int64_t nrows) {
  if (skip_if && skip_if[0]) return;  // the per-sample totals are being derived by k_row_fix instead
  // rows are strided over gridDim.y so that the launch that finds skip_if set costs a few thousand empty blocks,
  // not one per (chunk, row)
  const int64_t c0 = (int64_t)blockIdx.x * kRowChunk;
  __shared__ unsigned long long ss[kThreads / 32];
  __shared__ int sc[kThreads / 32];
  for (int64_t row = blockIdx.y; row < nrows; row += gridDim.y) {
  using C = Cells<T>;
  const T* p = depth + row * ld;
  typename C::Raw v[kRowUnroll];
  uint32_t m[kRowUnroll];
#pragma unroll
  for (int u = 0; u < kRowUnroll; ++u) {
    const int64_t j = c0 + ((int64_t)u * kThreads + threadIdx.x) * 4;
    if (j < b) {  // ld >= roundup4(b): the 16-byte load stays inside the row
      v[u] = C::load(p + j);
      if (j + 4 <= b) {
        m[u] = *reinterpret_cast<const uint32_t*>(valid + j);
      } else {
        m[u] = 0;
        for (int q = 0; q < 4 && j + q < b; ++q) m[u] |= (uint32_t)valid[j + q] << (8 * q);
      }
    } else {
      v[u] = C::zero();
      m[u] = 0;
    }
  }
  unsigned long long s = 0;
  int c = 0;
#pragma unroll
  for (int u = 0; u < kRowUnroll; ++u) {
    int hh[4];
    C::hund_exact(v[u], esc, row_base + row, c0 + ((int64_t)u * kThreads + threadIdx.x) * 4, hh);
    const int h0 = (m[u] & 0x000000ffu) ? hh[0] : 0;
    const int h1 = (m[u] & 0x0000ff00u) ? hh[1] : 0;
    const int h2 = (m[u] & 0x00ff0000u) ? hh[2] : 0;
    const int h3 = (m[u] & 0xff000000u) ? hh[3] : 0;
    s += (unsigned)(h0 + h1) + (unsigned)(h2 + h3);
    c += (h0 > 0) + (h1 > 0) + (h2 > 0) + (h3 > 0);
  }
  s = (unsigned long long)warp_sum((long long)s);
  c = warp_sum(c);
  if ((threadIdx.x & 31) == 0) {
    ss[threadIdx.x >> 5] = s;
    sc[threadIdx.x >> 5] = c;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long ts = 0;
    int tc = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) {
      ts += ss[w];
      tc += sc[w];
    }
    if (tc) {
      atomicAdd(&rowsum[row], ts);
      atomicAdd(&rowcnt[row], tc);
    }
  }
  __syncthreads();  // ss / sc are reused by the next row
  }
}

// ---------------------------------------------------------------------------------------------------------
// K3': per-sample sums over the VALID bins from the totals over all bins that K1 produced:
//   rowsum = rowsum_all - sum over bins that have data but failed the depth filter ("bad" bins).
// Reading single columns costs a 32-byte sector per 4-byte cell, so this beats the full pass K3 only while the
// bad bins are fewer than ~1/8 of all bins; k_bad_list decides ON THE DEVICE (no host round trip): it writes
// mode[0] = 1 (fix-up) or 0 (full pass); k_row_stats and k_row_fix both read it and one of them returns at once.
// ---------------------------------------------------------------------------------------------------------
__global__ void k_bad_list(const uint8_t* __restrict__ valid, const int* __restrict__ colcnt, int64_t b,
                           int* __restrict__ bad, int* __restrict__ nbad) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool is_bad = j < b && !valid[j] && colcnt[j] > 0;
  const unsigned m = __ballot_sync(0xffffffffu, is_bad);
  if (!m) return;
  int base = 0;
  if ((threadIdx.x & 31) == 0) base = atomicAdd(nbad, __popc(m));
  base = __shfl_sync(0xffffffffu, base, 0);
  if (is_bad) bad[base + __popc(m & ((1u << (threadIdx.x & 31)) - 1))] = (int)j;  // order is irrelevant: exact sums
}

__global__ void k_row_mode(const int* __restrict__ nbad, int64_t b, int* __restrict__ mode) {
  mode[0] = ((int64_t)nbad[0] * 8 <= b) ? 1 : 0;
}

template <typename T>
__global__ void __launch_bounds__(kThreads) k_row_fix(const T* __restrict__ depth, const Esc esc, int64_t ld,
                                                      const int* __restrict__ bad, const int* __restrict__ nbad,
                                                      const int* __restrict__ mode,
                                                      const unsigned long long* __restrict__ rowsum_all,
                                                      const int* __restrict__ rowcnt_all,
                                                      unsigned long long* __restrict__ rowsum,
                                                      int* __restrict__ rowcnt) {
  if (!mode[0]) return;
  const int64_t row = blockIdx.x;
  const int nb = nbad[0];
  const T* p = depth + row * ld;
  unsigned long long s = 0;
  int c = 0;
  for (int t = threadIdx.x; t < nb; t += kThreads) {
    const int h = Cells<T>::hund1(p + bad[t], esc, row, bad[t]);
    s += (unsigned)h;
    c += h > 0;
  }
  s = (unsigned long long)warp_sum((long long)s);
  c = warp_sum(c);
  __shared__ unsigned long long ss[kThreads / 32];
  __shared__ int sc[kThreads / 32];
  if ((threadIdx.x & 31) == 0) ss[threadIdx.x >> 5] = s, sc[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long ts = 0;
    int tc = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) ts += ss[w], tc += sc[w];
    rowsum[row] = rowsum_all[row] - ts;
    rowcnt[row] = rowcnt_all[row] - tc;
  }
}

__global__ void k_row_scales(const long long* __restrict__ rowsum, const int* __restrict__ rowcnt, int64_t n,
                             double* __restrict__ raw_mean, double* __restrict__ scale_q,
                             double* __restrict__ row_factor) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int c = rowcnt[i];
  const long long s = rowsum[i];
  double m = nan(""), q = nan(""), f = nan("");
  if (c > 0 && s > 0) {
    m = (double)s / (100.0 * (double)c);    // nanmean of the raw depths of the row
    q = printf_round(m, 100.0) / 100.0;     // float("%.2f" % m)
    f = (double)c / (double)s;              // y_ij = h_ij * f  ==  depth_ij / rowmean_i
  }
  if (raw_mean) raw_mean[i] = m;
  if (scale_q) scale_q[i] = q;
  if (row_factor) row_factor[i] = f;
}

// ---------------------------------------------------------------------------------------------------------
// K3'': per-sample means that sit EXACTLY on a %.2f tie.
//
// The reference prints np.nanmean(row) with %.2f.  k_row_scales computes the mean from exact integer sums, which is
// the correctly rounded value; numpy's is a pairwise floating-point sum and may be an ulp or two off.  The two print
// the same two decimals unless the exact mean is k + 0.005 -- then the ulp decides, and that is not rare for small
// windows (probability 1/count for an even count of bins: percent level at BASELINE config 1).  The distance of an
// exact mean to the nearest tie is 0 or at least 1/(2*count), far beyond the summation noise, so ONLY exact ties
// need numpy's own value: for those rows one thread replays numpy's pairwise summation (loops_utils.h.src,
// DOUBLE_pairwise_sum: blocks of <= 128 with 8 strided accumulators, halves split at a multiple of 8) over the
// row's valid bins in column order, NaN cells as 0.0, exactly as nansum does.
// ---------------------------------------------------------------------------------------------------------
template <typename T>
struct RowCursor {
  const T* row;
  const uint8_t* valid;
  int64_t j;
  Esc esc;
  int64_t i;
  __device__ double next() {  // value of the next valid bin: depth as the double the reference parsed, absent -> 0.0
    while (!valid[j]) ++j;
    const int h = Cells<T>::hund1(row + j, esc, i, j);
    ++j;
    return h ? __ddiv_rn((double)h, 100.0) : 0.0;
  }
};

template <typename T>
__device__ double numpy_pairwise_leaf(RowCursor<T>& c, int64_t n) {  // n <= 128
  if (n < 8) {
    double res = 0.0;
    for (int64_t i = 0; i < n; ++i) res = __dadd_rn(res, c.next());
    return res;
  }
  double r[8];
  for (int k = 0; k < 8; ++k) r[k] = c.next();
  int64_t i = 8;
  for (; i < n - (n % 8); i += 8)
    for (int k = 0; k < 8; ++k) r[k] = __dadd_rn(r[k], c.next());
  double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                         __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
  for (; i < n; ++i) res = __dadd_rn(res, c.next());
  return res;
}

// sum(n) = n <= 128 ? leaf(n) : sum(n2) + sum(n - n2), n2 = n/2 rounded down to a multiple of 8; the elements are
// consumed strictly in order, so the recursion runs over an explicit stack (depth <= log2(n))
template <typename T>
__device__ double numpy_pairwise_sum(RowCursor<T>& c, int64_t n) {
  struct Frame {
    int64_t n;
    double left;
    int phase;
  };
  Frame st[40];
  int sp = 0;
  st[0] = Frame{n, 0.0, 0};
  double ret = 0.0;
  while (sp >= 0) {
    Frame& f = st[sp];
    if (f.n <= 128) {
      ret = numpy_pairwise_leaf(c, f.n);
      --sp;
      continue;
    }
    int64_t n2 = f.n / 2;
    n2 -= n2 % 8;
    if (f.phase == 0) {
      f.phase = 1;
      st[++sp] = Frame{n2, 0.0, 0};
    } else if (f.phase == 1) {
      f.left = ret;
      f.phase = 2;
      const int64_t rest = f.n - n2;
      st[++sp] = Frame{rest, 0.0, 0};
    } else {
      ret = __dadd_rn(f.left, ret);
      --sp;
    }
  }
  return ret;
}

template <typename T>
__global__ void k_row_mean_ties(const T* __restrict__ depth, const Esc esc, int64_t n, int64_t b, int64_t ld,
                                const uint8_t* __restrict__ valid, const long long* __restrict__ rowsum,
                                const int* __restrict__ rowcnt, double* __restrict__ raw_mean,
                                double* __restrict__ scale_q) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const long long cnt = rowcnt[i], s = rowsum[i];
  if (cnt <= 0 || s <= 0 || (2 * s) % (2 * cnt) != cnt) return;  // mean in hundredths = s / cnt is not k + 1/2
  int64_t nvalid = 0;
  for (int64_t j = 0; j < b; ++j) nvalid += valid[j] != 0;
  RowCursor<T> c{depth + i * ld, valid, 0, esc, i};
  const double m = __ddiv_rn(__dadd_rn(0.0, numpy_pairwise_sum(c, nvalid)), (double)cnt);
  if (raw_mean) raw_mean[i] = m;
  if (scale_q) scale_q[i] = printf_round(m, 100.0) / 100.0;
}

// ---------------------------------------------------------------------------------------------------------
// K4: sd[j] = sum_i (y_ij - 1), sd2[j] = sum_i (y_ij - 1)^2 over present cells (shifted-data variance).
// Deterministic: with several row slabs each slab writes its own partial, reduced in slab order afterwards.
// ---------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kThreads) k_col_norm(const T* __restrict__ depth, const Esc esc, int64_t n, int64_t b,
                                                       int64_t ld, int64_t rows_per_slab,
                                                       const uint8_t* __restrict__ valid,
                                                       const double* __restrict__ row_factor,
                                                       double* __restrict__ out_sd, double* __restrict__ out_sd2) {
  using C = Wide<T>;
  constexpr int kN = C::kN;
  const int64_t j0 = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * kN;
  if (j0 >= b) return;
  bool ok[kN];
  bool any = false;
#pragma unroll
  for (int q = 0; q < kN; ++q) {
    ok[q] = (j0 + q < b) && valid[j0 + q];
    any |= ok[q];
  }
  double s1[kN], s2[kN];
#pragma unroll
  for (int q = 0; q < kN; ++q) s1[q] = 0.0, s2[q] = 0.0;
  if (any) {
    const int64_t r0 = (int64_t)blockIdx.y * rows_per_slab;
    const int64_t r1 = min(n, r0 + rows_per_slab);
    // rows are taken strictly in order per bin, whatever the storage: sd / sd2 are bit-identical for fp32 and uint16
    auto take = [&](const typename C::Raw& v, double f, int64_t row) {
      int h[kN];
      C::hund(v, esc, row, j0, h);
#pragma unroll
      for (int q = 0; q < kN; ++q) {
        if (h[q] > 0) {  // a row with no valid bin has f = NaN but then also no present cell
          const double d = fma(u32_to_double(h[q]), f, -1.0);
          s1[q] += d;
          s2[q] = fma(d, d, s2[q]);
        }
      }
    };
    const T* p = depth + r0 * ld + j0;
    int64_t i = r0;
    for (; i + 4 <= r1; i += 4) {
      typename C::Raw v[4];
      double f[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        v[u] = C::load(p + u * ld);
        f[u] = row_factor[i + u];
      }
      p += 4 * ld;
#pragma unroll
      for (int u = 0; u < 4; ++u) take(v[u], f[u], i + u);
    }
    for (; i < r1; ++i) {
      const typename C::Raw v = C::load(p);
      const double f = row_factor[i];
      p += ld;
      take(v, f, i);
    }
  }
  double* o1 = out_sd + (int64_t)blockIdx.y * b;
  double* o2 = out_sd2 + (int64_t)blockIdx.y * b;
#pragma unroll
  for (int q = 0; q < kN; ++q) {
    if (j0 + q < b) {
      o1[j0 + q] = ok[q] ? s1[q] : 0.0;
      o2[j0 + q] = ok[q] ? s2[q] : 0.0;
    }
  }
}

// K4 on the 2-byte matrix.  Same sums in the same order as k_col_norm (rows strictly in order per bin, same slabs), but
// built for latency instead of bandwidth: at 2 bytes per cell the pass is bound by how many loads are in flight (ncu on
// the generic kernel: 70 % of the stalls on the loads, DRAM at 36 %).  Four rows are always in flight -- a row's
// registers are reloaded four rows ahead as soon as it has been summed -- and the per-sample factors of a 512-row chunk
// sit in shared memory instead of registers, so four blocks fit an SM.
__global__ void __launch_bounds__(kThreads, 3) k_col_norm_u16(const uint16_t* __restrict__ depth, int64_t n, int64_t b,
                                                              int64_t ld, int64_t rows_per_slab,
                                                              const uint8_t* __restrict__ valid,
                                                              const double* __restrict__ row_factor,
                                                              double* __restrict__ out_sd, double* __restrict__ out_sd2) {
  __shared__ double s_f[512];
  using C = Wide<uint16_t>;
  const int64_t j0 = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * 8;
  const bool live = j0 < b;  // dead threads of the last block stay for the barriers
  unsigned okm = 0;
#pragma unroll
  for (int q = 0; q < 8; ++q)
    if (live && j0 + q < b && valid[j0 + q]) okm |= 1u << q;
  double s1[8], s2[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) s1[q] = 0.0, s2[q] = 0.0;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_slab;
  const int64_t r1 = min(n, r0 + rows_per_slab);
  auto take = [&](const uint4& v, double f) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const unsigned h = (q & 1) ? (w[q >> 1] >> 16) : (w[q >> 1] & 0xFFFFu);
      if (h > 0) {  // a row with no valid bin has f = NaN but then also no present cell
        const double d = fma(u32_to_double(h), f, -1.0);
        s1[q] += d;
        s2[q] = fma(d, d, s2[q]);
      }
    }
  };
  for (int64_t base = r0; base < r1; base += 512) {
    const int rows = (int)(r1 - base < 512 ? r1 - base : 512);
    __syncthreads();
    for (int t = threadIdx.x; t < rows; t += kThreads) s_f[t] = row_factor[base + t];
    __syncthreads();
    if (!okm) continue;
    const uint16_t* p = depth + base * ld + j0;
    int i = 0;
    uint4 v[4];  // rows i .. i+3; each register set is reloaded with row i+4+u as soon as row i+u has been summed
    if (rows >= 4) {
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = C::load(p + u * ld);
      p += 4 * ld;
    }
    for (; i + 8 <= rows; i += 4) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint4 cur = v[u];
        v[u] = C::load(p + u * ld);  // issued BEFORE the sums of row i+u (left after them, ptxas sinks all four to the loop end)
        take(cur, s_f[i + u]);
      }
      p += 4 * ld;
    }
    if (i + 4 <= rows) {
#pragma unroll
      for (int u = 0; u < 4; ++u) take(v[u], s_f[i + u]);
      i += 4;
    }
    for (; i < rows; ++i) {
      const uint4 x = C::load(p);
      p += ld;
      take(x, s_f[i]);
    }
  }
  if (!live) return;
  double* o1 = out_sd + (int64_t)blockIdx.y * b;
  double* o2 = out_sd2 + (int64_t)blockIdx.y * b;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    if (j0 + q < b) {
      o1[j0 + q] = (okm >> q & 1) ? s1[q] : 0.0;
      o2[j0 + q] = (okm >> q & 1) ? s2[q] : 0.0;
    }
  }
}

__global__ void k_reduce_slabs(const double* __restrict__ part1, const double* __restrict__ part2, int64_t b,
                               int slabs, double* __restrict__ sd, double* __restrict__ sd2) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= b) return;
  double a = 0, c = 0;
  for (int s = 0; s < slabs; ++s) {
    a += part1[(int64_t)s * b + j];
    c += part2[(int64_t)s * b + j];
  }
  sd[j] = a;
  sd2[j] = c;
}

__global__ void k_col_finalize(const double* __restrict__ sd, const double* __restrict__ sd2,
                               const int* __restrict__ colcnt, const uint8_t* __restrict__ valid, int64_t b,
                               int64_t n_total, double* __restrict__ mu, double* __restrict__ s2,
                               double* __restrict__ ratio) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= b) return;
  double m = nan(""), v = nan(""), r = nan("");
  const int c = colcnt[j];
  if (valid[j] && c > 0) {
    const double a = sd[j];
    const double dbar = a / (double)c;
    m = 1.0 + dbar;
    // sum (d - dbar)^2 = sd2 - sd*dbar ; the reference divides by N-1 with N = all rows (ddof=1, :446)
    v = fma(-a, dbar, sd2[j]) / (double)(n_total - 1);
    if (v < 0.0) v = 0.0;
    if (m > 0.0) r = 100.0 * v / m;
  }
  mu[j] = m;
  s2[j] = v;
  ratio[j] = r;
}

// ---------------------------------------------------------------------------------------------------------
// K7: transform + printf-exact quantisation + compaction to the selected bins (+ clipped fp16 operand,
// + exact row norms).  Block = 1024 bins x a slab of rows, processed in batches of 8 rows:
//   1. compute z in fp64 for the batch already in registers, round exactly, drop the 2-byte results into a
//      shared staging row at the compacted position (selected bins of one block are contiguous in the output),
//   2. issue the 8 row loads (16 B each) of the NEXT batch, so they fly during steps 3-4,
//   3. one __syncthreads (the staging buffer is double-buffered),
//   4. copy the staged rows out with ALIGNED 16-byte stores, (chunk, row) pairs spread over all threads
//      (element stores only on the two ragged ends, which belong to neighbouring blocks).
// ---------------------------------------------------------------------------------------------------------
constexpr int kTransformSlab = 256;
constexpr int kTB = 8;                        // rows per batch
constexpr int kStageRow = kColsPerBlock + 16;  // halves per staged row (+8 alignment shift, +8 dump slots)
constexpr int kDump = kStageRow - 1;          // where unselected bins drop their (ignored) result

// kBlocked = false: row-major output, element (r, p) at out[r*ld_out + p].
// kBlocked = true : k-blocked operand layout of the Gram kernel, element (r, p) at out[((p/64)*ld_out + r)*64 + p%64]
//                   (ld_out = rows per 64-bin slab); an aligned 8-element chunk never straddles two k-blocks.
// stage[r][shift + e] holds output element p0 + e of row row0 + r, shift = p0 % 8, so 16-byte chunk q of the
// staged row lines up with the aligned global chunk starting at element (p0 - shift) + 8q.
template <typename T, bool kBlocked>
__device__ __forceinline__ void copy_out_rows(const T* __restrict__ stage, T* __restrict__ out, int64_t ld_out,
                                              int64_t row0, int nrows, int p0, int len) {
  const int shift = p0 & 7;
  const int nchunk = (shift + len + 7) >> 3;
  const int lane = threadIdx.x & 31;
  // one warp per staged row (kTB == warps per block): 32 consecutive 16-byte chunks = 512 contiguous bytes per store
  for (int r = threadIdx.x >> 5; r < nrows; r += kThreads / 32) {
    const T* src = stage + r * kStageRow + 8 * lane;
    const int64_t pe0 = (int64_t)(p0 - shift) + 8 * lane;  // first output element of this lane's first chunk
    T* dst = kBlocked ? out + ((pe0 >> 6) * ld_out + row0 + r) * 64 + (pe0 & 63) : out + (row0 + r) * ld_out + pe0;
    const int64_t step = kBlocked ? 4 * ld_out * 64 : 256;  // 32 chunks = 256 elements = 4 k-blocks further on
    for (int q = lane; q < nchunk; q += 32, src += 256, dst += step) {
      const int e0 = 8 * q - shift;  // first element of this chunk relative to p0
      if (e0 >= 0 && e0 + 8 <= len) *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(src);
    }
    // the (at most two) ragged chunks of the row: lanes 0-7 take the elements of the first chunk, lanes 8-15 those
    // of the last one -- one 2-byte store per lane instead of one lane walking eight guarded stores
    if (lane < 16) {
      const int q = lane < 8 ? 0 : nchunk - 1;
      const int e = 8 * q - shift + (lane & 7);  // element relative to p0
      const bool ragged = (8 * q - shift < 0) || (8 * q - shift + 8 > len);
      if (ragged && e >= 0 && e < len && (lane < 8 || nchunk > 1)) {
        const int64_t pe = (int64_t)p0 + e;
        T* d = kBlocked ? out + ((pe >> 6) * ld_out + row0 + r) * 64 + (pe & 63) : out + (row0 + r) * ld_out + pe;
        *d = stage[r * kStageRow + shift + e];
      }
    }
  }
}

// Rare path of the quantiser: t = w*a100 landed exactly on k+0.5 (the exact residual decides, as printf would on the
// exact value) or |t| is beyond the int16 code range (saturate and count).
__device__ __noinline__ int quantise_slow(double w, double a100, unsigned long long* sat) {
  const double t = w * a100;
  double q = round_even(t);
  if (fabs(t - q) == 0.5) {
    const double r = fma(w, a100, -t);
    if (r > 0.0) q = t + 0.5;
    else if (r < 0.0) q = t - 0.5;
  }
  if (fabs(q) > 32766.0) {
    if (sat) atomicAdd(sat, 1ULL);
    q = q > 0 ? 32766.0 : -32766.0;
  }
  return double_to_int_exact(q);
}

__device__ __forceinline__ void sts_u16(uint32_t addr, int v) {
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"((unsigned short)v) : "memory");
}

// Fast path of one cell: z in fp64, rounded half-even through the magic-number add (no conversion pipe).
// Returns the hundredths code (garbage when `rare` comes back set: exactly on .5, or beyond the int16 range).
__device__ __forceinline__ int quantise_fast(int h, double f, double negmu, double a100, bool& rare, bool& neg) {
  const double w = fma(u32_to_double(h), f, negmu);
  const double t = w * a100;
  const double tm = t + kRoundMagic;  // low word = round-half-even(t) as int32
  const double d = t - (tm - kRoundMagic);
  rare |= (fabs(d) == 0.5) | ((__double2hiint(t) & 0x7fffffff) >= 0x40DFFF00);
  neg = __double2hiint(t) < 0;
  return __double2loint(tm);
}

// One batch of kRows rows for this thread's 4 bins.  kRows is a compile-time constant for the full batches (no row
// guards in the unrolled body); the ragged last batch of a slab takes the generic instantiation with nr < kTB.
template <bool kZq, bool kZh, bool kFull, typename T>
__device__ __forceinline__ void transform_batch(const typename Cells<T>::Raw (&v)[kTB], int nr,
                                                const double* __restrict__ sf, const double (&negmu)[4],
                                                const double (&a100)[4], const int (&zc)[4], const uint32_t (&aq)[4],
                                                const uint32_t (&ah)[4], unsigned* __restrict__ part, bool lane0,
                                                const T* __restrict__ colp, int64_t ld,
                                                unsigned long long* __restrict__ sat, const Esc& esc, int64_t row0,
                                                int64_t j0) {
  unsigned rare_rows = 0;
#pragma unroll
  for (int u = 0; u < kTB; ++u) {
    if (!kFull && u >= nr) break;
    const double f = sf[u];
    int hh[4];
    Cells<T>::hund(v[u], esc, row0 + u, j0, hh);
    bool rare = false;
    int vsq = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int h = hh[q];
      bool neg;
      int z = quantise_fast(h, f, negmu[q], a100[q], rare, neg);
      if (kZq) {
        int code = (z == 0 && neg) ? -32767 : z;  // prints as "-0.00"
        if (h == 0) code = -32768;                // absent cell (a row without valid bins has no present cell either)
        sts_u16(aq[q] + u * (kStageRow * 2), code);
      }
      if (kZh) {
        const int c = h != 0 ? max(-zc[q], min(zc[q], z)) : 0;
        sts_u16(ah[q] + u * (kStageRow * 2), (int)__half_as_ushort(__int2half_rn(c)));
        vsq += c * c;
      }
    }
    if (rare) rare_rows |= 1u << u;
    if (kZh) {
      const int tot = __reduce_add_sync(0xffffffffu, vsq);  // <= 128 * 2048^2 = 2^29
      if (lane0) part[u] = (unsigned)tot;
    }
  }
  // Rare rows (a value exactly on k+0.5, or outside the int16 code range): redo this thread's 4 cells of that row
  // from the (cache-resident) source with the exact slow quantiser and patch the staged values and the norm.
  while (rare_rows) {
    const int u = __ffs(rare_rows) - 1;
    rare_rows &= rare_rows - 1;
    const double f = sf[u];
    int hh[4];
    Cells<T>::hund(Cells<T>::load_cached(colp + (int64_t)u * ld), esc, row0 + u, j0, hh);
    int delta = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int h = hh[q];
      if (h == 0 || !(f == f) || a100[q] == 0.0) continue;  // absent cell / unselected bin: nothing was kept
      bool r2 = false, neg;
      const int zf = quantise_fast(h, f, negmu[q], a100[q], r2, neg);
      if (!r2) continue;
      const double w = fma(u32_to_double(h), f, negmu[q]);
      const int z = quantise_slow(w, a100[q], sat);
      if (kZq) sts_u16(aq[q] + u * (kStageRow * 2), (z == 0 && w * a100[q] < 0.0) ? -32767 : z);
      if (kZh) {
        const int c_old = max(-zc[q], min(zc[q], zf)), c = max(-zc[q], min(zc[q], z));
        sts_u16(ah[q] + u * (kStageRow * 2), (int)__half_as_ushort(__int2half_rn(c)));
        delta += c * c - c_old * c_old;
      }
    }
    if (kZh && delta) atomicAdd(&part[u], (unsigned)delta);  // after the warp's own store to part[u] (same warp: ordered)
  }
}

template <bool kZq, bool kZh, typename T>
// (2-byte cells need half the row registers: three blocks fit an SM without spills, -2 %; four spill and lose 17 %)
__global__ void __launch_bounds__(kThreads, (sizeof(T) == 2 && !kZq) ? 3 : 2) k_transform(const T* __restrict__ depth, const Esc esc, int64_t n, int64_t b,
                                                           int64_t ld, const double* __restrict__ row_factor,
                                                           const double* __restrict__ mu,
                                                           const int32_t* __restrict__ sel_pos,
                                                           const uint8_t* __restrict__ keep,
                                                           const double* __restrict__ info, int zclip,
                                                           int16_t* __restrict__ zq, int64_t ldq,
                                                           __half* __restrict__ zh, int64_t ldh,
                                                           unsigned long long* __restrict__ norms,
                                                           unsigned long long* __restrict__ sat) {
  __shared__ __align__(16) int16_t s_zq[kZq ? 2 * kTB * kStageRow : 8];
  __shared__ __align__(16) __half s_zh[kZh ? 2 * kTB * kStageRow : 8];
  __shared__ double s_f[kTransformSlab];
  __shared__ unsigned s_part[kZh ? (kThreads / 32) * kTransformSlab : 1];  // per-warp partial row norms
  __shared__ int s_lo, s_hi;
  const int64_t r0 = (int64_t)blockIdx.y * kTransformSlab;
  const int nrow = (int)(n - r0 < kTransformSlab ? n - r0 : (int64_t)kTransformSlab);
  for (int t = threadIdx.x; t < kTransformSlab; t += kThreads) s_f[t] = t < nrow ? row_factor[r0 + t] : nan("");
  if (threadIdx.x == 0) s_lo = 0x7fffffff, s_hi = -1;
  __syncthreads();
  const int64_t j0 = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * 4;
  int pos[4];
  double negmu[4], a100[4];  // t = (h*f - mu) * (100*zscale/sqrt(mu))
  int zc[4];                 // clip bound of the fp16 operand; 0 for bins the kNN filter drops
  int lo = 0x7fffffff, hi = -1;
  const double zscale = info[0];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    pos[q] = (j0 + q < b) ? sel_pos[j0 + q] : -1;
    negmu[q] = 0.0, a100[q] = 0.0, zc[q] = 0;  // unselected bins run with t = 0 into the dump slots
    if (pos[q] >= 0) {
      lo = min(lo, pos[q]), hi = max(hi, pos[q]);
      const double m = mu[j0 + q];
      negmu[q] = -m;
      a100[q] = 100.0 * (zscale / sqrt(m));  // selected bins have a finite ratio, hence mu > 0
      zc[q] = (keep ? keep[j0 + q] != 0 : true) ? zclip : 0;
    }
  }
  lo = __reduce_min_sync(0xffffffffu, lo);
  hi = __reduce_max_sync(0xffffffffu, hi);
  const bool any = hi >= 0;  // WARP-uniform (the batch loop contains warp collectives): some lane has a selected bin
  if ((threadIdx.x & 31) == 0 && hi >= 0) {
    atomicMin(&s_lo, lo);
    atomicMax(&s_hi, hi);
  }
  __syncthreads();
  const int p0 = s_lo, len = s_hi - s_lo + 1;  // this block's contiguous output range
  if (s_hi < 0) return;                        // no selected bin in these 1024 columns (block-uniform)
  const int shift = p0 & 7;
  uint32_t aq[4], ah[4];  // shared-memory byte addresses of this thread's 4 staging slots in row 0 of buffer 0
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int off = pos[q] >= 0 ? pos[q] - p0 + shift : kDump - q;
    aq[q] = (uint32_t)__cvta_generic_to_shared(s_zq) + 2 * off;
    ah[q] = (uint32_t)__cvta_generic_to_shared(s_zh) + 2 * off;
  }
  using C = Cells<T>;
  const T* colp = depth + (j0 < b ? j0 : 0) + r0 * ld;
  const int warp = threadIdx.x >> 5;
  const bool lane0 = (threadIdx.x & 31) == 0;
  constexpr uint32_t kBufBytes = kTB * kStageRow * 2;

  typename C::Raw v[kTB];
  if (any) {
#pragma unroll
    for (int u = 0; u < kTB; ++u) v[u] = u < nrow ? C::load(colp + (int64_t)u * ld) : C::zero();
  }
  int buf = 0;
  for (int rb = 0; rb < nrow; rb += kTB, buf ^= 1) {
    const int nr = min(kTB, nrow - rb);
    if (any) {
      unsigned* part = s_part + (kZh ? warp * kTransformSlab + rb : 0);
      const T* src = colp + (int64_t)rb * ld;
      if (nr == kTB)
        transform_batch<kZq, kZh, true, T>(v, nr, s_f + rb, negmu, a100, zc, aq, ah, part, lane0, src, ld, sat, esc, r0 + rb, j0);
      else
        transform_batch<kZq, kZh, false, T>(v, nr, s_f + rb, negmu, a100, zc, aq, ah, part, lane0, src, ld, sat, esc, r0 + rb, j0);
      // next batch: in flight during the barrier and the copy-out
      const int rn = rb + kTB;
      if (rn + kTB <= nrow) {
#pragma unroll
        for (int u = 0; u < kTB; ++u) v[u] = C::load(colp + (int64_t)(rn + u) * ld);
      } else {
#pragma unroll
        for (int u = 0; u < kTB; ++u)
          if (rn + u < nrow) v[u] = C::load(colp + (int64_t)(rn + u) * ld);
      }
    } else if (kZh) {
      for (int u = 0; u < nr; ++u)
        if (lane0) s_part[warp * kTransformSlab + rb + u] = 0;
    }
    __syncthreads();
    if (kZq) copy_out_rows<int16_t, false>(s_zq + (buf ? kTB * kStageRow : 0), zq, ldq, r0 + rb, nr, p0, len);
    if (kZh) copy_out_rows<__half, true>(s_zh + (buf ? kTB * kStageRow : 0), zh, ldh, r0 + rb, nr, p0, len);
    // the other staging buffer next time
    const uint32_t flip = buf ? (0u - kBufBytes) : kBufBytes;
#pragma unroll
    for (int q = 0; q < 4; ++q) aq[q] += kZq ? flip : 0u, ah[q] += kZh ? flip : 0u;
  }
  if (kZh) {
    __syncthreads();
    for (int t = threadIdx.x; t < nrow; t += kThreads) {
      unsigned long long tot = 0;
#pragma unroll
      for (int w = 0; w < kThreads / 32; ++w) tot += s_part[w * kTransformSlab + t];
      if (tot) atomicAdd(&norms[r0 + t], tot);
    }
  }
}

// zq (hundredths, from a step-4 file) -> clipped fp16 operand + exact norms.  One block per row.
__global__ void __launch_bounds__(kThreads) k_knn_prepare(const int16_t* __restrict__ zq, int64_t rw, int64_t ldq,
                                                          const uint8_t* __restrict__ keep, int zclip,
                                                          __half* __restrict__ zh, int64_t ldh,
                                                          unsigned long long* __restrict__ norms) {
  const int64_t i = blockIdx.x;
  unsigned long long acc = 0;
  const int64_t kdim = (rw + 63) / 64 * 64;  // zero padded to whole k-blocks
  for (int64_t j = threadIdx.x; j < kdim; j += kThreads) {
    int c = 0;
    if (j < rw) {
      const int v = zq[i * ldq + j];
      if (v > -32767 && (!keep || keep[j])) c = max(-zclip, min(zclip, v));
    }
    zh[((j >> 6) * ldh + i) * 64 + (j & 63)] = __int2half_rn(c);  // k-blocked, ldh = rows per slab
    acc += (unsigned)(c * c);
  }
  acc = (unsigned long long)warp_sum((long long)acc);
  __shared__ unsigned long long ss[kThreads / 32];
  if ((threadIdx.x & 31) == 0) ss[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t = 0;
    for (int w = 0; w < kThreads / 32; ++w) t += ss[w];
    norms[i] = t;
  }
}

// ---------------------------------------------------------------------------------------------------------
// Escape patches of the 2-byte matrix.  The streaming kernels above took every escape cell as the code 65535; these small
// kernels (one thread or warp per escape cell / column) put the true hundredths H in, so that every result equals the
// fp32 path's bit for bit.
// ---------------------------------------------------------------------------------------------------------
// K1 patch: integer sums, exact in any order
__global__ void k_esc_raw(const Esc esc, int64_t n, unsigned long long* __restrict__ colsum,
                          unsigned long long* __restrict__ rowsum_all) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= esc.n) return;
  const long long row = esc.key[e] / esc.b - esc.row0, col = esc.key[e] % esc.b;
  if (row < 0 || row >= n) return;
  const unsigned long long extra = (unsigned long long)(esc.h[e] - 65535);
  atomicAdd(&colsum[col], extra);
  atomicAdd(&rowsum_all[row], extra);
}

// K4 patch: the bins that hold an escape are summed again, rows strictly in order (as the streaming kernel does per bin),
// one warp per (bin, row slab): 32 rows are fetched at once, lane 0 adds them one after the other.
__global__ void k_esc_col_norm(const uint16_t* __restrict__ depth, const Esc esc, const int* __restrict__ cols, int64_t ncols,
                               int64_t n, int64_t b, int64_t ld, int64_t rows_per_slab, const uint8_t* __restrict__ valid,
                               const double* __restrict__ row_factor, double* __restrict__ out_sd,
                               double* __restrict__ out_sd2) {
  const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= ncols) return;
  const int64_t j = cols[w];
  if (j < 0 || j >= b || !valid[j]) return;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_slab, r1 = min(n, r0 + rows_per_slab);
  double s1 = 0.0, s2 = 0.0;
  for (int64_t base = r0; base < r1; base += 32) {
    const int64_t i = base + lane;
    int h = 0;
    double f = 0.0;
    if (i < r1) {
      h = Cells<uint16_t>::hund1(depth + i * ld + j, esc, i, j);
      f = row_factor[i];
    }
    const int cnt = (int)min((int64_t)32, r1 - base);
    for (int t = 0; t < cnt; ++t) {
      const int ht = __shfl_sync(0xffffffffu, h, t);
      const double ft = __shfl_sync(0xffffffffu, f, t);
      if (ht > 0) {
        const double d = fma(u32_to_double((unsigned)ht), ft, -1.0);
        s1 += d;
        s2 = fma(d, d, s2);
      }
    }
  }
  if (lane == 0) {
    out_sd[(int64_t)blockIdx.y * b + j] = s1;
    out_sd2[(int64_t)blockIdx.y * b + j] = s2;
  }
}

// K7 patch: one thread per escape cell recomputes the cell's code from H (the exact quantiser the rare path uses)
__global__ void k_esc_transform(const Esc esc, int64_t n, int64_t b, const double* __restrict__ row_factor,
                                const double* __restrict__ mu, const int32_t* __restrict__ sel_pos,
                                const uint8_t* __restrict__ keep, const double* __restrict__ info, int zclip,
                                int16_t* __restrict__ zq, int64_t ldq, __half* __restrict__ zh, int64_t ldh,
                                unsigned long long* __restrict__ norms, unsigned long long* __restrict__ sat) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= esc.n) return;
  const long long row = esc.key[e] / esc.b - esc.row0, col = esc.key[e] % esc.b;
  if (row < 0 || row >= n || col >= b) return;
  const int pos = sel_pos[col];
  if (pos < 0) return;
  const double f = row_factor[row];
  if (!(f == f)) return;
  const double m = mu[col], a100 = 100.0 * (info[0] / sqrt(m));
  unsigned long long sat_true = 0, sat_coded = 0;
  const double w_true = fma(u32_to_double((unsigned)esc.h[e]), f, -m), w_coded = fma(u32_to_double(65535u), f, -m);
  const int z_true = quantise_slow(w_true, a100, nullptr), z_coded = quantise_slow(w_coded, a100, nullptr);
  sat_true = fabs(round_even(w_true * a100)) > 32766.0, sat_coded = fabs(round_even(w_coded * a100)) > 32766.0;
  if (zq) zq[row * ldq + pos] = (int16_t)((z_true == 0 && w_true * a100 < 0.0) ? -32767 : z_true);
  if (zh) {
    const int zc = (keep ? keep[col] != 0 : true) ? zclip : 0;
    const int c_true = max(-zc, min(zc, z_true)), c_coded = max(-zc, min(zc, z_coded));
    if (c_true != c_coded) {
      zh[(((int64_t)pos >> 6) * ldh + row) * 64 + (pos & 63)] = __int2half_rn(c_true);
      atomicAdd(&norms[row], (unsigned long long)(long long)(c_true * c_true - c_coded * c_coded));
    }
  }
  if (sat && sat_true > sat_coded) atomicAdd(sat, 1ULL);
}

}  // namespace g2

using namespace g2;

static int check_matrix(const void* p, int64_t n, int64_t b, int64_t ld, int align = 16) {
  G2_REQUIRE(p != nullptr, GRIDB200_EINVAL, "depth matrix pointer is null");
  G2_REQUIRE(n > 0 && b > 0, GRIDB200_EINVAL, "empty matrix (n=%lld, b=%lld)", (long long)n, (long long)b);
  G2_REQUIRE(ld >= b && ld % 4 == 0, GRIDB200_EINVAL, "leading dimension %lld must be >= b=%lld and a multiple of 4",
             (long long)ld, (long long)b);
  G2_REQUIRE((reinterpret_cast<uintptr_t>(p) & (align - 1)) == 0, GRIDB200_EINVAL, "depth matrix must be %d-byte aligned", align);
  G2_REQUIRE(n < (1LL << 31) && b < (1LL << 40), GRIDB200_EINVAL, "matrix too large");
  return 0;
}

static int make_esc(Esc& e, const int64_t* key, const int32_t* h, int64_t n_esc, int64_t row0, int64_t b_key) {
  G2_REQUIRE(n_esc >= 0 && (n_esc == 0 || (key && h)) && row0 >= 0 && b_key > 0, GRIDB200_EINVAL, "bad escape table");
  e.key = reinterpret_cast<const long long*>(key), e.h = h, e.n = n_esc, e.row0 = row0, e.b = b_key;
  return 0;
}

extern "C" int gridb200_col_raw_stats(const float* depth, int64_t n, int64_t b, int64_t ld, int64_t* colsum,
                                      int32_t* colcnt, void* stream) {
  if (int e = check_matrix(depth, n, b, ld)) return e;
  G2_REQUIRE(colsum && colcnt, GRIDB200_EINVAL, "null output");
  cudaStream_t st = as_stream(stream);
  const int64_t cb = (b + kColsPerBlock - 1) / kColsPerBlock;
  const int64_t rows = pick_row_slab(n, cb, 64);
  const int64_t slabs = (n + rows - 1) / rows;
  G2_REQUIRE(slabs <= 65535, GRIDB200_EINVAL, "too many row slabs");
  if (slabs > 1) {
    G2_CUDA(cudaMemsetAsync(colsum, 0, b * sizeof(int64_t), st));
    G2_CUDA(cudaMemsetAsync(colcnt, 0, b * sizeof(int32_t), st));
  }
  k_col_raw<false, float><<<dim3((unsigned)cb, (unsigned)slabs), kThreads, 0, st>>>(
      depth, Esc{}, n, b, ld, rows, reinterpret_cast<unsigned long long*>(colsum), colcnt, nullptr, nullptr);
  G2_LAUNCHED();
  return 0;
}

template <typename T>
static int col_row_raw_impl(const T* depth, const Esc& esc, int64_t n, int64_t b, int64_t ld, int64_t* colsum,
                            int32_t* colcnt, int64_t* rowsum_all, int32_t* rowcnt_all, void* stream) {
  if (int e = check_matrix(depth, n, b, ld, 16)) return e;
  G2_REQUIRE(ld % Wide<T>::kN == 0, GRIDB200_EINVAL, "leading dimension must be a multiple of %d", Wide<T>::kN);
  G2_REQUIRE(colsum && colcnt && rowsum_all && rowcnt_all, GRIDB200_EINVAL, "null output");
  cudaStream_t st = as_stream(stream);
  const int64_t cb = (b + kThreads * Wide<T>::kN - 1) / (kThreads * Wide<T>::kN);
  // 2-byte cells: 3 blocks of 2048 bins per SM are resident, and with one slab every block walks all n rows -- 1417
  // equal blocks are 3.2 waves at config 2, a fifth of the machine idle in the last.  ~48 blocks per SM (row slabs of
  // >= 512 rows, a multiple of 32) make the last wave a sixteenth of the run; the sums are integers, so the slab
  // partials meet in exact atomics.
  int64_t rows = sizeof(T) == 2 ? pick_row_slab(n, cb, 512, 48) : pick_row_slab(n, cb, 64);
  if (sizeof(T) == 2 && rows < n) rows = (rows + 31) / 32 * 32;
  const int64_t slabs = (n + rows - 1) / rows;
  G2_REQUIRE(slabs <= 65535, GRIDB200_EINVAL, "too many row slabs");
  if (slabs > 1) {
    G2_CUDA(cudaMemsetAsync(colsum, 0, b * sizeof(int64_t), st));
    G2_CUDA(cudaMemsetAsync(colcnt, 0, b * sizeof(int32_t), st));
  }
  G2_CUDA(cudaMemsetAsync(rowsum_all, 0, n * sizeof(int64_t), st));
  G2_CUDA(cudaMemsetAsync(rowcnt_all, 0, n * sizeof(int32_t), st));
  if constexpr (sizeof(T) == 2) {
    k_col_raw_u16<<<dim3((unsigned)cb, (unsigned)slabs), kThreads, 0, st>>>(
        depth, n, b, ld, rows, reinterpret_cast<unsigned long long*>(colsum), colcnt,
        reinterpret_cast<unsigned long long*>(rowsum_all), rowcnt_all);
  } else {
    k_col_raw<true, T><<<dim3((unsigned)cb, (unsigned)slabs), kThreads, 0, st>>>(
        depth, esc, n, b, ld, rows, reinterpret_cast<unsigned long long*>(colsum), colcnt,
        reinterpret_cast<unsigned long long*>(rowsum_all), rowcnt_all);
  }
  G2_LAUNCHED();
  if (Cells<T>::kEsc && esc.n > 0) {
    k_esc_raw<<<(unsigned)((esc.n + 255) / 256), 256, 0, st>>>(esc, n, reinterpret_cast<unsigned long long*>(colsum),
                                                                reinterpret_cast<unsigned long long*>(rowsum_all));
    G2_LAUNCHED();
  }
  return 0;
}

extern "C" int gridb200_col_row_raw_stats(const float* depth, int64_t n, int64_t b, int64_t ld, int64_t* colsum,
                                          int32_t* colcnt, int64_t* rowsum_all, int32_t* rowcnt_all, void* stream) {
  return col_row_raw_impl<float>(depth, Esc{}, n, b, ld, colsum, colcnt, rowsum_all, rowcnt_all, stream);
}

extern "C" size_t gridb200_row_stats_fused_ws_bytes(int64_t b) { return (size_t)(b > 0 ? b : 0) * 4 + 256; }

template <typename T>
static int row_stats_fused_impl(const T* depth, const Esc& esc, int64_t n, int64_t b, int64_t ld, const uint8_t* valid,
                                const int32_t* colcnt, const int64_t* rowsum_all, const int32_t* rowcnt_all,
                                int64_t* rowsum, int32_t* rowcnt, void* ws, size_t ws_bytes, void* stream) {
  if (int e = check_matrix(depth, n, b, ld, 4 * (int)sizeof(T))) return e;
  G2_REQUIRE(valid && colcnt && rowsum_all && rowcnt_all && rowsum && rowcnt, GRIDB200_EINVAL, "null argument");
  G2_REQUIRE((reinterpret_cast<uintptr_t>(valid) & 3) == 0, GRIDB200_EINVAL, "valid flags must be 4-byte aligned");
  G2_REQUIRE(ws && ws_bytes >= gridb200_row_stats_fused_ws_bytes(b), GRIDB200_EINVAL,
             "workspace too small: need %zu bytes", gridb200_row_stats_fused_ws_bytes(b));
  G2_REQUIRE(b < (1LL << 31) && n <= 0x7fffffff, GRIDB200_EINVAL, "matrix too large");
  cudaStream_t st = as_stream(stream);
  int* flags = static_cast<int*>(ws);  // [0] = number of bad bins, [1] = mode
  int* bad = flags + 64;
  G2_CUDA(cudaMemsetAsync(flags, 0, 256, st));
  G2_CUDA(cudaMemsetAsync(rowsum, 0, n * sizeof(int64_t), st));
  G2_CUDA(cudaMemsetAsync(rowcnt, 0, n * sizeof(int32_t), st));
  k_bad_list<<<(unsigned)((b + 255) / 256), 256, 0, st>>>(valid, colcnt, b, bad, flags);
  G2_LAUNCHED();
  k_row_mode<<<1, 1, 0, st>>>(flags, b, flags + 1);
  G2_LAUNCHED();
  const int64_t chunks = (b + kRowChunk - 1) / kRowChunk;
  for (int64_t r0 = 0; r0 < n; r0 += 65535) {  // gridDim.y limit
    const int64_t rows = std::min<int64_t>(65535, n - r0);
    k_row_stats<T><<<dim3((unsigned)chunks, (unsigned)std::min<int64_t>(rows, kRowGridY)), kThreads, 0, st>>>(
        depth + r0 * ld, esc, r0, b, ld, valid, reinterpret_cast<unsigned long long*>(rowsum) + r0, rowcnt + r0, flags + 1,
        rows);
    G2_LAUNCHED();
  }
  k_row_fix<T><<<(unsigned)n, kThreads, 0, st>>>(depth, esc, ld, bad, flags, flags + 1,
                                                 reinterpret_cast<const unsigned long long*>(rowsum_all), rowcnt_all,
                                                 reinterpret_cast<unsigned long long*>(rowsum), rowcnt);
  G2_LAUNCHED();
  return 0;
}

extern "C" int gridb200_row_stats_fused(const float* depth, int64_t n, int64_t b, int64_t ld, const uint8_t* valid,
                                        const int32_t* colcnt, const int64_t* rowsum_all, const int32_t* rowcnt_all,
                                        int64_t* rowsum, int32_t* rowcnt, void* ws, size_t ws_bytes, void* stream) {
  return row_stats_fused_impl<float>(depth, Esc{}, n, b, ld, valid, colcnt, rowsum_all, rowcnt_all, rowsum, rowcnt, ws,
                                     ws_bytes, stream);
}

extern "C" int gridb200_depth_filter(const int64_t* colsum, const int32_t* colcnt, int64_t b, double min_depth,
                                     double max_depth, uint8_t* valid, void* stream) {
  G2_REQUIRE(colsum && colcnt && valid && b > 0, GRIDB200_EINVAL, "bad argument");
  k_depth_filter<<<(unsigned)((b + 255) / 256), 256, 0, as_stream(stream)>>>(
      reinterpret_cast<const long long*>(colsum), colcnt, b, min_depth, max_depth, valid);
  G2_LAUNCHED();
  return 0;
}

extern "C" int gridb200_row_stats(const float* depth, int64_t n, int64_t b, int64_t ld, const uint8_t* valid,
                                  int64_t* rowsum, int32_t* rowcnt, void* stream) {
  if (int e = check_matrix(depth, n, b, ld)) return e;
  G2_REQUIRE(valid && rowsum && rowcnt, GRIDB200_EINVAL, "null argument");
  G2_REQUIRE((reinterpret_cast<uintptr_t>(valid) & 3) == 0, GRIDB200_EINVAL, "valid flags must be 4-byte aligned");
  cudaStream_t st = as_stream(stream);
  G2_CUDA(cudaMemsetAsync(rowsum, 0, n * sizeof(int64_t), st));
  G2_CUDA(cudaMemsetAsync(rowcnt, 0, n * sizeof(int32_t), st));
  const int64_t chunks = (b + kRowChunk - 1) / kRowChunk;
  for (int64_t r0 = 0; r0 < n; r0 += 65535) {  // gridDim.y limit
    const int64_t rows = std::min<int64_t>(65535, n - r0);
    k_row_stats<float><<<dim3((unsigned)chunks, (unsigned)std::min<int64_t>(rows, kRowGridY)), kThreads, 0, st>>>(
        depth + r0 * ld, Esc{}, r0, b, ld, valid, reinterpret_cast<unsigned long long*>(rowsum) + r0, rowcnt + r0, nullptr,
        rows);
    G2_LAUNCHED();
  }
  return 0;
}

extern "C" int gridb200_row_scales(const int64_t* rowsum, const int32_t* rowcnt, int64_t n, double* raw_mean,
                                   double* scale_q, double* row_factor, void* stream) {
  G2_REQUIRE(rowsum && rowcnt && n > 0, GRIDB200_EINVAL, "bad argument");
  k_row_scales<<<(unsigned)((n + 255) / 256), 256, 0, as_stream(stream)>>>(
      reinterpret_cast<const long long*>(rowsum), rowcnt, n, raw_mean, scale_q, row_factor);
  G2_LAUNCHED();
  return 0;
}

template <typename T>
static int row_means_exact_impl(const T* depth, const Esc& esc, int64_t n, int64_t b, int64_t ld, const uint8_t* valid,
                                const int64_t* rowsum, const int32_t* rowcnt, double* raw_mean, double* scale_q,
                                void* stream) {
  if (int e = check_matrix(depth, n, b, ld, 4 * (int)sizeof(T))) return e;
  G2_REQUIRE(valid && rowsum && rowcnt, GRIDB200_EINVAL, "null argument");
  k_row_mean_ties<T><<<(unsigned)((n + 63) / 64), 64, 0, as_stream(stream)>>>(
      depth, esc, n, b, ld, valid, reinterpret_cast<const long long*>(rowsum), rowcnt, raw_mean, scale_q);
  G2_LAUNCHED();
  return 0;
}

extern "C" int gridb200_row_means_exact(const float* depth, int64_t n, int64_t b, int64_t ld, const uint8_t* valid,
                                        const int64_t* rowsum, const int32_t* rowcnt, double* raw_mean,
                                        double* scale_q, void* stream) {
  return row_means_exact_impl<float>(depth, Esc{}, n, b, ld, valid, rowsum, rowcnt, raw_mean, scale_q, stream);
}

static int64_t col_norm_slab_rows(int64_t n, int64_t b) {
  const int64_t cb = (b + kColsPerBlock - 1) / kColsPerBlock;
  if (cb >= (int64_t)kSMs * 4) {
    // Plenty of bin blocks -- but with ONE slab every block walks all n rows, and on the 2-byte matrix (2048 bins per
    // block, 3 blocks resident per SM) config 2 is 3.2 waves of equal blocks: a fifth of the machine idle in the last.
    // Row slabs of >= 256 rows for ~96 fp32-sized blocks per SM cut the tail to a few percent; the partials are reduced
    // in slab order (deterministic), at 16 bytes of workspace per bin and slab.
    const int64_t rows = pick_row_slab(n, cb, 256, 96);
    return rows >= n ? n : (rows + 31) / 32 * 32;
  }
  return pick_row_slab(n, cb, 32);
}

extern "C" size_t gridb200_col_norm_stats_ws_bytes(int64_t n, int64_t b) {
  if (n <= 0 || b <= 0) return 0;
  const int64_t rows = col_norm_slab_rows(n, b);
  const int64_t slabs = (n + rows - 1) / rows;
  return slabs > 1 ? (size_t)slabs * (size_t)b * 2 * sizeof(double) : 0;
}

template <typename T>
static int col_norm_patch(const T*, const Esc&, const int32_t*, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t,
                          const uint8_t*, const double*, double*, double*, cudaStream_t) {
  return 0;
}
template <>
int col_norm_patch<uint16_t>(const uint16_t* depth, const Esc& esc, const int32_t* cols, int64_t ncols, int64_t n, int64_t b,
                             int64_t ld, int64_t rows, int64_t slabs, const uint8_t* valid, const double* row_factor,
                             double* o1, double* o2, cudaStream_t st) {
  if (esc.n == 0) return 0;
  G2_REQUIRE(cols && ncols > 0, GRIDB200_EINVAL, "escape cells need the list of their bins (esc_cols)");
  k_esc_col_norm<<<dim3((unsigned)((ncols * 32 + 127) / 128), (unsigned)slabs), 128, 0, st>>>(
      depth, esc, cols, ncols, n, b, ld, rows, valid, row_factor, o1, o2);
  G2_LAUNCHED();
  return 0;
}

template <typename T>
static int col_norm_impl(const T* depth, const Esc& esc, const int32_t* esc_cols, int64_t n_esc_cols, int64_t n, int64_t b,
                         int64_t ld, const uint8_t* valid, const double* row_factor, double* sd, double* sd2, void* ws,
                         size_t ws_bytes, void* stream) {
  if (int e = check_matrix(depth, n, b, ld, 16)) return e;
  G2_REQUIRE(ld % Wide<T>::kN == 0, GRIDB200_EINVAL, "leading dimension must be a multiple of %d", Wide<T>::kN);
  G2_REQUIRE(valid && row_factor && sd && sd2, GRIDB200_EINVAL, "null argument");
  cudaStream_t st = as_stream(stream);
  const int64_t cb = (b + kThreads * Wide<T>::kN - 1) / (kThreads * Wide<T>::kN);
  const int64_t rows = col_norm_slab_rows(n, b);  // (the row slabs depend on b only: the same partial sums for either storage)
  const int64_t slabs = (n + rows - 1) / rows;
  G2_REQUIRE(slabs <= 65535, GRIDB200_EINVAL, "too many row slabs");
  if (slabs == 1) {
    if constexpr (sizeof(T) == 2)
      k_col_norm_u16<<<dim3((unsigned)cb, 1), kThreads, 0, st>>>(depth, n, b, ld, rows, valid, row_factor, sd, sd2);
    else
      k_col_norm<T><<<dim3((unsigned)cb, 1), kThreads, 0, st>>>(depth, esc, n, b, ld, rows, valid, row_factor, sd, sd2);
    G2_LAUNCHED();
    if (int e = col_norm_patch<T>(depth, esc, esc_cols, n_esc_cols, n, b, ld, rows, 1, valid, row_factor, sd, sd2, st)) return e;
  } else {
    G2_REQUIRE(ws && ws_bytes >= gridb200_col_norm_stats_ws_bytes(n, b), GRIDB200_EINVAL,
               "workspace too small: need %zu bytes", gridb200_col_norm_stats_ws_bytes(n, b));
    double* p1 = static_cast<double*>(ws);
    double* p2 = p1 + slabs * b;
    if constexpr (sizeof(T) == 2)
      k_col_norm_u16<<<dim3((unsigned)cb, (unsigned)slabs), kThreads, 0, st>>>(depth, n, b, ld, rows, valid, row_factor, p1, p2);
    else
      k_col_norm<T><<<dim3((unsigned)cb, (unsigned)slabs), kThreads, 0, st>>>(depth, esc, n, b, ld, rows, valid, row_factor,
                                                                               p1, p2);
    G2_LAUNCHED();
    if (int e = col_norm_patch<T>(depth, esc, esc_cols, n_esc_cols, n, b, ld, rows, slabs, valid, row_factor, p1, p2, st)) return e;
    k_reduce_slabs<<<(unsigned)((b + 255) / 256), 256, 0, st>>>(p1, p2, b, (int)slabs, sd, sd2);
    G2_LAUNCHED();
  }
  return 0;
}

extern "C" int gridb200_col_norm_stats(const float* depth, int64_t n, int64_t b, int64_t ld, const uint8_t* valid,
                                       const double* row_factor, double* sd, double* sd2, void* ws,
                                       size_t ws_bytes, void* stream) {
  return col_norm_impl<float>(depth, Esc{}, nullptr, 0, n, b, ld, valid, row_factor, sd, sd2, ws, ws_bytes, stream);
}

extern "C" int gridb200_col_finalize(const double* sd, const double* sd2, const int32_t* colcnt,
                                     const uint8_t* valid, int64_t b, int64_t n_total, double* mu, double* s2,
                                     double* ratio, void* stream) {
  G2_REQUIRE(sd && sd2 && colcnt && valid && mu && s2 && ratio && b > 0, GRIDB200_EINVAL, "bad argument");
  k_col_finalize<<<(unsigned)((b + 255) / 256), 256, 0, as_stream(stream)>>>(sd, sd2, colcnt, valid, b, n_total, mu,
                                                                            s2, ratio);
  G2_LAUNCHED();
  return 0;
}

template <typename T>
static int transform_impl(const T* depth, const Esc& esc, int64_t n, int64_t b, int64_t ld, const double* row_factor,
                          const double* mu, const int32_t* sel_pos, const uint8_t* keep, const double* info,
                          int32_t zclip, int16_t* zq, int64_t ldq, void* zh, int64_t ldh, int64_t* norms,
                          unsigned long long* sat, void* stream) {
  if (int e = check_matrix(depth, n, b, ld, 4 * (int)sizeof(T))) return e;
  G2_REQUIRE(row_factor && mu && sel_pos && info && sat, GRIDB200_EINVAL, "null argument");
  G2_REQUIRE(zq || zh, GRIDB200_EINVAL, "at least one of zq / zh must be requested");
  G2_REQUIRE(!zh || (norms && zclip > 0 && zclip <= 2048), GRIDB200_EINVAL,
             "zh needs norms and 0 < zclip <= 2048 (fp16 holds integers exactly up to 2048)");
  G2_REQUIRE(!zq || (ldq % 8 == 0 && (reinterpret_cast<uintptr_t>(zq) & 15) == 0), GRIDB200_EINVAL,
             "zq must be 16-byte aligned with ldq a multiple of 8");
  G2_REQUIRE(!zh || (ldh >= n && (reinterpret_cast<uintptr_t>(zh) & 15) == 0), GRIDB200_EINVAL,
             "zh must be 16-byte aligned with at least n rows per k-block slab");
  const int64_t cb = (b + kColsPerBlock - 1) / kColsPerBlock;
  const int64_t slabs = (n + kTransformSlab - 1) / kTransformSlab;
  G2_REQUIRE(slabs <= 65535, GRIDB200_EINVAL, "too many row slabs");
  dim3 grid((unsigned)cb, (unsigned)slabs);
  cudaStream_t st = as_stream(stream);
  __half* zhp = static_cast<__half*>(zh);
  unsigned long long* np = reinterpret_cast<unsigned long long*>(norms);
  // both outputs wanted (tests, small cohorts): two launches, so that one block's double-buffered staging stays
  // inside the 48 KB static shared-memory limit; the second launch does not count saturations again
  if (zq) {
    k_transform<true, false, T><<<grid, kThreads, 0, st>>>(depth, esc, n, b, ld, row_factor, mu, sel_pos, keep, info, zclip,
                                                           zq, ldq, zhp, ldh, np, sat);
    if (zh) G2_LAUNCHED();
  }
  if (zh)
    k_transform<false, true, T><<<grid, kThreads, 0, st>>>(depth, esc, n, b, ld, row_factor, mu, sel_pos, keep, info, zclip,
                                                           zq, ldq, zhp, ldh, np, zq ? nullptr : sat);
  G2_LAUNCHED();
  if (Cells<T>::kEsc && esc.n > 0) {
    k_esc_transform<<<(unsigned)((esc.n + 255) / 256), 256, 0, st>>>(esc, n, b, row_factor, mu, sel_pos, keep, info, zclip, zq,
                                                                      ldq, zhp, ldh, np, sat);
    G2_LAUNCHED();
  }
  return 0;
}

extern "C" int gridb200_transform(const float* depth, int64_t n, int64_t b, int64_t ld, const double* row_factor,
                                  const double* mu, const int32_t* sel_pos, const uint8_t* keep,
                                  const double* info, int32_t zclip, int16_t* zq, int64_t ldq, void* zh,
                                  int64_t ldh, int64_t* norms, unsigned long long* sat, void* stream) {
  return transform_impl<float>(depth, Esc{}, n, b, ld, row_factor, mu, sel_pos, keep, info, zclip, zq, ldq, zh, ldh, norms,
                               sat, stream);
}

// ---- the same five passes on the 2-byte matrix (uint16 hundredths, 0 = absent, 0xFFFF = see the escape table) ----------
extern "C" int gridb200_col_row_raw_stats_u16(const uint16_t* depth, int64_t n, int64_t b, int64_t ld,
                                              const int64_t* esc_key, const int32_t* esc_h, int64_t n_esc, int64_t esc_row0,
                                              int64_t esc_b, int64_t* colsum, int32_t* colcnt, int64_t* rowsum_all,
                                              int32_t* rowcnt_all, void* stream) {
  Esc esc;
  if (int e = make_esc(esc, esc_key, esc_h, n_esc, esc_row0, esc_b)) return e;
  return col_row_raw_impl<uint16_t>(depth, esc, n, b, ld, colsum, colcnt, rowsum_all, rowcnt_all, stream);
}

extern "C" int gridb200_row_stats_fused_u16(const uint16_t* depth, int64_t n, int64_t b, int64_t ld, const int64_t* esc_key,
                                            const int32_t* esc_h, int64_t n_esc, int64_t esc_row0, int64_t esc_b,
                                            const uint8_t* valid, const int32_t* colcnt, const int64_t* rowsum_all,
                                            const int32_t* rowcnt_all, int64_t* rowsum, int32_t* rowcnt, void* ws,
                                            size_t ws_bytes, void* stream) {
  Esc esc;
  if (int e = make_esc(esc, esc_key, esc_h, n_esc, esc_row0, esc_b)) return e;
  return row_stats_fused_impl<uint16_t>(depth, esc, n, b, ld, valid, colcnt, rowsum_all, rowcnt_all, rowsum, rowcnt, ws,
                                        ws_bytes, stream);
}

extern "C" int gridb200_row_means_exact_u16(const uint16_t* depth, int64_t n, int64_t b, int64_t ld, const int64_t* esc_key,
                                            const int32_t* esc_h, int64_t n_esc, int64_t esc_row0, int64_t esc_b,
                                            const uint8_t* valid, const int64_t* rowsum, const int32_t* rowcnt,
                                            double* raw_mean, double* scale_q, void* stream) {
  Esc esc;
  if (int e = make_esc(esc, esc_key, esc_h, n_esc, esc_row0, esc_b)) return e;
  return row_means_exact_impl<uint16_t>(depth, esc, n, b, ld, valid, rowsum, rowcnt, raw_mean, scale_q, stream);
}

extern "C" int gridb200_col_norm_stats_u16(const uint16_t* depth, int64_t n, int64_t b, int64_t ld, const int64_t* esc_key,
                                           const int32_t* esc_h, int64_t n_esc, int64_t esc_row0, int64_t esc_b,
                                           const int32_t* esc_cols, int64_t n_esc_cols, const uint8_t* valid, const double* row_factor, double* sd, double* sd2,
                                           void* ws, size_t ws_bytes, void* stream) {
  Esc esc;
  if (int e = make_esc(esc, esc_key, esc_h, n_esc, esc_row0, esc_b)) return e;
  return col_norm_impl<uint16_t>(depth, esc, esc_cols, n_esc_cols, n, b, ld, valid, row_factor, sd, sd2, ws, ws_bytes, stream);
}

extern "C" int gridb200_transform_u16(const uint16_t* depth, int64_t n, int64_t b, int64_t ld, const int64_t* esc_key,
                                      const int32_t* esc_h, int64_t n_esc, int64_t esc_row0, int64_t esc_b,
                                      const double* row_factor, const double* mu, const int32_t* sel_pos,
                                      const uint8_t* keep, const double* info, int32_t zclip, int16_t* zq, int64_t ldq,
                                      void* zh, int64_t ldh, int64_t* norms, unsigned long long* sat, void* stream) {
  Esc esc;
  if (int e = make_esc(esc, esc_key, esc_h, n_esc, esc_row0, esc_b)) return e;
  return transform_impl<uint16_t>(depth, esc, n, b, ld, row_factor, mu, sel_pos, keep, info, zclip, zq, ldq, zh, ldh, norms,
                                  sat, stream);
}

extern "C" int gridb200_knn_prepare(const int16_t* zq, int64_t n, int64_t rw, int64_t ldq, const uint8_t* keep,
                                    int32_t zclip, void* zh, int64_t ldh, int64_t* norms, void* stream) {
  G2_REQUIRE(zq && zh && norms && n > 0 && rw >= 0 && ldq >= rw && ldh >= n, GRIDB200_EINVAL, "bad argument");
  G2_REQUIRE(zclip > 0 && zclip <= 2048, GRIDB200_EINVAL, "0 < zclip <= 2048 required");
  G2_REQUIRE(n < (1LL << 31), GRIDB200_EINVAL, "too many rows");
  k_knn_prepare<<<(unsigned)n, kThreads, 0, as_stream(stream)>>>(zq, rw, ldq, keep, zclip, static_cast<__half*>(zh),
                                                                 ldh, reinterpret_cast<unsigned long long*>(norms));
  G2_LAUNCHED();
  return 0;
}
