// The two producers of the path's inputs that are data-parallel:
//
//   gridb200_region_coverage : overlap-weighted mean depth of a target region from decompressed mosdepth text
//                              (mosdepth.py:264-297), for a whole batch of sample files in one launch.
//   gridb200_count_reads     : the per-read filter of count_reads_in_region (count_reads.py:82-108) over columnar
//                              read records, binned into G target regions at once (the reference makes one pass per
//                              region; BASELINE config 5 needs 734 of them).
//
// Both are HBM-bound streaming kernels with exact integer results; the floating-point question the reference's
// `int(round(100 * (region_cov / covered_bp)))` raises is settled on the host from the exact integers (see
// grid_b200/utils/mosdepth.py).
#include "common.cuh"
#include "region_line.cuh"

namespace g2 {

// ------------------------------------------------------------------ region coverage ---------------------------

// One line: classify with the shared grammar (region_line.cuh) and accumulate.
__device__ __forceinline__ void rc_line(const char* text, long long n, long long i, const RegionParams& P,
                                        unsigned long long* out) {
  const LineResult r = rc_parse_line(text, n, i, P);
  if (r.status == RC_LINE_IRREGULAR) {
    atomicAdd(&out[RC_IRREGULAR], 1ULL);
  } else if (r.status == RC_LINE_COUNTS) {
    atomicAdd(&out[RC_WEIGHTED], (unsigned long long)r.h * (unsigned long long)r.ov);
    atomicAdd(&out[RC_COVERED], (unsigned long long)r.ov);
    atomicAdd(&out[RC_LINES], 1ULL);
  }
}

// Every thread owns 16 text bytes and handles the lines that START inside them.  The batch is the concatenation of
// the files' texts, each ending in a newline, so a file's first byte always follows a newline (or is byte 0).
__global__ void __launch_bounds__(256) k_region_coverage(const char* __restrict__ text, long long n,
                                                         const long long* __restrict__ file_off, int nfiles,
                                                         const __grid_constant__ RegionParams P,
                                                         unsigned long long* __restrict__ out) {
  const long long c0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 16;
  if (c0 >= n) return;
  unsigned starts = 0;
  if (c0 + 16 <= n && (reinterpret_cast<uintptr_t>(text) & 15) == 0) {
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
    const long long p = c0 + t;
    if (p >= n) break;
    int lo = 0, hi = nfiles;  // file f with file_off[f] <= p < file_off[f + 1]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (file_off[mid] <= p) lo = mid;
      else hi = mid;
    }
    rc_line(text, file_off[lo + 1], p, P, out + 4 * (long long)lo);
  }
}

// ------------------------------------------------------------------ read counting -----------------------------

constexpr int kReadsPerThread = 8;
constexpr int kReadsPerBlock = 256 * kReadsPerThread;
constexpr int kSmemIntervals = 4096;

struct ReadCols {
  const uint16_t* flag;
  const uint8_t* mapq;
  const int32_t* tid;
  const int32_t* next_tid;
  const int32_t* pos;
};

__device__ __forceinline__ int sample_of(long long r, const long long* __restrict__ sample_off, int ns) {
  int lo = 0, hi = ns;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (sample_off[mid] <= r) lo = mid;
    else hi = mid;
  }
  return lo;
}

// 128-/64-bit streaming loads that do not allocate in L1: the read columns are touched once, and L1 is better spent
// on the bucket directory and the boundary table (ncu: with plain loads the lookups cost 2.2 GB of L2->SM traffic
// on top of the 1.5 GB of columns).
__device__ __forceinline__ uint4 ld_stream_u4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 ld_stream_u2(const void* p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}

// hist[s, e] += reads of sample s that pass the filter and fall into elementary interval e of the region boundaries.
// A block takes kChunks x 2048 consecutive reads (one 16-byte load per column and thread where the column's element
// size allows); when they all belong to one sample -- nearly always -- the histogram is built in shared memory.
constexpr int kChunks = 4;

__global__ void __launch_bounds__(256) k_read_hist(ReadCols R, long long n_reads,
                                                   const long long* __restrict__ sample_off, int ns,
                                                   const uint32_t* __restrict__ flag_bits, int min_mapq,
                                                   const long long* __restrict__ bounds, int nb, BucketDir D,
                                                   uint32_t* __restrict__ hist) {
  __shared__ uint32_t sh[kSmemIntervals];
  __shared__ int s_first, s_last;
  const long long r0 = (long long)blockIdx.x * (kReadsPerBlock * kChunks);
  const long long r1 = r0 + kReadsPerBlock * kChunks < n_reads ? r0 + kReadsPerBlock * kChunks : n_reads;
  const int ne = nb + 1;
  if (threadIdx.x == 0) s_first = sample_of(r0, sample_off, ns);
  if (threadIdx.x == 32) s_last = sample_of(r1 - 1, sample_off, ns);
  const bool fits = ne <= kSmemIntervals;
  if (fits)
    for (int e = threadIdx.x; e < ne; e += blockDim.x) sh[e] = 0;
  __syncthreads();
  const bool local = fits && s_first == s_last;
  auto take = [&](unsigned flag, int mapq, int tid, int ntid, int pos, long long q) {
    // count_reads.py:96-100, then the position test as an interval lookup
    if (!read_passes(flag, mapq, tid, ntid, flag_bits, min_mapq)) return;
    if (tid < 0 || pos < 0) return;  // unplaced: no region contains it
    const int e = D.first ? interval_by_bucket(tid, pos, bounds, D)
                          : interval_of(((long long)tid << 32) | (long long)(unsigned)pos, bounds, nb);
    if (e <= 0 || e >= nb) return;  // no region spans the first or the last interval
    if (local) atomicAdd(&sh[e], 1u);
    else atomicAdd(&hist[(long long)sample_of(q, sample_off, ns) * ne + e], 1u);
  };
#pragma unroll 1
  for (int c = 0; c < kChunks; ++c) {
    const long long r = r0 + (long long)c * kReadsPerBlock + (long long)threadIdx.x * kReadsPerThread;
    if (r >= r1) break;
    if (r + kReadsPerThread <= r1) {
      const uint4 f = ld_stream_u4(R.flag + r);
      const uint2 m = ld_stream_u2(R.mapq + r);
      const uint4 t0 = ld_stream_u4(R.tid + r), t1 = ld_stream_u4(R.tid + r + 4);
      const uint4 n0 = ld_stream_u4(R.next_tid + r), n1 = ld_stream_u4(R.next_tid + r + 4);
      const uint4 p0 = ld_stream_u4(R.pos + r), p1 = ld_stream_u4(R.pos + r + 4);
      const uint32_t fw[4] = {f.x, f.y, f.z, f.w}, mw[2] = {m.x, m.y};
      const uint32_t tt[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
      const uint32_t nn[8] = {n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, n1.z, n1.w};
      const uint32_t pp[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
#pragma unroll
      for (int t = 0; t < kReadsPerThread; ++t)
        take((fw[t >> 1] >> (16 * (t & 1))) & 0xFFFFu, (int)((mw[t >> 2] >> (8 * (t & 3))) & 0xFFu), (int)tt[t], (int)nn[t],
             (int)pp[t], r + t);
    } else {
      for (long long q = r; q < r1; ++q) take(R.flag[q], R.mapq[q], R.tid[q], R.next_tid[q], R.pos[q], q);
    }
  }
  if (local) {
    __syncthreads();
    uint32_t* row = hist + (long long)s_first * ne;
    for (int e = threadIdx.x; e < ne; e += blockDim.x)
      if (sh[e]) atomicAdd(&row[e], sh[e]);
  }
}

// One block per sample: inclusive prefix over the sample's interval histogram (in place), then
// counts[s, g] = prefix[hi_g] - prefix[lo_g]  (region g = intervals lo_g + 1 .. hi_g).
__global__ void __launch_bounds__(256) k_region_counts(uint32_t* __restrict__ hist, int ne, const int32_t* __restrict__ lo,
                                                       const int32_t* __restrict__ hi, int g_regions,
                                                       long long* __restrict__ counts) {
  __shared__ uint32_t part[256];
  uint32_t* row = hist + (long long)blockIdx.x * ne;
  const int per = (ne + 255) / 256;
  const int a = threadIdx.x * per, b = min(ne, a + per);
  uint32_t s = 0;
  for (int e = a; e < b; ++e) s += row[e];
  part[threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0;
    for (int t = 0; t < 256; ++t) {
      const uint32_t v = part[t];
      part[t] = run;
      run += v;
    }
  }
  __syncthreads();
  uint32_t run = part[threadIdx.x];
  for (int e = a; e < b; ++e) {
    run += row[e];
    row[e] = run;
  }
  __syncthreads();
  for (int g = threadIdx.x; g < g_regions; g += blockDim.x) {
    const int l = lo[g], h = hi[g];
    counts[(long long)blockIdx.x * g_regions + g] = h > l ? (long long)(row[h] - row[l]) : 0;
  }
}

}  // namespace g2

extern "C" int gridb200_region_coverage(const char* text, int64_t nbytes, const int64_t* file_off, int32_t nfiles,
                                        const char* chrom, int32_t chrom_len, int64_t start, int64_t end,
                                        int64_t* out, void* stream) {
  using namespace g2;
  G2_REQUIRE(file_off && out && nfiles > 0 && nbytes >= 0, GRIDB200_EINVAL, "bad argument");
  G2_REQUIRE(text || nbytes == 0, GRIDB200_EINVAL, "null text");
  G2_REQUIRE(chrom && chrom_len >= 0 && chrom_len <= 64, GRIDB200_EINVAL, "chromosome name must be at most 64 bytes");
  G2_REQUIRE(start >= -(1LL << 40) && end <= (1LL << 40), GRIDB200_EDOMAIN, "region coordinates out of range");
  cudaStream_t st = as_stream(stream);
  RegionParams P;
  memset(&P, 0, sizeof(P));
  for (int t = 0; t < chrom_len; ++t) P.chrom[t] = chrom[t];
  P.chrom_len = chrom_len;
  P.start = start, P.end = end;
  G2_CUDA(cudaMemsetAsync(out, 0, (size_t)nfiles * 4 * sizeof(int64_t), st));
  if (nbytes > 0) {
    const long long chunks = (nbytes + 15) / 16;
    k_region_coverage<<<(unsigned)((chunks + 255) / 256), 256, 0, st>>>(
        text, nbytes, reinterpret_cast<const long long*>(file_off), nfiles, P, reinterpret_cast<unsigned long long*>(out));
    G2_LAUNCHED();
  }
  return 0;
}

extern "C" size_t gridb200_count_reads_ws_bytes(int32_t ns, int32_t nb) {
  return (size_t)(ns > 0 ? ns : 0) * (size_t)((nb > 0 ? nb : 0) + 1) * sizeof(uint32_t) + 256;
}

extern "C" int gridb200_count_reads(const uint16_t* flag, const uint8_t* mapq, const int32_t* tid, const int32_t* next_tid,
                                    const int32_t* pos, int64_t n_reads, const int64_t* sample_off, int32_t ns,
                                    const uint32_t* flag_bits, int32_t min_mapq, const int64_t* bounds, int32_t nb,
                                    const int32_t* lo, const int32_t* hi, int32_t g_regions, const int32_t* bucket_first,
                                    int32_t bucket_shift, int32_t buckets_per_tid, int32_t n_tid, int64_t* counts, void* ws,
                                    size_t ws_bytes, void* stream) {
  using namespace g2;
  G2_REQUIRE(sample_off && flag_bits && counts && ns > 0 && n_reads >= 0 && g_regions > 0 && nb >= 0, GRIDB200_EINVAL,
             "bad argument");
  G2_REQUIRE(n_reads == 0 || (flag && mapq && tid && next_tid && pos), GRIDB200_EINVAL, "null read column");
  G2_REQUIRE((nb == 0 || bounds) && lo && hi, GRIDB200_EINVAL, "null region table");
  G2_REQUIRE(((reinterpret_cast<uintptr_t>(flag) | reinterpret_cast<uintptr_t>(mapq) | reinterpret_cast<uintptr_t>(tid) |
               reinterpret_cast<uintptr_t>(next_tid) | reinterpret_cast<uintptr_t>(pos)) & 15) == 0,
             GRIDB200_EINVAL, "read columns must be 16-byte aligned");
  G2_REQUIRE(!bucket_first || (bucket_shift >= 0 && bucket_shift < 31 && buckets_per_tid > 0 && n_tid > 0), GRIDB200_EINVAL,
             "bad bucket directory");
  G2_REQUIRE(ws && ws_bytes >= gridb200_count_reads_ws_bytes(ns, nb), GRIDB200_EINVAL,
             "workspace too small: need %zu bytes", gridb200_count_reads_ws_bytes(ns, nb));
  cudaStream_t st = as_stream(stream);
  uint32_t* hist = static_cast<uint32_t*>(ws);
  const int ne = nb + 1;
  G2_CUDA(cudaMemsetAsync(hist, 0, (size_t)ns * ne * sizeof(uint32_t), st));
  if (n_reads > 0) {
    const long long per_block = (long long)kReadsPerBlock * kChunks;
    const long long blocks = (n_reads + per_block - 1) / per_block;
    G2_REQUIRE(blocks < (1LL << 31), GRIDB200_EINVAL, "too many reads in one call");
    ReadCols R{flag, mapq, tid, next_tid, pos};
    k_read_hist<<<(unsigned)blocks, 256, 0, st>>>(R, n_reads, reinterpret_cast<const long long*>(sample_off), ns, flag_bits,
                                                 min_mapq, reinterpret_cast<const long long*>(bounds), nb,
                                                 BucketDir{bucket_first, bucket_shift, buckets_per_tid, n_tid}, hist);
    G2_LAUNCHED();
  }
  k_region_counts<<<(unsigned)ns, 256, 0, st>>>(hist, ne, lo, hi, g_regions, reinterpret_cast<long long*>(counts));
  G2_LAUNCHED();
  return 0;
}
