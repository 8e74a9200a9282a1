// Order statistics and compaction over the per-bin ratio vector (length b, up to a few million):
//   - median + top_frac threshold of the finite ratios  (normalize_mosdepth.py:460-468, :479-499)
//   - sigma2_min of the %.3f-rounded ratios             (find_neighbors.py:128-175)
// Implemented as an 8-pass MSB radix select on order-preserving 64-bit keys (the last four passes on the few keys that
// still match a target after the fourth, collected on the way: k_sel_tail).  Each full pass is one multi-block
// kernel: warp-aggregated shared-memory histograms for up to 3 target ranks at once, merged with global
// atomics; the last block to finish picks the digit of every target.  No host round trip, no sort.
#include "common.cuh"

namespace g2 {

struct SelState {
  unsigned long long prefix[3];
  unsigned long long rank[3];
  unsigned long long count;
  double value[3];
  unsigned int done;
  int ntargets;
  int error;
  int pad;
  unsigned int ncand;      // keys collected for the tail (may exceed kCandCap: then `overflow` is set and nothing is used)
  unsigned int overflow;
  unsigned int tail_done;  // passes 4-7 were done on the collected keys: the full-vector passes return at once
};
constexpr int kCollectPass = 3;        // the pass that also collects the keys matching a target's 24-bit prefix
constexpr unsigned kCandCap = 1u << 16;  // keys the tail kernel will take (a few hundred are typical for 3e6 ratios)
constexpr size_t kStateBytes = 256;
constexpr size_t kHistBytes = 8 * 3 * 256 * sizeof(unsigned);
constexpr int kSelThreads = 256;
constexpr int kScanTile = 2048;

__device__ __forceinline__ unsigned long long key_of(double v) {
  const unsigned long long u = (unsigned long long)__double_as_longlong(v);
  return (u >> 63) ? ~u : (u | 0x8000000000000000ULL);
}
__device__ __forceinline__ double value_of(unsigned long long k) {
  const unsigned long long u = (k >> 63) ? (k & 0x7FFFFFFFFFFFFFFFULL) : ~k;
  return __longlong_as_double((long long)u);
}
// mode 0: candidates are the non-NaN values; mode 1: the finite values
__device__ __forceinline__ bool is_candidate(double v, int mode) { return mode == 0 ? (v == v) : isfinite(v); }

__global__ void k_sel_count(const double* __restrict__ vals, int64_t n, int mode, SelState* st) {
  int c = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    c += is_candidate(vals[i], mode);
  c = warp_sum(c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(&st->count, (unsigned long long)c);
}

// kind 0: ranks (cnt-1)/2, cnt/2 (median) and int(top_frac*cnt) (threshold).  kind 1: min(int(rw*(1-frac_r)), cnt-1)
__global__ void k_sel_ranks(SelState* st, int kind, double param, const double* info) {
  const long long cnt = (long long)st->count;
  st->error = 0;
  if (cnt == 0) {
    st->ntargets = 0;
    return;
  }
  if (kind == 0) {
    st->ntargets = 3;
    st->rank[0] = (unsigned long long)((cnt - 1) / 2);
    st->rank[1] = (unsigned long long)(cnt / 2);
    long long t = (long long)(param * (double)cnt);  // Python: int(top_frac * len(sorted_ratios))
    if (t >= cnt || t < 0) {
      st->error = 1;  // the reference's list index is out of range here
      t = cnt - 1;
    }
    st->rank[2] = (unsigned long long)t;
  } else {
    st->ntargets = 1;
    const long long rw = (long long)info[3];
    long long t = (long long)((double)rw * (1.0 - param));  // int(R * (1.0 - frac_r))
    if (t > cnt - 1) t = cnt - 1;
    if (t < 0) t = 0;
    st->rank[0] = (unsigned long long)t;
  }
  for (int t = 0; t < 3; ++t) st->prefix[t] = 0;
}

__global__ void __launch_bounds__(kSelThreads) k_sel_pass(const double* __restrict__ vals, int64_t n, int mode,
                                                          int pass, SelState* st, unsigned* __restrict__ hist_all,
                                                          unsigned long long* __restrict__ cand) {
  __shared__ unsigned sh[3][256];
  __shared__ bool s_last;
  const int nt = st->ntargets;
  if (nt == 0 || st->tail_done) return;
  const bool collect = cand != nullptr && pass == kCollectPass;
  for (int t = threadIdx.x; t < 3 * 256; t += kSelThreads) (&sh[0][0])[t] = 0;
  __syncthreads();
  const int shift = 56 - 8 * pass;
  unsigned long long pre[3];
  for (int t = 0; t < 3; ++t) pre[t] = st->prefix[t];
  // targets whose prefixes still agree share one histogram: alias[t] = the first target with t's prefix.  (The
  // ratios of one cohort share sign, exponent and often several mantissa bytes, and the two median ranks agree to
  // the last byte almost always -- counting once instead of three times is most of this kernel's time.)
  int alias[3] = {0, 1, 2};
  for (int t = 1; t < nt; ++t)
    for (int u = t - 1; u >= 0; --u)
      if (pass == 0 || (pre[u] >> (shift + 8)) == (pre[t] >> (shift + 8))) alias[t] = alias[u];
  const int64_t stride = (int64_t)gridDim.x * kSelThreads;
  const int64_t nround = ((n + stride - 1) / stride) * stride;  // keep warps converged for match_any
  for (int64_t i = (int64_t)blockIdx.x * kSelThreads + threadIdx.x; i < nround; i += stride) {
    const double v = i < n ? vals[i] : nan("");
    const bool is_cand = i < n && is_candidate(v, mode);
    const unsigned long long key = key_of(v);
    const unsigned digit = (unsigned)(key >> shift) & 0xffu;
    bool any_hit = false;
    for (int t = 0; t < nt; ++t) {
      if (alias[t] != t) continue;
      const bool hit = is_cand && (pass == 0 || (key >> (shift + 8)) == (pre[t] >> (shift + 8)));
      any_hit |= hit;
      const unsigned ballot = __ballot_sync(0xffffffffu, hit);
      if (!ballot) continue;
      if (hit) {
        const unsigned peers = __match_any_sync(ballot, digit);
        if ((threadIdx.x & 31) == __ffs(peers) - 1) {
          const unsigned cnt = __popc(peers);
          atomicAdd(&sh[t][digit], cnt);
        }
      }
    }
    if (collect) {  // the keys that still match some target's prefix: the last four passes run on this list alone
      const unsigned ballot = __ballot_sync(0xffffffffu, any_hit);
      if (ballot) {
        unsigned base = 0;
        const int lane = threadIdx.x & 31;
        if (lane == 0) base = atomicAdd(&st->ncand, (unsigned)__popc(ballot));
        base = __shfl_sync(0xffffffffu, base, 0);
        const unsigned slot = base + __popc(ballot & ((1u << lane) - 1u));
        if (any_hit) {
          if (slot < kCandCap) cand[slot] = key;
          else st->overflow = 1u;
        }
      }
    }
  }
  __syncthreads();
  unsigned* hist = hist_all + (size_t)pass * 3 * 256;
  for (int t = threadIdx.x; t < 3 * 256; t += kSelThreads) {
    const unsigned c = sh[alias[t >> 8]][t & 255];  // an aliased target receives a copy of its leader's counts
    if (c && (t >> 8) < nt) atomicAdd(&hist[t], c);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(&st->done, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // the complete histograms, fetched by the whole block (a single thread walking 255 dependent L2 loads cost
  // ~50 us per pass -- more than the counting itself)
  for (int t = threadIdx.x; t < 3 * 256; t += kSelThreads) (&sh[0][0])[t] = __ldcg(&hist[t]);
  __syncthreads();
  if (threadIdx.x < nt) {
    const int t = threadIdx.x;
    unsigned long long r = st->rank[t];
    const unsigned* h = sh[t];
    int d = 0;
    for (; d < 255; ++d) {
      const unsigned c = h[d];
      if (r < c) break;
      r -= c;
    }
    st->rank[t] = r;
    const unsigned long long p = pre[t] | ((unsigned long long)d << shift);
    st->prefix[t] = p;
    if (pass == 7) st->value[t] = value_of(p);
  }
  if (threadIdx.x == 0) st->done = 0;
}

// Passes kCollectPass+1 .. 7 on the collected keys, all in ONE block and one launch (each full-vector pass is a grid of
// 592 blocks that re-reads the 23 MB vector to count a few hundred survivors: ~25 us apiece).
__global__ void __launch_bounds__(1024) k_sel_tail(const unsigned long long* __restrict__ cand, SelState* st) {
  __shared__ unsigned sh[3][256];
  __shared__ unsigned long long s_pre[3], s_rank[3];
  const int nt = st->ntargets;
  if (nt == 0 || st->overflow) return;  // the full-vector passes take over
  const unsigned nc = st->ncand;
  if (threadIdx.x < 3) s_pre[threadIdx.x] = st->prefix[threadIdx.x], s_rank[threadIdx.x] = st->rank[threadIdx.x];
  for (int pass = kCollectPass + 1; pass < 8; ++pass) {
    for (int t = threadIdx.x; t < 3 * 256; t += 1024) (&sh[0][0])[t] = 0;
    __syncthreads();
    const int shift = 56 - 8 * pass;
    unsigned long long pre[3];
    for (int t = 0; t < 3; ++t) pre[t] = s_pre[t];
    int alias[3] = {0, 1, 2};
    for (int t = 1; t < nt; ++t)
      for (int u = t - 1; u >= 0; --u)
        if ((pre[u] >> (shift + 8)) == (pre[t] >> (shift + 8))) alias[t] = alias[u];
    for (unsigned i = threadIdx.x; i < nc; i += 1024) {
      const unsigned long long key = cand[i];
      const unsigned digit = (unsigned)(key >> shift) & 0xffu;
      for (int t = 0; t < nt; ++t)
        if (alias[t] == t && (key >> (shift + 8)) == (pre[t] >> (shift + 8))) atomicAdd(&sh[t][digit], 1u);
    }
    __syncthreads();
    if (threadIdx.x < nt) {
      const int t = threadIdx.x;
      unsigned long long r = s_rank[t];
      const unsigned* h = sh[alias[t]];
      int d = 0;
      for (; d < 255; ++d) {
        const unsigned c = h[d];
        if (r < c) break;
        r -= c;
      }
      s_rank[t] = r;
      s_pre[t] = pre[t] | ((unsigned long long)d << shift);
    }
    __syncthreads();
  }
  if (threadIdx.x < nt) {
    const int t = threadIdx.x;
    st->rank[t] = s_rank[t];
    st->prefix[t] = s_pre[t];
    st->value[t] = value_of(s_pre[t]);
  }
  if (threadIdx.x == 0) st->tail_done = 1u;
}

// zscale / threshold from the three order statistics
__global__ void k_norm_select_finish(SelState* st, double* info) {
  double zscale = 1.0, thr = nan("");
  if (st->ntargets) {
    const double med = (st->value[0] + st->value[1]) / 2.0;  // np.median: mean of the two middle values
    if (med > 0.0) zscale = 1.0 / sqrt(med / 100.0);
    thr = st->value[2];
  }
  info[0] = zscale;
  info[1] = thr;
  info[2] = (double)st->count;
  info[3] = st->error ? -1.0 : 0.0;  // completed by the scan
}

// flags -> exclusive scan -> compact positions.  flag = ratio > thr (NaN compares false).
__global__ void __launch_bounds__(1024) k_flag_count(const double* __restrict__ ratio, int64_t b,
                                                     const double* __restrict__ info, int* __restrict__ block_sums) {
  const double thr = info[1];
  const int64_t base = (int64_t)blockIdx.x * kScanTile;
  int c = 0;
  for (int e = threadIdx.x; e < kScanTile; e += 1024) {
    const int64_t j = base + e;
    if (j < b) c += ratio[j] > thr;
  }
  c = warp_sum(c);
  __shared__ int sw[32];
  if ((threadIdx.x & 31) == 0) sw[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x < 32) {
    int v = sw[threadIdx.x];
    v = warp_sum(v);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = v;
  }
}

__global__ void __launch_bounds__(1024) k_scan_block_sums(int* __restrict__ block_sums, int64_t nblocks, double* info) {
  __shared__ int sw[32];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int64_t base = 0; base < nblocks; base += 1024) {
    const int64_t i = base + threadIdx.x;
    const int v = i < nblocks ? block_sums[i] : 0;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, o);
      if ((threadIdx.x & 31) >= o) x += y;
    }
    if ((threadIdx.x & 31) == 31) sw[threadIdx.x >> 5] = x;
    __syncthreads();
    if (threadIdx.x < 32) {
      int w = sw[threadIdx.x];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, w, o);
        if (threadIdx.x >= o) w += y;
      }
      sw[threadIdx.x] = w;
    }
    __syncthreads();
    const int warp_off = (threadIdx.x >> 5) ? sw[(threadIdx.x >> 5) - 1] : 0;
    const int incl = x + warp_off + carry;
    if (i < nblocks) block_sums[i] = incl - v;  // exclusive
    __syncthreads();
    if (threadIdx.x == 1023) carry = incl;
    __syncthreads();
  }
  if (threadIdx.x == 0 && info[3] >= 0.0) info[3] = (double)carry;
}

__global__ void __launch_bounds__(1024) k_write_pos(const double* __restrict__ ratio, int64_t b,
                                                    const double* __restrict__ info,
                                                    const int* __restrict__ block_sums, int32_t* __restrict__ sel_pos) {
  const double thr = info[1];
  const int64_t base = (int64_t)blockIdx.x * kScanTile;
  // each thread owns 2 consecutive elements
  const int64_t j = base + 2 * (int64_t)threadIdx.x;
  const int f0 = (j < b) && (ratio[j] > thr);
  const int f1 = (j + 1 < b) && (ratio[j + 1] > thr);
  int x = f0 + f1;
  const int mine = x;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int y = __shfl_up_sync(0xffffffffu, x, o);
    if ((threadIdx.x & 31) >= o) x += y;
  }
  __shared__ int sw[32];
  if ((threadIdx.x & 31) == 31) sw[threadIdx.x >> 5] = x;
  __syncthreads();
  if (threadIdx.x < 32) {
    int w = sw[threadIdx.x];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, w, o);
      if (threadIdx.x >= o) w += y;
    }
    sw[threadIdx.x] = w;
  }
  __syncthreads();
  const int excl = x - mine + ((threadIdx.x >> 5) ? sw[(threadIdx.x >> 5) - 1] : 0) + block_sums[blockIdx.x];
  if (j < b) sel_pos[j] = f0 ? excl : -1;
  if (j + 1 < b) sel_pos[j + 1] = f1 ? excl + f0 : -1;
}

// ---- find_neighbors' region filter --------------------------------------------------------------------
__global__ void k_round3(const double* __restrict__ ratio, const int32_t* __restrict__ sel_pos, int64_t b,
                         double* __restrict__ rq) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= b) return;
  double v = nan("");
  if (sel_pos[j] >= 0) {
    const double r = ratio[j];
    v = isfinite(r) ? printf_round(r, 1000.0) / 1000.0 : r;  // float("%.3f" % r)
  }
  rq[j] = v;
}

__global__ void k_keep(const double* __restrict__ rq, const int32_t* __restrict__ sel_pos, int64_t b,
                       const SelState* st, double sigma2_max, uint8_t* __restrict__ keep,
                       unsigned long long* __restrict__ r_use, double* info2) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool none_finite = st->ntargets == 0;
  const double lo = none_finite ? nan("") : st->value[0];
  int k = 0;
  if (j < b) {
    if (sel_pos[j] >= 0) {
      const double v = rq[j];
      k = none_finite ? 1 : (isfinite(v) && v >= lo && v <= sigma2_max);
    }
    keep[j] = (uint8_t)k;
  }
  const int c = __syncthreads_count(k);  // one atomic per block: 90 000 warp-level adds on ONE address made this the
  if (threadIdx.x == 0 && c) atomicAdd(r_use, (unsigned long long)c);  // longest kernel of the selection (65 us)
  if (j == 0) info2[1] = lo;
}

__global__ void k_ruse_to_double(const unsigned long long* r_use, double* info2) { info2[0] = (double)*r_use; }

// kind 1 with frac_r >= 1: the wanted rank is int(Rw * (1 - frac_r)) clamped at 0 -- the MINIMUM of the finite values (the
// reference's default, find_neighbors.py:160-163).  One pass instead of the eight of the radix select.
__global__ void k_sel_min(const double* __restrict__ vals, int64_t n, SelState* st, unsigned long long* __restrict__ minkey) {
  unsigned long long m = ~0ULL;
  int c = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const double v = vals[i];
    if (isfinite(v)) {
      const unsigned long long k = key_of(v);
      m = k < m ? k : m;
      ++c;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long y = __shfl_xor_sync(0xffffffffu, m, o);
    m = y < m ? y : m;
  }
  c = warp_sum(c);
  if ((threadIdx.x & 31) == 0 && c) {
    atomicMin(minkey, m);
    atomicAdd(&st->count, (unsigned long long)c);
  }
}
__global__ void k_sel_min_finish(SelState* st, const unsigned long long* minkey) {
  st->error = 0;
  st->ntargets = st->count ? 1 : 0;
  if (st->count) st->value[0] = value_of(*minkey);
}

static int run_select_min(const double* vals, int64_t n, SelState* st, unsigned* hist, cudaStream_t stream) {
  G2_CUDA(cudaMemsetAsync(st, 0, kStateBytes, stream));
  G2_CUDA(cudaMemsetAsync(hist, 0xFF, sizeof(unsigned long long), stream));  // the running minimum key lives in hist[0..1]
  const int blocks = (int)std::min<int64_t>((n + kSelThreads * 8 - 1) / (kSelThreads * 8), (int64_t)kSMs * 4);
  k_sel_min<<<blocks, kSelThreads, 0, stream>>>(vals, n, st, reinterpret_cast<unsigned long long*>(hist));
  G2_LAUNCHED();
  k_sel_min_finish<<<1, 1, 0, stream>>>(st, reinterpret_cast<const unsigned long long*>(hist));
  G2_LAUNCHED();
  return 0;
}

// cand: room for kCandCap keys, or nullptr (all eight passes over the whole vector, as the allocation-free variant does)
static int run_select(const double* vals, int64_t n, int mode, int kind, double param, const double* info,
                      SelState* st, unsigned* hist, cudaStream_t stream, unsigned long long* cand = nullptr) {
  G2_CUDA(cudaMemsetAsync(st, 0, kStateBytes + kHistBytes, stream));
  const int blocks = (int)std::min<int64_t>((n + kSelThreads * 8 - 1) / (kSelThreads * 8), (int64_t)kSMs * 4);
  k_sel_count<<<blocks, kSelThreads, 0, stream>>>(vals, n, mode, st);
  G2_LAUNCHED();
  k_sel_ranks<<<1, 1, 0, stream>>>(st, kind, param, info);
  G2_LAUNCHED();
  for (int pass = 0; pass < 8; ++pass) {
    k_sel_pass<<<blocks, kSelThreads, 0, stream>>>(vals, n, mode, pass, st, hist, cand);
    G2_LAUNCHED();
    if (cand && pass == kCollectPass) {
      k_sel_tail<<<1, 1024, 0, stream>>>(cand, st);
      G2_LAUNCHED();
    }
  }
  return 0;
}

}  // namespace g2

using namespace g2;

// where the tail's key list lives inside a gridb200_select_ws_bytes(b) workspace
static unsigned long long* select_cand(void* ws, int64_t b) {
  const size_t nblk = (size_t)((b + kScanTile - 1) / kScanTile) + 1;
  size_t bytes = kStateBytes + kHistBytes + ((nblk * sizeof(int) + 255) / 256) * 256 + 256 + (size_t)b * sizeof(double);
  bytes = (bytes + 255) / 256 * 256;
  return reinterpret_cast<unsigned long long*>(static_cast<char*>(ws) + bytes);
}

extern "C" size_t gridb200_select_ws_bytes(int64_t b) {
  if (b <= 0) return 0;
  const size_t nblk = (size_t)((b + kScanTile - 1) / kScanTile) + 1;
  size_t bytes = kStateBytes + kHistBytes + ((nblk * sizeof(int) + 255) / 256) * 256 + 256;
  bytes += (size_t)b * sizeof(double);  // rounded ratios for the kNN filter
  bytes = (bytes + 255) / 256 * 256 + (size_t)kCandCap * sizeof(unsigned long long);  // keys of the select's tail
  return bytes;
}

extern "C" int gridb200_select(const double* ratio, int64_t b, double top_frac, int32_t* sel_pos, double* info,
                               void* ws, size_t ws_bytes, void* stream) {
  G2_REQUIRE(ratio && sel_pos && info && ws && b > 0, GRIDB200_EINVAL, "bad argument");
  G2_REQUIRE(ws_bytes >= gridb200_select_ws_bytes(b), GRIDB200_EINVAL, "workspace too small: need %zu bytes",
             gridb200_select_ws_bytes(b));
  G2_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 255) == 0, GRIDB200_EINVAL, "workspace must be 256-byte aligned");
  G2_REQUIRE(b < (1LL << 31), GRIDB200_EINVAL, "too many bins");
  cudaStream_t st = as_stream(stream);
  char* w = static_cast<char*>(ws);
  SelState* state = reinterpret_cast<SelState*>(w);
  unsigned* hist = reinterpret_cast<unsigned*>(w + kStateBytes);
  int* block_sums = reinterpret_cast<int*>(w + kStateBytes + kHistBytes);
  if (int e = run_select(ratio, b, /*mode=*/0, /*kind=*/0, top_frac, nullptr, state, hist, st, select_cand(ws, b))) return e;
  k_norm_select_finish<<<1, 1, 0, st>>>(state, info);
  G2_LAUNCHED();
  const int64_t nblk = (b + kScanTile - 1) / kScanTile;
  k_flag_count<<<(unsigned)nblk, 1024, 0, st>>>(ratio, b, info, block_sums);
  G2_LAUNCHED();
  k_scan_block_sums<<<1, 1024, 0, st>>>(block_sums, nblk, info);
  G2_LAUNCHED();
  k_write_pos<<<(unsigned)nblk, 1024, 0, st>>>(ratio, b, info, block_sums, sel_pos);
  G2_LAUNCHED();
  return 0;
}

extern "C" int gridb200_knn_col_filter(const double* ratio, const int32_t* sel_pos, int64_t b, double frac_r,
                                       double sigma2_max, uint8_t* keep, const double* info, double* info2,
                                       void* ws, size_t ws_bytes, void* stream) {
  G2_REQUIRE(ratio && sel_pos && keep && info && info2 && ws && b > 0, GRIDB200_EINVAL, "bad argument");
  G2_REQUIRE(ws_bytes >= gridb200_select_ws_bytes(b), GRIDB200_EINVAL, "workspace too small: need %zu bytes",
             gridb200_select_ws_bytes(b));
  G2_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 255) == 0, GRIDB200_EINVAL, "workspace must be 256-byte aligned");
  cudaStream_t st = as_stream(stream);
  char* w = static_cast<char*>(ws);
  SelState* state = reinterpret_cast<SelState*>(w);
  unsigned* hist = reinterpret_cast<unsigned*>(w + kStateBytes);
  const size_t nblk = (size_t)((b + kScanTile - 1) / kScanTile) + 1;
  char* after = w + kStateBytes + kHistBytes + ((nblk * sizeof(int) + 255) / 256) * 256;
  unsigned long long* r_use = reinterpret_cast<unsigned long long*>(after);
  double* rq = reinterpret_cast<double*>(after + 256);
  k_round3<<<(unsigned)((b + 255) / 256), 256, 0, st>>>(ratio, sel_pos, b, rq);
  G2_LAUNCHED();
  if (frac_r >= 1.0) {
    if (int e = run_select_min(rq, b, state, hist, st)) return e;
  } else if (int e = run_select(rq, b, /*mode=*/1, /*kind=*/1, frac_r, info, state, hist, st, select_cand(ws, b))) {
    return e;
  }
  G2_CUDA(cudaMemsetAsync(r_use, 0, sizeof(unsigned long long), st));
  k_keep<<<(unsigned)((b + 255) / 256), 256, 0, st>>>(rq, sel_pos, b, state, sigma2_max, keep, r_use, info2);
  G2_LAUNCHED();
  k_ruse_to_double<<<1, 1, 0, st>>>(r_use, info2);
  G2_LAUNCHED();
  return 0;
}

// ---- allocation-free variant (no counterpart in the reference) ------------
// Callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc) get
// gridb200_select with its scratch in a file-scope device array instead of a caller-provided workspace.  The array is
// sized for GRIDB200_SELECT_STATIC_MAX_BINS bins (selection state + radix histograms + one int per 2048-bin scan tile:
// about 33 KB); every GPU that loads the library holds it until exit.  In host code `g_select_scratch` is only a stand-in
// for the symbol: the device address comes from cudaGetSymbolAddress, once per device, with that device current.
// Every call shares the array (it keeps the last call's contents; what must start at zero is cleared at the start of a
// call, exactly as in gridb200_select), so calls that overlap on one GPU -- two streams -- race.
namespace g2 {
constexpr int64_t kSelectStaticMaxBins = GRIDB200_SELECT_STATIC_MAX_BINS;
constexpr size_t kSelectStaticBytes =
    kStateBytes + kHistBytes + (((size_t)(kSelectStaticMaxBins / kScanTile + 2) * sizeof(int) + 255) / 256) * 256;
static __device__ __align__(256) unsigned char g_select_scratch[kSelectStaticBytes];
}  // namespace g2

extern "C" int gridb200_select_static(const double* ratio, int64_t b, double top_frac, int32_t* sel_pos, double* info,
                                      void* stream) {
  G2_REQUIRE(ratio && sel_pos && info && b > 0, GRIDB200_EINVAL, "bad argument");
  G2_REQUIRE(b <= kSelectStaticMaxBins, GRIDB200_EINVAL,
             "gridb200_select_static: %lld bins exceed the built-in scratch (GRIDB200_SELECT_STATIC_MAX_BINS = %lld); use "
             "gridb200_select with a caller-provided workspace", (long long)b, (long long)kSelectStaticMaxBins);
  static std::atomic<unsigned char*> addr_of[64];  // device address of the scratch array, per device
  int dev = 0;
  G2_CUDA(cudaGetDevice(&dev));
  G2_REQUIRE(dev >= 0 && dev < 64, GRIDB200_ENODEV, "device index %d outside the per-device cache", dev);
  unsigned char* w = addr_of[dev].load();
  if (!w) {
    void* p = nullptr;
    G2_CUDA(cudaGetSymbolAddress(&p, g_select_scratch));
    G2_REQUIRE(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255) == 0, GRIDB200_ECUDA, "scratch symbol address unusable");
    w = static_cast<unsigned char*>(p);
    addr_of[dev].store(w);
  }
  cudaStream_t st = as_stream(stream);
  SelState* state = reinterpret_cast<SelState*>(w);
  unsigned* hist = reinterpret_cast<unsigned*>(w + kStateBytes);
  int* block_sums = reinterpret_cast<int*>(w + kStateBytes + kHistBytes);
  if (int e = run_select(ratio, b, /*mode=*/0, /*kind=*/0, top_frac, nullptr, state, hist, st)) return e;  // (clears state + hist)
  k_norm_select_finish<<<1, 1, 0, st>>>(state, info);
  G2_LAUNCHED();
  const int64_t nblk = (b + kScanTile - 1) / kScanTile;
  k_flag_count<<<(unsigned)nblk, 1024, 0, st>>>(ratio, b, info, block_sums);
  G2_LAUNCHED();
  k_scan_block_sums<<<1, 1024, 0, st>>>(block_sums, nblk, info);
  G2_LAUNCHED();
  k_write_pos<<<(unsigned)nblk, 1024, 0, st>>>(ratio, b, info, block_sums, sel_pos);
  G2_LAUNCHED();
  return 0;
}
