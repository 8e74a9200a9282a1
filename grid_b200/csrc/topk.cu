// k-best merge: d2 = ||q||^2 + ||r||^2 - 2 q.r in exact int64, self excluded, merged into running sorted lists
// ordered by (d2, global index).  One block per list; candidates are first filtered against the list's current k-th
// entry (only a candidate that beats it can enter), the survivors are appended behind the list in shared memory and a
// bitonic sort of the next power of two >= k + survivors puts the k best in front.  After the first block of a cohort
// the survivors are a small fraction of the candidates, so the sort shrinks from 8192 keys to about 2k.
//
// Three candidate sources share the kernel:
//   kRows  : list = query row i,      candidates = the row's entries of a Gram block            (gridb200_topk_merge)
//   kCols  : list = reference row j,  candidates = the COLUMN j of the block (its transpose)    (gridb200_topk_merge_t)
//   kLists : list = row i,            candidates = another k-best list for the same row          (gridb200_topk_merge_lists)
//   kCand  : list = row i,            candidates = the first min(cnt[i], cap) entries of the row's candidate buffer
//                                     filled by the fused Gram epilogue                            (gridb200_topk_merge_cand)
//   kRowsF32: list = query row i,     candidates = the row's entries of an fp32 Gram block (3xTF32 general path); the
//                                     squared distance is formed in fp64 and ordered through its bit pattern
//                                                                                                (gridb200_topk_merge_f32)
// kCols and kLists serve the half ring (grid_b200/ring.py): a rank that has multiplied its rows with a peer's shard
// extracts the PEER's k best from the transpose and sends those lists instead of the peer recomputing the block.
#include "common.cuh"

namespace g2 {

constexpr long long kInfD2 = 0x7fffffffffffffffLL;
constexpr int kTopkMaxKeys = 8192;
enum { kRows = 0, kCols = 1, kLists = 2, kCand = 3, kRowsF32 = 4 };

__device__ __forceinline__ bool pair_less(long long da, int ia, long long db, int ib) {
  return da < db || (da == db && ia < ib);
}

__global__ void k_topk_init(long long* __restrict__ d2, int* __restrict__ idx, int64_t total) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < total) {
    d2[i] = kInfD2;
    idx[i] = -1;
  }
}

struct TopkArgs {
  const long long* dot;      // kRows / kCols: Gram block, element (q, r) at dot[q * ldd + r]
  int64_t ldd;
  int64_t ncand;             // candidates per list
  const long long* norm_l;   // exact norms of the list owners (rows of kRows, columns of kCols)
  const long long* norm_c;   // ... of the candidates
  int64_t l_gid0, c_gid0;    // global sample ids of list 0 / candidate 0
  const long long* src_d2;   // kLists: candidate lists [nlists, src_k]
  const int* src_idx;
  int src_k;
  const int* src_cnt;        // kCand: entries appended to each row's buffer (may exceed src_k = its capacity)
  const float* dotf;         // kRowsF32: fp32 Gram block and fp64 norms
  const double* normd_l;
  const double* normd_c;
  int k;
  long long* best_d2;
  int* best_idx;
};

template <int kMode>
__global__ void __launch_bounds__(256) k_topk_merge(const TopkArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  long long* sd = reinterpret_cast<long long*>(smem_raw);
  int* si = reinterpret_cast<int*>(sd + kTopkMaxKeys);
  __shared__ int s_cnt;
  const int64_t row = blockIdx.x;
  const int k = a.k;
  if (threadIdx.x == 0) s_cnt = 0;
  for (int e = threadIdx.x; e < k; e += 256) {
    long long d = a.best_d2[row * k + e];
    int id = a.best_idx[row * k + e];
    if (id < 0) id = 0x7fffffff, d = kInfD2;
    sd[e] = d;
    si[e] = id;
  }
  __syncthreads();
  // the list is sorted: its last entry is the bar a candidate has to beat (an unfilled list ends in +inf)
  const long long tau_d = sd[k - 1];
  const int tau_i = si[k - 1];
  const int64_t self = a.l_gid0 + row;
  const long long nl = (kMode == kLists || kMode == kCand || kMode == kRowsF32) ? 0 : a.norm_l[row];
  const double nld = kMode == kRowsF32 ? a.normd_l[row] : 0.0;
  const int64_t ncand = kMode == kCand ? (int64_t)min(a.src_cnt[row], a.src_k) : a.ncand;
  for (int64_t j = threadIdx.x; j < ncand; j += 256) {
    long long d;
    int id;
    if (kMode == kLists || kMode == kCand) {
      d = a.src_d2[row * a.src_k + j];
      id = a.src_idx[row * a.src_k + j];
      if (id < 0) continue;
    } else if (kMode == kRowsF32) {
      const int64_t gid = a.c_gid0 + j;
      if (gid == self) continue;
      // a non-negative double orders like its bit pattern read as int64: the integer k-best machinery serves unchanged
      const double dd = fmax(0.0, nld + a.normd_c[j] - 2.0 * (double)a.dotf[row * a.ldd + j]);
      d = __double_as_longlong(dd);
      id = (int)gid;
    } else {
      const int64_t gid = a.c_gid0 + j;
      if (gid == self) continue;
      const long long dot = kMode == kRows ? a.dot[row * a.ldd + j] : a.dot[j * a.ldd + row];
      d = nl + a.norm_c[j] - 2 * dot;
      id = (int)gid;
    }
    if (pair_less(d, id, tau_d, tau_i)) {
      const int pos = k + atomicAdd(&s_cnt, 1);
      sd[pos] = d;
      si[pos] = id;
    }
  }
  __syncthreads();
  const int total = k + s_cnt;
  if (total == k) return;  // nothing beat the list
  int p2 = 2;
  while (p2 < total) p2 <<= 1;
  for (int e = total + threadIdx.x; e < p2; e += 256) sd[e] = kInfD2, si[e] = 0x7fffffff;
  __syncthreads();
  for (int size = 2; size <= p2; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = threadIdx.x; t < (p2 >> 1); t += 256) {
        const int lo = ((t / stride) * (stride << 1)) + (t % stride);
        const int hi = lo + stride;
        const bool up = ((lo & size) == 0);
        const long long da = sd[lo], db = sd[hi];
        const int ia = si[lo], ib = si[hi];
        if (pair_less(db, ib, da, ia) == up) {
          sd[lo] = db, sd[hi] = da;
          si[lo] = ib, si[hi] = ia;
        }
      }
      __syncthreads();
    }
  }
  for (int e = threadIdx.x; e < k; e += 256) {
    const long long d = sd[e];
    a.best_d2[row * k + e] = d;
    a.best_idx[row * k + e] = d == kInfD2 ? -1 : si[e];
  }
}

template <int kMode>
static int launch_topk(const TopkArgs& a, int64_t nlists, cudaStream_t st) {
  constexpr size_t smem = (size_t)kTopkMaxKeys * 12;
  static std::atomic<uint64_t> ready{0};  // one bit per device: the >48 KB shared-memory opt-in is per device
  int dev = 0;
  G2_CUDA(cudaGetDevice(&dev));
  if (!(ready.load() >> (dev & 63) & 1)) {
    G2_CUDA(cudaFuncSetAttribute(k_topk_merge<kMode>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ready.fetch_or(1ull << (dev & 63));
  }
  k_topk_merge<kMode><<<(unsigned)nlists, 256, smem, st>>>(a);
  G2_LAUNCHED();
  return 0;
}

}  // namespace g2

using namespace g2;

extern "C" int gridb200_topk_init(int64_t* best_d2, int32_t* best_idx, int64_t nq, int32_t k, void* stream) {
  G2_REQUIRE(best_d2 && best_idx && nq > 0 && k > 0, GRIDB200_EINVAL, "bad argument");
  const int64_t total = nq * k;
  k_topk_init<<<(unsigned)((total + 255) / 256), 256, 0, as_stream(stream)>>>(
      reinterpret_cast<long long*>(best_d2), best_idx, total);
  G2_LAUNCHED();
  return 0;
}

static int topk_dot(int mode, const int64_t* dot, int64_t ldd, int64_t nlists, int64_t ncand, const int64_t* norm_l,
                    const int64_t* norm_c, int64_t l_gid0, int64_t c_gid0, int32_t k, int64_t* best_d2, int32_t* best_idx,
                    void* stream) {
  G2_REQUIRE(dot && norm_l && norm_c && best_d2 && best_idx, GRIDB200_EINVAL, "null pointer");
  G2_REQUIRE(nlists > 0 && ncand > 0 && k > 0 && ldd >= (mode == kRows ? ncand : nlists), GRIDB200_EINVAL, "bad shape");
  G2_REQUIRE(ncand + k <= kTopkMaxKeys, GRIDB200_EINVAL, "candidates + k = %lld exceeds %d per call",
             (long long)(ncand + k), kTopkMaxKeys);
  G2_REQUIRE(c_gid0 + ncand < (1LL << 31) && l_gid0 + nlists < (1LL << 31), GRIDB200_EINVAL, "indices exceed int32");
  TopkArgs a{};
  a.dot = reinterpret_cast<const long long*>(dot), a.ldd = ldd, a.ncand = ncand;
  a.norm_l = reinterpret_cast<const long long*>(norm_l), a.norm_c = reinterpret_cast<const long long*>(norm_c);
  a.l_gid0 = l_gid0, a.c_gid0 = c_gid0, a.k = k;
  a.best_d2 = reinterpret_cast<long long*>(best_d2), a.best_idx = best_idx;
  return mode == kRows ? launch_topk<kRows>(a, nlists, as_stream(stream)) : launch_topk<kCols>(a, nlists, as_stream(stream));
}

extern "C" int gridb200_topk_merge(const int64_t* dot, int64_t ldd, int64_t nq, int64_t nr, const int64_t* normq,
                                   const int64_t* normr, int64_t q_gid0, int64_t r_gid0, int32_t k,
                                   int64_t* best_d2, int32_t* best_idx, void* stream) {
  return topk_dot(kRows, dot, ldd, nq, nr, normq, normr, q_gid0, r_gid0, k, best_d2, best_idx, stream);
}

extern "C" int gridb200_topk_merge_t(const int64_t* dot, int64_t ldd, int64_t nq, int64_t nr, const int64_t* normq,
                                     const int64_t* normr, int64_t q_gid0, int64_t r_gid0, int32_t k,
                                     int64_t* best_d2, int32_t* best_idx, void* stream) {
  // lists belong to the nr reference rows (columns of the block); the nq query rows are the candidates
  return topk_dot(kCols, dot, ldd, nr, nq, normr, normq, r_gid0, q_gid0, k, best_d2, best_idx, stream);
}

extern "C" int gridb200_topk_merge_lists(const int64_t* src_d2, const int32_t* src_idx, int64_t n, int32_t src_k,
                                         int32_t k, int64_t* best_d2, int32_t* best_idx, void* stream) {
  G2_REQUIRE(src_d2 && src_idx && best_d2 && best_idx, GRIDB200_EINVAL, "null pointer");
  G2_REQUIRE(n > 0 && k > 0 && src_k > 0 && src_k + k <= kTopkMaxKeys, GRIDB200_EINVAL, "bad shape");
  TopkArgs a{};
  a.src_d2 = reinterpret_cast<const long long*>(src_d2), a.src_idx = src_idx, a.src_k = src_k, a.ncand = src_k;
  a.k = k, a.best_d2 = reinterpret_cast<long long*>(best_d2), a.best_idx = best_idx;
  return launch_topk<kLists>(a, n, as_stream(stream));
}

extern "C" int gridb200_topk_merge_cand(const int64_t* cand_d2, const int32_t* cand_idx, const int32_t* cand_cnt, int64_t n,
                                        int32_t cap, int32_t k, int64_t* best_d2, int32_t* best_idx, void* stream) {
  G2_REQUIRE(cand_d2 && cand_idx && cand_cnt && best_d2 && best_idx, GRIDB200_EINVAL, "null pointer");
  G2_REQUIRE(n > 0 && k > 0 && cap > 0 && cap + k <= kTopkMaxKeys, GRIDB200_EINVAL, "bad shape (cap + k must be <= %d)",
             kTopkMaxKeys);
  TopkArgs a{};
  a.src_d2 = reinterpret_cast<const long long*>(cand_d2), a.src_idx = cand_idx, a.src_k = cap, a.src_cnt = cand_cnt;
  a.k = k, a.best_d2 = reinterpret_cast<long long*>(best_d2), a.best_idx = best_idx;
  return launch_topk<kCand>(a, n, as_stream(stream));
}

extern "C" int gridb200_topk_merge_f32(const float* dot, int64_t ldd, int64_t nq, int64_t nr, const double* normq,
                                       const double* normr, int64_t q_gid0, int64_t r_gid0, int32_t k, int64_t* best_d2,
                                       int32_t* best_idx, void* stream) {
  G2_REQUIRE(dot && normq && normr && best_d2 && best_idx, GRIDB200_EINVAL, "null pointer");
  G2_REQUIRE(nq > 0 && nr > 0 && k > 0 && ldd >= nr, GRIDB200_EINVAL, "bad shape");
  G2_REQUIRE(nr + k <= kTopkMaxKeys, GRIDB200_EINVAL, "candidates + k = %lld exceeds %d per call", (long long)(nr + k),
             kTopkMaxKeys);
  G2_REQUIRE(r_gid0 + nr < (1LL << 31) && q_gid0 + nq < (1LL << 31), GRIDB200_EINVAL, "indices exceed int32");
  TopkArgs a{};
  a.dotf = dot, a.ldd = ldd, a.ncand = nr, a.normd_l = normq, a.normd_c = normr, a.l_gid0 = q_gid0, a.c_gid0 = r_gid0;
  a.k = k, a.best_d2 = reinterpret_cast<long long*>(best_d2), a.best_idx = best_idx;
  return launch_topk<kRowsF32>(a, nq, as_stream(stream));
}
