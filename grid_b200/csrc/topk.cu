// k-best merge: d2 = ||q||^2 + ||r||^2 - 2 q.r in exact int64, self excluded, merged into running sorted lists
// ordered by (d2, global index).  One block per query row; bitonic sort of (k + nr) <= 8192 keys in shared memory.
#include "common.cuh"

namespace g2 {

constexpr long long kInfD2 = 0x7fffffffffffffffLL;

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

__global__ void __launch_bounds__(256) k_topk_merge(const long long* __restrict__ dot, int64_t ldd, int64_t nr,
                                                    const long long* __restrict__ normq,
                                                    const long long* __restrict__ normr, int64_t q_gid0,
                                                    int64_t r_gid0, int k, int p2, long long* __restrict__ best_d2,
                                                    int* __restrict__ best_idx) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  long long* sd = reinterpret_cast<long long*>(smem_raw);
  int* si = reinterpret_cast<int*>(sd + p2);
  const int64_t row = blockIdx.x;
  const long long nq = normq[row];
  const int64_t self = q_gid0 + row;
  for (int e = threadIdx.x; e < p2; e += 256) {
    long long d = kInfD2;
    int id = 0x7fffffff;
    if (e < k) {
      d = best_d2[row * k + e];
      id = best_idx[row * k + e];
      if (id < 0) id = 0x7fffffff, d = kInfD2;
    } else if (e - k < nr) {
      const int64_t j = e - k;
      const int64_t gid = r_gid0 + j;
      if (gid != self) {
        d = nq + normr[j] - 2 * dot[row * ldd + j];
        id = (int)gid;
      }
    }
    sd[e] = d;
    si[e] = id;
  }
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
    best_d2[row * k + e] = d;
    best_idx[row * k + e] = d == kInfD2 ? -1 : si[e];
  }
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

extern "C" int gridb200_topk_merge(const int64_t* dot, int64_t ldd, int64_t nq, int64_t nr, const int64_t* normq,
                                   const int64_t* normr, int64_t q_gid0, int64_t r_gid0, int32_t k,
                                   int64_t* best_d2, int32_t* best_idx, void* stream) {
  G2_REQUIRE(dot && normq && normr && best_d2 && best_idx, GRIDB200_EINVAL, "null pointer");
  G2_REQUIRE(nq > 0 && nr > 0 && k > 0 && ldd >= nr, GRIDB200_EINVAL, "bad shape");
  G2_REQUIRE(nr + k <= 8192, GRIDB200_EINVAL, "nr + k = %lld exceeds 8192 per call", (long long)(nr + k));
  G2_REQUIRE(r_gid0 + nr < (1LL << 31) && nq < (1LL << 31), GRIDB200_EINVAL, "indices exceed int32");
  int p2 = 2;
  while (p2 < nr + k) p2 <<= 1;
  const size_t smem = (size_t)p2 * 12;
  static bool attr_set = false;
  if (!attr_set) {
    G2_CUDA(cudaFuncSetAttribute(k_topk_merge, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192 * 12));
    attr_set = true;
  }
  k_topk_merge<<<(unsigned)nq, 256, smem, as_stream(stream)>>>(
      reinterpret_cast<const long long*>(dot), ldd, nr, reinterpret_cast<const long long*>(normq),
      reinterpret_cast<const long long*>(normr), q_gid0, r_gid0, k, p2, reinterpret_cast<long long*>(best_d2),
      best_idx);
  G2_LAUNCHED();
  return 0;
}
