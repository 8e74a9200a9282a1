// Step 7, haplotype-informed inference (grid/utils/hi_inference.py:179-256) -- SURVEY 8f rank 2.
//
// The reference is an IN-PLACE Gauss-Seidel iteration: every sweep visits the samples in file order and a sample's
// update sees the values its lower-numbered neighbours received earlier in the same sweep.  That order is kept
// exactly, in parallel, by level scheduling: level(i) = 1 + max level of the neighbour samples j < i that take part,
// so the samples of one level depend only on earlier levels.  A sample reads
//     hap[nb]  (this sweep's value)     when the neighbour's sample index is below its own,
//     old[nb]  (snapshot at sweep start) otherwise -- in the sequential order that neighbour has not been visited yet,
// which also removes every write-after-read hazard between samples of different levels.  One launch per level; the
// sums run in neighbour-list order with separately rounded multiply and add, so the doubles equal the reference's.
#include "common.cuh"

namespace g2 {

struct HapSums {
  double m[2];  // weighted neighbour mean per haplotype
  double s[2];  // weight sum (starts at 1e-9)
};

template <bool kSweep>
__device__ __forceinline__ HapSums hap_sums(int64_t i, const int64_t* __restrict__ nbr_ptr,
                                            const int32_t* __restrict__ nbr_hap, const double* __restrict__ nbr_w,
                                            const double* hap, const double* __restrict__ old) {
  HapSums r;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    double wsum = 1e-9, wval = 0.0;
    for (int64_t e = nbr_ptr[2 * i + h]; e < nbr_ptr[2 * i + h + 1]; ++e) {
      const int32_t nb = nbr_hap[e];
      const double v = (!kSweep || (nb >> 1) < i) ? hap[nb] : old[nb];
      if (v == v) {
        const double w = nbr_w[e];
        wsum = __dadd_rn(wsum, w);
        wval = __dadd_rn(wval, __dmul_rn(w, v));
      }
    }
    r.m[h] = __ddiv_rn(wval, wsum);
    r.s[h] = wsum;
  }
  return r;
}

__global__ void __launch_bounds__(128) k_hi_level(const double* __restrict__ irr, const int64_t* __restrict__ nbr_ptr,
                                                  const int32_t* __restrict__ nbr_hap, const double* __restrict__ nbr_w,
                                                  const int32_t* __restrict__ order, int64_t lo, int64_t hi,
                                                  double* hap, const double* __restrict__ old) {
  const int64_t t = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= hi) return;
  const int64_t i = order[t];
  const HapSums r = hap_sums<true>(i, nbr_ptr, nbr_hap, nbr_w, hap, old);
  const double d = __dadd_rn(r.m[0], r.m[1]);
  if (d > 0.0) {
    const double x = irr[i];
    hap[2 * i] = __ddiv_rn(__dmul_rn(x, r.m[0]), d);
    hap[2 * i + 1] = __ddiv_rn(__dmul_rn(x, r.m[1]), d);
  }
}

__global__ void __launch_bounds__(128) k_hi_impute(const int64_t* __restrict__ nbr_ptr,
                                                   const int32_t* __restrict__ nbr_hap, const double* __restrict__ nbr_w,
                                                   const double* __restrict__ hap, int64_t n, double mean_irr,
                                                   double* __restrict__ imp) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const HapSums r = hap_sums<false>(i, nbr_ptr, nbr_hap, nbr_w, hap, hap);
  imp[2 * i] = r.s[0] <= 1e-9 ? mean_irr / 2 : r.m[0];
  imp[2 * i + 1] = r.s[1] <= 1e-9 ? mean_irr / 2 : r.m[1];
}

}  // namespace g2

using namespace g2;

extern "C" int gridb200_hi_phase(const double* irr, int64_t n, const int64_t* nbr_ptr, const int32_t* nbr_hap,
                                 const double* nbr_w, const int32_t* order, const int64_t* level_ptr_host,
                                 int32_t n_levels, int32_t n_iters, double* hap, double* old_ws, void* stream) {
  G2_REQUIRE(irr && nbr_ptr && hap && old_ws && n > 0 && n < (1LL << 30), GRIDB200_EINVAL, "bad argument");
  G2_REQUIRE(n_levels >= 0 && n_iters >= 0 && (n_levels == 0 || (order && level_ptr_host)), GRIDB200_EINVAL,
             "bad level schedule");
  for (int l = 0; l < n_levels; ++l)
    G2_REQUIRE(level_ptr_host[l] <= level_ptr_host[l + 1] && level_ptr_host[l] >= 0 && level_ptr_host[l + 1] <= n,
               GRIDB200_EINVAL, "level_ptr must be non-decreasing within [0, n]");
  cudaStream_t st = as_stream(stream);
  for (int it = 0; it < n_iters && n_levels > 0; ++it) {
    G2_CUDA(cudaMemcpyAsync(old_ws, hap, (size_t)n * 2 * sizeof(double), cudaMemcpyDeviceToDevice, st));
    for (int l = 0; l < n_levels; ++l) {
      const int64_t lo = level_ptr_host[l], hi = level_ptr_host[l + 1];
      if (hi == lo) continue;
      k_hi_level<<<(unsigned)((hi - lo + 127) / 128), 128, 0, st>>>(irr, nbr_ptr, nbr_hap, nbr_w, order, lo, hi, hap, old_ws);
      G2_LAUNCHED();
    }
  }
  return 0;
}

extern "C" int gridb200_hi_impute(const int64_t* nbr_ptr, const int32_t* nbr_hap, const double* nbr_w, const double* hap,
                                  int64_t n, double mean_irr, double* imp, void* stream) {
  G2_REQUIRE(nbr_ptr && hap && imp && n > 0, GRIDB200_EINVAL, "bad argument");
  k_hi_impute<<<(unsigned)((n + 127) / 128), 128, 0, as_stream(stream)>>>(nbr_ptr, nbr_hap, nbr_w, hap, n, mean_irr, imp);
  G2_LAUNCHED();
  return 0;
}
