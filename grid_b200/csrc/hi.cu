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
#include <cooperative_groups.h>
#include <stdlib.h>

#include "common.cuh"

namespace g2 {

struct HapSums {
  double m[2];  // weighted neighbour mean per haplotype
  double s[2];  // weight sum (starts at 1e-9)
};

// kCg: values are read with ld.global.cg (L2) -- inside the persistent sweep kernel they were written by other SMs
// earlier in the same launch, and L1 is not coherent between SMs.
template <bool kSweep, bool kCg = false>
__device__ __forceinline__ HapSums hap_sums(int64_t i, const int64_t* __restrict__ nbr_ptr,
                                            const int32_t* __restrict__ nbr_hap, const double* __restrict__ nbr_w,
                                            const double* hap, const double* old) {
  HapSums r;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    double wsum = 1e-9, wval = 0.0;
    for (int64_t e = nbr_ptr[2 * i + h]; e < nbr_ptr[2 * i + h + 1]; ++e) {
      const int32_t nb = nbr_hap[e];
      const double* src = (!kSweep || (nb >> 1) < i) ? hap + nb : old + nb;
      const double v = kCg ? __ldcg(src) : *src;
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

// All sweeps and levels in ONE cooperative launch: a grid barrier replaces the kernel boundary between levels
// (2 800 launches at N = 5e4, 100 sweeps -> 2 900 barriers of a few microseconds).
__global__ void __launch_bounds__(256) k_hi_sweeps(const double* __restrict__ irr, const int64_t* __restrict__ nbr_ptr,
                                                   const int32_t* __restrict__ nbr_hap, const double* __restrict__ nbr_w,
                                                   const int32_t* __restrict__ order, const int64_t* __restrict__ level_ptr,
                                                   int n_levels, int n_iters, int64_t n, double* hap, double* old) {
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, gsize = (int64_t)gridDim.x * blockDim.x;
  for (int it = 0; it < n_iters; ++it) {
    for (int64_t x = gtid; x < 2 * n; x += gsize) old[x] = __ldcg(hap + x);
    grid.sync();
    for (int l = 0; l < n_levels; ++l) {
      const int64_t lo = level_ptr[l], hi = level_ptr[l + 1];
      for (int64_t t = lo + gtid; t < hi; t += gsize) {
        const int64_t i = order[t];
        const HapSums r = hap_sums<true, true>(i, nbr_ptr, nbr_hap, nbr_w, hap, old);
        const double d = __dadd_rn(r.m[0], r.m[1]);
        if (d > 0.0) {
          const double x = irr[i];
          hap[2 * i] = __ddiv_rn(__dmul_rn(x, r.m[0]), d);
          hap[2 * i + 1] = __ddiv_rn(__dmul_rn(x, r.m[1]), d);
        }
      }
      grid.sync();
    }
  }
}

// One relaxation pass of level(i) = 1 + max level(j) over neighbour samples j < i that take part (0 if none), in place.
// Values only ever rise towards the fixed point, so passes may overlap freely; *changed is set when anything moved.
__global__ void __launch_bounds__(256) k_hi_level_pass(const int64_t* __restrict__ nbr_ptr,
                                                       const int32_t* __restrict__ nbr_hap, int64_t n, int32_t* level,
                                                       int32_t* __restrict__ changed) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t mine = level[i];
  if (mine < 0) return;
  int32_t m = -1;
  for (int64_t e = nbr_ptr[2 * i]; e < nbr_ptr[2 * i + 2]; ++e) {
    const int64_t j = nbr_hap[e] >> 1;
    if (j < i) m = max(m, __ldcg(level + j));
  }
  if (m + 1 != mine) {
    level[i] = m + 1;
    *changed = 1;
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
  // One cooperative launch when the device supports it and the caller's scratch also holds the level table
  // (old_ws: 2n doubles, then n_levels + 1 int64); GRIDB200_HI_SCHED=levels keeps one launch per level (A/B aid).
  const char* sched = getenv("GRIDB200_HI_SCHED");
  const bool per_level = sched && sched[0] == 'l';
  int dev = 0, coop = 0, sms = 0, per_sm = 0;
  G2_CUDA(cudaGetDevice(&dev));
  G2_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
  G2_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  G2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_hi_sweeps, 256, 0));
  if (coop && per_sm > 0 && !per_level && n_levels > 0 && n_iters > 0) {
    int64_t* level_ptr_dev = reinterpret_cast<int64_t*>(old_ws + 2 * n);
    G2_CUDA(cudaMemcpyAsync(level_ptr_dev, level_ptr_host, (size_t)(n_levels + 1) * sizeof(int64_t),
                            cudaMemcpyHostToDevice, st));
    const int64_t want = (2 * n + 255) / 256;  // the snapshot copy is the widest phase
    const unsigned grid = (unsigned)(want < (int64_t)sms * per_sm ? want : (int64_t)sms * per_sm);
    void* args[] = {&irr, &nbr_ptr, &nbr_hap, &nbr_w, &order, &level_ptr_dev, &n_levels, &n_iters, &n, &hap, &old_ws};
    const cudaError_t e = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(k_hi_sweeps), dim3(grid), dim3(256), args, 0, st);
    if (e == cudaSuccess) {
      g_launches.fetch_add(1, std::memory_order_relaxed);
      return 0;
    }
    // the grid cannot be made co-resident right now (e.g. the device is shared): nothing has run; one launch per level
    (void)cudaGetLastError();
  }
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

extern "C" int gridb200_hi_level_pass(const int64_t* nbr_ptr, const int32_t* nbr_hap, int64_t n, int32_t* level,
                                      int32_t* changed, void* stream) {
  G2_REQUIRE(nbr_ptr && level && changed && n > 0 && n < (1LL << 30), GRIDB200_EINVAL, "bad argument");
  k_hi_level_pass<<<(unsigned)((n + 255) / 256), 256, 0, as_stream(stream)>>>(nbr_ptr, nbr_hap, n, level, changed);
  G2_LAUNCHED();
  return 0;
}
