// Step 6: neighbour-normalised diploid copy number (compute_dipcn.py:62-88), batched over G regions.
#include "common.cuh"

namespace g2 {

// t[i,g] = counts[i,g] / scale_q[i]
__global__ void k_dipcn_scaled(const int* __restrict__ counts, int64_t n, int64_t g, int64_t ldc,
                               const double* __restrict__ scale_q, double* __restrict__ t) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n * g) return;
  const int64_t i = e / g, c = e % g;
  t[e] = (double)counts[i * ldc + c] / scale_q[i];
}

// One block per query sample.  Warp 0 walks the neighbour list in rank order and keeps the first n_nbr
// neighbours that have a count (the others do not use up the budget); then threads fan out over regions and
// add the contributors IN LIST ORDER, as the reference's loop does.
__global__ void __launch_bounds__(128) k_dipcn(const double* __restrict__ t, int64_t g,
                                               const uint8_t* __restrict__ has_count,
                                               const int* __restrict__ nbr_idx, int k, int64_t q_gid0, int n_nbr,
                                               double* __restrict__ out, int64_t ldo) {
  extern __shared__ int s_list[];
  __shared__ int s_cnt;
  const int64_t row = blockIdx.x;
  const int64_t self = q_gid0 + row;
  if (threadIdx.x < 32) {
    int cnt = 0;
    for (int base = 0; base < k && cnt < n_nbr; base += 32) {
      const int e = base + threadIdx.x;
      const int id = e < k ? nbr_idx[row * k + e] : -1;
      const bool ok = id >= 0 && has_count[id];
      const unsigned bal = __ballot_sync(0xffffffffu, ok);
      const int pos = cnt + __popc(bal & ((1u << threadIdx.x) - 1));
      if (ok && pos < n_nbr) s_list[pos] = id;
      cnt += __popc(bal);
    }
    if (threadIdx.x == 0) s_cnt = min(cnt, n_nbr);
  }
  __syncthreads();
  const int cnt = s_cnt;
  const bool have = has_count[self] && cnt > 0;
  for (int64_t c = threadIdx.x; c < g; c += 128) {
    double v = nan("");
    if (have) {
      double tot = 0.0;
      for (int e = 0; e < cnt; ++e) tot += t[(int64_t)s_list[e] * g + c];
      v = t[self * g + c] / (tot / (double)cnt);
    }
    out[row * ldo + c] = v;
  }
}

// The exon rule (compute_dipcn_dir/compute_diploid_cn.py:8-63): the list is cut to its first n_nbr entries BEFORE
// neighbours without counts are dropped (they use up a slot); per region only neighbours with count > 0 and
// scale > 0 contribute; a query with a zero count, a non-positive scale or no contributor is NaN (omitted).
__global__ void __launch_bounds__(128) k_dipcn_exon(const int* __restrict__ counts, int64_t ldc,
                                                    const double* __restrict__ scale_q, const double* __restrict__ t,
                                                    int64_t g, const uint8_t* __restrict__ has_count,
                                                    const int* __restrict__ nbr_idx, int k, int64_t q_gid0, int n_nbr,
                                                    double* __restrict__ out, int64_t ldo) {
  extern __shared__ int s_list[];
  __shared__ int s_cnt;
  const int64_t row = blockIdx.x;
  const int64_t self = q_gid0 + row;
  const int lim = min(k, n_nbr);
  if (threadIdx.x < 32) {
    int cnt = 0;
    for (int base = 0; base < lim; base += 32) {
      const int e = base + threadIdx.x;
      const int id = e < lim ? nbr_idx[row * k + e] : -1;
      const bool ok = id >= 0 && has_count[id] && scale_q[id] > 0.0;
      const unsigned bal = __ballot_sync(0xffffffffu, ok);
      if (ok) s_list[cnt + __popc(bal & ((1u << threadIdx.x) - 1))] = id;
      cnt += __popc(bal);
    }
    if (threadIdx.x == 0) s_cnt = cnt;
  }
  __syncthreads();
  const int cnt = s_cnt;
  const bool have = has_count[self] && scale_q[self] > 0.0;
  for (int64_t c = threadIdx.x; c < g; c += 128) {
    double v = nan("");
    if (have && counts[self * ldc + c] != 0) {
      double tot = 0.0;
      int num = 0;
      for (int e = 0; e < cnt; ++e) {
        const int64_t j = s_list[e];
        if (counts[j * ldc + c] > 0) tot += t[j * g + c], ++num;
      }
      if (num > 0) {
        const double mean = tot / (double)num;
        if (mean > 0.0) v = t[self * g + c] / mean;
      }
    }
    out[row * ldo + c] = v;
  }
}

}  // namespace g2

using namespace g2;

extern "C" int gridb200_dipcn(const int32_t* counts, int64_t n_all, int64_t g, int64_t ldc, const uint8_t* has_count,
                              const double* scale_q, const int32_t* nbr_idx, int64_t nq, int32_t k, int64_t q_gid0,
                              int32_t n_nbr, int32_t rule, double* out, int64_t ldo, double* t_ws, void* stream) {
  G2_REQUIRE(counts && has_count && scale_q && nbr_idx && out && t_ws, GRIDB200_EINVAL, "null pointer");
  G2_REQUIRE(n_all > 0 && g > 0 && nq > 0 && k > 0 && n_nbr > 0 && ldc >= g && ldo >= g, GRIDB200_EINVAL, "bad shape");
  G2_REQUIRE(q_gid0 >= 0 && q_gid0 + nq <= n_all, GRIDB200_EINVAL, "query rows outside the cohort");
  G2_REQUIRE(rule == GRIDB200_DIPCN_PIPELINE || rule == GRIDB200_DIPCN_EXON, GRIDB200_EINVAL, "unknown dipCN rule %d", rule);
  n_nbr = n_nbr < k ? n_nbr : k;  // a list holds k entries: a larger budget changes nothing (and sizes the shared list)
  G2_REQUIRE(n_nbr <= 12000, GRIDB200_EINVAL, "min(n_nbr, k) = %d too large for the shared-memory list", n_nbr);
  cudaStream_t st = as_stream(stream);
  const int64_t cells = n_all * g;
  k_dipcn_scaled<<<(unsigned)((cells + 255) / 256), 256, 0, st>>>(counts, n_all, g, ldc, scale_q, t_ws);
  G2_LAUNCHED();
  if (rule == GRIDB200_DIPCN_EXON)
    k_dipcn_exon<<<(unsigned)nq, 128, (size_t)n_nbr * sizeof(int), st>>>(counts, ldc, scale_q, t_ws, g, has_count, nbr_idx,
                                                                         k, q_gid0, n_nbr, out, ldo);
  else
    k_dipcn<<<(unsigned)nq, 128, (size_t)n_nbr * sizeof(int), st>>>(t_ws, g, has_count, nbr_idx, k, q_gid0, n_nbr, out,
                                                                    ldo);
  G2_LAUNCHED();
  return 0;
}
