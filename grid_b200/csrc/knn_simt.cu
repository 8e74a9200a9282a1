// Exact Gram block on plain SIMT integer units: the cross-check for the tcgen05 kernel (impl = 1) and the
// definition of what gridb200_gram computes: dot[i,j] (+)= sum_k q[i,k]*r[j,k], operands fp16 integers.
#include "common.cuh"

namespace g2 {

constexpr int kTile = 64;
constexpr int kKT = 32;

__global__ void __launch_bounds__(256) k_gram_simt(const __half* __restrict__ q, int64_t nq, int64_t ldq,
                                                   const __half* __restrict__ r, int64_t nr, int64_t ldr,
                                                   int64_t kdim, long long* __restrict__ dot, int64_t ldd,
                                                   int accumulate) {
  __shared__ int sq[kTile][kKT + 1];
  __shared__ int sr[kTile][kKT + 1];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int64_t i0 = (int64_t)blockIdx.y * kTile, j0 = (int64_t)blockIdx.x * kTile;
  long long acc[4][4];
  int part[4][4];
#pragma unroll
  for (int u = 0; u < 4; ++u)
#pragma unroll
    for (int v = 0; v < 4; ++v) acc[u][v] = 0, part[u][v] = 0;
  int tiles = 0;
  for (int64_t k0 = 0; k0 < kdim; k0 += kKT) {
    for (int e = threadIdx.x; e < kTile * kKT; e += 256) {
      const int row = e / kKT, kk = e % kKT;
      const int64_t k = k0 + kk;
      // k-blocked operands: element (row, k) at ((k/64)*slab + row)*64 + k%64
      sq[row][kk] = (i0 + row < nq && k < kdim) ? __half2int_rn(q[((k >> 6) * ldq + i0 + row) * 64 + (k & 63)]) : 0;
      sr[row][kk] = (j0 + row < nr && k < kdim) ? __half2int_rn(r[((k >> 6) * ldr + j0 + row) * 64 + (k & 63)]) : 0;
    }
    __syncthreads();
#pragma unroll 8
    for (int kk = 0; kk < kKT; ++kk) {
      int a[4], bb[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) a[u] = sq[ty * 4 + u][kk], bb[u] = sr[tx * 4 + u][kk];
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 4; ++v) part[u][v] += a[u] * bb[v];
    }
    __syncthreads();
    if (++tiles == 8) {  // 8*32 products of at most 2048^2 stay below 2^31
      tiles = 0;
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 4; ++v) acc[u][v] += part[u][v], part[u][v] = 0;
    }
  }
#pragma unroll
  for (int u = 0; u < 4; ++u)
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      const int64_t i = i0 + ty * 4 + u, j = j0 + tx * 4 + v;
      if (i < nq && j < nr) {
        const long long val = acc[u][v] + part[u][v];
        long long* o = dot + i * ldd + j;
        *o = accumulate ? *o + val : val;
      }
    }
}

}  // namespace g2

int gridb200_gram_tc(const void* zq, int64_t nq, int64_t ldq, const void* zr, int64_t nr, int64_t ldr,
                     int64_t kdim, int32_t vmax, int64_t* dot, int64_t ldd, int accumulate, void* stream);

extern "C" int gridb200_gram(const void* zq, int64_t nq, int64_t ldq, const void* zr, int64_t nr, int64_t ldr,
                             int64_t kdim, int32_t vmax, int64_t* dot, int64_t ldd, int accumulate, int impl,
                             void* stream) {
  using namespace g2;
  G2_REQUIRE(zq && zr && dot, GRIDB200_EINVAL, "null pointer");
  G2_REQUIRE(nq > 0 && nr > 0 && kdim > 0 && ldd >= nr, GRIDB200_EINVAL, "bad shape");
  G2_REQUIRE(ldq >= nq && ldr >= nr && kdim % 64 == 0, GRIDB200_EINVAL,
             "kdim must be a multiple of 64 and each k-block slab must hold its rows (got kdim=%lld nq=%lld slabq=%lld "
             "nr=%lld slabr=%lld)",
             (long long)kdim, (long long)nq, (long long)ldq, (long long)nr, (long long)ldr);
  G2_REQUIRE(((reinterpret_cast<uintptr_t>(zq) | reinterpret_cast<uintptr_t>(zr)) & 15) == 0, GRIDB200_EINVAL,
             "operands must be 16-byte aligned");
  G2_REQUIRE(vmax > 0 && vmax <= 2048, GRIDB200_EDOMAIN, "vmax=%d outside (0, 2048]: fp16 cannot hold it exactly",
             vmax);
  if (impl == 1) {
    dim3 grid((unsigned)((nr + kTile - 1) / kTile), (unsigned)((nq + kTile - 1) / kTile));
    k_gram_simt<<<grid, 256, 0, as_stream(stream)>>>(static_cast<const __half*>(zq), nq, ldq,
                                                     static_cast<const __half*>(zr), nr, ldr, kdim,
                                                     reinterpret_cast<long long*>(dot), ldd, accumulate);
    G2_LAUNCHED();
    return 0;
  }
  G2_REQUIRE(impl == 0, GRIDB200_EINVAL, "unknown impl %d", impl);
  return gridb200_gram_tc(zq, nq, ldq, zr, nr, ldr, kdim, vmax, dot, ldd, accumulate, stream);
}
