// General (unquantised) Gram block on Blackwell tensor cores with 3xTF32 split precision:
//     dot[i,j] = sum_k Q[i,k] * R[j,k]          for ARBITRARY fp32 operands
//
// north_star's neighbour search, step (c).  The pipeline's own operands are quantised (the %.2f hand-off) and take the
// exact fp16 kernel of gram_tc.cu; this kernel serves callers that hold unrounded fp32 z-scores.  tcgen05's kind::tf32
// reads only the upper 19 bits of an fp32 operand, so every operand is split once into
//     hi = tf32(x)   (round to nearest, low 13 mantissa bits cleared)      lo = x - hi   (exact in fp32)
// and three tensor-core products are accumulated in the same fp32 TMEM accumulator:
//     Q.R  ~=  Qhi.Rhi + Qhi.Rlo + Qlo.Rhi          (the dropped Qlo.Rlo term is 2^-22 relative)
// which restores fp32-level accuracy of every product (ranking then matches an fp32 reference; ties within 1e-6 rel).
//
// Operand layout: k-blocked like the fp16 operand, 32 floats (128 B) per row and k-block:
//     element (row i, k) at base[((k/32)*slab + i)*32 + k%32],     two arrays per operand (hi, lo).
// One CTA pair (cluster 2x1x1, cta_group::2) per 256 x 256 tile, persistent round robin over tiles, whole K per tile (no
// split-K: the fp32 result of a tile is summed in one fixed order -- deterministic).  Per k-block and CTA four TMA boxes
// (Qhi, Qlo, Rhi, Rlo: 4 x 16 KB), 12 MMAs (4 K-steps of 8 x 3 products).  TMEM double-buffered in chunks of `ck`
// k-blocks; the epilogue warps add each drained chunk into fp32 registers (chunked summation keeps the accumulation error
// near sqrt(K/ck) instead of K) and store the tile.
//   warp 0  TMA producer + TMEM allocator        warp 1  MMA issuer (leader)        warps 4-11  epilogue
#include <cuda.h>

#include <algorithm>

#include "common.cuh"
#include "tc_common.cuh"

namespace g2 {
namespace tf {

using namespace g2::tc;

constexpr int BK = 32, UMMA_K = 8;          // fp32 elements per 128-byte swizzle row / per MMA
constexpr int STAGES = 3;
constexpr int TM = 256, TN = 256, CTA_ROWS = 128;
constexpr int BOX_BYTES = CTA_ROWS * BK * 4;             // 16 KB
constexpr int STAGE_BYTES = 4 * BOX_BYTES;               // Qhi, Qlo, Rhi, Rlo
constexpr int ACC_STAGES = 2, TMEM_COLS = ACC_STAGES * TN;
constexpr int NUM_THREADS = 384, EPI_WARP0 = 4, EPI_WARPS = 8, EPI_COLS = TN / 2;
constexpr int REGS_CTRL = 40, REGS_EPI = 232;
constexpr int CHUNK_KB = 2;   // k-blocks (64 bins) accumulated in TMEM between drains: the tensor core TRUNCATES when it adds to
                              // an fp32 accumulator, a bias that grows with the partial sum; short chunks keep the partial sums small
                              // and the long sum runs in registers with round-to-nearest adds
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
// kind::tf32 instruction descriptor: D = f32 (bit 4), A = B = TF32 (2 at bits 7 and 10), K-major, N>>3 at 17, M>>4 at 24
constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);

struct Work {
  int tiles_m, tiles_n, ntiles, kb_total, ck;
  int64_t nq, nr, ldd;
};

__device__ __forceinline__ void tma_load_box(uint32_t dst, const CUtensorMap* map, uint32_t leader_bar, int row, int kb) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
      "l"(map), "r"(leader_bar), "r"(0), "r"(row), "r"(kb)
      : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
k_gram_tf32x3_pair(const __grid_constant__ CUtensorMap map_qh, const __grid_constant__ CUtensorMap map_ql,
                   const __grid_constant__ CUtensorMap map_rh, const __grid_constant__ CUtensorMap map_rl, const Work w,
                   float* __restrict__ dot) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = base + STAGES * STAGE_BYTES;
  const uint32_t full0 = bars, empty0 = bars + 8 * STAGES, tfull0 = bars + 16 * STAGES, tempty0 = tfull0 + 8 * ACC_STAGES;
  const uint32_t tmem_slot = tempty0 + 8 * ACC_STAGES;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair_id = blockIdx.x >> 1, npairs = gridDim.x >> 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full0 + 8 * s, 1);   // the leader's expect_tx arrive; both CTAs' TMA bytes complete_tx on the leader
      mbar_init(empty0 + 8 * s, 1);  // one multicast commit
    }
    for (int a = 0; a < ACC_STAGES; ++a) {
      mbar_init(tfull0 + 8 * a, 1);
      mbar_init(tempty0 + 8 * a, 2 * EPI_WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp < EPI_WARP0) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(REGS_CTRL));
  if (warp == 0) {
    // ===================================== TMA producer (both CTAs) =========================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = pair_id; item < w.ntiles; item += npairs) {
        const int tm = item / w.tiles_n, tn = item - tm * w.tiles_n;
        const int qrow = tm * TM + (int)rank * CTA_ROWS, rrow = tn * TN + (int)rank * CTA_ROWS;
        for (int kb = 0; kb < w.kb_total; ++kb) {
          mbar_wait(empty0 + 8 * stage, phase ^ 1);
          const uint32_t leader_full = map_to_cta(full0 + 8 * stage, 0);
          if (rank == 0) mbar_expect_tx(full0 + 8 * stage, 2 * STAGE_BYTES);
          const uint32_t sa = base + stage * STAGE_BYTES;
          tma_load_box(sa, &map_qh, leader_full, qrow, kb);
          tma_load_box(sa + BOX_BYTES, &map_ql, leader_full, qrow, kb);
          tma_load_box(sa + 2 * BOX_BYTES, &map_rh, leader_full, rrow, kb);
          tma_load_box(sa + 3 * BOX_BYTES, &map_rl, leader_full, rrow, kb);
          if (++stage == STAGES) stage = 0, phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ===================================== MMA issuer (leader only) =========================
    if (rank == 0 && lane == 0) {
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int item = pair_id; item < w.ntiles; item += npairs) {
        for (int c0 = 0; c0 < w.kb_total; c0 += w.ck) {
          const int c1 = min(w.kb_total, c0 + w.ck);
          mbar_wait(tempty0 + 8 * acc, acc_phase ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + (uint32_t)(acc * TN);
          for (int kb = c0; kb < c1; ++kb) {
            mbar_wait(full0 + 8 * stage, phase);
            tc_fence_after();
            const uint32_t sa = base + stage * STAGE_BYTES;
            const uint64_t qh = smem_desc(sa), ql = smem_desc(sa + BOX_BYTES);
            const uint64_t rh = smem_desc(sa + 2 * BOX_BYTES), rl = smem_desc(sa + 3 * BOX_BYTES);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {  // +32 bytes (2 x 16 B) per K8 step inside the swizzle atom
              tc_mma_pair_tf32(d_tmem, qh + 2 * k, rh + 2 * k, IDESC, (kb > c0 || k > 0) ? 1u : 0u);
              tc_mma_pair_tf32(d_tmem, qh + 2 * k, rl + 2 * k, IDESC, 1u);
              tc_mma_pair_tf32(d_tmem, ql + 2 * k, rh + 2 * k, IDESC, 1u);
            }
            tc_commit_pair(empty0 + 8 * stage);
            if (++stage == STAGES) stage = 0, phase ^= 1;
          }
          tc_commit_pair(tfull0 + 8 * acc);
          if (++acc == ACC_STAGES) acc = 0, acc_phase ^= 1;
        }
      }
    }
  } else if (warp >= EPI_WARP0) {
    // ===================================== epilogue (both CTAs) =============================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(REGS_EPI));
    const int quarter = warp & 3, half = (warp - EPI_WARP0) >> 2;
    float accf[EPI_COLS];
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int item = pair_id; item < w.ntiles; item += npairs) {
      const int tm = item / w.tiles_n, tn = item - tm * w.tiles_n;
      const int64_t row = (int64_t)tm * TM + rank * CTA_ROWS + quarter * 32 + lane;
      const int64_t col0 = (int64_t)tn * TN + half * EPI_COLS;
#pragma unroll
      for (int i = 0; i < EPI_COLS; ++i) accf[i] = 0.f;
      for (int c0 = 0; c0 < w.kb_total; c0 += w.ck) {
        mbar_wait(tfull0 + 8 * acc, acc_phase);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * TN + half * EPI_COLS);
        uint32_t va[32], vb[32];
        tmem_ld32(taddr, va);
        tmem_ld_wait32(va);
        tmem_ld32(taddr + 32, vb);
#pragma unroll
        for (int i = 0; i < 32; ++i) accf[i] += __uint_as_float(va[i]);
        tmem_ld_wait32(vb);
        tmem_ld32(taddr + 64, va);
#pragma unroll
        for (int i = 0; i < 32; ++i) accf[32 + i] += __uint_as_float(vb[i]);
        tmem_ld_wait32(va);
        tmem_ld32(taddr + 96, vb);
#pragma unroll
        for (int i = 0; i < 32; ++i) accf[64 + i] += __uint_as_float(va[i]);
        tmem_ld_wait32(vb);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(map_to_cta(tempty0 + 8 * acc, 0));
#pragma unroll
        for (int i = 0; i < 32; ++i) accf[96 + i] += __uint_as_float(vb[i]);
        if (++acc == ACC_STAGES) acc = 0, acc_phase ^= 1;
      }
      if (row < w.nq) {
        float* out = dot + row * w.ldd + col0;
#pragma unroll
        for (int i = 0; i < EPI_COLS; ++i)
          if (col0 + i < w.nr) out[i] = accf[i];
      }
    }
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

// row-major fp32 x[n, ldx] -> k-blocked hi / lo arrays [kdim/32, slab, 32] (zero padded to kdim) + fp64 row norms.
// One block per row; the norm is reduced in a fixed order (deterministic).
__global__ void __launch_bounds__(256) k_split_tf32(const float* __restrict__ x, int64_t k, int64_t ldx, int64_t kdim,
                                                    float* __restrict__ hi, float* __restrict__ lo, int64_t slab,
                                                    double* __restrict__ norms) {
  const int64_t row = blockIdx.x;
  double acc = 0.0;
  for (int64_t j = threadIdx.x; j < kdim; j += 256) {
    const float v = j < k ? x[row * ldx + j] : 0.f;
    uint32_t hb;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hb) : "f"(v));   // round to nearest: 10 explicit mantissa bits kept
    const float h = __uint_as_float(hb & 0xFFFFE000u);
    const int64_t at = ((j >> 5) * slab + row) * 32 + (j & 31);
    hi[at] = h;
    lo[at] = v - h;                                           // exact: the low 13 mantissa bits of v
    acc += (double)v * (double)v;
  }
  acc = warp_sum(acc);
  __shared__ double part[8];
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0 && norms) {
    double t = 0.0;
#pragma unroll
    for (int wv = 0; wv < 8; ++wv) t += part[wv];
    norms[row] = t;
  }
}

static int make_map(CUtensorMap* map, const void* ptr, int64_t rows, int64_t nkb, int64_t slab) {
  typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeTiledFn enc = nullptr;
  if (!enc) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      enc = reinterpret_cast<EncodeTiledFn>(p);
  }
  G2_REQUIRE(enc != nullptr, GRIDB200_ENODEV, "cuTensorMapEncodeTiled is not available from this driver");
  const cuuint64_t dims[3] = {(cuuint64_t)BK, (cuuint64_t)rows, (cuuint64_t)nkb};
  const cuuint64_t strides[2] = {(cuuint64_t)BK * 4, (cuuint64_t)slab * BK * 4};
  const cuuint32_t box[3] = {(cuuint32_t)BK, (cuuint32_t)CTA_ROWS, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  G2_REQUIRE(r == CUDA_SUCCESS, GRIDB200_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

}  // namespace tf
}  // namespace g2

using namespace g2;

extern "C" int gridb200_split_tf32(const float* x, int64_t n, int64_t k, int64_t ldx, float* hi, float* lo, int64_t slab,
                                   double* norms, void* stream) {
  G2_REQUIRE(x && hi && lo, GRIDB200_EINVAL, "null pointer");
  G2_REQUIRE(n > 0 && k > 0 && ldx >= k && slab >= n, GRIDB200_EINVAL, "bad shape");
  const int64_t kdim = (k + tf::BK - 1) / tf::BK * tf::BK;
  G2_REQUIRE(n < (1LL << 31), GRIDB200_EINVAL, "too many rows");
  tf::k_split_tf32<<<(unsigned)n, 256, 0, as_stream(stream)>>>(x, k, ldx, kdim, hi, lo, slab, norms);
  G2_LAUNCHED();
  return 0;
}

extern "C" int gridb200_gram_f32(const float* qhi, const float* qlo, int64_t nq, int64_t slabq, const float* rhi,
                                 const float* rlo, int64_t nr, int64_t slabr, int64_t kdim, float* dot, int64_t ldd,
                                 void* stream) {
  using namespace g2::tf;
  G2_REQUIRE(qhi && qlo && rhi && rlo && dot, GRIDB200_EINVAL, "null pointer");
  G2_REQUIRE(nq > 0 && nr > 0 && kdim > 0 && kdim % BK == 0 && slabq >= nq && slabr >= nr && ldd >= nr, GRIDB200_EINVAL,
             "bad shape (kdim must be a multiple of %d)", BK);
  G2_REQUIRE(nq < (1LL << 31) && nr < (1LL << 31) && kdim < (1LL << 31), GRIDB200_EINVAL, "shape exceeds int32");
  G2_REQUIRE(((reinterpret_cast<uintptr_t>(qhi) | reinterpret_cast<uintptr_t>(qlo) | reinterpret_cast<uintptr_t>(rhi) |
               reinterpret_cast<uintptr_t>(rlo)) & 15) == 0, GRIDB200_EINVAL, "operands must be 16-byte aligned");
  cudaStream_t st = as_stream(stream);
  static std::atomic<int> sms_of[64];
  int dev = 0;
  G2_CUDA(cudaGetDevice(&dev));
  G2_REQUIRE(dev >= 0 && dev < 64, GRIDB200_ENODEV, "device index %d outside the per-device cache", dev);
  int sms = sms_of[dev].load();
  if (!sms) {
    cudaDeviceProp prop;
    G2_CUDA(cudaGetDeviceProperties(&prop, dev));
    G2_REQUIRE(prop.major == 10, GRIDB200_ENODEV, "tcgen05 path needs an sm_100 device (found sm_%d%d)", prop.major, prop.minor);
    G2_CUDA(cudaFuncSetAttribute(k_gram_tf32x3_pair, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    sms = prop.multiProcessorCount;
    sms_of[dev].store(sms);
  }
  const int64_t nkb = kdim / BK;
  CUtensorMap mqh, mql, mrh, mrl;
  if (int e = make_map(&mqh, qhi, nq, nkb, slabq)) return e;
  if (int e = make_map(&mql, qlo, nq, nkb, slabq)) return e;
  if (int e = make_map(&mrh, rhi, nr, nkb, slabr)) return e;
  if (int e = make_map(&mrl, rlo, nr, nkb, slabr)) return e;
  Work w;
  w.nq = nq, w.nr = nr, w.ldd = ldd;
  w.tiles_m = (int)((nq + TM - 1) / TM), w.tiles_n = (int)((nr + TN - 1) / TN);
  w.ntiles = w.tiles_m * w.tiles_n;
  w.kb_total = (int)nkb;
  w.ck = (int)std::min<int64_t>(CHUNK_KB, nkb);
  const int pairs = std::max(1, std::min(sms / 2, w.ntiles));
  k_gram_tf32x3_pair<<<2 * pairs, NUM_THREADS, SMEM_BYTES, st>>>(mqh, mql, mrh, mrl, w, dot);
  G2_LAUNCHED();
  return 0;
}
