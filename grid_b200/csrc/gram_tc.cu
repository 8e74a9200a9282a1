// Exact Gram block on Blackwell tensor cores:  dot[i,j] += sum_k Q[i,k] * R[j,k]
//
// Operands are fp16 holding small integers (clipped z-scores in hundredths, |v| <= vmax), so every product is
// exact and an fp32 accumulator stays an exact integer while its magnitude is below 2^24.  The K loop is cut into
// chunks of ck k-blocks with ck*64*vmax^2 < 2^24; each chunk accumulates in TMEM (tcgen05.mma, fp32), is drained
// with tcgen05.ld, converted to int32 and added into per-thread int32 registers; those are flushed to the int64
// result with red.global.add before they can overflow.  Nothing is ever rounded: the result is the exact integer
// dot product, whatever K is.
//
// Operand layout ("k-blocked"): element (row i, bin k) lives at  base + ((k/64)*slab + i)*64 + k%64  (fp16), i.e.
// every 64-bin k-block is one contiguous slab of `slab` rows x 128 B.  A 128-row operand box of one k-block is a
// single contiguous 16 KB read (the row-major layout this replaces cost one DRAM page and one TLB entry per row).
//
// One CTA PAIR (cluster 2x1x1, tcgen05 cta_group::2) per 256 x 256 output tile, persistent over work items =
// (tile, K split).  Each CTA stages its own 128 query rows and 128 of the 256 reference rows per k-block; the
// leader issues M=256 N=256 K=16 MMAs; each CTA's TMEM holds its 128 x 256 half of the accumulator (two stages:
// chunk c+1 accumulates while chunk c drains).
//   warp 0      TMA producer (both CTAs; transaction bytes land on the LEADER's full barrier, on which only the
//               leader arrives) + TMEM allocator
//               + L2 pacing (below)
//   warp 1      MMA issuer (leader CTA only); commits are multicast to both CTAs' barriers
//   warps 4-11  epilogue: lane quarter = warp%4, column half = (warp-4)/4, 128 int32 accumulators per thread
//
// L2 pacing.  K is millions of bins long, so free-running pairs drift apart along K until nothing a pair loads is
// still in L2 when the next pair wants it (ncu, first version: 335 GB of DRAM reads for a 16.7 GB operand, 33 %
// L2 hits).  The producers therefore publish their k-block step in a global table and do not run more than
// `window` steps ahead of the slowest pair, which keeps the union of all in-flight operand slabs inside L2.
// Pacing is a performance hint only: a wait that exceeds its time-out is abandoned and pacing is switched off for
// the rest of that CTA's life, so no forward-progress assumption about other CTAs is ever needed.
//
// Symmetric mode (Q and R are the same matrix): only tiles on or above the diagonal are computed; a small
// transpose kernel then copies them below the diagonal (dot[j,i] = dot[i,j]).
#include <cuda.h>
#include <stdlib.h>

#include <algorithm>

#include "common.cuh"
#include "tc_common.cuh"

namespace g2 {
namespace tc {

constexpr int BK = 64, UMMA_K = 16;
constexpr int STAGES = 6;
constexpr int TM = 256, TN = 256;           // tile of the pair
constexpr int CTA_ROWS = 128;               // query rows / reference rows staged per CTA
constexpr int A_BYTES = CTA_ROWS * BK * 2, B_BYTES = CTA_ROWS * BK * 2, STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int ACC_STAGES = 2;
constexpr int TMEM_COLS = ACC_STAGES * TN;  // 512: all of TMEM
constexpr int NUM_THREADS = 384;            // warpgroup 0 = {TMA, MMA, 2 idle warps}, warpgroups 1-2 = epilogue
constexpr int EPI_WARP0 = 4, EPI_WARPS = 8;
constexpr int REGS_CTRL = 40, REGS_EPI = 232;  // setmaxnreg: 128*40 + 256*232 = 64512 <= 65536
constexpr int EPI_COLS = TN / 2;            // 128 columns per epilogue thread
constexpr int CHECK_EVERY = 16;             // chunks between int32 overflow checks
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*alignment slack*/ + 256 /*barriers*/;
constexpr int PACE = 16;                    // k-block steps between pacing checks
constexpr long long PACE_TIMEOUT = 400000;  // cycles a producer will wait for the slowest pair (~0.2 ms)
constexpr int MAX_PAIRS = 128;              // static schedule: CTA pairs of one launch
constexpr int PACE_SLOTS = 192;             // progress table entries (static: pair id; dynamic: the leader's %smid)
constexpr int CLC_SLOTS = 2;                // work-stealing responses in flight per pair
constexpr int CLC_CONSUMERS = 2 * (1 + EPI_WARPS) + 2;  // both producers + all epilogue warps + MMA lane + scheduler

// kind::f16 instruction descriptor: D=f32 (bit 4), A=B=f16 (0), both K-major, N>>3 at bit 17, M>>4 at bit 24
constexpr uint32_t IDESC = (1u << 4) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);

// progress table of the pacing scheme: (launch epoch << 32) | k-block offset inside the current item
__device__ unsigned long long g_pace[PACE_SLOTS];

struct Work {
  int tiles_m, tiles_n, ntiles, ksplits, kb_total, kb_per_split, ck, group_m;
  int symmetric, window, pace_slots;
  unsigned epoch;
  int64_t nq, nr, ldd;
};

// Fused k-best epilogue (gridb200_gram_topk): instead of writing the block of dot products, a finished tile forms
// d2 = |q|^2 + |r|^2 - 2 q.r in int64 and keeps only what can still enter a list -- d2 <= tau of its query row, or
// <= tau of its reference row (the half ring wants the block read both ways) -- appending (d2, id) to small per-row
// candidate buffers.  The distances themselves never reach memory.
struct Fused {
  const long long* norm_q;   // exact norms of the nq query rows / nr reference rows of this block
  const long long* norm_r;
  const long long* tau_q;    // current k-th best d2 of every query row (element i at tau_q[i * tau_q_stride]); +inf = take all
  const long long* tau_r;    // ... of every reference row; nullptr = no lists for the reference rows
  int64_t tau_q_stride, tau_r_stride;
  long long* cq_d2;          // candidates of the query rows: [nq, cap] (d2, id), cq_cnt[nq] = appended so far (may exceed cap)
  int* cq_id;
  int* cq_cnt;
  long long* cr_d2;          // candidates of the reference rows: [nr, cap]
  int* cr_id;
  int* cr_cnt;
  int cap;
  int64_t q_gid0, r_gid0;    // global sample ids of query row 0 / reference row 0
  long long* scratch;        // [SMs][128][256] int64: private carries of the tile a CTA is working on
};

// acc += (int)x for an integer-valued float |x| < 2^24, two ways, so that the work is split between two pipes:
//   add_exact_xu : F2I (XU pipe, 16 lanes/clk/SM -- ncu showed this pipe 82 % busy when it did all the work)
//   add_exact_fma: FFMA/FADD/IMAD/IADD3 only.  With M = 1.5*2^23: t = x/4096 + M has bits(M) + hi, hi = rint(x/4096);
//                  lo = x - 4096*hi is exact and |lo| <= 2048, so bits(lo + M) = bits(M) + lo; x = 4096*hi + lo.
__device__ __forceinline__ void add_exact_xu(int& acc, uint32_t xbits) { acc += __float2int_rn(__uint_as_float(xbits)); }
__device__ __forceinline__ void add_exact_fma(int& acc, uint32_t xbits) {
  constexpr float M = 12582912.0f;                                // 0x4B400000
  constexpr uint32_t K = 0u - (0x4B400000u * 4096u + 0x4B400000u);  // removes both biases (mod 2^32)
  const float x = __uint_as_float(xbits);
  const float t = __fmaf_rn(x, 0.000244140625f, M);
  const float hf = __fadd_rn(t, -M);
  const float lo = __fmaf_rn(hf, -4096.0f, x);
  const float u = __fadd_rn(lo, M);
  acc = (int)((uint32_t)acc + (__float_as_uint(t) << 12) + (__float_as_uint(u) + K));
}
template <int kBase>
__device__ __forceinline__ void drain32(int* acc32, const uint32_t* v) {
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    if (i & 1) add_exact_fma(acc32[kBase + i], v[i]);
    else add_exact_xu(acc32[kBase + i], v[i]);
  }
}

// item -> (k split, tile m, tile n).  K-split major, so that the pairs of one wave stream the same K window.
// General mode: tiles walked in bands of group_m tile-rows so concurrent tiles share operand rows.
// Symmetric mode: the tiles_m*(tiles_m+1)/2 tiles with tn >= tm, row by row.
__device__ __forceinline__ void decode_item(const Work& w, int item, int& ks, int& tm, int& tn) {
  ks = item / w.ntiles;
  int t = item - ks * w.ntiles;
  if (w.symmetric) {
    int rowlen = w.tiles_n;
    tm = 0;
    while (t >= rowlen) t -= rowlen, --rowlen, ++tm;
    tn = tm + t;
    return;
  }
  const int per_group = w.group_m * w.tiles_n;
  const int g = t / per_group;
  const int first_m = g * w.group_m;
  const int gsize = min(w.tiles_m - first_m, w.group_m);
  const int r = t - g * per_group;
  tm = first_m + r % gsize;
  tn = r / gsize;
}

// Whole warp 0.  Publish this pair's offset inside its item, then hold the producer until every pair of its cohort
// is within `window` k-blocks.  Items all have the same K length and the item order is K-split major, so pairs at
// the same offset stream the same bins.  A pair further behind than 4 windows is out of phase (it joined late: the
// dynamic schedule starts clusters whenever SMs become free) and is not waited for.
__device__ __forceinline__ void pace(const Work& w, int slot, uint32_t rank, int lane, unsigned off, bool& enabled) {
  if (rank == 0 && lane == 0)
    *reinterpret_cast<volatile unsigned long long*>(&g_pace[slot]) = ((unsigned long long)w.epoch << 32) | off;
  if (!enabled || off < (unsigned)w.window) return;
  const unsigned need = off - (unsigned)w.window;
  const unsigned behind = off > 4u * (unsigned)w.window ? off - 4u * (unsigned)w.window : 0u;
  const long long t0 = clock64();
  for (;;) {
    unsigned m = 0xffffffffu;
    for (int p = lane; p < w.pace_slots; p += 32) {
      const unsigned long long v = *reinterpret_cast<volatile unsigned long long*>(&g_pace[p]);
      // another epoch = no pair of this launch on that slot (yet, or any more): nothing to wait for
      const unsigned s = (unsigned)(v >> 32) == w.epoch ? (unsigned)v : 0xffffffffu;
      if (s >= behind) m = min(m, s);
    }
    m = __reduce_min_sync(0xffffffffu, m);
    if (m >= need) return;
    if (clock64() - t0 > PACE_TIMEOUT) enabled = false;  // decided by lane 0 below
    enabled = __shfl_sync(0xffffffffu, (int)enabled, 0) != 0;
    if (!enabled) return;
    __nanosleep(200);
  }
}

// The item after `item` for one role of a pair.
//   static  : round robin over the pairs of the launch.
//   dynamic : the next answer of the cluster-launch-control ring (filled by the leader's scheduler warp).  Every
//             consumer of an answer arrives on the LEADER's `empty` barrier of that slot; kWarp = a whole warp
//             reads the answer and lane 0 arrives for it.
struct Cursor {
  int slot = 0;
  uint32_t phase = 0;
};
template <bool kDynamic, bool kWarp>
__device__ __forceinline__ int next_item(int item, int npairs, int nitems, uint32_t clc_full0, uint32_t clc_empty0,
                                         uint32_t clc_resp0, Cursor& c, int lane) {
  if (!kDynamic) {
    item += npairs;
    return item < nitems ? item : -1;
  }
  mbar_wait(clc_full0 + 8 * c.slot, c.phase);
  const int nxt = clc_decode(clc_resp0 + 16 * c.slot);
  if (kWarp) __syncwarp();
  if (!kWarp || lane == 0) mbar_arrive_cluster(map_to_cta(clc_empty0 + 8 * c.slot, 0));
  if (++c.slot == CLC_SLOTS) c.slot = 0, c.phase ^= 1;
  return nxt;
}

template <bool kDynamic, bool kFused>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
k_gram_tc_pair(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_r, const Work w,
               long long* __restrict__ dot, const Fused f) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // 128B-swizzle atoms need 1024-byte alignment
  const uint32_t bars = base + STAGES * STAGE_BYTES;
  const uint32_t full0 = bars, empty0 = bars + 8 * STAGES, tfull0 = bars + 16 * STAGES, tempty0 = tfull0 + 8 * ACC_STAGES;
  const uint32_t tmem_slot = tempty0 + 8 * ACC_STAGES;
  const uint32_t clc_full0 = tmem_slot + 8, clc_empty0 = clc_full0 + 8 * CLC_SLOTS, sched_go = clc_empty0 + 8 * CLC_SLOTS;
  const uint32_t clc_resp0 = bars + 192;  // 16-byte aligned (bars is 1024-aligned)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  // static: this pair takes items pair_id, pair_id + npairs, ...; dynamic: one cluster was launched per item, a
  // running pair starts with its own and then cancels not-yet-started clusters and does their items itself
  const int pair_id = blockIdx.x >> 1, npairs = gridDim.x >> 1;
  const int nitems = w.ntiles * w.ksplits;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full0 + 8 * s, 1);   // the leader's expect_tx arrive; both CTAs' TMA bytes complete_tx on the leader
      mbar_init(empty0 + 8 * s, 1);  // one multicast commit
    }
    for (int a = 0; a < ACC_STAGES; ++a) {
      mbar_init(tfull0 + 8 * a, 1);                // one multicast commit
      mbar_init(tempty0 + 8 * a, 2 * EPI_WARPS);   // every epilogue warp of both CTAs (leader's copy is used)
    }
    for (int c = 0; c < CLC_SLOTS; ++c) {
      mbar_init(clc_full0 + 8 * c, 1);               // the scheduler's expect_tx arrive (+ 16 answer bytes)
      mbar_init(clc_empty0 + 8 * c, CLC_CONSUMERS);  // leader's copy is used
    }
    mbar_init(sched_go, 1);
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
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_q) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&map_r) : "memory");
    }
    int stage = 0;
    uint32_t phase = 0;
    bool pacing = w.window > 0 && npairs > 1;
    const int pace_slot = kDynamic ? (int)sm_id() : pair_id;  // one CTA per SM: the leader's SM names the pair
    Cursor cur;
    for (int item = pair_id; item >= 0 && item < nitems;
         item = next_item<kDynamic, true>(item, npairs, nitems, clc_full0, clc_empty0, clc_resp0, cur, lane)) {
      if (kDynamic && rank == 0 && lane == 0) mbar_arrive_cluster(map_to_cta(sched_go, 0));  // fetch the item after this
      int ks, tm, tn;
      decode_item(w, item, ks, tm, tn);
      const int kb0 = ks * w.kb_per_split, kb1 = min(w.kb_total, kb0 + w.kb_per_split);
      const int qrow = tm * TM + (int)rank * CTA_ROWS, rrow = tn * TN + (int)rank * CTA_ROWS;
      for (int kb = kb0; kb < kb1; ++kb) {
        if ((kb - kb0) % PACE == 0) {
          pace(w, pace_slot, rank, lane, (unsigned)(kb - kb0), pacing);
          __syncwarp();
        }
        if (lane == 0) {
          mbar_wait(empty0 + 8 * stage, phase ^ 1);
          const uint32_t leader_full = map_to_cta(full0 + 8 * stage, 0);
          // Only the leader arrives (expecting the bytes of both CTAs).  The peer's bytes may land before that
          // arrive -- the transaction count just goes negative for a moment; the phase cannot complete early
          // because the leader's arrival is still pending, and it cannot be a phase late because the peer only
          // refills a slot after the multicast `empty` commit that follows the previous phase's completion.
          if (rank == 0) mbar_expect_tx(full0 + 8 * stage, 2 * STAGE_BYTES);
          const uint32_t sa = base + stage * STAGE_BYTES;
          tma_load_pair(sa, &map_q, leader_full, qrow, kb);
          tma_load_pair(sa + A_BYTES, &map_r, leader_full, rrow, kb);
        }
        if (++stage == STAGES) stage = 0, phase ^= 1;
      }
    }
    if (rank == 0 && lane == 0)  // finished: never hold anybody back
      *reinterpret_cast<volatile unsigned long long*>(&g_pace[pace_slot]) = ((unsigned long long)w.epoch << 32) | 0xffffffffu;
  } else if (warp == 1) {
    // ===================================== MMA issuer (leader only) =========================
    if (rank == 0 && lane == 0) {
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      Cursor cur;
      for (int item = pair_id; item >= 0 && item < nitems;
           item = next_item<kDynamic, false>(item, npairs, nitems, clc_full0, clc_empty0, clc_resp0, cur, 0)) {
        int ks, tm, tn;
        decode_item(w, item, ks, tm, tn);
        const int kb0 = ks * w.kb_per_split, kb1 = min(w.kb_total, kb0 + w.kb_per_split);
        for (int c0 = kb0; c0 < kb1; c0 += w.ck) {
          const int c1 = min(kb1, c0 + w.ck);
          mbar_wait(tempty0 + 8 * acc, acc_phase ^ 1);  // both CTAs' epilogues have drained this TMEM stage
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + (uint32_t)(acc * TN);
          for (int kb = c0; kb < c1; ++kb) {
            mbar_wait(full0 + 8 * stage, phase);
            tc_fence_after();
            const uint32_t sa = base + stage * STAGE_BYTES;
            const uint64_t da = smem_desc(sa), db = smem_desc(sa + A_BYTES);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)  // +32 bytes (2 x 16 B) per K16 step inside the swizzle atom
              tc_mma_pair(d_tmem, da + 2 * k, db + 2 * k, IDESC, (kb > c0 || k > 0) ? 1u : 0u);
            tc_commit_pair(empty0 + 8 * stage);  // smem slot is free (in both CTAs) once these MMAs have read it
            if (++stage == STAGES) stage = 0, phase ^= 1;
          }
          tc_commit_pair(tfull0 + 8 * acc);  // chunk complete in TMEM
          if (++acc == ACC_STAGES) acc = 0, acc_phase ^= 1;
        }
      }
    }
  } else if (warp == 2) {
    // ===================================== work-stealing scheduler (leader only, dynamic) ===
    if (kDynamic && rank == 0 && lane == 0) {
      Cursor cur;
      uint32_t go_phase = 0;
      for (;;) {
        mbar_wait(sched_go, go_phase);  // the producer has started an item: steal the one after it, no further ahead
        go_phase ^= 1;
        mbar_wait(clc_empty0 + 8 * cur.slot, cur.phase ^ 1);  // every role of both CTAs has read the old answer
        mbar_expect_tx(clc_full0 + 8 * cur.slot, 16);
        mbar_expect_tx_cluster(map_to_cta(clc_full0 + 8 * cur.slot, 1), 16);
        clc_try_cancel(clc_resp0 + 16 * cur.slot, clc_full0 + 8 * cur.slot);
        if (next_item<true, false>(0, 0, 0, clc_full0, clc_empty0, clc_resp0, cur, 0) < 0) break;
      }
    }
  } else if (warp >= EPI_WARP0) {
    // ===================================== epilogue (both CTAs) =============================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(REGS_EPI));
    const int quarter = warp & 3, half = (warp - EPI_WARP0) >> 2;
    int acc32[EPI_COLS];
#pragma unroll
    for (int i = 0; i < EPI_COLS; ++i) acc32[i] = 0;
    int acc = 0, since_check = 0;
    uint32_t acc_phase = 0;
    Cursor cur;
    // fused: carries of this thread's 128 elements live in a private slab of global memory (L2), [element][thread] so
    // that a warp's accesses coalesce; a bit per element says whether its slot has been written for the current tile
    uint32_t carried[EPI_COLS / 32];
    long long* scr = nullptr;
    if (kFused) scr = f.scratch + (size_t)sm_id() * (CTA_ROWS * TN) + (threadIdx.x - EPI_WARP0 * 32);
    for (int item = pair_id; item >= 0 && item < nitems;
         item = next_item<kDynamic, true>(item, npairs, nitems, clc_full0, clc_empty0, clc_resp0, cur, lane)) {
      int ks, tm, tn;
      decode_item(w, item, ks, tm, tn);
      const int kb0 = ks * w.kb_per_split, kb1 = min(w.kb_total, kb0 + w.kb_per_split);
      const int64_t row = (int64_t)tm * TM + rank * CTA_ROWS + quarter * 32 + lane;
      const int64_t col0 = (int64_t)tn * TN + half * EPI_COLS;
      long long* out = dot + row * w.ldd + col0;           // dot[row, col0 + i]
      const bool row_ok = row < w.nq;
      if (kFused) {
#pragma unroll
        for (int g = 0; g < EPI_COLS / 32; ++g) carried[g] = 0;
      }
      for (int c0 = kb0; c0 < kb1; c0 += w.ck) {
        mbar_wait(tfull0 + 8 * acc, acc_phase);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * TN + half * EPI_COLS);
        uint32_t va[32], vb[32];
        tmem_ld32(taddr, va);
        tmem_ld_wait32(va);
        tmem_ld32(taddr + 32, vb);
        drain32<0>(acc32, va);
        tmem_ld_wait32(vb);
        tmem_ld32(taddr + 64, va);
        drain32<32>(acc32, vb);
        tmem_ld_wait32(va);
        tmem_ld32(taddr + 96, vb);
        drain32<64>(acc32, va);
        tmem_ld_wait32(vb);
        // everything is in registers: hand the TMEM stage back to the MMA warp
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(map_to_cta(tempty0 + 8 * acc, 0));
        drain32<96>(acc32, vb);
        if (++acc == ACC_STAGES) acc = 0, acc_phase ^= 1;
        if (++since_check == CHECK_EVERY) {  // keep |acc32| + 16 more chunks (< 2^24 each) below 2^31
          since_check = 0;
#pragma unroll
          for (int i = 0; i < EPI_COLS; ++i) {
            if ((unsigned)(acc32[i] + 0x70000000) > 0xE0000000u) {
              if (kFused) {
                const bool had = (carried[i >> 5] >> (i & 31)) & 1u;
                scr[i * (CTA_ROWS * 2)] = (had ? scr[i * (CTA_ROWS * 2)] : 0LL) + (long long)acc32[i];
                carried[i >> 5] |= 1u << (i & 31);
              } else if (row_ok && col0 + i < w.nr) {
                red_add_s64(out + i, (long long)acc32[i]);
              }
              acc32[i] = 0;
            }
          }
        }
      }
      since_check = 0;
      if (kFused) {
        // the tile is complete in registers (+ carries): distances, threshold tests, candidate appends
        const long long nq_i = row_ok ? f.norm_q[row] : 0;
        const long long tq = row_ok ? f.tau_q[row * f.tau_q_stride] : -1;   // -1: nothing passes
        const int64_t gid_q = f.q_gid0 + row;
        const bool want_r = f.tau_r != nullptr;
#pragma unroll
        for (int i = 0; i < EPI_COLS; ++i) {
          long long dotv = (long long)acc32[i];
          if ((carried[i >> 5] >> (i & 31)) & 1u) dotv += scr[i * (CTA_ROWS * 2)];
          acc32[i] = 0;
          const int64_t col = col0 + i;
          if (col < w.nr) {  // warp-uniform
            const long long nr_j = __ldg(f.norm_r + col);
            const long long d2 = nq_i + nr_j - 2 * dotv;
            const int64_t gid_r = f.r_gid0 + col;
            const bool live = row_ok && gid_q != gid_r;
            if (live && d2 <= tq) {
              const int slot = atomicAdd(f.cq_cnt + row, 1);
              if (slot < f.cap) f.cq_d2[row * f.cap + slot] = d2, f.cq_id[row * f.cap + slot] = (int)gid_r;
            }
            if (want_r) {
              const long long tr = __ldg(f.tau_r + col * f.tau_r_stride);
              if (live && d2 <= tr) {
                const int slot = atomicAdd(f.cr_cnt + col, 1);
                if (slot < f.cap) f.cr_d2[col * f.cap + slot] = d2, f.cr_id[col * f.cap + slot] = (int)gid_q;
              }
            }
          }
        }
      } else {
#pragma unroll
        for (int i = 0; i < EPI_COLS; ++i) {
          if (row_ok && col0 + i < w.nr && acc32[i] != 0) red_add_s64(out + i, (long long)acc32[i]);
          acc32[i] = 0;
        }
      }
    }
  }
  tc_fence_before();
  cluster_sync_all();  // the peer's TMEM and barriers stay alive until the leader's last MMA/commit has landed
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

// dot[j,i] = dot[i,j] for every element of the 256 x 256 tiles strictly above the tile diagonal (symmetric mode).
// Block = one 32 x 32 patch, transposed through shared memory so both the read and the write are coalesced.
__global__ void __launch_bounds__(256) k_mirror_upper(long long* __restrict__ dot, int64_t n, int64_t ldd) {
  __shared__ long long t[32][33];
  const int bi = blockIdx.y, bj = blockIdx.x;           // 32-row / 32-column patch indices
  if ((bj * 32) / TN <= (bi * 32) / TM) return;         // on or below the tile diagonal: already complete
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int r = ty; r < 32; r += 8) {
    const int64_t i = (int64_t)bi * 32 + r, j = (int64_t)bj * 32 + tx;
    if (i < n && j < n) t[r][tx] = dot[i * ldd + j];
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int64_t j = (int64_t)bj * 32 + r, i = (int64_t)bi * 32 + tx;
    if (i < n && j < n) dot[j * ldd + i] = t[tx][r];
  }
}

// ---- host side ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 3-D view of a k-blocked operand: (bin in block: 64) x (row: rows) x (k-block: nkb); rows beyond `rows` read as 0
static int make_map(CUtensorMap* map, const void* ptr, int64_t rows, int64_t nkb, int64_t slab) {
  EncodeTiledFn enc = get_encode();
  G2_REQUIRE(enc != nullptr, GRIDB200_ENODEV, "cuTensorMapEncodeTiled is not available from this driver");
  const cuuint64_t dims[3] = {(cuuint64_t)BK, (cuuint64_t)rows, (cuuint64_t)nkb};
  const cuuint64_t strides[2] = {(cuuint64_t)BK * 2, (cuuint64_t)slab * BK * 2};
  const cuuint32_t box[3] = {(cuuint32_t)BK, (cuuint32_t)CTA_ROWS, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  G2_REQUIRE(r == CUDA_SUCCESS, GRIDB200_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

}  // namespace tc
}  // namespace g2

// Upper bound on the CTA pairs the Gram kernel occupies (0 = no bound, the default).  A bound selects the static
// schedule: a persistent grid of that many pairs with a round-robin item partition, which needs every pair to be
// co-resident.  Without a bound the kernel uses cluster launch control (one cluster per item, running clusters
// cancel and absorb the ones that have not started), which adapts to whatever SMs other kernels leave free.
static std::atomic<int> g_max_pairs{0};
extern "C" int gridb200_gram_set_max_pairs(int pairs) {
  G2_REQUIRE(pairs >= 0, GRIDB200_EINVAL, "pairs must be >= 0");
  g_max_pairs.store(pairs);
  return 0;
}

// Shared launcher.  fused == nullptr: dot += Q.R^T as exact int64 (gridb200_gram).  fused given: no dot matrix at all,
// one work item per tile over the whole K (the int32 -> int64 carry never leaves the pair), k-best filter in the epilogue.
static int launch_gram(const void* zq, int64_t nq, int64_t slabq, const void* zr, int64_t nr, int64_t slabr,
                       int64_t kdim, int32_t vmax, int64_t* dot, int64_t ldd, int accumulate, void* stream,
                       const g2::tc::Fused* fused) {
  using namespace g2;
  using namespace g2::tc;
  // largest chunk (in 64-wide k-blocks) whose partial sums stay strictly below 2^24
  const int64_t ck = ((1LL << 24) - 1) / ((int64_t)vmax * vmax * BK);
  G2_REQUIRE(ck >= 1, GRIDB200_EDOMAIN,
             "vmax=%d: one 64-wide k-block can exceed 2^24 in fp32; the tensor-core path needs vmax <= 511 (use impl=1)",
             vmax);
  G2_REQUIRE(nq < (1LL << 31) && nr < (1LL << 31) && kdim < (1LL << 31), GRIDB200_EINVAL, "shape exceeds int32");
  cudaStream_t st = as_stream(stream);
  // SM count and the >48 KB shared-memory opt-in are per DEVICE: cached per device index, not per process
  static std::atomic<int> sms_of[64];
  int dev = 0;
  G2_CUDA(cudaGetDevice(&dev));
  G2_REQUIRE(dev >= 0 && dev < 64, GRIDB200_ENODEV, "device index %d outside the per-device cache", dev);
  int sms = sms_of[dev].load();
  if (!sms) {
    cudaDeviceProp prop;
    G2_CUDA(cudaGetDeviceProperties(&prop, dev));
    G2_REQUIRE(prop.major == 10, GRIDB200_ENODEV, "tcgen05 path needs an sm_100 device (found sm_%d%d)", prop.major,
               prop.minor);
    G2_CUDA(cudaFuncSetAttribute(k_gram_tc_pair<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    G2_CUDA(cudaFuncSetAttribute(k_gram_tc_pair<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    G2_CUDA(cudaFuncSetAttribute(k_gram_tc_pair<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    G2_CUDA(cudaFuncSetAttribute(k_gram_tc_pair<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    G2_REQUIRE(prop.multiProcessorCount <= PACE_SLOTS, GRIDB200_ENODEV, "more SMs than pacing slots");
    sms = prop.multiProcessorCount;
    sms_of[dev].store(sms);
  }
  const int64_t nkb = kdim / BK;
  CUtensorMap mq, mr;
  if (int e = make_map(&mq, zq, nq, nkb, slabq)) return e;
  if (int e = make_map(&mr, zr, nr, nkb, slabr)) return e;

  static int no_sym = -1, no_pace = -1, static_sched = 0;  // debugging switches
  if (no_sym < 0) {
    const char* sch = getenv("GRIDB200_GRAM_SCHED");  // "static": round-robin item partition instead of work stealing
    static_sched = (sch && sch[0] == 's') ? 1 : 0;
    const char* e = getenv("GRIDB200_GRAM_NOSYM");
    no_sym = (e && e[0] == '1') ? 1 : 0;
    e = getenv("GRIDB200_GRAM_NOPACE");
    no_pace = (e && e[0] == '1') ? 1 : 0;
  }
  int workers = std::min(sms / 2, MAX_PAIRS);
  const int cap = g_max_pairs.load();
  if (cap) workers = std::max(1, std::min(workers, cap));
  const bool dynamic = !static_sched && !cap;
  Work w;
  w.nq = nq, w.nr = nr, w.ldd = ldd;
  w.symmetric = (!fused && !no_sym && zq == zr && nq == nr && slabq == slabr) ? 1 : 0;
  w.tiles_m = (int)((nq + TM - 1) / TM);
  w.tiles_n = (int)((nr + TN - 1) / TN);
  w.ntiles = w.symmetric ? w.tiles_m * (w.tiles_m + 1) / 2 : w.tiles_m * w.tiles_n;
  w.kb_total = (int)nkb;
  w.ck = (int)std::min<int64_t>(ck, w.kb_total);
  static int group_m_env = -1;  // GRIDB200_GRAM_GROUP_M: tile rows per band of the general (non-symmetric) walk
  if (group_m_env < 0) {
    const char* e = getenv("GRIDB200_GRAM_GROUP_M");
    group_m_env = (e && atoi(e) > 0) ? atoi(e) : 0;
  }
  w.group_m = group_m_env ? group_m_env : 4;
  // K splits: the shortest launch under the model  rounds(items over the pairs) x (k-blocks per split + ITEM_COST).
  // ITEM_COST (in k-block times, ~0.3 us each) is what an item costs beyond its MMAs: 128 uncoalesced 64-bit
  // reductions per epilogue thread during which the next item's MMAs run out of TMEM stages, the pipeline refill and
  // the pacing hand-over -- about 60 us, measured as the intercept of launch time against K (3.3 ms at K/8 where
  // 19.4/8 = 2.4 ms would be proportional).  Wave efficiency alone picked 13 splits for the 91 tiles of N = 3202
  // (1183 items); with the item cost 4 splits win: -20 % at K = 326 k bins (one rank's share at 8 GPUs).
  constexpr int ITEM_COST = 192;
  const int chunks = (w.kb_total + w.ck - 1) / w.ck;
  int best_s = 1;
  double best_cost = 0.0;
  for (int s = 1; !fused && s <= 64 && s * 4 <= std::max(chunks, 4); ++s) {
    const int cps = (chunks + s - 1) / s;          // chunks per split
    const int real_s = (chunks + cps - 1) / cps;   // splits actually needed
    const long long items = (long long)w.ntiles * real_s;
    const double cost = (double)((items + workers - 1) / workers) * (double)(cps * w.ck + ITEM_COST);
    if (s == 1 || cost < best_cost - 1e-9) best_cost = cost, best_s = real_s;
  }
  static int splits_env = -1;  // GRIDB200_GRAM_SPLITS: force the number of K splits (measurement switch)
  if (splits_env < 0) {
    const char* e = getenv("GRIDB200_GRAM_SPLITS");
    splits_env = (e && atoi(e) > 0) ? atoi(e) : 0;
  }
  if (splits_env && !fused) best_s = std::max(1, std::min(splits_env, chunks));
  const int cps = (chunks + best_s - 1) / best_s;
  w.kb_per_split = cps * w.ck;
  w.ksplits = (w.kb_total + w.kb_per_split - 1) / w.kb_per_split;
  // pacing window: the operand rows one wave touches, times `window` k-blocks, should stay near 32 MB
  {
    const int64_t wave_rows = w.symmetric ? nq
                                          : std::min<int64_t>(nq, (int64_t)w.group_m * TM) +
                                                std::min<int64_t>(nr, (int64_t)(workers / w.group_m + 1) * TN);
    static int64_t win_mb = 0;  // GRIDB200_GRAM_WINDOW_MB: operand bytes the pairs of one wave may be spread over
    if (!win_mb) {
      const char* e = getenv("GRIDB200_GRAM_WINDOW_MB");
      win_mb = (e && atoi(e) > 0) ? atoi(e) : 32;
    }
    int64_t win = (win_mb << 20) / (wave_rows * BK * 2);
    win = std::max<int64_t>(2 * PACE, std::min<int64_t>(win, 16 * PACE)) / PACE * PACE;
    w.window = no_pace ? 0 : (int)win;
  }
  static std::atomic<unsigned> epoch{0};
  w.epoch = epoch.fetch_add(1) + 1;
  if (!fused && !accumulate) G2_CUDA(cudaMemset2DAsync(dot, ldd * sizeof(int64_t), 0, nr * sizeof(int64_t), nq, st));
  const long long items = (long long)w.ntiles * w.ksplits;
  G2_REQUIRE(items < (1LL << 30), GRIDB200_EINVAL, "too many work items");
  const int nwork = (int)std::min<long long>(items, workers);
  w.pace_slots = dynamic ? sms : nwork;
  long long* dotp = reinterpret_cast<long long*>(dot);
  const Fused f0{};
  // dynamic: one cluster per item; the clusters that get SMs cancel the rest and take their items over
  if (fused && dynamic) k_gram_tc_pair<true, true><<<2 * (unsigned)items, NUM_THREADS, SMEM_BYTES, st>>>(mq, mr, w, dotp, *fused);
  else if (fused) k_gram_tc_pair<false, true><<<2 * nwork, NUM_THREADS, SMEM_BYTES, st>>>(mq, mr, w, dotp, *fused);
  else if (dynamic) k_gram_tc_pair<true, false><<<2 * (unsigned)items, NUM_THREADS, SMEM_BYTES, st>>>(mq, mr, w, dotp, f0);
  else k_gram_tc_pair<false, false><<<2 * nwork, NUM_THREADS, SMEM_BYTES, st>>>(mq, mr, w, dotp, f0);
  G2_LAUNCHED();
  if (w.symmetric && w.tiles_m > 1) {
    const unsigned nb = (unsigned)((nq + 31) / 32);
    k_mirror_upper<<<dim3(nb, nb), 256, 0, st>>>(reinterpret_cast<long long*>(dot), nq, ldd);
    G2_LAUNCHED();
  }
  return 0;
}

int gridb200_gram_tc(const void* zq, int64_t nq, int64_t slabq, const void* zr, int64_t nr, int64_t slabr,
                     int64_t kdim, int32_t vmax, int64_t* dot, int64_t ldd, int accumulate, void* stream) {
  return launch_gram(zq, nq, slabq, zr, nr, slabr, kdim, vmax, dot, ldd, accumulate, stream, nullptr);
}

extern "C" size_t gridb200_gram_topk_ws_bytes(void) {
  return (size_t)g2::tc::PACE_SLOTS * g2::tc::CTA_ROWS * g2::tc::TN * sizeof(long long);  // one carry slab per SM
}

extern "C" int gridb200_gram_topk(const void* zq, int64_t nq, int64_t slabq, const void* zr, int64_t nr, int64_t slabr,
                                  int64_t kdim, int32_t vmax, const int64_t* normq, const int64_t* normr, int64_t q_gid0,
                                  int64_t r_gid0, const int64_t* tau_q, int64_t tau_q_stride, const int64_t* tau_r,
                                  int64_t tau_r_stride, int32_t cap, int64_t* cq_d2, int32_t* cq_idx, int32_t* cq_cnt,
                                  int64_t* cr_d2, int32_t* cr_idx, int32_t* cr_cnt, void* ws, size_t ws_bytes,
                                  void* stream) {
  using namespace g2;
  G2_REQUIRE(zq && zr && normq && normr && tau_q && cq_d2 && cq_idx && cq_cnt && ws, GRIDB200_EINVAL, "null pointer");
  G2_REQUIRE(!tau_r || (cr_d2 && cr_idx && cr_cnt), GRIDB200_EINVAL, "tau_r given without candidate buffers for the reference rows");
  G2_REQUIRE(nq > 0 && nr > 0 && kdim > 0 && kdim % 64 == 0 && slabq >= nq && slabr >= nr, GRIDB200_EINVAL, "bad shape");
  G2_REQUIRE(cap > 0 && tau_q_stride > 0 && (!tau_r || tau_r_stride > 0), GRIDB200_EINVAL, "bad capacity / stride");
  G2_REQUIRE(ws_bytes >= gridb200_gram_topk_ws_bytes(), GRIDB200_EINVAL, "workspace too small");
  G2_REQUIRE(((reinterpret_cast<uintptr_t>(zq) | reinterpret_cast<uintptr_t>(zr)) & 15) == 0, GRIDB200_EINVAL,
             "operands must be 16-byte aligned");
  G2_REQUIRE(q_gid0 + nq < (1LL << 31) && r_gid0 + nr < (1LL << 31), GRIDB200_EINVAL, "indices exceed int32");
  tc::Fused f{};
  f.norm_q = reinterpret_cast<const long long*>(normq), f.norm_r = reinterpret_cast<const long long*>(normr);
  f.tau_q = reinterpret_cast<const long long*>(tau_q), f.tau_r = reinterpret_cast<const long long*>(tau_r);
  f.tau_q_stride = tau_q_stride, f.tau_r_stride = tau_r_stride;
  f.cq_d2 = reinterpret_cast<long long*>(cq_d2), f.cq_id = cq_idx, f.cq_cnt = cq_cnt;
  f.cr_d2 = reinterpret_cast<long long*>(cr_d2), f.cr_id = cr_idx, f.cr_cnt = cr_cnt;
  f.cap = cap, f.q_gid0 = q_gid0, f.r_gid0 = r_gid0, f.scratch = reinterpret_cast<long long*>(ws);
  return launch_gram(zq, nq, slabq, zr, nr, slabr, kdim, vmax, nullptr, 0, 1, stream, &f);
}
