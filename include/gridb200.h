/*
 * gridb200.h -- C ABI of libgridb200.so: the B200 (sm_100a) implementation of GRiD's depth-inference
 * hot path (reference steps 4-6: normalize_mosdepth -> find_neighbors -> compute_dipcn).
 *
 * The reference (caterer-z-t/GRiD) is pure Python and has no FFI; the boundary this library replaces is the
 * body of three Python step functions.  Each entry point below cites the reference lines whose arithmetic
 * it takes over (paths relative to /root/reference/grid/utils/).  The Python host in grid_b200/ keeps the
 * reference's function names, config keys and file formats and calls these through ctypes; see
 * INTEGRATION.md for the binding a reference maintainer would add.
 *
 * Conventions
 *   - every function returns 0 on success, a negative GRIDB200_E* code on failure; the message is available
 *     from gridb200_last_error() (thread-local, library-owned, valid until the next failing call).
 *   - all pointers are DEVICE pointers unless the name ends in _host; the caller owns every buffer; the
 *     library allocates nothing on the device; scratch space is passed in as `ws` (size from *_ws_bytes).
 *   - `stream` is a cudaStream_t passed as void*; all work is asynchronous on that stream.
 *   - matrices are row-major with an explicit leading dimension in ELEMENTS; rows are samples.
 *   - depth matrix: fp32, 0.0f = bin absent for that sample (the reference keeps only depth > 0,
 *     normalize_mosdepth.py:274-278), values are multiples of 0.01 below 41943.04 (mosdepth prints %.2f).
 *   - "hundredths": integer n standing for the decimal n/100, i.e. the value a %.2f text field carries.
 *   - no C++ exception crosses this boundary.
 */
#ifndef GRIDB200_H
#define GRIDB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define GRIDB200_API __attribute__((visibility("default")))
#else
#define GRIDB200_API
#endif

#define GRIDB200_OK 0
#define GRIDB200_EINVAL (-1)   /* bad argument (null pointer, misaligned leading dimension, size out of range) */
#define GRIDB200_ECUDA (-2)    /* CUDA runtime / driver error */
#define GRIDB200_EDOMAIN (-3)  /* input outside the documented exact domain */
#define GRIDB200_ENODEV (-4)   /* no sm_100 device / driver entry point unavailable */

#define GRIDB200_ZQ_NA ((int16_t)-32768)       /* "NA" cell in a hundredths matrix */
#define GRIDB200_ZQ_NEG_ZERO ((int16_t)-32767) /* negative value that prints as "-0.00"; numerically 0 */
#define GRIDB200_ZQ_LIMIT 32766                /* |hundredths| saturates here (counted in *sat) */

GRIDB200_API const char* gridb200_last_error(void);
GRIDB200_API const char* gridb200_version(void);
/* Number of kernels this library has launched in the calling process (bench.py's gpu_launches). */
GRIDB200_API uint64_t gridb200_launch_count(void);

/* ------------------------------------------------------------------ step 4: normalize_mosdepth ----------- */

/* Per-bin raw sums over the local rows: colsum[j] = sum_i round(100*depth[i,j]), colcnt[j] = #{i: depth>0}.
 * Replaces pass 1, compute_population_mean_depths (normalize_mosdepth.py:218-301) once the text is parsed.
 * Buffers are overwritten.  Row-sharded multi-GPU: all-reduce(sum) both outputs afterwards. */
GRIDB200_API int gridb200_col_raw_stats(const float* depth, int64_t n, int64_t b, int64_t ld,
                           int64_t* colsum, int32_t* colcnt, void* stream);

/* gridb200_col_raw_stats plus, from the same pass over the matrix, the per-sample totals over ALL bins
 * (rowsum_all[i] = sum_j round(100*depth[i,j]), rowcnt_all[i] = #present); feeds gridb200_row_stats_fused. */
GRIDB200_API int gridb200_col_row_raw_stats(const float* depth, int64_t n, int64_t b, int64_t ld,
                               int64_t* colsum, int32_t* colcnt, int64_t* rowsum_all, int32_t* rowcnt_all,
                               void* stream);

/* valid[j] = 1 iff colcnt>0 and min_depth <= colsum/(100*colcnt) <= max_depth (normalize_mosdepth.py:81-83). */
GRIDB200_API int gridb200_depth_filter(const int64_t* colsum, const int32_t* colcnt, int64_t b,
                          double min_depth, double max_depth, uint8_t* valid, void* stream);

/* Per-sample sums over the valid bins: rowsum[i] = sum_j round(100*depth[i,j]), rowcnt[i] = #present.
 * This is np.nanmean(mat, axis=1) (normalize_mosdepth.py:120 and :440) in exact integer form.
 * Column-sharded multi-GPU: all-reduce(sum) both outputs afterwards. */
GRIDB200_API int gridb200_row_stats(const float* depth, int64_t n, int64_t b, int64_t ld, const uint8_t* valid,
                       int64_t* rowsum, int32_t* rowcnt, void* stream);

/* Same result as gridb200_row_stats, derived from the totals of gridb200_col_row_raw_stats: the bins that have
 * data but failed the depth filter are subtracted column by column.  When those bins are more than 1/8 of all
 * bins (a strided column read costs a 32-byte sector per cell) the full pass runs instead; the choice is made
 * on the device, without a host round trip.  ws: gridb200_row_stats_fused_ws_bytes(b). */
GRIDB200_API size_t gridb200_row_stats_fused_ws_bytes(int64_t b);
GRIDB200_API int gridb200_row_stats_fused(const float* depth, int64_t n, int64_t b, int64_t ld, const uint8_t* valid,
                             const int32_t* colcnt, const int64_t* rowsum_all, const int32_t* rowcnt_all,
                             int64_t* rowsum, int32_t* rowcnt, void* ws, size_t ws_bytes, void* stream);

/* raw_mean[i] = rowsum/(100*rowcnt) (the "scale" column, NaN for an empty row); scale_q[i] = that value
 * after the %.2f text hand-off (normalize_mosdepth.py:552-554 -> find_neighbors.py:111); row_factor[i] =
 * rowcnt/rowsum, so that depth_ij/rowmean_i == hundredths_ij * row_factor[i] (normalize_mosdepth.py:440-442).
 * Any output may be null. */
GRIDB200_API int gridb200_row_scales(const int64_t* rowsum, const int32_t* rowcnt, int64_t n,
                        double* raw_mean, double* scale_q, double* row_factor, void* stream);

/* After gridb200_row_scales: rows whose exact mean lies exactly on a %.2f tie (k + 0.005) get numpy's own value --
 * np.nanmean's pairwise floating-point sum over the row's valid bins, replayed operation by operation -- because there
 * the last ulp decides which way the reference's "%.2f" goes (normalize_mosdepth.py:120, :548).  No other row can
 * differ: a non-tie mean is at least 1/(2*count) hundredths from a tie.  Needs the whole row (all bins) on this GPU. */
GRIDB200_API int gridb200_row_means_exact(const float* depth, int64_t n, int64_t b, int64_t ld, const uint8_t* valid,
                             const int64_t* rowsum, const int32_t* rowcnt, double* raw_mean, double* scale_q,
                             void* stream);

/* Per-bin sums of the row-normalised matrix y = depth/rowmean, shifted by 1: sd[j] = sum_i (y-1),
 * sd2[j] = sum_i (y-1)^2 over present cells of valid bins (normalize_mosdepth.py:445-446 in one pass).
 * ws: gridb200_col_norm_stats_ws_bytes(n,b).  Row-sharded multi-GPU: all-reduce(sum) sd and sd2. */
GRIDB200_API size_t gridb200_col_norm_stats_ws_bytes(int64_t n, int64_t b);
GRIDB200_API int gridb200_col_norm_stats(const float* depth, int64_t n, int64_t b, int64_t ld, const uint8_t* valid,
                            const double* row_factor,
                            double* sd, double* sd2, void* ws, size_t ws_bytes, void* stream);

/* mu = 1 + sd/cnt; s2 = (sd2 - sd^2/cnt)/(n_total-1); ratio = 100*s2/mu where mu > 0 else NaN; NaN for
 * invalid bins (normalize_mosdepth.py:445-451).  n_total = number of sample rows in the whole cohort. */
GRIDB200_API int gridb200_col_finalize(const double* sd, const double* sd2, const int32_t* colcnt, const uint8_t* valid,
                          int64_t b, int64_t n_total, double* mu, double* s2, double* ratio, void* stream);

/* Median rescale and variance selection on the FULL ratio vector (normalize_mosdepth.py:460-468, :479-499):
 *   zscale = 1/sqrt(median(finite ratio)/100) if that median > 0 else 1
 *   thr = sorted(finite)[int(top_frac*count)]; sel_pos[j] = rank of j among {ratio > thr} else -1
 * info (device, 4 doubles): [0]=zscale [1]=thr [2]=count of finite ratios [3]=number selected (Rw).
 * status GRIDB200_EDOMAIN is reported through info[3] = -1 when int(top_frac*count) == count (the reference
 * raises IndexError there).  ws: gridb200_select_ws_bytes(b). */
GRIDB200_API size_t gridb200_select_ws_bytes(int64_t b);
GRIDB200_API int gridb200_select(const double* ratio, int64_t b, double top_frac, int32_t* sel_pos, double* info,
                    void* ws, size_t ws_bytes, void* stream);

/* gridb200_select without a workspace argument.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  The scratch (selection state, radix histograms, scan-tile sums: about 33 KB) is a file-scope
 * `static __device__` array inside the library, held by every GPU that loads it until exit; the entry point calls no
 * cudaMalloc / cudaFree / allocating routine.  Same inputs, outputs and arithmetic as gridb200_select.
 * Limit: b <= GRIDB200_SELECT_STATIC_MAX_BINS (larger -> GRIDB200_EINVAL with a message naming the limit).
 * Every call on a GPU shares the one array: calls that overlap on the GPU (different streams) race; serialise them. */
#define GRIDB200_SELECT_STATIC_MAX_BINS 4194304
GRIDB200_API int gridb200_select_static(const double* ratio, int64_t b, double top_frac, int32_t* sel_pos, double* info,
                           void* stream);

/* find_neighbors' region filter on the %.3f-rounded ratios of the selected bins (find_neighbors.py:128-175):
 * keep[j] = 1 iff sel_pos[j] >= 0 and finite(rq) and rq >= sigma2_min and rq <= sigma2_max, where
 * rq = float("%.3f" % ratio[j]) and sigma2_min = sorted(finite rq)[min(int(Rw*(1-frac_r)), n_finite-1)];
 * no finite rq -> every selected bin is kept.  `info` is gridb200_select's output (Rw is read on the device);
 * info2 (device, 2 doubles): [0]=R_use [1]=sigma2_min.  ws: gridb200_select_ws_bytes(b). */
GRIDB200_API int gridb200_knn_col_filter(const double* ratio, const int32_t* sel_pos, int64_t b, double frac_r,
                            double sigma2_max, uint8_t* keep, const double* info, double* info2,
                            void* ws, size_t ws_bytes, void* stream);

/* The transform pass (normalize_mosdepth.py:454-470 + the %.2f hand-off :552-554 + find_neighbors.py:57-58):
 *   z = ((depth/rowmean) - mu_j) / sqrt(mu_j) * zscale          fp64, then printf-exact rounding to hundredths
 *   zq[i, sel_pos[j]] = hundredths(z)  (GRIDB200_ZQ_NA where absent, GRIDB200_ZQ_NEG_ZERO for "-0.00";
 *                                       saturates at +-GRIDB200_ZQ_LIMIT, counted in *sat)
 *   zh(i, sel_pos[j]) = fp16(clip(hundredths(z), +-zclip))  with NA -> 0 and 0 where keep[j] == 0
 *   norms[i] += sum_j zh(i,j)^2 (exact int64; caller zeroes it)
 * zq and zh are optional (either may be null).  zq is row-major with ldq in elements.  zh is the K-BLOCKED
 * operand of gridb200_gram: element (row i, compact bin p) lives at zh[((p/64)*zh_slab + i)*64 + p%64], i.e.
 * each 64-bin k-block is one contiguous slab of zh_slab (>= n) rows x 128 bytes; the pad bins [Rw, ceil64(Rw))
 * of the last k-block must be zeroed by the caller.  `info` is gridb200_select's output (zscale is read on
 * the device). */
GRIDB200_API int gridb200_transform(const float* depth, int64_t n, int64_t b, int64_t ld,
                       const double* row_factor, const double* mu,
                       const int32_t* sel_pos, const uint8_t* keep, const double* info, int32_t zclip,
                       int16_t* zq, int64_t ldq, void* zh, int64_t zh_slab, int64_t* norms,
                       unsigned long long* sat, void* stream);

/* ------------------------------------------------------------------ step 5: find_neighbors --------------- */

/* zq (as read back from a step-4 file) -> tensor-core operand: clip to +-zclip, NA -> 0, zero dropped bins,
 * fp16 in the k-blocked layout described at gridb200_transform (all ceil64(rw) bins are written, pad = 0);
 * norms[i] = exact sum of squares (overwritten) (find_neighbors.py:57-65). keep may be null. */
GRIDB200_API int gridb200_knn_prepare(const int16_t* zq, int64_t n, int64_t rw, int64_t ldq, const uint8_t* keep,
                         int32_t zclip, void* zh, int64_t zh_slab, int64_t* norms, void* stream);

/* Exact Gram block: dot[i,j] (+)= sum_k q[i,k]*r[j,k] over kdim bins, operands fp16 holding integers with
 * |v| <= vmax (<= 2048).  This is the -2*X.Y^T term of sklearn's EuclideanArgKmin64, which the reference
 * reaches through NearestNeighbors.kneighbors (find_neighbors.py:207-213).
 *   impl 0: tcgen05 tensor cores + TMA (sm_100a); fp32 TMEM accumulation over chunks short enough to stay
 *           below 2^24 (exact), drained into int64.        impl 1: plain SIMT integer kernel (cross-check).
 * Operands are K-BLOCKED (see gridb200_transform): element (i,k) at base[((k/64)*slab + i)*64 + k%64]; a block
 * of rows starting at row q0 of a larger operand is passed as base + q0*64 with the same slab.  Base pointers
 * 16-byte aligned, kdim a multiple of 64 (zero padded), slabq >= nq, slabr >= nr.
 * When both operands are the same matrix (same pointer, rows and slab) impl 0 computes only the tiles on or
 * above the diagonal and mirrors them.  accumulate != 0 adds to dot (bin-sharded multi-GPU, K-blocked calls). */
GRIDB200_API int gridb200_gram(const void* zq_f16, int64_t nq, int64_t slabq, const void* zr_f16, int64_t nr, int64_t slabr,
                  int64_t kdim, int32_t vmax, int64_t* dot, int64_t ldd, int accumulate, int impl,
                  void* stream);

/* Cap the number of CTA pairs (2 SMs each) the tensor-core Gram kernel occupies; 0 = all SMs (default).  The kernel
 * is persistent with a static work partition, so a caller that overlaps it with another kernel (the row-sharded ring
 * overlaps it with NCCL's send/recv kernel) must leave that kernel its SMs.  Process-wide. */
GRIDB200_API int gridb200_gram_set_max_pairs(int pairs);

/* d2[i,j] = normq[i] + normr[j] - 2*dot[i,j]; the row's own sample (q_gid0+i == r_gid0+j) is skipped
 * (find_neighbors.py:218-219); merge into the running k-best lists, ascending by (d2, global index) -- the
 * order sklearn's heap + sort yields when no two candidates tie, lower index first among exact ties.
 * best_d2 [nq,k] / best_idx [nq,k] must be initialised with gridb200_topk_init before the first block.
 * nr + k <= 8192 per call.  Candidates are filtered against the list's current k-th entry before anything is sorted. */
GRIDB200_API int gridb200_topk_init(int64_t* best_d2, int32_t* best_idx, int64_t nq, int32_t k, void* stream);
GRIDB200_API int gridb200_topk_merge(const int64_t* dot, int64_t ldd, int64_t nq, int64_t nr,
                        const int64_t* normq, const int64_t* normr, int64_t q_gid0, int64_t r_gid0,
                        int32_t k, int64_t* best_d2, int32_t* best_idx, void* stream);
/* The same block read through its TRANSPOSE: the lists best_d2/best_idx [nr,k] belong to the nr REFERENCE rows (global
 * ids r_gid0..) and the nq query rows (q_gid0..) are the candidates.  The squared distance is symmetric, so a rank that
 * has multiplied its rows with a peer's shard can hand the peer its k best instead of the peer recomputing the block
 * (the half ring of grid_b200/ring.py; same find_neighbors.py:207-227 semantics per list).  nq + k <= 8192 per call. */
GRIDB200_API int gridb200_topk_merge_t(const int64_t* dot, int64_t ldd, int64_t nq, int64_t nr,
                          const int64_t* normq, const int64_t* normr, int64_t q_gid0, int64_t r_gid0,
                          int32_t k, int64_t* best_d2, int32_t* best_idx, void* stream);
/* Merge candidate lists src_d2/src_idx [n, src_k] (entries with idx < 0 are empty; no sample may already be in the
 * destination list of its row) into the running lists best_d2/best_idx [n, k].  src_k + k <= 8192. */
GRIDB200_API int gridb200_topk_merge_lists(const int64_t* src_d2, const int32_t* src_idx, int64_t n, int32_t src_k,
                              int32_t k, int64_t* best_d2, int32_t* best_idx, void* stream);

/* The Gram block and the k-best filter in ONE kernel: the distances never reach memory.  Same operands and exact
 * arithmetic as gridb200_gram (impl 0; one work item per 256 x 256 tile over the whole K, the int32 -> int64 carry stays
 * with the CTA pair), but a finished tile forms d2 = normq[i] + normr[j] - 2*dot in int64 in the epilogue and appends
 * (d2, id) only where it can still enter a list: to the candidate buffer of query row i when d2 <= tau_q[i*tau_q_stride]
 * (tau = the row's current k-th best d2; +inf takes everything), and -- when tau_r is given: the half ring reads every
 * block both ways -- to that of reference row j when d2 <= tau_r[j*tau_r_stride].  The row's own sample is skipped by
 * global id.  This is sklearn's ArgKmin keeping per-row heaps while it walks the distance chunks, instead of a stored
 * N x N matrix (find_neighbors.py:207-213).  cq_* / cr_*: [rows, cap] buffers + per-row counters (zeroed by the caller;
 * a counter above cap means candidates were dropped: the caller must redo that block through gridb200_gram +
 * gridb200_topk_merge).  gridb200_topk_merge_cand then merges a buffer into the running lists.  ws: carry slabs,
 * gridb200_gram_topk_ws_bytes() bytes.  vmax <= 511. */
GRIDB200_API size_t gridb200_gram_topk_ws_bytes(void);
GRIDB200_API int gridb200_gram_topk(const void* zq_f16, int64_t nq, int64_t slabq, const void* zr_f16, int64_t nr,
                       int64_t slabr, int64_t kdim, int32_t vmax, const int64_t* normq, const int64_t* normr,
                       int64_t q_gid0, int64_t r_gid0, const int64_t* tau_q, int64_t tau_q_stride,
                       const int64_t* tau_r, int64_t tau_r_stride, int32_t cap,
                       int64_t* cq_d2, int32_t* cq_idx, int32_t* cq_cnt,
                       int64_t* cr_d2, int32_t* cr_idx, int32_t* cr_cnt, void* ws, size_t ws_bytes, void* stream);
/* Merge the first min(cand_cnt[i], cap) entries of each row's candidate buffer [n, cap] into the running lists [n, k]
 * (cap + k <= 8192). */
GRIDB200_API int gridb200_topk_merge_cand(const int64_t* cand_d2, const int32_t* cand_idx, const int32_t* cand_cnt,
                             int64_t n, int32_t cap, int32_t k, int64_t* best_d2, int32_t* best_idx, void* stream);

/* ---- general fp32 path (north_star's 3xTF32): unquantised operands ----------------------------------------------
 * The pipeline's operands are quantised by the %.2f hand-off and take the exact fp16 kernel above.  A caller that holds
 * UNROUNDED fp32 z-scores (the matrix find_neighbors.py:57-65 would feed sklearn without the text round trip) uses these:
 * gridb200_split_tf32 splits a row-major fp32 matrix x[n, ldx] into tf32 hi and lo parts (x = hi + lo exactly, hi has
 * 10 mantissa bits) in the k-blocked layout [kdim/32, slab, 32] (kdim = k rounded up to 32, zero padded) and returns the
 * fp64 row norms; gridb200_gram_f32 forms dot = Qhi.Rhi + Qhi.Rlo + Qlo.Rhi on tcgen05 (kind::tf32, fp32 TMEM
 * accumulators, chunked summation) -- the -2*X.Y^T term of EuclideanArgKmin at fp32 accuracy; gridb200_topk_merge_f32 forms
 * d2 = normq[i] + normr[j] - 2*dot in fp64 (clamped at 0) and merges into the running lists, whose best_d2 entries then
 * hold the BIT PATTERNS of those doubles (a non-negative double orders like its bits read as int64). */
GRIDB200_API int gridb200_split_tf32(const float* x, int64_t n, int64_t k, int64_t ldx, float* hi, float* lo,
                        int64_t slab, double* norms, void* stream);
GRIDB200_API int gridb200_gram_f32(const float* qhi, const float* qlo, int64_t nq, int64_t slabq,
                      const float* rhi, const float* rlo, int64_t nr, int64_t slabr, int64_t kdim,
                      float* dot, int64_t ldd, void* stream);
GRIDB200_API int gridb200_topk_merge_f32(const float* dot, int64_t ldd, int64_t nq, int64_t nr,
                            const double* normq, const double* normr, int64_t q_gid0, int64_t r_gid0,
                            int32_t k, int64_t* best_d2, int32_t* best_idx, void* stream);

/* ------------------------------------------------------------------ step 6: compute_dipcn ---------------- */

/* out[i,g] = (counts[q0+i,g]/scale[q0+i]) / mean over contributing neighbours j of counts[j,g]/scale[j]; NaN where
 * the reference omits the sample.  counts [n_all, g] int32 (read counts are integers and stay so), scale_q [n_all] =
 * the %.2f-rounded scales, nbr_idx [nq,k] global indices (-1 = empty slot).  t_ws: n_all*g doubles of scratch.
 * rule = which of the reference's two skip rules applies:
 *   GRIDB200_DIPCN_PIPELINE  compute_dipcn.py:62-88 (what `grid wgs` runs): the first n_nbr neighbours THAT HAVE a
 *                            count contribute (absent ones do not use up the budget); query omitted without a count
 *                            or without a contributor.
 *   GRIDB200_DIPCN_EXON      compute_dipcn_dir/compute_diploid_cn.py:8-63: the list is cut to n_nbr entries first;
 *                            per region only neighbours with count > 0 and scale > 0 contribute; query omitted on a
 *                            zero count, a non-positive scale or no contributor. */
#define GRIDB200_DIPCN_PIPELINE 0
#define GRIDB200_DIPCN_EXON 1
GRIDB200_API int gridb200_dipcn(const int32_t* counts, int64_t n_all, int64_t g, int64_t ldc, const uint8_t* has_count,
                   const double* scale_q, const int32_t* nbr_idx, int64_t nq, int32_t k, int64_t q_gid0,
                   int32_t n_nbr, int32_t rule, double* out, int64_t ldo, double* t_ws, void* stream);

/* ------------------------------------------------------------------ synthetic cohorts (bench / tests) ---- */

/* Device twin of grid_b200/synth.py:depth_hundredths -- fills depth[n, b] (fp32, ld) for global rows
 * row0.. and bins col0..; c/rc are per-row, bcol/rb per-bin parameter vectors (computed by the host). */
GRIDB200_API int gridb200_synth_depth(uint64_t seed, int64_t n, int64_t b, int64_t row0, int64_t col0,
                         const double* c, const double* rc, const double* bcol, const double* rb,
                         float* depth, int64_t ld, void* stream);

/* ------------------------------------------------------------------ ingest (mosdepth text) --------------- */

/* Parse one sample's decompressed `chrom\tstart\tend\tdepth\n` text on the device and write row `depth_row`
 * [b] of the packed matrix (the line filters of normalize_mosdepth.py:262-285/:321-352): chromosome PREFIX match
 * (chrom_len = 0 matches all), >= 4 tab fields, depth > 0, bin touching the closed window [win_lo, win_hi]
 * (win_lo < 0: no window), no masked kb in [start/1000, end/1000] (mask_bits: bit kb set = masked, mask_kb bits,
 * may be null).  Column = (start - bin0) / 1000 for bins with start aligned to bin0 and 0 < end - start <= 1000;
 * a bin repeated in the file keeps its LAST depth (:285).  Bins never seen are written as 0.0f (absent).
 * stats (device, 8 x int64): [0]=kept lines [1]=irregular lines (outside the strict grammar digits[.dd] /
 * unaligned / out of range: skipped here -- the caller re-parses such a file on the host to keep the
 * reference's Python int()/float() semantics) [2]=lines matching the prefix [3]=masked lines [4]=lines whose
 * chromosome field is longer than the prefix (the reference's startswith quirk: with a mask the caller must
 * check these, because the mask is looked up by the line's own chromosome).
 * ws: gridb200_ingest_ws_bytes(b). */
GRIDB200_API size_t gridb200_ingest_ws_bytes(int64_t b);
GRIDB200_API int gridb200_ingest_text(const char* text, int64_t nbytes, const char* chrom_prefix, int32_t chrom_len,
                         int64_t win_lo, int64_t win_hi, const uint32_t* mask_bits, int64_t mask_kb,
                         int64_t bin0, int64_t b, float* depth_row, int64_t* stats,
                         void* ws, size_t ws_bytes, void* stream);

/* Packed host format of the depth matrix: uint16 hundredths (0 = absent), 2 bytes per cell over PCIe instead of 4.
 * A cell that does not fit (depth >= 655.35) holds 0xFFFF and its value travels in an escape list
 * (esc_cell = global_row * b + bin, esc_val = the fp32 depth).  pack: device fp32 rows [row0, row0+n) -> codes;
 * *esc_count (device, caller-zeroed, accumulates across calls) may exceed esc_cap -- then re-run with larger arrays.
 * unpack: codes -> the fp32 matrix the other entry points take, bit-identical to ingesting the text
 * (h / 100 correctly rounded), then escape cells of these rows are overwritten.  n <= 65535 rows per call. */
GRIDB200_API int gridb200_depth_pack_u16(const float* depth, int64_t n, int64_t b, int64_t ld, uint16_t* out, int64_t ldo,
                            int64_t row0, int64_t* esc_cell, float* esc_val, int64_t esc_cap, int64_t* esc_count,
                            void* stream);
GRIDB200_API int gridb200_depth_unpack_u16(const uint16_t* src, int64_t n, int64_t b, int64_t lds, float* depth,
                              int64_t ld, const int64_t* esc_cell, const float* esc_val, int64_t esc_count,
                              int64_t row0, void* stream);

/* The same with `bits` (9..15) per cell: a 30x cohort's hundredths sit below 2^13, and the end-to-end step is bound by
 * PCIe.  A row is a sequence of groups of 32 cells, a group `bits` little-endian 32-bit words (cell c at bit c*bits);
 * the all-ones code is the escape; rows are ldw words apart, ldw >= gridb200_depth_packed_row_words(b, bits)
 * (whole groups, 16-byte aligned).  Unpacking yields the same bit-identical fp32 matrix.
 * row_q [n] / col_q [b] (both or neither; NULL = absolute codes): RELATIVE codes.  Coverage is to first order
 * (sample's depth) x (bin's bias), so with pred(i, j) = (row_q[i] * col_q[j] + 2048) >> 12 -- col_q = the bin's mean
 * hundredths (<= 65535), row_q = 4096 x the sample's mean over the cohort mean (< 32768), integers so that packer and
 * unpacker agree exactly -- a present cell travels as h - pred + 2^(bits-1); 0 stays "absent", all-ones the escape
 * (residual out of range, or h >= 65535).  A 30x cohort then needs 11 bits where the absolute value needs 13. */
GRIDB200_API int64_t gridb200_depth_packed_row_words(int64_t b, int32_t bits);
GRIDB200_API int gridb200_depth_pack_bits(const float* depth, int64_t n, int64_t b, int64_t ld, int32_t bits,
                             const int32_t* row_q, const int32_t* col_q, uint32_t* out, int64_t ldw, int64_t row0,
                             int64_t* esc_cell, float* esc_val, int64_t esc_cap, int64_t* esc_count, void* stream);
GRIDB200_API int gridb200_depth_unpack_bits(const uint32_t* src, int64_t n, int64_t b, int64_t ldw, int32_t bits,
                               const int32_t* row_q, const int32_t* col_q, float* depth, int64_t ld,
                               const int64_t* esc_cell, const float* esc_val, int64_t esc_count, int64_t row0,
                               void* stream);

/* ---- the depth matrix in 2 bytes per cell ON THE DEVICE ------------------------------------------------------------
 * uint16 hundredths, 0 = absent, rows ld elements apart (ld % 4 == 0, 8-byte aligned rows): the three step-4 passes then
 * read half the bytes of the fp32 matrix.  The code 0xFFFF is an escape: the cell's true hundredths (depth >= 655.35;
 * < 2^22 like everywhere) are in a table sorted by key (esc_row0 + row) * esc_b + bin -- esc_key [n_esc] ascending,
 * esc_h [n_esc].  The streaming kernels take an escape cell as 65535 and small patch kernels (one thread or warp per
 * escape cell / bin; gridb200_col_norm_stats_u16 also wants esc_cols [n_esc_cols], the distinct bins of the table)
 * put the true values in, so every result is bit-identical to the fp32 entry point of the same name (same reference
 * lines) and the hot loops carry no escape test.  gridb200_depth_pack_u16 produces the
 * matrix from fp32, gridb200_depth_unpack_bits_u16 from the bit-packed host format (its escape code becomes 0xFFFF). */
GRIDB200_API int gridb200_depth_unpack_bits_u16(const uint32_t* src, int64_t n, int64_t b, int64_t ldw, int32_t bits,
                                   const int32_t* row_q, const int32_t* col_q, uint16_t* out, int64_t ldo,
                                   void* stream);
GRIDB200_API int gridb200_col_row_raw_stats_u16(const uint16_t* depth, int64_t n, int64_t b, int64_t ld,
                                   const int64_t* esc_key, const int32_t* esc_h, int64_t n_esc, int64_t esc_row0,
                                   int64_t esc_b, int64_t* colsum, int32_t* colcnt, int64_t* rowsum_all,
                                   int32_t* rowcnt_all, void* stream);
GRIDB200_API int gridb200_row_stats_fused_u16(const uint16_t* depth, int64_t n, int64_t b, int64_t ld,
                                 const int64_t* esc_key, const int32_t* esc_h, int64_t n_esc, int64_t esc_row0,
                                 int64_t esc_b, const uint8_t* valid, const int32_t* colcnt, const int64_t* rowsum_all,
                                 const int32_t* rowcnt_all, int64_t* rowsum, int32_t* rowcnt, void* ws, size_t ws_bytes,
                                 void* stream);
GRIDB200_API int gridb200_row_means_exact_u16(const uint16_t* depth, int64_t n, int64_t b, int64_t ld,
                                 const int64_t* esc_key, const int32_t* esc_h, int64_t n_esc, int64_t esc_row0,
                                 int64_t esc_b, const uint8_t* valid, const int64_t* rowsum, const int32_t* rowcnt,
                                 double* raw_mean, double* scale_q, void* stream);
GRIDB200_API int gridb200_col_norm_stats_u16(const uint16_t* depth, int64_t n, int64_t b, int64_t ld,
                                const int64_t* esc_key, const int32_t* esc_h, int64_t n_esc, int64_t esc_row0,
                                int64_t esc_b, const int32_t* esc_cols, int64_t n_esc_cols, const uint8_t* valid,
                                const double* row_factor, double* sd, double* sd2, void* ws, size_t ws_bytes, void* stream);
GRIDB200_API int gridb200_transform_u16(const uint16_t* depth, int64_t n, int64_t b, int64_t ld,
                           const int64_t* esc_key, const int32_t* esc_h, int64_t n_esc, int64_t esc_row0, int64_t esc_b,
                           const double* row_factor, const double* mu, const int32_t* sel_pos, const uint8_t* keep,
                           const double* info, int32_t zclip, int16_t* zq, int64_t ldq, void* zh, int64_t ldh,
                           int64_t* norms, unsigned long long* sat, void* stream);

/* ------------------------------------------------------------------ step-4 hand-off text --- */

/* The N x Rw body of the step-4 file, on the device.  Codes are the int16 hundredths gridb200_transform writes
 * (-32768 = "NA", -32767 = "-0.00").  Replaces the per-cell f-string of write_normalized_output
 * (normalize_mosdepth.py:548-552) and the per-token float() of read_normalized_data (find_neighbors.py:107-121).
 *
 * gridb200_format_z_measure: row_off[0..rows] (device) = byte offset of every row's text in the batch; row i is
 *   "z\tz\t...z\n" (the host writes "id\tscale\t" in front).  rows <= 65535 per call.
 * gridb200_format_z_write:   fills text[0 .. row_off[rows]) (text_cap >= that; 8 bytes per cell always suffice).
 *   `ws` must be the workspace the measure call filled.  ws: gridb200_format_z_ws_bytes(rows, rw). */
GRIDB200_API size_t gridb200_format_z_ws_bytes(int64_t rows, int64_t rw);
GRIDB200_API int gridb200_format_z_measure(const int16_t* zq, int64_t rows, int64_t rw, int64_t ldq, int64_t* row_off,
                              void* ws, size_t ws_bytes, void* stream);
GRIDB200_API int gridb200_format_z_write(const int16_t* zq, int64_t rows, int64_t rw, int64_t ldq,
                            const int64_t* row_off, const void* ws, char* text, int64_t text_cap, void* stream);

/* Inverse: line i of the batch is text[line_off[i], line_off[i+1]) including its newline (line_off: device,
 * rows+1 entries; the host finds line ends with memchr while it inflates).  Fields skip_fields.. of every line are
 * decoded into zq[i, 0..rw).  Accepted cell grammar: NA | nan | [-]d{1,3}.dd with |value| <= 327.66 ("-0.00" reads
 * back as 0, as float() does).  stats[0] = cells decoded, stats[1] = tokens outside the grammar, stats[2] = lines
 * whose field count is not skip_fields + rw: when either is non-zero the caller must re-read the file on the host
 * (the reference's float() accepts more spellings than the writer produces).
 * ws: gridb200_parse_z_ws_bytes(rows, max_line_bytes), max_line_bytes >= the longest line. */
GRIDB200_API size_t gridb200_parse_z_ws_bytes(int64_t rows, int64_t max_line_bytes);
GRIDB200_API int gridb200_parse_z_rows(const char* text, const int64_t* line_off, int64_t rows, int64_t max_line_bytes,
                          int32_t skip_fields, int64_t rw, int16_t* zq, int64_t ldq, int64_t* stats,
                          void* ws, size_t ws_bytes, void* stream);

/* The body of the step-5 neighbours file (save_neighbors, find_neighbors.py:231-267) from the k-best lists in HBM:
 * row i = token[row0+i] ( '\t' token[j] '\t' dist )* '\n' over the entries j = idx[i][0..] up to the first negative
 * index, dist = "%.2f" % (sqrt(d2/1e4)^2 / (2 * r_use)) with d2 the exact integer squared distance in units of 1e-4
 * (r_use == 0 counts as 1, :256-257).  token[s] = tok[tok_off[s] .. tok_off[s+1]) is the host-made "id\tscale" text of
 * sample s (n_tok samples; ids are arbitrary strings).  idx / d2 point at the first row of the batch, row stride ld.
 * measure: row_off[0..rows] (device) = byte offsets of the rows in the batch text; rows <= 65535 per call.
 * write:   fills text[0 .. row_off[rows]).  ws: gridb200_format_nbrs_ws_bytes(rows). */
GRIDB200_API size_t gridb200_format_nbrs_ws_bytes(int64_t rows);
GRIDB200_API int gridb200_format_nbrs_measure(const char* tok, const int64_t* tok_off, int64_t n_tok, const int32_t* idx,
                                 const int64_t* d2, int64_t rows, int32_t k, int64_t ld, int64_t row0,
                                 int64_t r_use, int64_t* row_off, void* ws, size_t ws_bytes, void* stream);
GRIDB200_API int gridb200_format_nbrs_write(const char* tok, const int64_t* tok_off, int64_t n_tok, const int32_t* idx,
                               const int64_t* d2, int64_t rows, int32_t k, int64_t ld, int64_t row0,
                               int64_t r_use, const int64_t* row_off, char* text, int64_t text_cap, void* stream);

/* Reading side (load_neighbors, compute_dipcn.py:105-152) for files step 5 wrote: every "nbr_id\tnbr_scale" of such a
 * file is byte for byte the "id\tscale" head of one of its lines, so neighbours become row indices by hashing that
 * token and comparing bytes.
 * gridb200_hash_tokens: hash[s] (int64) of token s = tok[tok_off[s] .. tok_off[s+1]); the caller sorts them (they must
 *   be distinct) and passes hash_sorted / hash_row (row of each sorted hash).
 * gridb200_parse_nbrs_rows: line i of the batch is text[line_off[i], line_off[i+1]) including its newline (as for
 *   gridb200_parse_z_rows).  idx[i][e] (row stride ld) = row index of entry e, -1 beyond the line's entries (columns
 *   0..k_cap-1 are overwritten); n_entries[i] = number of entries, -1 for a line whose field count is not 2 + 3m or
 *   exceeds k_cap.  stats (device, 4 x int64): [0] entries resolved, [1] entries whose token is no line's head,
 *   [2] irregular lines -- when [1] or [2] is non-zero the caller re-reads the file with the reference's rules.
 *   The distance field is not looked at (the reference ignores it too).  ws: gridb200_parse_z_ws_bytes(rows,
 *   max_line_bytes). */
GRIDB200_API int gridb200_hash_tokens(const char* tok, const int64_t* tok_off, int64_t n, int64_t* hash, void* stream);
GRIDB200_API int gridb200_parse_nbrs_rows(const char* text, const int64_t* line_off, int64_t rows, int64_t max_line_bytes,
                             const char* tok, const int64_t* tok_off, const int64_t* hash_sorted,
                             const int32_t* hash_row, int64_t n_tok, int32_t k_cap, int32_t* idx, int64_t ld,
                             int32_t* n_entries, int64_t* stats, void* ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------ step 7: hi_inference --- */

/* Iterative haplotype phasing of hi_inference.py:179-233, in the reference's in-place (Gauss-Seidel) order.
 * Haplotype h of sample i is index 2i+h.  Neighbours in CSR form: entries nbr_ptr[x] .. nbr_ptr[x+1] of nbr_hap
 * (haplotype indices) / nbr_w (weights) belong to haplotype x, in the reference's list order.
 * hap[2n] in/out: NaN for samples that do not take part, IRR/2 for the others on entry.
 * order / level_ptr_host: the level schedule -- order[level_ptr[l] .. level_ptr[l+1]) are the samples of level l, where
 *   level(i) = 1 + max level over neighbour samples j < i that take part (0 if none); level_ptr_host is a HOST
 *   array of n_levels+1 entries (the call launches one kernel per level and sweep).
 * old_ws: device scratch of 2n doubles followed by n_levels + 1 int64 (the level table, for the single cooperative
 *   launch that runs all sweeps with a grid barrier between levels).  Results equal the reference's doubles bit for bit. */
GRIDB200_API int gridb200_hi_phase(const double* irr, int64_t n, const int64_t* nbr_ptr, const int32_t* nbr_hap,
                      const double* nbr_w, const int32_t* order, const int64_t* level_ptr_host, int32_t n_levels,
                      int32_t n_iters, double* hap, double* old_ws, void* stream);

/* One in-place relaxation pass of the level schedule gridb200_hi_phase needs to keep the visiting order of the reference's
 * sweep (`for i in range(N)`, hi_inference.py:186-233) while running samples in parallel: level[i] (>= 0 for samples that take part, -1 otherwise; start
 * from 0 / -1) becomes 1 + max level over neighbour samples j < i that take part (0 if none); *changed (device) is set to
 * 1 when any level moved.  Repeat until a pass leaves *changed at 0 (at most depth + 1 passes). */
GRIDB200_API int gridb200_hi_level_pass(const int64_t* nbr_ptr, const int32_t* nbr_hap, int64_t n, int32_t* level,
                           int32_t* changed, void* stream);

/* imp[2i+h] = weighted mean of the phased neighbours of haplotype 2i+h after the last sweep, mean_irr/2 when it
 * has none (hi_inference.py:236-256). */
GRIDB200_API int gridb200_hi_impute(const int64_t* nbr_ptr, const int32_t* nbr_hap, const double* nbr_w,
                       const double* hap, int64_t n, double mean_irr, double* imp, void* stream);

/* ------------------------------------------------------------------ the path's producers */

/* Overlap-weighted depth of the region [start, end) on `chrom` for a BATCH of samples: the body of
 * compute_region_coverage (mosdepth.py:264-297) without its float arithmetic.  `text` is the concatenation of the
 * samples' decompressed `chrom\tstart\tend\tdepth\n` files, EACH ENDING IN A NEWLINE; file f is
 * text[file_off[f], file_off[f+1]) (file_off: device, nfiles+1 entries).  Chromosome test: exact equality (:290).
 * out (device, [nfiles][4] int64, overwritten): [0] = sum over overlapping lines of hundredths(depth) * overlap_bp,
 * [1] = sum of overlap_bp, [2] = number of overlapping lines, [3] = lines outside the strict grammar
 * (digits \t digits \t digits[.dd]; blank lines; leading blanks; < 4 fields) on ANY chromosome -- the reference
 * converts every line before it looks at the chromosome, so the caller replays such a file with Python semantics.
 * The caller turns [0], [1], [2] into int(round(100 * (region_cov / covered_bp))): exact quotient, except within
 * the rounding noise of the reference's sequential fp64 sum around a .5 tie (grid_b200/utils/mosdepth.py). */
GRIDB200_API int gridb200_region_coverage(const char* text, int64_t nbytes, const int64_t* file_off, int32_t nfiles,
                             const char* chrom, int32_t chrom_len, int64_t start, int64_t end,
                             int64_t* out, void* stream);

/* The read filter of count_reads_in_region (count_reads.py:95-105) over columnar read records, for a batch of
 * samples and G regions at once.  Read r of sample s: sample_off[s] <= r < sample_off[s+1] (device, ns+1 entries).
 * A read is counted when flag_bits has bit `flag` set (65536 bits = 2048 words: the exact-value test
 * `read.flag in proper_flags` -- 8 KB that stay in L1), mapq >= min_mapq, tid == next_tid, flag & 0x400 == 0 (duplicate), flag & 0x100 == 0
 * (secondary), and start_g <= pos < end_g on the region's reference.
 * Regions arrive as a sorted table of distinct boundary keys, bounds[nb] with key = (tid << 32) | position, and per
 * region the indices lo_g / hi_g of its start / end key in that table (hi_g <= lo_g: empty region, count 0).
 * bucket_first (device, optional): a directory over bounds -- entry tid * buckets_per_tid + (pos >> bucket_shift) =
 * number of boundary keys below (tid << 32) | ((pos >> bucket_shift) << bucket_shift), n_tid * buckets_per_tid + 1
 * entries, buckets_per_tid << bucket_shift beyond every region end; a read's interval then costs one load plus a
 * search among the boundaries of its own bucket.  NULL: plain binary search over bounds.
 * counts (device, [ns][g_regions] int64, overwritten).  Columns must be 16-byte aligned.
 * ws: gridb200_count_reads_ws_bytes(ns, nb).  15 bytes per read of compulsory traffic. */
GRIDB200_API size_t gridb200_count_reads_ws_bytes(int32_t ns, int32_t nb);
GRIDB200_API int gridb200_count_reads(const uint16_t* flag, const uint8_t* mapq, const int32_t* tid, const int32_t* next_tid,
                         const int32_t* pos, int64_t n_reads, const int64_t* sample_off, int32_t ns,
                         const uint32_t* flag_bits, int32_t min_mapq, const int64_t* bounds, int32_t nb,
                         const int32_t* lo, const int32_t* hi, int32_t g_regions, const int32_t* bucket_first,
                         int32_t bucket_shift, int32_t buckets_per_tid, int32_t n_tid, int64_t* counts,
                         void* ws, size_t ws_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GRIDB200_H */
