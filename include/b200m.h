/*
 * b200m.h -- C-ABI of the B200-native pairwise descriptor matcher.
 *
 * This is the drop-in boundary for ONE hot path of daovietanh190499/3D_Reconstruction:
 * "two descriptor sets in -> (query index, train index, distance) match lists out".
 * The reference has no FFI of its own; the interface each entry point stands in for is cited
 * as reference file:line (paths relative to the reference repository root):
 *
 *   matching.py:122            matches01 = matcher({'image0': feats0, 'image1': feats1})
 *   matching.py:123-128        idx0 / idx1 extraction from matches01['matches']
 *   lightglue/lightglue.py:218-227   MatchAssignment.forward  (dense M x N similarity)
 *   lightglue/lightglue.py:235-254   filter_matches           (row/col arg-best + mutual check)
 *   lightglue/lightglue.py:417-426   match list packing, ascending query index
 *   (named by BASELINE.json, third-party)  cv2.BFMatcher(NORM_L2).knnMatch(k=2) + Lowe ratio + cross-check
 *
 * Conventions
 *   - every pointer except `pairs_host`-style arguments is a DEVICE pointer unless its name ends in _host;
 *   - the library never allocates or frees caller-visible memory: outputs and the scratch
 *     workspace are caller-allocated (size it with b200m_workspace_bytes);
 *   - all compute entry points are asynchronous on `stream` (a cudaStream_t passed as void*);
 *   - return value: 0 = OK, negative = error (B200M_E_*), text via b200m_last_error_string();
 *   - there is no CPU fallback and no other-architecture fallback: on a machine without an
 *     sm_100 GPU every compute entry point returns B200M_E_NO_DEVICE.
 *
 * Descriptor bank layout (HBM): uint8 [n_img, k_stride, 128], row-major, k_stride a multiple
 * of 256; image i owns rows [0, k_per_image[i]) of its slab, the remaining rows are padding.
 * b200m_bank_prepare_u8 zeroes the padding and writes the int32 norm table
 * (squared norms [n_img, k_stride], padding rows = B200M_PAD_NORM, + per-32-row minima).
 */
#ifndef B200M_H_
#define B200M_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200M_VERSION 120          /* 0.2.0 */
#define B200M_DESC_DIM 128         /* SIFT / DISK descriptor length handled by the u8 path */
#define B200M_TILE_Q 128           /* query rows per UMMA M-tile (two per work item) */
#define B200M_TILE_T 256           /* train rows per tile (UMMA N) */
#define B200M_PAD_NORM (1 << 23)   /* > 128*255^2 = 8,323,200: marks a padding row */

enum {
  B200M_OK = 0,
  B200M_E_BAD_ARG = -1,      /* null pointer, k_stride not a multiple of 256, negative sizes ... */
  B200M_E_BAD_DIM = -2,      /* descriptor length not supported */
  B200M_E_WORKSPACE = -3,    /* workspace too small */
  B200M_E_NO_DEVICE = -4,    /* no sm_100 device / driver entry point missing */
  B200M_E_CUDA = -5          /* a CUDA call failed; see b200m_last_error_string() */
};

int b200m_version(void);
const char* b200m_last_error_string(void);

/* 0 when the current CUDA device can run the sm_100a kernels, B200M_E_NO_DEVICE otherwise. */
int b200m_device_check(void);

/* Number of SMs the sweep kernel will be launched on (persistent grid); <0 on error. */
int b200m_sm_count(void);

/* Elements (int32) of the norm table of a bank: n_img*k_stride squared norms followed by
 * n_img*k_stride/32 minima over 32 consecutive rows (used by the threshold-reject epilogue). */
int64_t b200m_norm_table_elems(int32_t n_img, int64_t k_stride);

/* Zero the padding rows of `bank` and fill the norm table (exact int32 squared L2 norms,
 * B200M_PAD_NORM on padding rows, then the per-32-row minima).
 * Replaces the per-pair repacking at matching.py:107-120 with a one-off pass. */
int b200m_bank_prepare_u8(uint8_t* bank, int32_t n_img, int64_t k_stride,
                          const int32_t* k_per_image, int32_t* norm_table, void* stream);

/* Norm-sorted copy of the bank for the default pipeline (the chunk sweep): per image the rows
 * ordered by (squared norm, row), so that the 32 rows of a chunk have nearly equal norms and the
 * epilogue can rank whole chunks by max(row.col) alone.
 *   sorted_bank  uint8 [n_img, k_stride, 128]   (also used as sort scratch while the call runs)
 *   sort_table   int32 [b200m_sort_table_elems]: perm (sorted position -> row) | inv (row -> sorted
 *                position) | sorted norms | per-chunk minimum norm | per-image largest chunk spread
 * Needs the norm table of b200m_bank_prepare_u8. The original bank stays in use (exact fallback,
 * mutual check, outputs are in original row numbers). */
int64_t b200m_sort_table_elems(int32_t n_img, int64_t k_stride);
int b200m_bank_sort_u8(const uint8_t* bank, const int32_t* norm_table, const int32_t* k_per_image,
                       int32_t n_img, int64_t k_stride,
                       uint8_t* sorted_bank, int32_t* sort_table, void* stream);

/* The same two passes over the images [img_lo, img_lo + img_count) only; table layouts stay those of the WHOLE bank
 * (n_img images). Lets a caller prepare the bank while it is still arriving -- block by block behind the
 * host->device copies of one process, or slice by slice on the ranks of a multi-GPU job, each preparing the images
 * it uploaded before an all-gather distributes bank and tables (the per-pair host->device repacking of
 * matching.py:107-120, overlapped and sharded). Ranges are independent of each other. */
int b200m_bank_prepare_range_u8(uint8_t* bank, int32_t n_img, int64_t k_stride,
                                const int32_t* k_per_image, int32_t* norm_table,
                                int32_t img_lo, int32_t img_count, void* stream);
int b200m_bank_sort_range_u8(const uint8_t* bank, const int32_t* norm_table, const int32_t* k_per_image,
                             int32_t n_img, int64_t k_stride,
                             uint8_t* sorted_bank, int32_t* sort_table,
                             int32_t img_lo, int32_t img_count, void* stream);

/* Scratch bytes needed to process `n_pairs` pairs of images with slab stride `k_stride`.
 * Pairs are processed in internal batches, so this saturates at a fixed batch size. */
size_t b200m_workspace_bytes(int64_t k_stride, int64_t n_pairs);

/* cv2.BFMatcher(NORM_L2).knnMatch(query, train, k=2) for every (query_img, train_img) pair.
 *   pairs          int32 [n_pairs, 2]  (query image, train image), device
 *   out_train_idx  int32 [n_pairs, k_stride, 2]  best / second-best train row, -1 = none
 *   out_dist       float [n_pairs, k_stride, 2]  float32 sqrt of the exact int32 squared distance
 * Rows >= k_per_image[query_img] are left untouched. Ties -> lowest train index (cv2 rule). */
int b200m_knn2_u8(const uint8_t* bank, const int32_t* norm_table, const int32_t* k_per_image,
                  int32_t n_img, int64_t k_stride,
                  const int32_t* pairs, int64_t n_pairs,
                  int32_t* out_train_idx, float* out_dist,
                  void* workspace, size_t workspace_bytes, void* stream);

/* The same knnMatch(k=2) table on the chunk pipeline (needs the outputs of b200m_bank_sort_u8): one chunk sweep on the
 * norm-sorted bank keeps the three smallest chunk bounds of every query row, the two winning 64-column chunks are
 * evaluated exactly (rows bucketed by chunk, gathered into a swizzled operand tile, tcgen05 int8 MMA against the chunk;
 * option "knn_eval" = 1: dp4a), rows whose runner-up is not provably ahead of every other column go through the exact
 * top-2 sweep. Bit-identical outputs, about 2.3x the throughput of b200m_knn2_u8 (43 k vs 19 k pairs/s at 16,384 rows). */
int b200m_knn2_sorted_u8(const uint8_t* bank, const int32_t* norm_table,
                         const uint8_t* sorted_bank, const int32_t* sort_table,
                         const int32_t* k_per_image, int32_t n_img, int64_t k_stride,
                         const int32_t* pairs, int64_t n_pairs,
                         int32_t* out_train_idx, float* out_dist,
                         void* workspace, size_t workspace_bytes, void* stream);

/* The full matching rule of the hot path for every pair:
 *   2-NN in L2  ->  keep q if dist1 < ratio * dist2 (float32 distances, double compare,
 *   exactly `m.distance < ratio * n.distance` on cv2 DMatch)  ->  if mutual != 0 keep (q,t) only
 *   when q is the nearest query of t (lowest-index ties, cv2 crossCheck rule).
 * ratio <= 0 disables the ratio test.  A query with a single train candidate has no second
 * neighbour and is dropped by the ratio test (cv2: the `for m, n in knn` unpack needs two).
 *   out_query_idx / out_train_idx / out_dist   [n_pairs, cap], ascending query index
 *   out_count  int32 [n_pairs]: matches FOUND; if > cap only the first cap were stored and the
 *              caller must treat the pair as overflowed (never silently truncated). -1: the pair-list
 *              entry names an image outside [0, n_img) (such entries are skipped, never dereferenced;
 *              k_per_image entries are clamped to [0, k_stride] by every kernel).
 *   sorted_bank / sort_table: outputs of b200m_bank_sort_u8, or both NULL (then the slower
 *              pipeline 1 runs on the original bank alone). Results are identical either way. */
int b200m_match_pairs_u8(const uint8_t* bank, const int32_t* norm_table,
                         const uint8_t* sorted_bank, const int32_t* sort_table,
                         const int32_t* k_per_image, int32_t n_img, int64_t k_stride,
                         const int32_t* pairs, int64_t n_pairs,
                         double ratio, int mutual,
                         int32_t* out_query_idx, int32_t* out_train_idx, float* out_dist,
                         int32_t* out_count, int64_t cap,
                         void* workspace, size_t workspace_bytes, void* stream);

/* Float path (SuperPoint / DISK style descriptors, lightglue/lightglue.py:223-250 in raw-similarity
 * form): similarity = inner product, bf16 operands with fp32 accumulation on the tensor cores.
 *   bank_bf16  bf16 [n_img, k_stride, dim], dim 128 or 256, rows >= k_per_image[i] ignored
 * A query row q matches train row t when t is q's arg-max similarity, (mutual != 0) q is t's
 * arg-max similarity (lowest index on exact ties, both ways), and (use_threshold != 0)
 * similarity > score_threshold. Outputs as b200m_match_pairs_u8, out_score = the fp32 similarity.
 * Not bit-comparable with an fp32 reference: the stated tolerance is index agreement (tests/). */
size_t b200m_workspace_bytes_bf16(int64_t k_stride, int64_t n_pairs);
int b200m_match_pairs_bf16(const void* bank_bf16, const int32_t* k_per_image, int32_t n_img,
                           int64_t k_stride, int32_t dim,
                           const int32_t* pairs, int64_t n_pairs,
                           float score_threshold, int use_threshold, int mutual,
                           int32_t* out_query_idx, int32_t* out_train_idx, float* out_score,
                           int32_t* out_count, int64_t cap,
                           void* workspace, size_t workspace_bytes, void* stream);

/* The reference matcher's assignment head without the M x N matrices: log double-softmax of the
 * similarity plus matchability (sigmoid_log_double_softmax, lightglue/lightglue.py:196-208, on
 * sim = mdesc0 . mdesc1 of MatchAssignment.forward :218-227) followed by filter_matches (:235-254)
 * and the match list (:417-426).
 *   mdesc_bf16    bf16 [n_img, k_stride, dim]: final_proj(desc) / dim^0.25 (:220-222), dim 128 or 256
 *   matchability  float32 [n_img, k_stride]: the logits z = matchability(desc) (:224-225)
 * Two tensor sweeps per pair (option "float_sweep" = 0, default): pass 1 the log-sum-exp of every row AND every
 * column of sim from one sweep (all exponentials measured from one reference point, the product of the longest row
 * norms; sums that leave the fp32 range around it are detected on the device and the batch is redone by the
 * two-direction sweep with per-row maxima), pass 2 row arg-max + column maxima of
 *   score[i,j] = 2 sim[i,j] - lse_row[i] - lse_col[j] + logsigmoid(z0[i]) + logsigmoid(z1[j]).
 * ("float_sweep" = 1: one sweep per direction and pass, four per pair.)
 * A row i matches j when j = arg-max_j score[i,:], i = arg-max_i score[:,j] and (use_threshold != 0)
 * exp(score) > filter_threshold. out_score = exp(score) (= matching_scores0 of the reference);
 * other outputs as b200m_match_pairs_u8. With use_threshold == 0 the list holds every mutual pair,
 * from which matches0/1 and matching_scores0/1 of the reference follow exactly (host side).
 * Workspace footprint, with B = pairs per internal batch (min(n_pairs, 2^22 / k_stride) unless "batch" is set):
 *   24 * B * k_stride bytes (row results and log-sum-exp tables) + 1 MiB (largest row norm per image)
 *   + with "float_sweep" = 0 only: 4 * B * (k_stride / 256) * k_stride bytes of column partial sums -- quadratic in
 *     k_stride: 128 MiB at k_stride = 2,048 (B = 2,048), 1 GiB at k_stride = 16,384 (B = 256); lower "batch" to shrink it.
 * The size depends on the "float_sweep" and "batch" options as they are when b200m_workspace_bytes_assign is called:
 * a workspace sized under one setting and used under another is refused with B200M_E_WORKSPACE if it is too small. */
size_t b200m_workspace_bytes_assign(int64_t k_stride, int64_t n_pairs);
int b200m_assign_pairs_bf16(const void* mdesc_bf16, const float* matchability, const int32_t* k_per_image,
                            int32_t n_img, int64_t k_stride, int32_t dim,
                            const int32_t* pairs, int64_t n_pairs,
                            float filter_threshold, int use_threshold,
                            int32_t* out_query_idx, int32_t* out_train_idx, float* out_score,
                            int32_t* out_count, int64_t cap,
                            void* workspace, size_t workspace_bytes, void* stream);

/* Visual-word assignment of the BoW pair scheduler, `vq(descriptors, codebook)` at matching.py:26-28:
 * out_code[i] = arg-min over codebook rows of the L2 distance to obs[i] (first minimum on ties),
 * out_dist[i] = that distance (may be NULL). fp64 arithmetic in scipy's form (|x|^2 + |c|^2 - 2 x.c).
 *   obs float32 [n_obs, dim] (dim <= 256), codebook float64 [n_codes, dim],
 *   code_sqr_scratch float64 [n_codes] (caller-allocated scratch). */
int b200m_vq_f32(const float* obs, int64_t n_obs, int32_t dim, const double* codebook, int32_t n_codes,
                 int32_t* out_code, double* out_dist, double* code_sqr_scratch, void* stream);

/* b200m_vq_f32 for callers that cannot provide device scratch. (No counterpart in the reference: the reference never
 * manages device memory itself.)
 *
 * b200m_vq_f32 needs a float64 [n_codes] scratch array that the caller allocates on the device. Some hosts of this
 * library have no way to do that: a plugin that is handed its input and output device buffers by an engine which owns
 * the memory pool, a binding (ctypes / cgo / JNI stub) that links this library but not the CUDA runtime, a service
 * that runs inside a fixed memory budget where cudaMalloc after start-up is forbidden, or code running while another
 * component holds the device in a state where allocation is not allowed. For them the squared codebook norms live in
 * an array that is part of the library image itself: B200M_VQ_STATIC_MAX_CODES float64 values (32 KiB) in device
 * global memory, held by every GPU that loads the library until the process exits.
 *
 * Same arguments (minus the scratch pointer), same kernels and bit-identical outputs as b200m_vq_f32. Differences:
 *   - n_codes > B200M_VQ_STATIC_MAX_CODES is refused with B200M_E_BAD_ARG (use b200m_vq_f32 with your own scratch);
 *   - the call allocates and frees nothing, on the device or through any library. The first call on each device looks
 *     up the device address of the built-in array (once per device, with that device current); make that first call
 *     before starting any stream capture;
 *   - all calls on one device share the one array. It keeps whatever the previous call left there; nothing in it has
 *     to start at zero, because each call rewrites entries [0, n_codes) before it reads them. Two calls that can run on
 *     the GPU at the same time (different streams, or different host threads on different streams) race on it:
 *     serialise them, or give all but one of them b200m_vq_f32 with private scratch. */
#define B200M_VQ_STATIC_MAX_CODES 4096
int b200m_vq_f32_static(const float* obs, int64_t n_obs, int32_t dim, const double* codebook, int32_t n_codes,
                        int32_t* out_code, double* out_dist, void* stream);

/* Batched geometric verification of match lists, the step after the matcher in the reference
 * (matching.py:133-142): cv2.findEssentialMat(k0, k1, K, method=RANSAC, prob=0.999, threshold=1)
 * followed by cv2.recoverPose(E, k0[mask], k1[mask], K).
 *   matches_xyxy  float32 [total, 4] (16-byte aligned): (x0, y0, x1, y1) of every match of every pair in
 *                 NORMALISED image coordinates (pixel / focal for the reference's K = diag(f, f, 1))
 *   offsets       int64 [n_pairs + 1]: matches of pair p are rows [offsets[p], offsets[p+1])
 *   iters         RANSAC hypotheses per pair (rounded up to a multiple of 128; cv2's maxIters is 1000)
 *   threshold_norm  Sampson threshold in normalised units (threshold_px / focal); max_depth: recoverPose's
 *                 distanceThresh (<= 0 -> 50)
 * Per pair: out_E [9] (row-major, unit Frobenius norm; zeros when the pair has < 5 matches or no model),
 * out_mask [total] / out_inliers: Sampson inliers of the best model; out_R [9], out_t [3] (unit), out_pose_mask /
 * out_pose_inliers: the (R, t) candidate with the most inliers in front of both cameras and those inliers.
 * Five-point minimal solver (Nister) per hypothesis; deterministic for a given seed; parity with cv2 is
 * statistical (RANSAC draws differ), see tests/test_gpu_geometry.py. */
size_t b200m_workspace_bytes_ransac(int64_t n_pairs, int32_t iters);
int b200m_essential_ransac_f32(const float* matches_xyxy, const int64_t* offsets, int64_t n_pairs, int32_t iters,
                               float threshold_norm, uint64_t seed, float max_depth,
                               double* out_E, uint8_t* out_mask, int32_t* out_inliers,
                               double* out_R, double* out_t, uint8_t* out_pose_mask, int32_t* out_pose_inliers,
                               void* workspace, size_t workspace_bytes, void* stream);
/* One minimal solve (test hook): x0, x1 float64 [5, 2] device -> out_E float64 [10, 9], *out_n solutions. */
int b200m_five_point_f64(const double* x0, const double* x1, double* out_E, int32_t* out_n, void* stream);

/* Tuning / introspection knobs (process-global, read at launch time).
 *   "epilogue"  0 = exact insertion of every element (debug), 1 = threshold-reject (default),
 *               2 = MMA-only ceiling measurement (results invalid)
 *   "pipeline"  2 = chunk sweep on the norm-sorted bank + exact refine / fallback / verify (default),
 *               1 = top-1 sweep + lazy ratio / mutual verify sweeps over the candidates,
 *               0 = exact top-2 sweep + lazy mutual verify sweep
 *   "kernel"    5 = chunk sweep with N256 MMAs / two accumulators (default), 3 = N128 / four accumulators
 *               (4 / 6, the slower cta_group::2 variants of csrc/experimental/, are refused unless the library
 *               was built with -DB200M_EXPERIMENTAL)
 *   "chunk"     columns per chunk of the chunk sweep: 64 (default, kernel 5 only) or 32
 *   "grid"      number of persistent CTAs, 0 = one per SM (default)
 *   "batch"     pairs per internal batch, 0 = default
 *   "float_sweep" 0 = float path as ONE tensor sweep per pair and pass: column maxima / column log-sum-exp fused into the
 *               row sweep (default), 1 = one sweep per direction
 *   "refine"    pipeline 2, from chunk bounds to exact rows: 0 = one fused kernel (default), 1 = streaming
 *               classification + one warp per row whose winning chunk has to be evaluated (pays on banks where many rows
 *               need it)
 *   "knn_eval"  b200m_knn2_sorted_u8, exact evaluation of the two winning chunks: 0 = tensor cores over gathered
 *               rows (default), 1 = dp4a
 *   "profile"   1 = time sweep launches with CUDA events (see b200m_profile_read) */
int b200m_set_option(const char* name, int64_t value);
int64_t b200m_get_option(const char* name);

/* [n_pairs, cap] slots + counts -> flat lists. offsets int64 [n_pairs] = exclusive scan of
 * min(count, cap) (device; the caller computes it, e.g. with a cumsum). This is the device-side
 * form of the per-pair idx0 / idx1 arrays matching.py:128 pulls back to the host. */
int b200m_compact_matches(const int32_t* query_idx, const int32_t* train_idx, const float* dist,
                          const int32_t* count, const int64_t* offsets, int64_t n_pairs, int64_t cap,
                          int32_t* flat_query_idx, int32_t* flat_train_idx, float* flat_dist, void* stream);

/* Counters of the last b200m_match_pairs_u8 call that used `workspace` (blocks until `stream` has drained):
 * out_host[0..5] = query rows seen by the refine step | rows whose winning chunk was evaluated exactly | rows sent
 * to the exact fallback sweep | candidates surviving the Lowe test | rows whose Lowe test needed the verify sweep |
 * matches emitted. Throughput of the default pipeline depends on these fractions. */
int b200m_match_stats_read(const void* workspace, int64_t* out_host, int n, void* stream);

/* With option "profile" = 1 every sweep launch is bracketed by CUDA events on its stream.
 * Reads (and resets) the accumulated device time of the first sweeps and of the verify sweeps,
 * the number of first-sweep launches, and the number of kernels launched by the library.
 * Blocks until the recorded events have completed. */
int b200m_profile_read(double* first_sweep_ms, double* verify_sweep_ms, int64_t* first_sweeps, int64_t* launches);
/* Call BEFORE b200m_profile_read: device ms per kernel class since the last read,
 * out_ms[0..5] = first sweep, verify sweep, fallback sweep, refine, select, emit. */
int b200m_profile_detail(double* out_ms, int n);

/* Diagnostics: raw 128 x 256 int32 dot-product tile through the same TMA + tcgen05 path
 * (used by the self-test to validate descriptors/swizzle independently of the epilogue). */
int b200m_debug_dot_tile_u8(const uint8_t* a_rows /*[128,128]*/, const uint8_t* b_rows /*[256,128]*/,
                            int32_t* out /*[128,256]*/, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200M_H_ */
