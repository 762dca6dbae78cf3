/*
 * sfm_match.h -- C ABI of the B200 pairwise-matching library (libsfm_match.so).
 *
 * This is the drop-in boundary for the `preprocess` hot path of Sovik-Ghosh/SFM.
 * The reference has no FFI layer (it is pure Python over OpenCV); each entry
 * point below names the reference call site it replaces (paths relative to the
 * reference repo root).  Host language above this ABI is Python (ctypes, see
 * sfm_b200/lib.py and INTEGRATION.md).
 *
 * Conventions
 *  - every function returns 0 on success, non-zero on failure;
 *    sfm_last_error(ctx) gives the message.  No C++ exceptions cross the ABI.
 *  - `d_` pointers are DEVICE memory owned by the caller (PyTorch allocates),
 *    `h_` pointers are HOST memory owned by the caller.  The library never
 *    frees caller memory and only keeps the pointers passed to sfm_set_images.
 *  - `stream` is a cudaStream_t passed as void*; all work is asynchronous on
 *    that stream; nothing synchronises unless stated.
 *  - one context per GPU / rank, used from one host thread.
 */
#ifndef SFM_MATCH_H_
#define SFM_MATCH_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define SFM_API __attribute__((visibility("default")))
#else
#define SFM_API
#endif

#define SFM_DESC_DIM 128          /* SIFT descriptor length, bytes per row */
#define SFM_QBLOCK 256            /* query rows per work item of the kNN kernel */
#define SFM_TTILE 128             /* train rows per tensor-core tile */
#define SFM_FILTER_STATIC_MAX_PAIRS 16384   /* most pairs per sfm_filter_matches_static call */

typedef struct sfm_ctx sfm_ctx;

/* status codes written per pair by sfm_verify_pairs; they name the reference's
 * skip gates in ImageMatcher.process_image_pair (utils/find_matches.py:268,
 * :274, :284, :290). */
enum {
    SFM_PAIR_OK = 0,
    SFM_PAIR_TOO_FEW_FEATURES = 1, /* :268, and Nt<2 (knnMatch rows of length 1 raise at :151) */
    SFM_PAIR_TOO_FEW_MATCHES = 2,  /* :274  len(matches) < 5 */
    SFM_PAIR_NO_MODEL = 3,         /* :284  F is None / not 3x3 */
    SFM_PAIR_QUALITY_FAIL = 4      /* :290  verify_match_quality false */
};

/* Parameters of RANSAC + verification.  Defaults (sfm_default_params) are the
 * constants the reference hard-codes. */
typedef struct sfm_verify_params {
    double ransac_thresh;   /* 3.0   utils/find_matches.py:282 */
    double confidence;      /* 0.99  cv2.findFundamentalMat default */
    int32_t max_iters;      /* 1000  cv2.findFundamentalMat default (rounded up to a multiple of 32) */
    int32_t min_matches;    /* 5     utils/find_matches.py:274 */
    uint64_t seed;          /* hypothesis seed; results are a pure function of (seed, pair inputs) */
    double verify_thresh;   /* 3.0   utils/find_matches.py:157 */
    int32_t min_inliers;    /* 15    utils/find_matches.py:203 */
    int32_t lo_iters;       /* 2     inlier-refit rounds on the best hypothesis */
    double min_ratio;       /* 0.3 */
    double max_error;       /* 2.0 */
    double min_spread;      /* 20    utils/find_matches.py:185 */
    int32_t score_mode;     /* 3.  bit 0: hypotheses scored by the symmetric epipolar error instead of
                               Sampson; bit 1: the inlier refits use the symmetric error (the
                               measure geometric_verification applies, :165-172).  Sampson scoring
                               (0 / 2) leaves the SURVEY 8c envelope on the 148 golden pairs (4-6
                               inlier-bound failures per seed, tests/ransac_tune.py) and is no faster */
    int32_t min_work;       /* 65536: floor on hypotheses x matches before the adaptive stop may end the
                               search, for pairs of 15..256 matches (tapering to none at 384; at most
                               4096 hypotheses); 0 disables */
} sfm_verify_params;

/* Per-pair metrics, the dict returned by ImageMatcher.geometric_verification
 * (utils/find_matches.py:190-201) plus RANSAC bookkeeping. */
typedef struct sfm_pair_metrics {
    int32_t total_matches;
    int32_t inliers;             /* sum(inlier_mask) */
    float inlier_ratio;          /* informative; hosts recompute inliers/total in double */
    float reprojection_error;    /* mean symmetric error over inliers, +inf if none */
    float symmetric_error;       /* mean symmetric error over all matches */
    int32_t well_distributed;    /* 0/1 */
    int32_t ransac_inliers;      /* support of the best hypothesis under the RANSAC threshold */
    int32_t ransac_iters;        /* hypotheses evaluated */
} sfm_pair_metrics;

/* ---- context ---------------------------------------------------------- */
SFM_API int sfm_create(int device_ordinal, sfm_ctx** out);
SFM_API void sfm_destroy(sfm_ctx* ctx);
SFM_API const char* sfm_last_error(sfm_ctx* ctx);     /* ctx may be NULL: last create error */
SFM_API void sfm_default_params(sfm_verify_params* p);
SFM_API int sfm_abi_version(void);

/* ---- image set -------------------------------------------------------- */
/* Register the feature set (replaces the per-pair detect_features calls,
 * utils/find_matches.py:265-266; extraction itself stays on the host).
 *   d_desc     u8  [total_rows, 128] packed row-major, 16-byte aligned
 *   d_kps      f32 [total_rows, 2]   keypoint (x, y) per descriptor row; may be NULL
 *                                     if sfm_verify_pairs is not used
 *   h_offsets  i64 [n_images + 1]    row range of image i is [off[i], off[i+1])
 * Launches the descriptor-norm kernel on `stream`.  The pointers must stay
 * valid until the next sfm_set_images / sfm_destroy. */
SFM_API int sfm_set_images(sfm_ctx* ctx, const uint8_t* d_desc, const float* d_kps,
                   const int64_t* h_offsets, int32_t n_images, void* stream);

/* ---- K1: brute-force L2 kNN, k = 2 ------------------------------------ */
/* Replaces cv2.BFMatcher(NORM_L2).knnMatch(desc1, desc2, k=2)
 * (utils/find_matches.py:144-147, L2 per SURVEY.md 0.2) for a batch of
 * DIRECTED pairs (query image, train image).
 *   h_pairs    i32 [n_pairs, 2]
 *   h_row_off  i64 [n_pairs + 1]  first output row of pair p; pair p owns
 *                                  n_query(p) consecutive rows
 *   d_knn      i32 [total_rows, 4] per query row: (idx1, idx2, d1^2, d2^2),
 *              exact integers, ascending distance, lowest train index wins
 *              ties; idx2 = -1 and d2^2 = INT32_MAX when the train image has
 *              fewer than 2 rows (idx1 too when it has none). */
SFM_API int sfm_knn2(sfm_ctx* ctx, const int32_t* h_pairs, int32_t n_pairs,
             const int64_t* h_row_off, int32_t* d_knn, void* stream);

/* ---- K2: ratio test (+ mutual-NN) and order-preserving compaction ------ */
/* Replaces the ratio loop utils/find_matches.py:150-153: keep a query row iff
 * float64(f32sqrt(d1^2)) < ratio * float64(f32sqrt(d2^2)); with d_knn_rev
 * non-NULL additionally iff the best train row's own nearest query row
 * (d_knn_rev, computed by sfm_knn2 on the reversed pairs with offsets
 * h_row_off_rev) is this query row (README.md:108; == BFMatcher crossCheck).
 * Output arena uses the forward row offsets: pair p may write up to
 * n_query(p) matches starting at h_row_off[p], ascending queryIdx:
 *   d_mq, d_mt  i32 query / train row index within their image
 *   d_md2       i32 squared distance of the kept match
 *   d_count     i32 [n_pairs] */
SFM_API int sfm_filter_matches(sfm_ctx* ctx, const int32_t* h_pairs, int32_t n_pairs,
                       const int64_t* h_row_off, const int32_t* d_knn,
                       const int64_t* h_row_off_rev, const int32_t* d_knn_rev,
                       double ratio, int32_t* d_mq, int32_t* d_mt, int32_t* d_md2,
                       int32_t* d_count, void* stream);

/* sfm_filter_matches without allocation.  No counterpart in the reference: added
 * for callers that may not call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same rule, same outputs and same argument meaning as
 * sfm_filter_matches with d_knn_rev = NULL (ratio test only), bit for bit.  The per-call metadata
 * (16 bytes per pair) is staged in a file-scope __device__ array of SFM_FILTER_STATIC_MAX_PAIRS
 * records (256 KiB, held by every GPU that loads the library until the process exits) instead of
 * buffers that grow with cudaMalloc / cudaMallocHost: the call makes no cudaMalloc, cudaFree,
 * cudaMallocHost or other allocating call.  n_pairs beyond the maximum is rejected (return 1,
 * sfm_last_error names the limit): split the batch, or use sfm_filter_matches.
 * The array is shared by every call on the device.  Each call rewrites the records it reads, so
 * nothing has to be cleared between calls; calls that OVERLAP ON THE GPU (different streams, or
 * threads on one device) race on it -- order them on one stream or with an event.  The first call
 * on a device resolves the array's address (cudaGetSymbolAddress; with lazy module loading this
 * loads the kernel image): a call with n_pairs = 0 does only that and is the warm-up to make
 * before graph capture.  The metadata copy reads pageable host memory, so the call returns after
 * the stream's earlier work has been reached (it does not synchronise on its own kernel). */
SFM_API int sfm_filter_matches_static(sfm_ctx* ctx, const int32_t* h_pairs, int32_t n_pairs,
                              const int64_t* h_row_off, const int32_t* d_knn, double ratio,
                              int32_t* d_mq, int32_t* d_mt, int32_t* d_md2, int32_t* d_count,
                              void* stream);

/* ---- K3 + K4: RANSAC fundamental matrix + geometric verification ------ */
/* Mutual nearest-neighbour check on match lists that have passed the ratio test (README.md:108 of the
 * reference, "mutual nearest neighbor check"; the reference's code path has none).  Match i of pair p
 * = (i, j) is kept iff its query row is the nearest row of image i to its train descriptor (lowest
 * index on ties, as cv2.BFMatcher(crossCheck=True) decides).  Sparse form: only the train descriptors
 * of the surviving matches are matched back against image i, not all n_j rows.
 * h_count: host copy of d_count (the caller fetches it; this is the one synchronisation the check needs).
 * The lists d_mq / d_mt / d_md2 are compacted in place (ascending query index kept), d_count rewritten.
 * Fails when a descriptor norm is beyond the folded kernel's range (then use sfm_knn2 on the reversed
 * pairs + sfm_filter_matches with d_knn_rev). */
SFM_API int sfm_mutual_filter(sfm_ctx* ctx, const int32_t* h_pairs, int32_t n_pairs, const int64_t* h_row_off,
                              const int32_t* h_count, int32_t* d_mq, int32_t* d_mt, int32_t* d_md2,
                              int32_t* d_count, void* stream);

/* Replaces, per pair: the point gather utils/find_matches.py:278-279,
 * cv2.findFundamentalMat(pts1, pts2, FM_RANSAC, 3.0) (:282),
 * geometric_verification (:157-201) and verify_match_quality (:203-214).
 * Inputs are the arenas written by sfm_filter_matches.  Outputs:
 *   d_pts1, d_pts2  f32 [arena, 2]  gathered keypoints (rows of pair p at h_row_off[p])
 *   d_F             f64 [n_pairs, 9] row-major, rank 2, F[2][2] = 1 when not tiny
 *   d_mask          u8  [arena]     inlier_mask of geometric_verification
 *   d_sym           f32 [arena]     symmetric_errors
 *   d_metrics       [n_pairs], d_status i32 [n_pairs] (SFM_PAIR_*) */
SFM_API int sfm_verify_pairs(sfm_ctx* ctx, const int32_t* h_pairs, int32_t n_pairs,
                     const int64_t* h_row_off, const int32_t* d_mq, const int32_t* d_mt,
                     const int32_t* d_count, const sfm_verify_params* params,
                     float* d_pts1, float* d_pts2, double* d_F, uint8_t* d_mask,
                     float* d_sym, sfm_pair_metrics* d_metrics, int32_t* d_status,
                     void* stream);

/* geometric_verification alone for a caller-supplied F (parity tests of
 * utils/find_matches.py:157-214 against the golden fundamental/*.npz):
 *   h_off i64 [n_pairs+1] point ranges, d_pts1/d_pts2 f32 [total,2], d_F f64 [n_pairs,9] */
SFM_API int sfm_verify_given_F(sfm_ctx* ctx, int32_t n_pairs, const int64_t* h_off,
                       const float* d_pts1, const float* d_pts2, const double* d_F,
                       const sfm_verify_params* params, uint8_t* d_mask, float* d_sym,
                       sfm_pair_metrics* d_metrics, int32_t* d_status, void* stream);

/* RANSAC + verification on caller-supplied correspondences (no gather):
 * same outputs as sfm_verify_pairs. */
SFM_API int sfm_ransac_points(sfm_ctx* ctx, int32_t n_pairs, const int64_t* h_off,
                      const float* d_pts1, const float* d_pts2,
                      const sfm_verify_params* params, double* d_F, uint8_t* d_mask,
                      float* d_sym, sfm_pair_metrics* d_metrics, int32_t* d_status,
                      void* stream);

/* ---- result packing ---------------------------------------------------- */
/* The arenas above are sized for the worst case (one match per query row).
 * Before the device->host copy, gather the `d_count[p]` valid entries of every
 * kept pair (all pairs, or only those with d_status == SFM_PAIR_OK when
 * only_ok != 0) into dense arrays -- what save_pair_data
 * (utils/find_matches.py:312-327) writes for passing pairs.
 *   d_pack_off i64 [n_pairs + 1] (out) dense offsets; [n_pairs] = total kept
 * Any source pointer may be NULL (that array is skipped); each destination
 * must hold as many elements as its source arena. */
SFM_API int sfm_pack_results(sfm_ctx* ctx, int32_t n_pairs, const int64_t* h_row_off,
                     const int32_t* d_count, const int32_t* d_status, int32_t only_ok,
                     const int32_t* d_mq, const int32_t* d_mt, const int32_t* d_md2,
                     const uint8_t* d_mask, const float* d_sym, const float* d_pts1,
                     const float* d_pts2, int64_t* d_pack_off, int32_t* o_mq, int32_t* o_mt,
                     int32_t* o_md2, uint8_t* o_mask, float* o_sym, float* o_pts1, float* o_pts2,
                     void* stream);

/* ---- 2-D radius matching of track points ----------------- */
/* The dense distance test inside StructureFromMotion.find_2d3d_matches
 * (utils/sfm_reconstruction.py:211-213):
 *     distances = np.linalg.norm(valid_track_points[:, None] - other_img_pts, axis=2)
 *     matches = np.where(distances < MATCHING_THRESHOLD)
 * for n_seg segments at once (one segment = the tracks that contain the other
 * image x the points of one image pair, :181-207).
 *   h_track_off i64 [n_seg+1]  track rows of segment s = [off[s], off[s+1])
 *   d_tracks    f64 [rows,2]   valid_track_points of all segments
 *   h_pt_off    i64 [n_seg+1]  point ranges
 *   d_pts       f64 [n_pts,2]  other_img_pts (the f32 files widened exactly)
 *   threshold                  MATCHING_THRESHOLD (:14); the test is the f64
 *                              expression above, decision for decision
 *   h_seg_off   i64 [n_seg+1]  (out) match ranges: segment s owns entries
 *                              [h_seg_off[s], h_seg_off[s+1]); [n_seg] = total
 *   d_out_track, d_out_pt  i32 [cap] (out) the two arrays np.where returns, per
 *                              segment in its order (track-major, point ascending);
 *                              indices are local to the segment
 * Pass d_out_track = d_out_pt = NULL to size the result (h_seg_off only).
 * If the total h_seg_off[n_seg] exceeds cap nothing is written to the output
 * arrays (the call still returns 0): call again with room.  Synchronises `stream`. */
SFM_API int sfm_radius_match(sfm_ctx* ctx, int32_t n_seg, const int64_t* h_track_off,
                     const double* d_tracks, const int64_t* h_pt_off, const double* d_pts,
                     double threshold, int64_t* h_seg_off, int32_t* d_out_track,
                     int32_t* d_out_pt, int64_t cap, void* stream);

/* ---- measurement helpers ---------------------------------------------- */
/* Number of kernel launches issued by this context since creation. */
SFM_API int64_t sfm_launch_count(sfm_ctx* ctx);
/* Tensor-pipe peak probe: runs an MMA-only tcgen05 kind::i8 loop on every SM
 * for `iters` 128x128x128 tile products per CTA and returns the elapsed
 * milliseconds (CUDA events) in *ms and the int8 op count in *ops. */
SFM_API int sfm_probe_int8_peak(sfm_ctx* ctx, int32_t iters, float* ms, double* ops, void* stream);
/* FP64 issue-rate probe (the bound of sfm_radius_match): eight independent DFMA chains per thread on every
 * SM; *instr = thread-level f64 instructions executed, *ms = elapsed. */
SFM_API int sfm_probe_fp64_rate(sfm_ctx* ctx, int32_t iters, float* ms, double* instr, void* stream);
/* Same with 128 x tile_n x 128 tile products (tile_n = 64, 128 or 256; the peak probe uses 256):
 * shows how the shared-memory operand reads of narrower tiles hold the tensor pipe back. */
SFM_API int sfm_probe_int8_rate(sfm_ctx* ctx, int32_t iters, int32_t tile_n, float* ms, double* ops,
                        void* stream);
/* Debug: raw u8 x u8 dot products of one 128x128 tile pair through the same
 * TMA + tcgen05 path (rows q_row.. of the registered set against t_row..),
 * written to d_out i32 [128,128]. */
SFM_API int sfm_debug_dot_tile(sfm_ctx* ctx, int64_t q_row, int64_t t_row, int32_t* d_out, void* stream);

/* Debug: folded accumulators of one 128 x 256 tile through the wide kernel's path
 * (sfm_b200/csrc/knn_wide.cu: N = 256 MMAs, fold operand as SWIZZLE_32B rows):
 * a.b + 2016029 - floor(||b||^2 / 2) for query rows q_row.. against descriptor rows t_row..,
 * written to d_out i32 [128,256]. */
SFM_API int sfm_debug_wide_tile(sfm_ctx* ctx, int64_t q_row, int64_t t_row, int32_t* d_out, void* stream);
/* Debug: the same folded accumulators for a 128 x 128 tile through the TS path
 * (sfm_b200/csrc/knn_ts.cu: query rows written to tensor memory with tcgen05.st, tcgen05.mma with
 * the A operand in TMEM), written to d_out i32 [128,128]. */
SFM_API int sfm_debug_ts_tile(sfm_ctx* ctx, int64_t q_row, int64_t t_row, int32_t* d_out, void* stream);
/* Debug: the same folded accumulators for a 256 x 128 tile through the CTA-pair path
 * (sfm_b200/csrc/knn_pair.cu: a cluster of two CTAs, tcgen05.mma.cta_group::2 with M = 256, N = 128,
 * each CTA staging its 128 query rows and 64 of the 128 train rows), written to d_out i32 [256,128]. */
SFM_API int sfm_debug_pair_tile(sfm_ctx* ctx, int64_t q_row, int64_t t_row, int32_t* d_out, void* stream);
/* Cycle counters the pair kernel's first epilogue warp leaves behind when probe flag 512 is set
 * (main loop, resolve, recompute, merge + write, recomputed rows, work items, -, -). */
SFM_API int sfm_debug_pair_counters(sfm_ctx* ctx, int64_t* h_out8, void* stream);
/* Issue-rate probe of 128 x tile_n x 32 u8 MMAs (tile_n = 64, 96, 128, 192 or 256) in groups of
 * five, accumulators rotating over 384 / tile_n TMEM stages, no data movement.  ts != 0: A operand
 * in tensor memory, else in shared memory; commit != 0: one tcgen05.commit per group.
 * *ms = elapsed (CUDA events), *ops = int8 operations issued. */
SFM_API int sfm_debug_ts_rate_probe(sfm_ctx* ctx, int32_t groups, int32_t tile_n, int32_t ts, int32_t commit,
                            float* ms, double* ops, void* stream);
/* TMEM read-rate probe: every SM runs `warps` warps issuing `iters` tcgen05.ld each
 * (variant 0: 32x32b.x32 + wait each; 1: .x32, wait every 4; 2: .x64 + wait each;
 * 3: 16x256b.x8, wait every 4); *h_cycles = slowest warp's SM cycles (synchronises). */
SFM_API int sfm_debug_tmem_ld_probe(sfm_ctx* ctx, int32_t warps, int32_t variant, int32_t iters,
                            int64_t* h_cycles, void* stream);
/* Issue-pattern probe: `steps` steps of 2 accumulators x k_per_acc 128x128x32 u8 MMAs on every
 * SM, no data movement.  flags bit 0: commit per step; bit 1: consumer warps hand the TMEM stage
 * back (needs bit 0 and consumers >= 1); bit 2: one extra commit per step.  *ms = elapsed. */
SFM_API int sfm_debug_mma_handshake_probe(sfm_ctx* ctx, int32_t steps, int32_t k_per_acc, int32_t flags,
                                  int32_t consumers, float* ms, void* stream);
/* Select the kNN kernel variant used by sfm_knn2: 0 automatic (wide folded kernel, N = 256
 * MMAs with the norm term folded into the contraction, unless a registered row has
 * ||b||^2 > 4,032,059, then the generic variant), 1 generic only, 5 wide folded kernel only,
 * 12 CTA-pair folded kernel (knn_pair.cu: cta_group::2, four accumulator stages), 15 wide kernel
 * with the warp-uniform skip of the unit bookkeeping (measured slower; kept for the A/B), 16 wide
 * kernel with the lean unit tracker (fewer ALU instructions, same speed; A/B), 17 ping-pong kernel (the two accumulators alternate
 * over train tiles of a 128-row query block; measured slower; kept for the A/B);
 * timing probes whose output is NOT a kNN: 6 / 13 / 18 max tree only (wide / pair / ping-pong),
 * 7 / 14 / 19 TMA + MMA pipeline without an epilogue. */
SFM_API int sfm_debug_set_knn_mode(sfm_ctx* ctx, int32_t mode);
/* mbarrier suspend-time hints (ns; 0 = spin) of the kNN kernel's producer/MMA warps and of
 * its epilogue warps.  Tuning knob; results do not depend on it. */
SFM_API int sfm_debug_set_knn_waits(sfm_ctx* ctx, int32_t wait_ns, int32_t epi_wait_ns);
/* Probe switches of the kNN kernel.  Timing probes (results become meaningless): bit 0 the
 * producer skips the train-tile TMA loads (generic and pair kernels), bit 1 the MMA warp skips the tcgen05.mma
 * issue, bit 2 the wide kernel runs without its tie-list and fix-up passes, bit 5 without its parking
 * stores, bit 6 (ping-pong kernel) without the end-of-pass work, bit 8 the second row block's epilogue
 * warps of the wide kernel do no work, bit 10 the wide kernel issues only the first of a tile's eight
 * TMEM loads per thread.  Bit 4: CTA 0 of the wide / ping-pong kernel counts cycles in
 * its MMA warp and its first epilogue warp; sfm_debug_pair_counters reads them: [0] epilogue loop,
 * [1] of it waiting for a full accumulator, [2] for tcgen05.ld, [3] at the end of work items, [4] MMA
 * warp's loop, [5] of it waiting for train tiles, [6] for an empty accumulator, [7] tiles counted.
 * (Bits 4, 5, 6, 8, 10 select a separate instantiation of the kernel: compiled into the default one the
 * counters cost 15 %.)  Test switches (results stay exact): bit 3 shrinks the fix-up list to 4
 * entries, so its overflow scan is exercised; bit 7 makes the wide and ping-pong kernels recompute
 * tied units inside the kernel instead of deferring them to the tie list; bit 9 shrinks the tie list
 * to 2 entries (rows beyond it go to the fix-up pass). */
SFM_API int sfm_debug_set_knn_probe_flags(sfm_ctx* ctx, int32_t flags);
/* Hand-over latency probe (one CTA): n_mma tcgen05.mma (128 x tile_n x 32, tile_n = 128 or 256) + commit by
 * one warp, a second warp waits and answers through a plain mbarrier.  h_out4 = mean cycles of: issue,
 * issue start -> other warp woken, its arrive -> issuer woken, issue start -> issuer itself woken.
 * trail > 0: that many more MMAs (another accumulator) are issued right after the commit. */
SFM_API int sfm_debug_handover_probe(sfm_ctx* ctx, int32_t iters, int32_t n_mma, int32_t tile_n, int32_t trail,
                             int64_t* h_out4, void* stream);
/* Reads (synchronising `stream`) whether the registered set exceeded the fold range. */
SFM_API int sfm_debug_fold_flag(sfm_ctx* ctx, int32_t* h_flag, void* stream);

/* Statistics: query rows the folded kNN kernel recomputed by brute force since the last
 * sfm_set_images (rows whose candidate tracking met an exact tie; see knn_wide.cu).  Synchronises. */
SFM_API int sfm_debug_brute_rows(sfm_ctx* ctx, int32_t* h_rows, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SFM_MATCH_H_ */
