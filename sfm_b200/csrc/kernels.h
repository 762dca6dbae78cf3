// Internal launch interface between api.cu and the kernel translation units.
#pragma once
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>

#include "../../include/sfm_match.h"

namespace sfm {

// One unit of kNN work: a block of up to SFM_QBLOCK query rows of one image
// against every row of one train image.
struct __align__(16) WorkItem {
    int32_t q_row;    // first query row (row index into the packed descriptor array)
    int32_t q_cnt;    // valid query rows in this block (1..256)
    int32_t t_row;    // first row of the train image
    int32_t t_cnt;    // rows of the train image
    int32_t t_pad;    // offset of the train image in the 128-padded column-norm array
    int32_t x_row;    // unused (kept for the 32-byte layout)
    int64_t out_row;  // first output row of this block in d_knn
};
static_assert(sizeof(WorkItem) == 32, "WorkItem must be 32 bytes");

// ---- knn.cu (generic kNN kernel: the fallback for norms beyond the folded kernels' range; norms; work items)
void launch_norms(const uint8_t* d_desc, int64_t total_rows, const int64_t* d_img_off,
                  const int64_t* d_pad_off, int32_t n_images, int32_t* d_norms, int32_t* d_norms_pad,
                  int32_t* d_flags, cudaStream_t st, int64_t* launches);
void launch_build_work(const int32_t* d_pairs, int32_t n_pairs, const int64_t* d_blk_prefix,
                       const int64_t* d_row_off, const int64_t* d_img_off, const int64_t* d_pad_off,
                       WorkItem* d_work, cudaStream_t st, int64_t* launches);
struct KnnTuning {
    uint32_t wait_ns = 0;         // mbarrier suspend hint of the producer / MMA warps (0 = spin; measured best)
    uint32_t epi_wait_ns = 0;     // same for the epilogue warps' accumulator-ready wait
    uint32_t probe_flags = 0;     // timing probes: bit 0 no B-tile TMA traffic, bit 1 no MMA issue
};
// only_if_flag != 0: exits at once unless flags[0] says a norm is beyond the folded kernels' range.
int launch_knn2(const CUtensorMap& tmap, const WorkItem* d_work, int64_t n_work, const int32_t* d_norms,
                const int32_t* d_norms_pad, const int32_t* d_flags, int only_if_flag, const KnnTuning& tn,
                int32_t* d_knn, int num_sms, cudaStream_t st, int64_t* launches);
int launch_debug_dot(const CUtensorMap& tmap, int32_t q_row, int32_t t_row, int32_t* d_out, cudaStream_t st,
                     int64_t* launches);
int launch_mma_probe(int32_t iters, int32_t probe_n, int num_sms, cudaStream_t st, int64_t* launches);

// ---- knn_wide.cu (N = 256 folded kernel; fold operands as u8 [128 + R, 32])
void launch_bias32(const int32_t* d_norms, int64_t total_rows, uint8_t* d_bias32, int32_t* d_flags,
                   cudaStream_t st, int64_t* launches);
// Query rows that are not rows of the registered descriptor array (the mutual check matches the train
// descriptors of the surviving matches back against the query image): their own tensor map (128-row
// boxes), descriptor rows and squared norms; WorkItem::q_row indexes THIS array.
struct KnnQuery {
    const CUtensorMap* tmap;
    const uint8_t* desc;
    const int32_t* norms;
};
int launch_knn2_wide(const CUtensorMap& tmap, const CUtensorMap& tmap_x, const CUtensorMap& tmap_ax,
                     const uint8_t* d_desc, const WorkItem* d_work, int64_t n_work, const int32_t* d_norms,
                     const int32_t* d_norms_pad,
                     int32_t* d_flags, int variant, int force, const struct KnnTuning& tn, int32_t* d_fix_list,
                     int32_t fix_cap, int32_t* d_knn, int num_sms, cudaStream_t st, int64_t* launches,
                     const KnnQuery* query = nullptr, int32_t* d_tie_list = nullptr, int32_t tie_cap = 0);
int launch_debug_wide(const CUtensorMap& tmap, const CUtensorMap& tmap_x, const CUtensorMap& tmap_ax,
                      int32_t q_row, int32_t t_row, int32_t* d_out, cudaStream_t st, int64_t* launches);

// ---- knn_ts.cu (A operand in TMEM; probes and debug tile)
int launch_debug_ts(const CUtensorMap& tmap, const CUtensorMap& tmap_ax, const uint8_t* d_desc,
                    const uint8_t* d_bias32, int32_t q_row, int32_t t_row, int32_t* d_out, cudaStream_t st,
                    int64_t* launches);
// Fix-up pass shared by the folded kernels (knn_wide.cu): rows flagged kFlagged are recomputed by brute force.
int launch_knn_fix(const WorkItem* d_work, int64_t n_work, const uint8_t* d_desc, const int32_t* d_norms,
                   const int32_t* d_flags, const int32_t* d_fix_list, int32_t fix_cap, int32_t* d_knn, int num_sms,
                   cudaStream_t st, int64_t* launches);
// Norm parities of the padded column space as a bitmap (32 columns per word; knn_pair.cu).
void launch_parity_bits(const int32_t* d_norms_pad, int64_t pad_total, uint32_t* d_par_bits, cudaStream_t st,
                        int64_t* launches);
// ---- knn_pair.cu (cta_group::2: a cluster of two CTAs per 256 query rows, four 128-column accumulator stages)
// tmap_a = descriptor rows in 128-row boxes, tmap_b = the same array in 64-row boxes, tmap_x = fold operand rows
// (32 B) in 64-row boxes, tmap_ax = the same array in a 128-row box (constant rows).
int launch_knn2_pair(const CUtensorMap& tmap_a, const CUtensorMap& tmap_b, const CUtensorMap& tmap_x,
                     const CUtensorMap& tmap_ax, const uint8_t* d_desc, const WorkItem* d_work, int64_t n_work,
                     const int32_t* d_norms, const int32_t* d_norms_pad, const uint32_t* d_par_bits,
                     int32_t* d_flags, int variant, int force, const struct KnnTuning& tn, int32_t* d_fix_list,
                     int32_t fix_cap, int32_t* d_knn, int num_sms, cudaStream_t st, int64_t* launches);
int launch_debug_pair(const CUtensorMap& tmap_a, const CUtensorMap& tmap_b, const CUtensorMap& tmap_x,
                      const CUtensorMap& tmap_ax, int32_t q_row, int32_t t_row, int32_t* d_out, cudaStream_t st,
                      int64_t* launches);
int launch_handover_probe(int iters, int n_mma, int tile_n, int trail, long long* d_out, cudaStream_t st, int64_t* launches);
int launch_ts_rate_probe(int groups, int tile_n, int ts, int commit, int num_sms, cudaStream_t st,
                         int64_t* launches);

// ---- probes.cu
int launch_tmem_ld_probe(int warps, int variant, int iters, int num_sms, long long* d_out, cudaStream_t st,
                         int64_t* launches);

// FP64 DFMA issue-rate probe: num_sms * 4 CTAs x 512 threads x 8 chains x iters instructions.
int launch_fp64_probe(int iters, int num_sms, double* d_out, cudaStream_t st, int64_t* launches);

int launch_mma_handshake_probe(int steps, int k_per_acc, int flags, int consumers, int num_sms,
                               cudaStream_t st, int64_t* launches);

// ---- filter.cu
// Mutual check, sparse form: the train descriptors (and norms) of every surviving match, gathered to
// q_desc / q_norms rows m_off[p] + i.
void launch_mutual_gather(const int32_t* d_pairs, int32_t n_pairs, const int64_t* d_img_off, const int64_t* d_row_off,
                          const int64_t* d_m_off, const int32_t* d_mt, const uint8_t* d_desc, const int32_t* d_norms,
                          uint8_t* d_q_desc, int32_t* d_q_norms, cudaStream_t st, int64_t* launches);
// Keeps match i of pair p iff the nearest row of the QUERY image to its train descriptor
// (d_knn_rev[m_off[p] + i].x) is its query row; compacts the lists in place and rewrites d_count.
void launch_mutual_compact(int32_t n_pairs, const int64_t* d_row_off, const int64_t* d_m_off, const int4* d_knn_rev,
                           int32_t* d_mq, int32_t* d_mt, int32_t* d_md2, int32_t* d_count, cudaStream_t st,
                           int64_t* launches);
void launch_filter(const int32_t* d_pairs, int32_t n_pairs, const int64_t* d_img_off,
                   const int64_t* d_row_off, const int4* d_knn, const int64_t* d_row_off_rev,
                   const int4* d_knn_rev, double ratio, int32_t* d_mq, int32_t* d_mt,
                   int32_t* d_md2, int32_t* d_count, cudaStream_t st, int64_t* launches);

// ---- filter_static.cu: the ratio filter with its metadata in a file-scope __device__ array (no allocation).
// h_img_off: the registered image offsets (host).  Returns cudaErrorInvalidValue beyond
// SFM_FILTER_STATIC_MAX_PAIRS; n_pairs == 0 only resolves the array's address for `device` (must be current).
cudaError_t launch_filter_static(int device, const int32_t* h_pairs, int32_t n_pairs, const int64_t* h_img_off,
                                 const int64_t* h_row_off, const int4* d_knn, double ratio, int32_t* d_mq,
                                 int32_t* d_mt, int32_t* d_md2, int32_t* d_count, cudaStream_t st,
                                 int64_t* launches);

struct PackArrays {
    const int32_t *mq, *mt, *md2;
    const uint8_t* mask;
    const float *sym, *pts1, *pts2;
    int32_t *o_mq, *o_mt, *o_md2;
    uint8_t* o_mask;
    float *o_sym, *o_pts1, *o_pts2;
};
// d_status may be NULL when only_ok == 0.
void launch_pack(int32_t n_pairs, const int64_t* d_row_off, const int32_t* d_count,
                 const int32_t* d_status, int only_ok, int64_t* d_pack_off, const PackArrays& a,
                 cudaStream_t st, int64_t* launches);

// ---- radius.cu (2-D radius matching of track points)
constexpr int kRadiusRows = 256;   // tracks per block, one per thread
constexpr int kRadiusKeep = 4;     // hits per track the counting pass keeps for the fill pass
struct RadiusBlock {
    long long row0;    // first track row (global, over all segments)
    long long pt0;     // first point of the segment (global)
    int32_t n_rows;    // tracks of this block (<= kRadiusRows)
    int32_t n_pts;     // points of the segment
    int32_t track0;    // index of row0 inside its segment
    int32_t pad;
};
// Counts per track row into d_row_cnt (+ the first kRadiusKeep point indices of each row into
// d_first[row * kRadiusKeep]), block totals into d_block_cnt[n_blocks], then their exclusive scan
// into d_block_off[n_blocks + 1].
void launch_radius_count(const RadiusBlock* d_blocks, int32_t n_blocks, const double* d_tracks,
                         const double* d_pts, double s_lim, long long* d_row_cnt, int32_t* d_first,
                         long long* d_block_cnt, long long* d_block_off, cudaStream_t st, int64_t* launches);
// d_out[i] = d_off[d_at[i]]: the match offsets at the segment boundaries.
void launch_gather_offsets(const long long* d_off, const long long* d_at, int32_t n, long long* d_out,
                           cudaStream_t st, int64_t* launches);
// (track index in its segment, point index in its segment) of every match, row-major per segment.
void launch_radius_fill(const RadiusBlock* d_blocks, int32_t n_blocks, const double* d_tracks,
                        const double* d_pts, double s_lim, const long long* d_row_cnt, const int32_t* d_first,
                        const long long* d_block_off, long long cap, int32_t* d_out_track, int32_t* d_out_pt,
                        cudaStream_t st, int64_t* launches);

// ---- ransac.cu
void launch_gather_points(const int32_t* d_pairs, int32_t n_pairs, const int64_t* d_img_off,
                          const int64_t* d_row_off, const int32_t* d_mq, const int32_t* d_mt,
                          const int32_t* d_count, const float* d_kps, float* d_pts1, float* d_pts2,
                          cudaStream_t st, int64_t* launches);
// d_cnt may be NULL: then the count of pair p is d_off[p+1]-d_off[p].
// d_F_in non-NULL: skip RANSAC and verify against the supplied F.
// d_pair_ids (nullable) i32 [n_pairs,2]: image ids keying the hypothesis RNG.
int launch_ransac_verify(int32_t n_pairs, const int64_t* d_off, const int32_t* d_cnt,
                         const int32_t* d_pair_ids, const float* d_pts1, const float* d_pts2, const double* d_F_in,
                         const sfm_verify_params& prm, double* d_F, uint8_t* d_mask, float* d_sym,
                         sfm_pair_metrics* d_metrics, int32_t* d_status, cudaStream_t st,
                         int64_t* launches);

}  // namespace sfm
