// 2-D radius matching of track points against the points of an image pair.
//
// Replaces the dense distance matrix of the reference's find_2d3d_matches
// (utils/sfm_reconstruction.py:211-213):
//     distances = np.linalg.norm(valid_track_points[:, None] - other_img_pts, axis=2)
//     matches = np.where(distances < MATCHING_THRESHOLD)
// for a whole batch of segments (one segment = the tracks x points of one image pair) at once.
// The result is the list np.where yields: (track, point) index pairs in row-major order, per
// segment.  Nothing is materialised but that list: a counting pass that also keeps the first
// kRadiusKeep hits of every track and sums the counts per block, a scan of the block totals, and a
// fill pass that scans its own tracks' counts on top of the block offset and copies the
// kept hits to their place (a track matches a handful of points at most); only blocks that hold a
// track with more hits than were kept repeat the comparison for those tracks.
//
// Exactness.  numpy evaluates sqrt(dx*dx + dy*dy) < thr in f64 with every operation rounded
// once.  sqrt_rn is monotone, so the test equals  dx*dx + dy*dy < s_lim  with
// s_lim = min{ s : sqrt_rn(s) >= thr }, which the host finds by bisection over the f64 bit
// patterns (api.cu).  The device does two subtractions, two products, one sum and one compare,
// all explicitly round-to-nearest without FMA contraction: the decisions are bit-identical.
// Rounding is monotone, so rn(dx*dx) >= s_lim already decides "no hit" after three operations.
//
// Layout: tracks f64 [rows, 2] and points f64 [n_pts, 2] concatenated over the segments; block b
// owns kRadiusRows consecutive tracks of one segment (host-built block list), one track per
// thread, and walks the segment's points through shared memory in tiles of kRadiusTile (every
// thread reads the same point: a broadcast).  FP64-pipe bound: 3 f64 instructions per (track,
// point), 7 for the few inside the x band; the bytes are negligible (each point is read from L2
// once per 256 tracks).

#include <cstdint>
#include <cuda_runtime.h>

#include "kernels.h"

namespace sfm {
namespace {

constexpr int kRadiusTile = 1024;   // points staged per pass: 16 KiB of double2

template <bool FILL>
__global__ void __launch_bounds__(kRadiusRows)
radius_kernel(const RadiusBlock* __restrict__ blocks, const double2* __restrict__ tracks,
              const double2* __restrict__ pts, double s_lim, long long* __restrict__ row_cnt,
              int32_t* __restrict__ first, long long* __restrict__ block_cnt,
              const long long* __restrict__ block_off, long long cap, int32_t* __restrict__ out_track,
              int32_t* __restrict__ out_pt) {
    __shared__ double2 tile[kRadiusTile];
    __shared__ long long warp_tot[kRadiusRows / 32];
    const RadiusBlock b = blocks[blockIdx.x];
    const int local = static_cast<int>(threadIdx.x);
    const bool live = local < b.n_rows;
    const long long row = b.row0 + local;
    const double2 t = live ? tracks[row] : make_double2(0.0, 0.0);
    long long n = 0;
    long long dst = 0;
    bool scan = live;
    if (FILL) {
        // offsets of this block's tracks: exclusive scan of their counts on top of the block's offset
        const long long cnt = live ? row_cnt[row] : 0;
        const int lane = local & 31, warp = local >> 5;
        long long x = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const long long y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= d) x += y;
        }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        long long before = block_off[blockIdx.x] + (x - cnt);
        for (int w = 0; w < warp; ++w) before += warp_tot[w];
        dst = before;
        scan = cnt > kRadiusKeep;
        if (!scan)
            for (int i = 0; i < static_cast<int>(cnt); ++i)
                if (dst + i < cap) {
                    out_track[dst + i] = b.track0 + local;
                    out_pt[dst + i] = first[row * kRadiusKeep + i];
                }
        if (!__syncthreads_or(scan)) return;
    }
    for (int p0 = 0; p0 < b.n_pts; p0 += kRadiusTile) {
        const int np = min(kRadiusTile, b.n_pts - p0);
        __syncthreads();
        for (int i = local; i < np; i += kRadiusRows) tile[i] = pts[b.pt0 + p0 + i];
        __syncthreads();
        if (!scan) continue;
#pragma unroll 4
        for (int i = 0; i < np; ++i) {
            const double2 q = tile[i];
            const double dx = __dsub_rn(t.x, q.x);
            const double sx = __dmul_rn(dx, dx);
            // rn(sx + sy) >= sx for sy >= 0: a point outside the band in x alone is decided here
            if (!(sx < s_lim)) continue;
            const double dy = __dsub_rn(t.y, q.y);
            const double s = __dadd_rn(sx, __dmul_rn(dy, dy));
            if (s < s_lim) {
                if (FILL) {
                    if (dst < cap) {
                        out_track[dst] = b.track0 + local;
                        out_pt[dst] = p0 + i;
                    }
                    ++dst;
                } else {
                    if (n < kRadiusKeep) first[row * kRadiusKeep + n] = p0 + i;
                    ++n;
                }
            }
        }
    }
    if (!FILL) {
        if (live) row_cnt[row] = n;
        // block total for the scan over blocks
        long long x = n;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) x += __shfl_xor_sync(0xffffffffu, x, d);
        __syncthreads();
        if ((local & 31) == 0) warp_tot[local >> 5] = x;
        __syncthreads();
        if (local == 0) {
            long long tot = 0;
            for (int w = 0; w < kRadiusRows / 32; ++w) tot += warp_tot[w];
            block_cnt[blockIdx.x] = tot;
        }
    }
}

// Exclusive scan of n block totals into n + 1 offsets by one block (one value per 256 tracks);
// off[n] = total.
constexpr int kScanThreads = 1024;

__global__ void __launch_bounds__(kScanThreads) scan_kernel(const long long* __restrict__ cnt, long long n,
                                                            long long* __restrict__ off) {
    __shared__ long long warp_sum[kScanThreads / 32];
    __shared__ long long carry_s;
    const int tid = static_cast<int>(threadIdx.x), lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (long long base = 0; base < n; base += kScanThreads) {
        const long long i = base + tid;
        const long long v = i < n ? cnt[i] : 0;
        long long x = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const long long y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= d) x += y;
        }
        if (lane == 31) warp_sum[warp] = x;
        __syncthreads();
        if (warp == 0) {
            long long w = warp_sum[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const long long y = __shfl_up_sync(0xffffffffu, w, d);
                if (lane >= d) w += y;
            }
            warp_sum[lane] = w;   // inclusive over the warps
        }
        __syncthreads();
        const long long carry = carry_s;
        const long long before = carry + (warp ? warp_sum[warp - 1] : 0) + (x - v);
        if (i < n) off[i] = before;
        __syncthreads();
        if (tid == kScanThreads - 1) carry_s = carry + warp_sum[kScanThreads / 32 - 1];
        __syncthreads();
    }
    if (tid == 0) off[n] = carry_s;
}

__global__ void gather_offsets_kernel(const long long* __restrict__ row_off, const long long* __restrict__ at,
                                      int n, long long* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = row_off[at[i]];
}

}  // namespace

void launch_gather_offsets(const long long* d_row_off, const long long* d_at, int32_t n, long long* d_out,
                           cudaStream_t st, int64_t* launches) {
    if (n <= 0) return;
    gather_offsets_kernel<<<(n + 255) / 256, 256, 0, st>>>(d_row_off, d_at, n, d_out);
    ++*launches;
}

void launch_radius_count(const RadiusBlock* d_blocks, int32_t n_blocks, const double* d_tracks,
                         const double* d_pts, double s_lim, long long* d_row_cnt, int32_t* d_first,
                         long long* d_block_cnt, long long* d_block_off, cudaStream_t st, int64_t* launches) {
    if (n_blocks > 0) {
        radius_kernel<false><<<n_blocks, kRadiusRows, 0, st>>>(
            d_blocks, reinterpret_cast<const double2*>(d_tracks), reinterpret_cast<const double2*>(d_pts), s_lim,
            d_row_cnt, d_first, d_block_cnt, nullptr, 0, nullptr, nullptr);
        ++*launches;
    }
    scan_kernel<<<1, kScanThreads, 0, st>>>(d_block_cnt, n_blocks, d_block_off);
    ++*launches;
}

void launch_radius_fill(const RadiusBlock* d_blocks, int32_t n_blocks, const double* d_tracks,
                        const double* d_pts, double s_lim, const long long* d_row_cnt, const int32_t* d_first,
                        const long long* d_block_off, long long cap, int32_t* d_out_track, int32_t* d_out_pt,
                        cudaStream_t st, int64_t* launches) {
    if (n_blocks <= 0) return;
    radius_kernel<true><<<n_blocks, kRadiusRows, 0, st>>>(
        d_blocks, reinterpret_cast<const double2*>(d_tracks), reinterpret_cast<const double2*>(d_pts), s_lim,
        const_cast<long long*>(d_row_cnt), const_cast<int32_t*>(d_first), nullptr, d_block_off, cap, d_out_track,
        d_out_pt);
    ++*launches;
}

}  // namespace sfm
