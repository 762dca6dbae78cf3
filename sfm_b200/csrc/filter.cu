// K2: Lowe ratio test (+ optional mutual-NN check) and order-preserving
// compaction of the surviving matches.
//
// Replaces the Python loop of reference utils/find_matches.py:150-153
//     for m, n in matches: if m.distance < 0.75 * n.distance: good.append(m)
// where m.distance / n.distance are float32 square roots of the exact integer
// squared distances and the comparison is evaluated in double precision; that
// rule is reproduced bit for bit (correctly rounded f32 sqrt, widened, one
// double multiply).  Mutual-NN (README.md:108, == BFMatcher crossCheck=True):
// the best train row's own nearest query row must be this query row.
//
// HBM-bound integer pass: one CTA per pair streams 16 B per query row with
// 128-bit loads and writes <= 12 B per kept match; a block-wide ballot scan
// keeps ascending queryIdx order.
#include "kernels.h"

namespace sfm {

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ bool ratio_pass(int d1, int d2, double ratio) {
    const float s1 = __fsqrt_rn(static_cast<float>(d1));   // exact ints < 2^24 -> exact floats
    const float s2 = __fsqrt_rn(static_cast<float>(d2));
    return static_cast<double>(s1) < ratio * static_cast<double>(s2);
}

// A block walks its pair in tiles of 4 x 256 query rows: every thread loads four rows (row = tile + 256 k + tid:
// each load is one coalesced 4 KiB warp request), the 32 per-(k, warp) survivor counts go to shared memory, and
// after ONE barrier every warp scans them itself with shuffles -- the running base lives in a register, the
// count arrays alternate between tiles.  (Round 1 took three barriers per 256 rows: 34 % of the HBM peak.)
constexpr int kRowsPerThread = 4;

__global__ void __launch_bounds__(kThreads)
filter_kernel(const int32_t* __restrict__ pairs, int n_pairs, const long long* __restrict__ img_off,
              const long long* __restrict__ row_off, const int4* __restrict__ knn,
              const long long* __restrict__ row_off_rev, const int4* __restrict__ knn_rev,
              double ratio, int32_t* __restrict__ mq, int32_t* __restrict__ mt,
              int32_t* __restrict__ md2, int32_t* __restrict__ count) {
    __shared__ int wc[2][32];
    static_assert(kRowsPerThread * (kThreads / 32) == 32, "one count per lane of the scanning warp");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int it = 0;
    for (int p = blockIdx.x; p < n_pairs; p += gridDim.x) {
        const int qi = pairs[2 * p], ti = pairs[2 * p + 1];
        const int nq = static_cast<int>(img_off[qi + 1] - img_off[qi]);
        const int nt = static_cast<int>(img_off[ti + 1] - img_off[ti]);
        const long long r0 = row_off[p];
        const int4* rev = knn_rev ? knn_rev + row_off_rev[p] : nullptr;
        int base = 0;
        for (int q0 = 0; q0 < nq; q0 += kRowsPerThread * kThreads, ++it) {
            int4 r[kRowsPerThread];
            bool keep[kRowsPerThread];
            unsigned bal[kRowsPerThread];
#pragma unroll
            for (int k = 0; k < kRowsPerThread; ++k) {
                const int q = q0 + k * kThreads + threadIdx.x;
                r[k] = make_int4(-1, -1, 0, 0);
                if (q < nq && nt >= 2) r[k] = __ldg(knn + r0 + q);
            }
#pragma unroll
            for (int k = 0; k < kRowsPerThread; ++k) {
                const int q = q0 + k * kThreads + threadIdx.x;
                keep[k] = r[k].x >= 0 && r[k].y >= 0 && ratio_pass(r[k].z, r[k].w, ratio);
                if (keep[k] && rev) keep[k] = __ldg(&rev[r[k].x].x) == q;
                bal[k] = __ballot_sync(0xffffffffu, keep[k]);
                if (lane == 0) wc[it & 1][k * (kThreads / 32) + warp] = __popc(bal[k]);
            }
            __syncthreads();
            const int x = wc[it & 1][lane];
            int incl = x;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += y;
            }
            const int excl = incl - x;
            const int total = __shfl_sync(0xffffffffu, incl, 31);
#pragma unroll
            for (int k = 0; k < kRowsPerThread; ++k) {
                const int before = __shfl_sync(0xffffffffu, excl, k * (kThreads / 32) + warp);
                if (keep[k]) {
                    const long long o = r0 + base + before + __popc(bal[k] & ((1u << lane) - 1u));
                    mq[o] = q0 + k * kThreads + threadIdx.x;
                    mt[o] = r[k].x;
                    md2[o] = r[k].z;
                }
            }
            base += total;
        }
        if (threadIdx.x == 0) count[p] = base;
    }
}


// ---------------------------------------------------------------------------
// Mutual nearest-neighbour check, sparse form (README.md:108 of the reference; oracle/knn.py
// match_features(mutual=True)).  Only the matches that survive the ratio test need the reverse
// question "is q the nearest row of image i to descriptor t of image j?", so instead of a second
// exhaustive pass j -> i (all n_j rows) the train descriptors of the survivors -- a few hundred per
// pair -- are gathered and matched against image i: +9 % tensor work instead of +100 % on the 8k sets.
__global__ void __launch_bounds__(256)
mutual_gather_kernel(const int32_t* __restrict__ pairs, int n_pairs, const long long* __restrict__ img_off,
                     const long long* __restrict__ row_off, const long long* __restrict__ m_off,
                     const int32_t* __restrict__ mt, const uint4* __restrict__ desc, const int32_t* __restrict__ norms,
                     uint4* __restrict__ q_desc, int32_t* __restrict__ q_norms) {
    const int sub = threadIdx.x & 7, grp = threadIdx.x >> 3;   // 8 threads copy one 128-byte row
    for (int p = blockIdx.x; p < n_pairs; p += gridDim.x) {
        const long long t0 = img_off[pairs[2 * p + 1]];
        const long long r0 = row_off[p], o0 = m_off[p];
        const int m = static_cast<int>(m_off[p + 1] - o0);
        for (int i = grp; i < m; i += 32) {
            const long long src = t0 + __ldg(mt + r0 + i);
            q_desc[(o0 + i) * 8 + sub] = __ldg(desc + src * 8 + sub);
            if (sub == 0) q_norms[o0 + i] = __ldg(norms + src);
        }
    }
}

__global__ void __launch_bounds__(kThreads)
mutual_compact_kernel(int n_pairs, const long long* __restrict__ row_off, const long long* __restrict__ m_off,
                      const int4* __restrict__ knn_rev, int32_t* __restrict__ mq, int32_t* __restrict__ mt,
                      int32_t* __restrict__ md2, int32_t* __restrict__ count) {
    __shared__ int warp_tot[kThreads / 32];
    __shared__ int base_s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int p = blockIdx.x; p < n_pairs; p += gridDim.x) {
        const long long r0 = row_off[p], o0 = m_off[p];
        const int m = static_cast<int>(m_off[p + 1] - o0);
        if (threadIdx.x == 0) base_s = 0;
        __syncthreads();
        // in place: entry i only ever moves to a position <= i, and a block of kThreads entries is read
        // (registers) before any of it is written
        for (int i0 = 0; i0 < m; i0 += kThreads) {
            const int i = i0 + threadIdx.x;
            int q = 0, t = 0, d = 0;
            bool keep = false;
            if (i < m) {
                q = mq[r0 + i];
                t = mt[r0 + i];
                d = md2[r0 + i];
                keep = __ldg(&knn_rev[o0 + i].x) == q;
            }
            const unsigned bal = __ballot_sync(0xffffffffu, keep);
            if (lane == 0) warp_tot[warp] = __popc(bal);
            __syncthreads();
            int before = base_s, total = 0;
#pragma unroll
            for (int w = 0; w < kThreads / 32; ++w) {
                const int c = warp_tot[w];
                if (w < warp) before += c;
                total += c;
            }
            if (keep) {
                const long long o = r0 + before + __popc(bal & ((1u << lane) - 1u));
                mq[o] = q;
                mt[o] = t;
                md2[o] = d;
            }
            __syncthreads();
            if (threadIdx.x == 0) base_s += total;
            __syncthreads();
        }
        if (threadIdx.x == 0) count[p] = base_s;
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------
// Result packing: the per-pair arenas are sized for the worst case (one match
// per query row); before the device->host copy the kept entries of the kept
// pairs are gathered into dense arrays.  Exclusive scan of the kept counts by
// one CTA, then one CTA per pair copies its slice with coalesced accesses.
__global__ void __launch_bounds__(1024)
pack_scan_kernel(const int32_t* __restrict__ count, const int32_t* __restrict__ status, int only_ok,
                 int n_pairs, long long* __restrict__ pack_off) {
    __shared__ long long part[1024];
    const int per = (n_pairs + 1023) / 1024;
    const int lo = threadIdx.x * per, hi = min(n_pairs, lo + per);
    long long s = 0;
    for (int p = lo; p < hi; ++p) s += (!only_ok || status[p] == SFM_PAIR_OK) ? count[p] : 0;
    part[threadIdx.x] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const long long v = threadIdx.x >= o ? part[threadIdx.x - o] : 0;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    long long run = part[threadIdx.x] - s;
    for (int p = lo; p < hi; ++p) {
        pack_off[p] = run;
        run += (!only_ok || status[p] == SFM_PAIR_OK) ? count[p] : 0;
    }
    if (threadIdx.x == 1023) pack_off[n_pairs] = part[1023];
}

__global__ void pack_gather_kernel(int n_pairs, const long long* __restrict__ row_off,
                                   const long long* __restrict__ pack_off, PackArrays a) {
    for (int p = blockIdx.x; p < n_pairs; p += gridDim.x) {
        const long long src = row_off[p], dst = pack_off[p];
        const int n = static_cast<int>(pack_off[p + 1] - dst);
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            if (a.mq) a.o_mq[dst + i] = a.mq[src + i];
            if (a.mt) a.o_mt[dst + i] = a.mt[src + i];
            if (a.md2) a.o_md2[dst + i] = a.md2[src + i];
            if (a.mask) a.o_mask[dst + i] = a.mask[src + i];
            if (a.sym) a.o_sym[dst + i] = a.sym[src + i];
            if (a.pts1) reinterpret_cast<float2*>(a.o_pts1)[dst + i] = reinterpret_cast<const float2*>(a.pts1)[src + i];
            if (a.pts2) reinterpret_cast<float2*>(a.o_pts2)[dst + i] = reinterpret_cast<const float2*>(a.pts2)[src + i];
        }
    }
}

}  // namespace

void launch_mutual_gather(const int32_t* d_pairs, int32_t n_pairs, const int64_t* d_img_off, const int64_t* d_row_off,
                          const int64_t* d_m_off, const int32_t* d_mt, const uint8_t* d_desc, const int32_t* d_norms,
                          uint8_t* d_q_desc, int32_t* d_q_norms, cudaStream_t st, int64_t* launches) {
    if (n_pairs <= 0) return;
    const int grid = n_pairs < 148 * 8 ? n_pairs : 148 * 8;
    mutual_gather_kernel<<<grid, 256, 0, st>>>(d_pairs, n_pairs, reinterpret_cast<const long long*>(d_img_off),
                                               reinterpret_cast<const long long*>(d_row_off),
                                               reinterpret_cast<const long long*>(d_m_off), d_mt,
                                               reinterpret_cast<const uint4*>(d_desc), d_norms,
                                               reinterpret_cast<uint4*>(d_q_desc), d_q_norms);
    ++*launches;
}

void launch_mutual_compact(int32_t n_pairs, const int64_t* d_row_off, const int64_t* d_m_off, const int4* d_knn_rev,
                           int32_t* d_mq, int32_t* d_mt, int32_t* d_md2, int32_t* d_count, cudaStream_t st,
                           int64_t* launches) {
    if (n_pairs <= 0) return;
    const int grid = n_pairs < 148 * 8 ? n_pairs : 148 * 8;
    mutual_compact_kernel<<<grid, kThreads, 0, st>>>(n_pairs, reinterpret_cast<const long long*>(d_row_off),
                                                     reinterpret_cast<const long long*>(d_m_off), d_knn_rev, d_mq, d_mt,
                                                     d_md2, d_count);
    ++*launches;
}

void launch_filter(const int32_t* d_pairs, int32_t n_pairs, const int64_t* d_img_off,
                   const int64_t* d_row_off, const int4* d_knn, const int64_t* d_row_off_rev,
                   const int4* d_knn_rev, double ratio, int32_t* d_mq, int32_t* d_mt,
                   int32_t* d_md2, int32_t* d_count, cudaStream_t st, int64_t* launches) {
    if (n_pairs <= 0) return;
    const int grid = n_pairs < 148 * 8 ? n_pairs : 148 * 8;
    filter_kernel<<<grid, kThreads, 0, st>>>(
        d_pairs, n_pairs, reinterpret_cast<const long long*>(d_img_off),
        reinterpret_cast<const long long*>(d_row_off), d_knn,
        reinterpret_cast<const long long*>(d_row_off_rev), d_knn_rev, ratio, d_mq, d_mt, d_md2,
        d_count);
    ++*launches;
}

void launch_pack(int32_t n_pairs, const int64_t* d_row_off, const int32_t* d_count,
                 const int32_t* d_status, int only_ok, int64_t* d_pack_off, const PackArrays& a,
                 cudaStream_t st, int64_t* launches) {
    if (n_pairs <= 0) return;
    pack_scan_kernel<<<1, 1024, 0, st>>>(d_count, d_status, only_ok, n_pairs,
                                         reinterpret_cast<long long*>(d_pack_off));
    ++*launches;
    const int grid = n_pairs < 148 * 8 ? n_pairs : 148 * 8;
    pack_gather_kernel<<<grid, 256, 0, st>>>(n_pairs, reinterpret_cast<const long long*>(d_row_off),
                                             reinterpret_cast<const long long*>(d_pack_off), a);
    ++*launches;
}

}  // namespace sfm
