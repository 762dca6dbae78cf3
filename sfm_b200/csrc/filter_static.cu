// K2 without allocation: the Lowe ratio test + order-preserving compaction of filter.cu
// (reference utils/find_matches.py:150-153) for callers that may not call cudaMalloc (CUDA graph
// capture around the call site, tight serving loops, harnesses that forbid it).  No counterpart in
// the reference.
//
// sfm_filter_matches stages its per-call metadata (pair list, row offsets) in device buffers that
// grow with cudaMalloc and through pinned slots that grow with cudaMallocHost.  Here the metadata
// of up to SFM_FILTER_STATIC_MAX_PAIRS pairs lives in ONE file-scope __device__ array (16 bytes
// per pair, 256 KiB, held by every GPU that loads the library until the process exits) and is
// staged from a file-scope host array of the same size: the call allocates nothing on either side.
//
// Sharing: every call on a device writes the same array.  A call rewrites the n_pairs records its
// kernel reads and reads nothing else, so no part of the array has to start at zero and a smaller
// batch after a larger one never sees the larger one's tail.  Calls that overlap ON THE GPU (two
// streams, or two threads on one device) race on the array: the second copy may land while the
// first kernel still reads.  The host side is serialised by a mutex; the device side is the
// caller's to order (one stream, or an event between the calls).
#include <mutex>

#include "kernels.h"

namespace sfm {

namespace {

constexpr int kThreads = 256;
constexpr int kRowsPerThread = 4;
constexpr int kMaxDevices = 64;

struct PairRec {
    long long row_off;   // first arena row of the pair (kNN rows in, match lists out)
    int nq, nt;          // rows of the query / train image
};
static_assert(sizeof(PairRec) == 16, "16 bytes per pair");

// In host code this name is only a handle for cudaGetSymbolAddress / cudaMemcpyToSymbol: it is NOT a
// device address (see scratch_for_device below).
static __device__ __align__(256) PairRec g_pair_rec[SFM_FILTER_STATIC_MAX_PAIRS];

PairRec g_host_rec[SFM_FILTER_STATIC_MAX_PAIRS];   // pageable staging, guarded by g_mu
PairRec* g_dev_ptr[kMaxDevices];                    // device address of g_pair_rec, per device ordinal
std::mutex g_mu;

// Set-up, once per device, with that device current: the device address of the scratch array.
cudaError_t scratch_for_device(int device, PairRec** out) {
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    if (!g_dev_ptr[device]) {
        int cur = -1;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess) return e;
        if (cur != device) return cudaErrorInvalidDevice;
        PairRec* p = nullptr;
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_pair_rec);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_dev_ptr[device] = p;
    }
    *out = g_dev_ptr[device];
    return cudaSuccess;
}

__device__ __forceinline__ bool ratio_pass(int d1, int d2, double ratio) {
    const float s1 = __fsqrt_rn(static_cast<float>(d1));
    const float s2 = __fsqrt_rn(static_cast<float>(d2));
    return static_cast<double>(s1) < ratio * static_cast<double>(s2);
}

// Same tiling as filter_kernel (filter.cu): 4 x 256 rows per tile, one barrier per tile, the 32
// per-(k, warp) counts scanned by every warp with shuffles.  Per-pair metadata comes from `rec`.
__global__ void __launch_bounds__(kThreads)
filter_static_kernel(const PairRec* __restrict__ rec, int n_pairs, const int4* __restrict__ knn, double ratio,
                     int32_t* __restrict__ mq, int32_t* __restrict__ mt, int32_t* __restrict__ md2,
                     int32_t* __restrict__ count) {
    __shared__ int wc[2][32];
    static_assert(kRowsPerThread * (kThreads / 32) == 32, "one count per lane of the scanning warp");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int it = 0;
    for (int p = blockIdx.x; p < n_pairs; p += gridDim.x) {
        const PairRec pr = rec[p];
        const bool live = pr.nt >= 2;
        int base = 0;
        for (int q0 = 0; q0 < pr.nq; q0 += kRowsPerThread * kThreads, ++it) {
            int4 r[kRowsPerThread];
            unsigned bal[kRowsPerThread];
            bool keep[kRowsPerThread];
#pragma unroll
            for (int k = 0; k < kRowsPerThread; ++k) {
                const int q = q0 + k * kThreads + threadIdx.x;
                r[k] = make_int4(-1, -1, 0, 0);
                if (live && q < pr.nq) r[k] = __ldg(knn + pr.row_off + q);
            }
#pragma unroll
            for (int k = 0; k < kRowsPerThread; ++k) {
                keep[k] = r[k].x >= 0 && r[k].y >= 0 && ratio_pass(r[k].z, r[k].w, ratio);
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
                    const long long o = pr.row_off + base + before + __popc(bal[k] & ((1u << lane) - 1u));
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

}  // namespace

cudaError_t launch_filter_static(int device, const int32_t* h_pairs, int32_t n_pairs, const int64_t* h_img_off,
                                 const int64_t* h_row_off, const int4* d_knn, double ratio, int32_t* d_mq,
                                 int32_t* d_mt, int32_t* d_md2, int32_t* d_count, cudaStream_t st,
                                 int64_t* launches) {
    if (n_pairs < 0 || n_pairs > SFM_FILTER_STATIC_MAX_PAIRS) return cudaErrorInvalidValue;
    std::lock_guard<std::mutex> lock(g_mu);
    PairRec* d_rec = nullptr;
    cudaError_t e = scratch_for_device(device, &d_rec);
    if (e != cudaSuccess || n_pairs == 0) return e;
    for (int32_t p = 0; p < n_pairs; ++p) {
        const int32_t qi = h_pairs[2 * p], ti = h_pairs[2 * p + 1];
        g_host_rec[p].row_off = h_row_off[p];
        g_host_rec[p].nq = static_cast<int>(h_img_off[qi + 1] - h_img_off[qi]);
        g_host_rec[p].nt = static_cast<int>(h_img_off[ti + 1] - h_img_off[ti]);
    }
    // From pageable memory the copy has consumed g_host_rec when the call returns (the driver stages it),
    // so the next call may refill the array; it is ordered on `st` before the kernel.
    e = cudaMemcpyAsync(d_rec, g_host_rec, sizeof(PairRec) * static_cast<size_t>(n_pairs), cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return e;
    const int grid = n_pairs < 148 * 8 ? n_pairs : 148 * 8;
    filter_static_kernel<<<grid, kThreads, 0, st>>>(d_rec, n_pairs, d_knn, ratio, d_mq, d_mt, d_md2, d_count);
    ++*launches;
    return cudaGetLastError();
}

}  // namespace sfm
