// K1, TS form (A operand in tensor memory) -- probes and debug tile only.  A kNN kernel on this form
// (query rows resident in TMEM, three 128-column accumulator stages, 32-column units) was built and
// measured in round 2: bit-exact, but 1735 TOP/s against 2205 for the wide kernel on the same box -- its
// MMA pipeline alone reached 3065 TOP/s (wide: 3470) because three stages of 320 tensor cycles do not cover
// the ~800-cycle hand-over round trip, and one issuing warp needs ~400 cycles of serial work per group.
// The kernel was removed again; what it taught went into knn_pair.cu (four stages, two issuing warps).
//
// tcgen05.mma can take its A operand from tensor memory ("TS" form): then only the B tile is read
// from shared memory per instruction (N = 128: 4 KiB per 64 cycles = 64 B/clk instead of the
// 128 B/clk an SS-mode 128x128x32 needs, which holds that shape at 74 % of the tensor pipe --
// profiles/r01_probe_mma.json).  With the 256 query rows of a work item resident in TMEM
// (2 x 32 columns + 8 columns for the constant fold-step operand) the remaining 440 columns hold
// three 128-column (or four 96-column) accumulator stages instead of two 256-column ones: a
// finer MMA -> epilogue -> MMA hand-over at the full tensor rate.
#include <climits>

#include "kernels.h"
#include "ptx.cuh"

namespace sfm {

namespace {

constexpr int kRowBytes = SFM_DESC_DIM;
constexpr int kUmmaK = 32;
constexpr int kKSteps = kRowBytes / kUmmaK;   // 4
constexpr int kTileM = 128;

// K-major operand whose rows are 32 bytes (one UMMA K step), SWIZZLE_32B: 8-row groups of 256 bytes.
__device__ __forceinline__ uint64_t make_kmajor_sw32_desc_ts(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
    d |= static_cast<uint64_t>(1) << 16;
    d |= static_cast<uint64_t>(256 >> 4) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(6) << 61;
    return d;
}

// TMEM column map shared by the probe, the debug tile and the kernel.
constexpr uint32_t kAccCols = 384;          // accumulator stages: 3 x 128 or 4 x 96
constexpr uint32_t kAColBase = kAccCols;    // query block a: columns [384 + 32 a, +32)
constexpr uint32_t kAxCol = kAccCols + 64;  // constant fold-step operand: 8 columns
constexpr uint32_t kTmemCols = 512;

// Debug: folded accumulators of one 128 x 128 tile through the TS path: A rows by tcgen05.st from
// registers, B tile + fold operand by TMA, four K steps + the fold step with A in TMEM.
__global__ void __launch_bounds__(128, 1)
debug_ts_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_ax,
                const uint8_t* __restrict__ desc, const uint8_t* __restrict__ bias32, int q_row, int t_row,
                int32_t* __restrict__ out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    uint8_t* sb = base;                         // 128 x 128 B, SWIZZLE_128B
    uint8_t* sx = sb + kTileM * kRowBytes;      // 128 x 32 B, SWIZZLE_32B
    uint64_t* bar_ld = reinterpret_cast<uint64_t*>(sx + kTileM * kUmmaK);
    uint64_t* bar_mma = bar_ld + 1;
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar_mma + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(bar_ld, 1);
        mbar_init(bar_mma, 1);
        fence_barrier_init();
    }
    if (warp == 0) {
        tmem_alloc(slot, kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *slot;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);
    // A: this thread's query row, 32 words -> 32 TMEM columns of its lane
    {
        const uint4* src = reinterpret_cast<const uint4*>(desc + static_cast<size_t>(q_row + warp * 32 + lane) * kRowBytes);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            uint32_t v[16];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint4 w = __ldg(src + 4 * h + i);
                v[4 * i] = w.x; v[4 * i + 1] = w.y; v[4 * i + 2] = w.z; v[4 * i + 3] = w.w;
            }
            tmem_st_32x16(lane_base + kAColBase + 16 * h, v);
        }
        uint32_t c[8];
        const uint4* cs = reinterpret_cast<const uint4*>(bias32);   // rows [0,128) of the fold array are all the same
        const uint4 c0 = __ldg(cs), c1 = __ldg(cs + 1);
        c[0] = c0.x; c[1] = c0.y; c[2] = c0.z; c[3] = c0.w; c[4] = c1.x; c[5] = c1.y; c[6] = c1.z; c[7] = c1.w;
        tmem_st_32x8(lane_base + kAxCol, c);
        tmem_st_wait();
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) {
        tc_fence_after();
        mbar_arrive_expect_tx(bar_ld, kTileM * kRowBytes + kTileM * kUmmaK);
        tma_load_2d(sb, &tmap, bar_ld, 0, t_row);
        tma_load_2d(sx, &tmap_ax, bar_ld, 0, kTileM + t_row);
        mbar_wait(bar_ld, 0);
        tc_fence_after();
        const uint64_t bd = make_kmajor_sw128_desc(smem_u32(sb));
        constexpr uint32_t idesc = make_idesc_u8(kTileM, 128);
        for (int k = 0; k < kKSteps; ++k)
            mma_i8_ts(tmem_base, tmem_base + kAColBase + 8 * k, bd + uint64_t(k) * (kUmmaK >> 4), idesc, k > 0);
        mma_i8_ts(tmem_base, tmem_base + kAxCol, make_kmajor_sw32_desc_ts(smem_u32(sx)), idesc, 1);
        mma_commit(bar_mma);
    }
    __syncwarp();
    mbar_wait(bar_mma, 0);
    tc_fence_after();
    const int row = warp * 32 + lane;
    for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(lane_base + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) out[row * 128 + c * 32 + j] = static_cast<int32_t>(v[j]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// Issue-rate probe: `groups` groups of 5 MMAs (128 x N x 32, u8), accumulators rotating over
// 384 / N stages, no data movement (operands are whatever TMEM / shared memory hold).
//   ts = 1: A from TMEM (columns 384..455), 0: A from shared memory.
//   commit = 1: one tcgen05.commit per group to a barrier nobody waits on.
template <int N>
__global__ void __launch_bounds__(128, 1) ts_rate_probe_kernel(int groups, int ts, int commit) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    constexpr int kTile = 128 * 128;
    uint8_t* sa = base;              // 2 A tiles
    uint8_t* sb = base + 2 * kTile;  // 4 x 128 rows of B
    uint64_t* bar = reinterpret_cast<uint64_t*>(base + 6 * kTile);
    uint64_t* other = bar + 1;
    uint32_t* slot = reinterpret_cast<uint32_t*>(other + 1);
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 6 * kTile / 4; i += blockDim.x)
        reinterpret_cast<uint32_t*>(base)[i] = 0x01010101u * (i & 3);
    fence_proxy_async();
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        mbar_init(other, 1);
        fence_barrier_init();
    }
    if (warp == 0) {
        tmem_alloc(slot, kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *slot;
    constexpr uint32_t idesc = make_idesc_u8(128, N);
    constexpr int kStages = N <= 128 ? static_cast<int>(kAccCols) / N : 1;
    if (warp == 0) {
        int stage = 0;
        for (int g = 0; g < groups; ++g) {
            const int a = g & 1;
            const int bsel = N <= 128 ? ((g >> 1) & 3) : ((g >> 1) & 1) * 2;
            const uint64_t bd = make_kmajor_sw128_desc(smem_u32(sb + bsel * kTile));
            const uint64_t ad = make_kmajor_sw128_desc(smem_u32(sa + a * kTile));
            const uint32_t d = tmem_base + stage * N;
            const uint32_t ta = tmem_base + kAColBase + 32 * a;
            if (elect_one()) {
                if (ts) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) mma_i8_ts(d, ta + 8 * k, bd + uint64_t(k) * 2, idesc, k > 0);
                    mma_i8_ts(d, tmem_base + kAxCol, bd, idesc, 1);
                } else {
#pragma unroll
                    for (int k = 0; k < 5; ++k) mma_i8_ss(d, ad + uint64_t(k & 3) * 2, bd + uint64_t(k & 3) * 2, idesc, k > 0);
                }
                if (commit) mma_commit(other);
            }
            __syncwarp();
            if (++stage == kStages) stage = 0;
        }
        if (elect_one()) mma_commit(bar);
        __syncwarp();
        mbar_wait(bar, 0);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}



// Hand-over latency probe (one CTA): warp 0 issues `n_mma` MMAs (128 x N x 32, A in TMEM) and a
// tcgen05.commit to `full`; warp 1 waits for it, takes a time stamp and arrives on `empty`; warp 0 waits
// for that and takes a time stamp.  out[0] = mean cycles issue start -> issue done (commit issued),
// out[1] = issue start -> warp 1 past its wait, out[2] = warp 1's arrive -> warp 0 past its wait,
// out[3] = same-thread variant: issue start -> warp 0 itself past a wait on `full` (iters rounds each).
template <int N>
__global__ void __launch_bounds__(64, 1) handover_probe_kernel(int iters, int n_mma, int trail, long long* __restrict__ out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    uint8_t* sb = base;   // 256 rows x 128 B of B (contents irrelevant)
    uint64_t* full = reinterpret_cast<uint64_t*>(base + 256 * 128);
    uint64_t* empty = full + 1;
    uint32_t* slot = reinterpret_cast<uint32_t*>(empty + 1);
    volatile long long* ts = reinterpret_cast<volatile long long*>(slot + 4);   // [0] seen by warp 1, [1] arrive time
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 256 * 128 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(base)[i] = 0x01010101u;
    fence_proxy_async();
    if (threadIdx.x == 0) {
        mbar_init(full, 1);
        mbar_init(empty, 1);
        fence_barrier_init();
    }
    if (warp == 0) {
        tmem_alloc(slot, kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *slot;
    constexpr uint32_t idesc = make_idesc_u8(128, N);
    constexpr uint32_t idesc_t = make_idesc_u8(128, 128);
    const uint64_t bd = make_kmajor_sw128_desc(smem_u32(sb));
    long long s_issue = 0, s_seen = 0, s_back = 0, s_self = 0;
    if (warp == 0) {
        uint32_t ph = 0;
        for (int it = 0; it < 2 * iters; ++it) {
            const bool cross = it < iters;
            const long long t0 = clock64();
            if (elect_one()) {
                for (int k = 0; k < n_mma; ++k)
                    mma_i8_ts(tmem_base, tmem_base + kAColBase + 8 * (k & 3), bd + uint64_t(k & 3) * 2, idesc, k > 0);
                mma_commit(full);
                for (int k = 0; k < trail; ++k)   // later MMAs into another accumulator: must not delay `full`
                    mma_i8_ts(tmem_base + 128, tmem_base + kAColBase + 8 * (k & 3), bd + uint64_t(k & 3) * 2, idesc_t, k > 0);
            }
            __syncwarp();
            const long long t1 = clock64();
            if (cross) {
                mbar_wait(empty, ph);
                const long long t2 = clock64();
                s_issue += t1 - t0;
                s_seen += ts[0] - t0;
                s_back += t2 - ts[1];
            } else {
                mbar_wait(full, ph);
                const long long t2 = clock64();
                s_self += t2 - t0;
                if (lane == 0) mbar_arrive(empty);   // keeps warp 1's phases in step
                mbar_wait(empty, ph);
            }
            ph ^= 1;
        }
        if (lane == 0) {
            out[0] = s_issue / iters;
            out[1] = s_seen / iters;
            out[2] = s_back / iters;
            out[3] = s_self / iters;
        }
    } else {
        uint32_t ph = 0;
        for (int it = 0; it < 2 * iters; ++it) {
            if (it < iters) {
                mbar_wait(full, ph);
                if (lane == 0) {
                    ts[0] = clock64();
                    ts[1] = clock64();
                    __threadfence_block();
                    mbar_arrive(empty);
                }
                __syncwarp();
            }
            ph ^= 1;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

}  // namespace

int launch_handover_probe(int iters, int n_mma, int tile_n, int trail, long long* d_out, cudaStream_t st, int64_t* launches) {
    const size_t smem = 256 * 128 + 256 + 1024;
    cudaError_t e = cudaSuccess;
    if (tile_n == 256) {
        e = cudaFuncSetAttribute(handover_probe_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e == cudaSuccess) handover_probe_kernel<256><<<1, 64, smem, st>>>(iters, n_mma, trail, d_out);
    } else {
        e = cudaFuncSetAttribute(handover_probe_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e == cudaSuccess) handover_probe_kernel<128><<<1, 64, smem, st>>>(iters, n_mma, trail, d_out);
    }
    ++*launches;
    return static_cast<int>(e != cudaSuccess ? e : cudaGetLastError());
}

int launch_debug_ts(const CUtensorMap& tmap, const CUtensorMap& tmap_ax, const uint8_t* d_desc,
                    const uint8_t* d_bias32, int32_t q_row, int32_t t_row, int32_t* d_out, cudaStream_t st,
                    int64_t* launches) {
    const size_t smem = kTileM * kRowBytes + kTileM * kUmmaK + 64 + 1024;
    cudaError_t e = cudaFuncSetAttribute(debug_ts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    debug_ts_kernel<<<1, 128, smem, st>>>(tmap, tmap_ax, d_desc, d_bias32, q_row, t_row, d_out);
    ++*launches;
    return static_cast<int>(cudaGetLastError());
}

int launch_ts_rate_probe(int groups, int tile_n, int ts, int commit, int num_sms, cudaStream_t st,
                         int64_t* launches) {
    const size_t smem = 6 * 128 * 128 + 128 + 1024;
    cudaError_t e = cudaSuccess;
#define SFM_TS_PROBE(NN)                                                                                   \
    if (tile_n == NN) {                                                                                    \
        e = cudaFuncSetAttribute(ts_rate_probe_kernel<NN>, cudaFuncAttributeMaxDynamicSharedMemorySize,    \
                                 static_cast<int>(smem));                                                  \
        if (e != cudaSuccess) return static_cast<int>(e);                                                  \
        ts_rate_probe_kernel<NN><<<num_sms, 128, smem, st>>>(groups, ts, commit);                          \
    }
    SFM_TS_PROBE(64)
    SFM_TS_PROBE(96)
    SFM_TS_PROBE(128)
    SFM_TS_PROBE(192)
    SFM_TS_PROBE(256)
#undef SFM_TS_PROBE
    ++*launches;
    return static_cast<int>(cudaGetLastError());
}

}  // namespace sfm
