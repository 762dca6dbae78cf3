// K1: exhaustive 128-d L2 kNN (k = 2) on the 5th-gen tensor cores.
//
// Replaces cv2.BFMatcher(NORM_L2).knnMatch(desc1, desc2, k=2)
// (reference utils/find_matches.py:144-147).  SIFT descriptors are integers
// 0..255, so ||a-b||^2 = ||a||^2 + ||b||^2 - 2 a.b is computed exactly:
// a.b as u8 x u8 -> s32 tcgen05.mma tiles accumulated in TMEM.  The top-2
// selection is fused into the TMEM epilogue, so the distance matrix never
// reaches HBM.
//
// One persistent CTA per SM, warp-specialised:
//   warps 0-7  epilogue: thread <-> one query row (TMEM lane); running
//              (best, second) in registers over all train tiles
//   warp 8     TMA producer (A block: 2 x 128 query rows, B ring: 128 train rows)
//   warp 9     TMEM allocator + single-thread tcgen05.mma issuer
// A work item is 256 query rows x one whole train image; the two 128x128
// accumulators of a train tile form one TMEM stage (256 columns), two stages
// are in flight so the epilogue of tile t overlaps the MMAs of tile t+1.
//
// This is the GENERIC variant: keys ||b||^2 - 2 a.b are formed per element from a padded column-norm
// array (4.2 thread instructions per element, 822 TOP/s).  It is the fallback for descriptor sets whose
// norms exceed the range of the folded kernels (knn_wide.cu, knn_pair.cu: the norm term rides in a fifth
// K step, the epilogue is a max tree); sfm_set_images raises a device flag for such sets, both kernels are
// launched and the one that does not apply exits at its first instruction.  This file also owns the
// norms, the work-item builder and the plain dot-product debug tile.
//
// Tie rule (== OpenCV): columns are visited in ascending train
// index and a candidate replaces an incumbent only on strict comparison, so
// the lowest index wins ties for both the best and the second best.
#include <climits>

#include "kernels.h"
#include "ptx.cuh"

namespace sfm {

namespace {

constexpr int kTileM = 128;                         // UMMA M
constexpr int kATiles = SFM_QBLOCK / kTileM;        // 2 accumulators per stage
constexpr int kTileN = SFM_TTILE;                   // UMMA N
constexpr int kRowBytes = SFM_DESC_DIM;             // K extent in bytes (u8)
constexpr int kTileBytes = kTileM * kRowBytes;      // 16 KiB (A tile == B tile)
constexpr int kABlockBytes = kATiles * kTileBytes;  // 32 KiB
constexpr int kAStages = 2;
constexpr int kBStages = 4;
constexpr int kTmemStages = 2;
constexpr int kTmemColsPerStage = kATiles * kTileN;  // 256
constexpr int kTmemCols = kTmemStages * kTmemColsPerStage;  // 512 = all of TMEM
constexpr int kEpiWarps = 8;                         // one thread per query row
constexpr int kThreads = (kEpiWarps + 2) * 32;
constexpr int kUmmaK = 32;                           // bytes of K per tcgen05.mma kind::i8
constexpr int kKSteps = kRowBytes / kUmmaK;          // 4
constexpr uint32_t kIdesc = make_idesc_u8(kTileM, kTileN);
constexpr int kPadNorm = 1 << 30;                    // column norm of a padding column
constexpr size_t kSmemData = size_t(kAStages) * kABlockBytes + size_t(kBStages) * kTileBytes;
constexpr int kNumBars = 2 * kAStages + 2 * kBStages + 4 * kTmemStages;
constexpr size_t kSmemBytes = kSmemData + kNumBars * 8 + 16 + 1024;  // + alignment slack
static_assert(kSmemBytes <= 232448, "shared memory budget");

struct Pipe {
    int stage = 0;
    uint32_t phase = 0;
    int n;
    __device__ explicit Pipe(int n_) : n(n_) {}
    __device__ void advance() {
        if (++stage == n) {
            stage = 0;
            phase ^= 1;
        }
    }
};

// ------------------------------------------------------------------ generic epilogue
struct Top2 {
    int d1, d2, i1, i2;
};

__device__ __forceinline__ void top2_insert(Top2& s, int key, int col) {
    if (key < s.d2) {
        if (key < s.d1) {
            s.d2 = s.d1;
            s.i2 = s.i1;
            s.d1 = key;
            s.i1 = col;
        } else {
            s.d2 = key;
            s.i2 = col;
        }
    }
}

// 32 accumulator columns of one row: keys = ||b||^2 - 2 a.b ; a group of 8
// keys is scanned one by one only when its minimum beats the current second best.
__device__ __forceinline__ void top2_chunk(Top2& s, const uint32_t (&acc)[32], const int4 (&nb)[8],
                                           int col0) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        int k[8];
        const int4 n0 = nb[2 * g], n1 = nb[2 * g + 1];
        k[0] = n0.x - 2 * static_cast<int>(acc[8 * g + 0]);
        k[1] = n0.y - 2 * static_cast<int>(acc[8 * g + 1]);
        k[2] = n0.z - 2 * static_cast<int>(acc[8 * g + 2]);
        k[3] = n0.w - 2 * static_cast<int>(acc[8 * g + 3]);
        k[4] = n1.x - 2 * static_cast<int>(acc[8 * g + 4]);
        k[5] = n1.y - 2 * static_cast<int>(acc[8 * g + 5]);
        k[6] = n1.z - 2 * static_cast<int>(acc[8 * g + 6]);
        k[7] = n1.w - 2 * static_cast<int>(acc[8 * g + 7]);
        const int m = min(__vimin3_s32(__vimin3_s32(k[0], k[1], k[2]), __vimin3_s32(k[3], k[4], k[5]),
                                       k[6]),
                          k[7]);
        if (m < s.d2) {
#pragma unroll
            for (int e = 0; e < 8; ++e) top2_insert(s, k[e], col0 + 8 * g + e);
        }
    }
}

struct SmemLayout {
    uint8_t* a;
    uint8_t* b;
    uint64_t* a_full;
    uint64_t* a_empty;
    uint64_t* b_full;
    uint64_t* b_empty;
    uint64_t* t_full;
    uint64_t* t_empty;
    uint32_t* tmem_slot;
    __device__ explicit SmemLayout(uint8_t* raw) {
        uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) &
                                                   ~static_cast<uintptr_t>(1023));
        a = base;
        b = a + size_t(kAStages) * kABlockBytes;
        uint64_t* bars = reinterpret_cast<uint64_t*>(base + kSmemData);
        a_full = bars;
        a_empty = a_full + kAStages;
        b_full = a_empty + kAStages;
        b_empty = b_full + kBStages;
        t_full = b_empty + kBStages;
        t_empty = t_full + 2 * kTmemStages;   // [stage][accumulator]
        tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2 * kTmemStages);
    }
};

// `flags[0]` != 0: some registered row's norm is beyond what the folded kernels' fifth K step can
// encode.  only_if_flag: run only then (the folded kernel launched beside this one handles the rest).
__global__ void __launch_bounds__(kThreads, 1)
knn2_kernel(const __grid_constant__ CUtensorMap tmap, const WorkItem* __restrict__ work, long long n_work,
            const int32_t* __restrict__ norms, const int32_t* __restrict__ norms_pad,
            const int32_t* __restrict__ flags, int only_if_flag, uint32_t wait_ns, uint32_t epi_wait_ns,
            uint32_t probe_flags, int4* __restrict__ out) {
    constexpr int kProducerWarp = kEpiWarps;
    constexpr int kMmaWarp = kEpiWarps + 1;
    if (only_if_flag && *static_cast<volatile const int32_t*>(flags) == 0) return;

    extern __shared__ uint8_t smem_raw[];
    SmemLayout sm(smem_raw);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == kProducerWarp && lane == 0) {
        tma_prefetch_desc(&tmap);
        for (int i = 0; i < kAStages; ++i) {
            mbar_init(&sm.a_full[i], 1);
            mbar_init(&sm.a_empty[i], 1);
        }
        for (int i = 0; i < kBStages; ++i) {
            mbar_init(&sm.b_full[i], 1);
            mbar_init(&sm.b_empty[i], 1);
        }
        for (int i = 0; i < 2 * kTmemStages; ++i) {   // [stage][accumulator]
            mbar_init(&sm.t_full[i], 1);
            mbar_init(&sm.t_empty[i], kEpiWarps / kATiles);
        }
        fence_barrier_init();
    }
    if (warp == kMmaWarp) {
        tmem_alloc(sm.tmem_slot, kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *sm.tmem_slot;

    // Producer and MMA warps run their loops warp-convergent (all 32 lanes wait and keep the
    // pipeline state) and only the TMA / tcgen05 issue is predicated on one elected lane: the
    // descriptors then live in uniform registers and each tcgen05.mma is a single predicated
    // instruction.  (Under `if (lane == 0)` ptxas wraps every UTCIMMA in an R2UR + elect loop
    // and the issuing thread, not the tensor pipe, paces the kernel.)
    if (warp == kProducerWarp) {
        Pipe ap(kAStages), bp(kBStages);
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int q_row = w0.x, t_row = w0.z, t_cnt = w0.w;
            mbar_wait_suspend(&sm.a_empty[ap.stage], ap.phase ^ 1, wait_ns);
            if (elect_one()) {
                mbar_arrive_expect_tx(&sm.a_full[ap.stage], kABlockBytes);
                uint8_t* a_dst = sm.a + size_t(ap.stage) * kABlockBytes;
#pragma unroll
                for (int a = 0; a < kATiles; ++a)
                    tma_load_2d(a_dst + a * kTileBytes, &tmap, &sm.a_full[ap.stage], 0, q_row + a * kTileM);
            }
            __syncwarp();
            ap.advance();
            const int n_tiles = (t_cnt + kTileN - 1) / kTileN;
            for (int t = 0; t < n_tiles; ++t) {
                mbar_wait_suspend(&sm.b_empty[bp.stage], bp.phase ^ 1, wait_ns);
                if (elect_one()) {
                    if (probe_flags & 1u) {   // probe: no B traffic
                        mbar_arrive(&sm.b_full[bp.stage]);
                    } else {
                        mbar_arrive_expect_tx(&sm.b_full[bp.stage], kTileBytes);
                        tma_load_2d(sm.b + size_t(bp.stage) * kTileBytes, &tmap, &sm.b_full[bp.stage], 0,
                                    t_row + t * kTileN);
                    }
                }
                __syncwarp();
                bp.advance();
            }
        }
    } else if (warp == kMmaWarp) {
        Pipe ap(kAStages), bp(kBStages), tp(kTmemStages);
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int n_tiles = (w0.w + kTileN - 1) / kTileN;
            mbar_wait_suspend(&sm.a_full[ap.stage], ap.phase, wait_ns);
            const uint64_t adesc0 = make_kmajor_sw128_desc(smem_u32(sm.a + size_t(ap.stage) * kABlockBytes));
            for (int t = 0; t < n_tiles; ++t) {
                mbar_wait_suspend(&sm.b_full[bp.stage], bp.phase, wait_ns);
                const uint64_t bdesc0 = make_kmajor_sw128_desc(smem_u32(sm.b + size_t(bp.stage) * kTileBytes));
                const uint32_t d0 = tmem_base + tp.stage * kTmemColsPerStage;
#pragma unroll
                for (int a = 0; a < kATiles; ++a) {
                    // each 128-row accumulator is handed over on its own barrier pair, so the
                    // epilogue of accumulator 0 overlaps the MMAs of accumulator 1
                    mbar_wait_suspend(&sm.t_empty[tp.stage * kATiles + a], tp.phase ^ 1, wait_ns);
                    tc_fence_after();
                    const uint32_t d = d0 + a * kTileN;
                    const uint64_t adesc = adesc0 + uint64_t(a) * (kTileBytes >> 4);
                    if (elect_one()) {
                        if (!(probe_flags & 2u)) {   // (probe bit 1: no MMA issue)
#pragma unroll
                            for (int k = 0; k < kKSteps; ++k)
                                mma_i8_ss(d, adesc + uint64_t(k) * (kUmmaK >> 4),
                                          bdesc0 + uint64_t(k) * (kUmmaK >> 4), kIdesc, k > 0);
                        }
                        mma_commit(&sm.t_full[tp.stage * kATiles + a]);   // accumulator ready for the epilogue
                        if (a == kATiles - 1) mma_commit(&sm.b_empty[bp.stage]);   // B slot reusable once the MMAs retire
                    }
                    __syncwarp();
                }
                bp.advance();
                tp.advance();
            }
            if (elect_one()) mma_commit(&sm.a_empty[ap.stage]);
            __syncwarp();
            ap.advance();
        }
    } else {
        // ------------------------------------------------ epilogue warps
        const int quad = warp & 3;          // TMEM lane quadrant this warp may read
        const int acc = (warp >> 2) & 1;    // which 128-row accumulator of the stage
        const int row = acc * kTileM + quad * 32 + lane;
        const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * kTileN;
        Pipe tp(kTmemStages);
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int4 w1 = __ldg(reinterpret_cast<const int4*>(work + w) + 1);
            const int q_row = w0.x, q_cnt = w0.y, t_cnt = w0.w, t_pad = w1.x;
            const long long out_row =
                static_cast<long long>(static_cast<uint32_t>(w1.z)) | (static_cast<long long>(w1.w) << 32);
            const bool valid = row < q_cnt;
            const int na = valid ? __ldg(norms + q_row + row) : 0;
            const int n_tiles = (t_cnt + kTileN - 1) / kTileN;
            const int32_t* nbase = norms_pad + t_pad;
            Top2 s{INT_MAX, INT_MAX, -1, -1};
            for (int t = 0; t < n_tiles; ++t) {
                uint64_t* full = &sm.t_full[tp.stage * kATiles + acc];
                uint64_t* empty = &sm.t_empty[tp.stage * kATiles + acc];
                mbar_wait_suspend(full, tp.phase, epi_wait_ns);
                tc_fence_after();
                const uint32_t taddr = lane_addr + tp.stage * kTmemColsPerStage;
                uint32_t v[2][32];
                tmem_ld_32x32(taddr, v[0]);
#pragma unroll
                for (int c = 0; c < kTileN / 32; ++c) {
                    const int col0 = t * kTileN + c * 32;
                    int4 nb[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        nb[j] = __ldg(reinterpret_cast<const int4*>(nbase) + col0 / 4 + j);
                    tmem_ld_wait();
                    if (c + 1 < kTileN / 32) {
                        tmem_ld_32x32(taddr + (c + 1) * 32, v[(c + 1) & 1]);
                    } else {
                        // every column of this thread is in registers: hand the accumulator back
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(empty);
                    }
                    top2_chunk(s, v[c & 1], nb, col0);
                }
                tp.advance();
            }
            if (valid) {
                // padding columns (index >= t_cnt) can only surface when t_cnt < 2
                if (s.i1 >= t_cnt) s.i1 = -1;
                if (s.i2 >= t_cnt) s.i2 = -1;
                const int d1 = s.i1 < 0 ? INT_MAX : s.d1 + na;
                const int d2 = s.i2 < 0 ? INT_MAX : s.d2 + na;
                out[out_row + row] = make_int4(s.i1, s.i2, d1, d2);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ---------------------------------------------------------------------------
// Debug: one 128x128 tile of raw dot products through the same TMA/UMMA path.
__global__ void __launch_bounds__(128, 1)
debug_dot_kernel(const __grid_constant__ CUtensorMap tmap, int q_row, int t_row, int32_t* __restrict__ out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    uint8_t* sa = base;
    uint8_t* sb = base + kTileBytes;
    uint64_t* bar_ld = reinterpret_cast<uint64_t*>(base + 2 * kTileBytes);
    uint64_t* bar_mma = bar_ld + 1;
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar_mma + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(bar_ld, 1);
        mbar_init(bar_mma, 1);
        fence_barrier_init();
    }
    if (warp == 0) {
        tmem_alloc(slot, 128);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *slot;
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(bar_ld, 2 * kTileBytes);
        tma_load_2d(sa, &tmap, bar_ld, 0, q_row);
        tma_load_2d(sb, &tmap, bar_ld, 0, t_row);
        mbar_wait(bar_ld, 0);
        tc_fence_after();
        const uint64_t ad = make_kmajor_sw128_desc(smem_u32(sa));
        const uint64_t bd = make_kmajor_sw128_desc(smem_u32(sb));
        for (int k = 0; k < kKSteps; ++k)
            mma_i8_ss(tmem_base, ad + uint64_t(k) * (kUmmaK >> 4), bd + uint64_t(k) * (kUmmaK >> 4),
                      kIdesc, k > 0);
        mma_commit(bar_mma);
    }
    __syncwarp();
    mbar_wait(bar_mma, 0);
    tc_fence_after();
    const int row = warp * 32 + lane;
    for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(warp * 32) << 16) + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) out[row * 128 + c * 32 + j] = static_cast<int32_t>(v[j]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 128);
    }
}

// ---------------------------------------------------------------------------
// Tensor-pipe peak probe: MMA-only loop (operands are whatever is in shared
// memory; the integer pipe's rate does not depend on the data).
__global__ void __launch_bounds__(128, 1) mma_probe_kernel(int iters, int probe_n) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    uint8_t* sa = base;                   // 128 rows
    uint8_t* sb = base + kTileBytes;      // 256 rows
    uint64_t* bar = reinterpret_cast<uint64_t*>(base + 3 * kTileBytes);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 3 * kTileBytes / 4; i += blockDim.x)
        reinterpret_cast<uint32_t*>(base)[i] = 0x01010101u * (i & 3);
    fence_proxy_async();
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        fence_barrier_init();
    }
    if (warp == 0) {
        tmem_alloc(slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *slot;
    if (threadIdx.x == 0) {
        const uint64_t ad = make_kmajor_sw128_desc(smem_u32(sa));
        const uint64_t bd = make_kmajor_sw128_desc(smem_u32(sb));
        const uint32_t idesc = make_idesc_u8(128, probe_n);
        for (int it = 0; it < iters; ++it) {
            const uint32_t d = tmem_base + (it & 1) * probe_n;
#pragma unroll
            for (int k = 0; k < kKSteps; ++k)
                mma_i8_ss(d, ad + uint64_t(k) * (kUmmaK >> 4), bd + uint64_t(k) * (kUmmaK >> 4), idesc,
                          k > 0);
        }
        mma_commit(bar);
        mbar_wait(bar, 0);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------
// ||row||^2 for every descriptor row: one warp per row, 4 bytes per lane.
__global__ void norms_kernel(const uint8_t* __restrict__ desc, long long total_rows,
                             int32_t* __restrict__ norms) {
    const long long row = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= total_rows) return;
    const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(desc + row * kRowBytes) + lane);
    uint32_t s = __dp4a(w, w, 0u);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) norms[row] = static_cast<int32_t>(s);
}

// Column-norm array padded per image to a multiple of 128 entries; padding
// columns get a norm no real key can reach, so they never enter a top-2.
__global__ void pad_norms_kernel(const int32_t* __restrict__ norms,
                                 const long long* __restrict__ img_off,
                                 const long long* __restrict__ pad_off, int n_images,
                                 int32_t* __restrict__ norms_pad) {
    const int img = blockIdx.y;
    if (img >= n_images) return;
    const long long r0 = img_off[img], n = img_off[img + 1] - r0;
    const long long p0 = pad_off[img], np = pad_off[img + 1] - p0;
    for (long long c = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; c < np;
         c += static_cast<long long>(gridDim.x) * blockDim.x)
        norms_pad[p0 + c] = c < n ? norms[r0 + c] : kPadNorm;
}

__global__ void build_work_kernel(const int32_t* __restrict__ pairs, int n_pairs,
                                  const long long* __restrict__ blk_prefix,
                                  const long long* __restrict__ row_off,
                                  const long long* __restrict__ img_off,
                                  const long long* __restrict__ pad_off, WorkItem* __restrict__ work) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    const int qi = pairs[2 * p], ti = pairs[2 * p + 1];
    const long long q0 = img_off[qi], nq = img_off[qi + 1] - q0;
    const long long t0 = img_off[ti], nt = img_off[ti + 1] - t0;
    const long long w0 = blk_prefix[p], nb = blk_prefix[p + 1] - w0;
    for (long long b = 0; b < nb; ++b) {
        WorkItem it;
        it.q_row = static_cast<int32_t>(q0 + b * SFM_QBLOCK);
        const long long left = nq - b * SFM_QBLOCK;
        it.q_cnt = static_cast<int32_t>(left < SFM_QBLOCK ? left : SFM_QBLOCK);
        it.t_row = static_cast<int32_t>(t0);
        it.t_cnt = static_cast<int32_t>(nt);
        it.t_pad = static_cast<int32_t>(pad_off[ti]);
        it.x_row = 0;
        it.out_row = row_off[p] + b * SFM_QBLOCK;
        work[w0 + b] = it;
    }
}

}  // namespace

void launch_norms(const uint8_t* d_desc, int64_t total_rows, const int64_t* d_img_off,
                  const int64_t* d_pad_off, int32_t n_images, int32_t* d_norms, int32_t* d_norms_pad,
                  int32_t* d_flags, cudaStream_t st, int64_t* launches) {
    if (total_rows > 0) {
        const int threads = 256;
        const long long blocks = (total_rows * 32 + threads - 1) / threads;
        norms_kernel<<<static_cast<unsigned>(blocks), threads, 0, st>>>(d_desc, total_rows, d_norms);
        ++*launches;
    }
    cudaMemsetAsync(d_flags, 0, 4 * sizeof(int32_t), st);
    // images are processed in slabs of 65535 (grid.y limit)
    for (int i0 = 0; i0 < n_images; i0 += 65535) {
        const int cnt = n_images - i0 < 65535 ? n_images - i0 : 65535;
        dim3 grid(8, cnt);
        pad_norms_kernel<<<grid, 256, 0, st>>>(d_norms, reinterpret_cast<const long long*>(d_img_off) + i0,
                                               reinterpret_cast<const long long*>(d_pad_off) + i0, cnt,
                                               d_norms_pad);
        ++*launches;
    }
}

void launch_build_work(const int32_t* d_pairs, int32_t n_pairs, const int64_t* d_blk_prefix,
                       const int64_t* d_row_off, const int64_t* d_img_off, const int64_t* d_pad_off,
                       WorkItem* d_work, cudaStream_t st, int64_t* launches) {
    if (n_pairs <= 0) return;
    build_work_kernel<<<(n_pairs + 127) / 128, 128, 0, st>>>(
        d_pairs, n_pairs, reinterpret_cast<const long long*>(d_blk_prefix),
        reinterpret_cast<const long long*>(d_row_off), reinterpret_cast<const long long*>(d_img_off),
        reinterpret_cast<const long long*>(d_pad_off), d_work);
    ++*launches;
}

// only_if_flag != 0: the kernel exits at once unless a registered norm is beyond the folded kernels' range.
int launch_knn2(const CUtensorMap& tmap, const WorkItem* d_work, int64_t n_work, const int32_t* d_norms,
                const int32_t* d_norms_pad, const int32_t* d_flags, int only_if_flag, const KnnTuning& tn,
                int32_t* d_knn, int num_sms, cudaStream_t st, int64_t* launches) {
    if (n_work <= 0) return 0;
    const unsigned grid = static_cast<unsigned>(n_work < num_sms ? n_work : num_sms);
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(knn2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(kSmemBytes));
        if (e != cudaSuccess) return static_cast<int>(e);
        configured = true;
    }
    knn2_kernel<<<grid, kThreads, kSmemBytes, st>>>(tmap, d_work, n_work, d_norms, d_norms_pad, d_flags, only_if_flag,
                                                    tn.wait_ns, tn.epi_wait_ns, tn.probe_flags,
                                                    reinterpret_cast<int4*>(d_knn));
    ++*launches;
    return static_cast<int>(cudaGetLastError());
}

int launch_debug_dot(const CUtensorMap& tmap, int32_t q_row, int32_t t_row, int32_t* d_out, cudaStream_t st,
                     int64_t* launches) {
    const size_t smem = 2 * kTileBytes + 64 + 1024;
    cudaError_t e = cudaFuncSetAttribute(debug_dot_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    debug_dot_kernel<<<1, 128, smem, st>>>(tmap, q_row, t_row, d_out);
    ++*launches;
    return static_cast<int>(cudaGetLastError());
}

int launch_mma_probe(int32_t iters, int32_t probe_n, int num_sms, cudaStream_t st, int64_t* launches) {
    const size_t smem = 3 * kTileBytes + 64 + 1024;
    cudaError_t e = cudaFuncSetAttribute(mma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    mma_probe_kernel<<<num_sms, 128, smem, st>>>(iters, probe_n);
    ++*launches;
    return static_cast<int>(cudaGetLastError());
}

}  // namespace sfm
