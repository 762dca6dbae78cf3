// K1, wide variant: the folded kNN kernel of knn.cu with 128x256x32 tcgen05.mma
// instructions (replaces cv2.BFMatcher(NORM_L2).knnMatch(k=2), reference
// utils/find_matches.py:144-147; same exact integer arithmetic and tie rule).
//
// Why: with N = 128 an SS-mode tcgen05.mma reads 8 KiB of operands from shared
// memory per 64 cycles -- exactly the 128 B/clk the SM has -- and the measured
// issue pattern reaches 74 % of the tensor pipe (profiles/r01_probe_mma.json,
// hs_k5_f3: 3366 of 4567 TOP/s); with N = 256 it is 12 KiB per 128 cycles and
// the same pattern runs at 100 %.
//
// Layout differences to knn.cu:
//  * a train tile is 256 rows (two 128-row TMA boxes, contiguous in shared
//    memory); TMEM holds two accumulators of 256 columns (query rows 0..127 and
//    128..255 of the work item), each with its own full/empty barrier pair;
//  * the fold-step operand (digits of C - floor(||b||^2/2)) is a plain u8
//    [128 + R, 32] array parallel to the descriptor rows, loaded per train tile
//    as a 256 x 32-byte SWIZZLE_32B box next to the B tile (one barrier);
//  * candidate units are 16 columns: fewer unit events per row (they cost far
//    more than the max tree) and a 15-input tree per
//    compare.
#include <climits>

#include "kernels.h"
#include "ptx.cuh"

namespace sfm {

namespace {

constexpr int kTileM = 128;
constexpr int kATiles = SFM_QBLOCK / kTileM;         // 2 accumulators
constexpr int kTileN = 256;                          // UMMA N = train rows per tile
constexpr int kRowBytes = SFM_DESC_DIM;
constexpr int kATileBytes = kTileM * kRowBytes;      // 16 KiB
constexpr int kABlockBytes = kATiles * kATileBytes;  // 32 KiB
constexpr int kBTileBytes = kTileN * kRowBytes;      // 32 KiB
constexpr int kUmmaK = 32;
constexpr int kKSteps = kRowBytes / kUmmaK;          // 4
constexpr int kXTileBytes = kTileN * kUmmaK;         // 8 KiB: fold operand of one train tile
constexpr int kAxBytes = kTileM * kUmmaK;            // 4 KiB: constant fold-step A operand
constexpr int kAStages = 2;
constexpr int kBStages = 2;                          // (B tile + X tile) per stage
constexpr int kStageBytes = kBTileBytes + kXTileBytes;
constexpr int kTmemCols = kATiles * kTileN;          // 512
constexpr int kEpiWarps = 16;                        // 2 accumulators x 4 lane quadrants x 2 column halves
constexpr int kThreads = (kEpiWarps + 2) * 32;
constexpr int kColsPerThread = kTileN / 2;           // 128 columns of every tile per thread
constexpr uint32_t kIdesc = make_idesc_u8(kTileM, kTileN);
constexpr int kNegV = -(1 << 28);
constexpr int kFoldC = 255 * 31 * 255 + 254;

constexpr int kUnit = 16;
constexpr uint32_t kSlotRow = kEpiWarps * 32 * 16;          // bytes between a thread's uint4s
constexpr uint32_t kSlotStride = (kUnit / 4) * kSlotRow;    // bytes per slot
constexpr size_t kSlotBytes = 2 * size_t(kSlotStride);      // 64 KiB
constexpr size_t kXchgBytes = 2 * SFM_QBLOCK * 16 + kEpiWarps * 32 * 4;
constexpr size_t kSmemData = size_t(kAStages) * kABlockBytes + size_t(kBStages) * kStageBytes + kAxBytes +
                             kSlotBytes + kXchgBytes;
constexpr int kNumBars = 2 * kAStages + 2 * kBStages + 2 * kATiles + 1;
constexpr size_t kSmemBytes = kSmemData + kNumBars * 8 + 16 + 1024;
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

// K-major operand whose rows are 32 bytes (one UMMA K step), written by TMA with
// SWIZZLE_32B: 8-row groups of 256 bytes.  Layout type 6 = SWIZZLE_32B.
__device__ __forceinline__ uint64_t make_kmajor_sw32_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
    d |= static_cast<uint64_t>(1) << 16;
    d |= static_cast<uint64_t>(256 >> 4) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(6) << 61;
    return d;
}

// ---- folded epilogue state (see knn.cu for the derivation; units are 16 columns here)
struct Cand {
    int m1, m2;   // units: largest / second largest unit maximum.   exact: k1 / entry threshold on v
    int c1, c2;   // units: first column of those units (-1: none).  exact: i1 / i2
    int s1;       // units: shared-memory slot (0/1) of unit c1.     exact: 2
    int x;        //                                                  exact: k2
    int thr;      // entry threshold on a unit maximum
};

struct Top2K {
    int k1, k2, i1, i2;
};

__device__ __forceinline__ void top2k_insert(Top2K& s, int key, int col) {
    if (key > s.k2) {
        if (key > s.k1) {
            s.k2 = s.k1;
            s.i2 = s.i1;
            s.k1 = key;
            s.i1 = col;
        } else {
            s.k2 = key;
            s.i2 = col;
        }
    }
}

__device__ __forceinline__ bool key_before(int ka, int ia, int kb, int ib) {
    return ka > kb || (ka == kb && ia < ib);
}

// Exact top-2 of the (at most) two parked units, visited in ascending column order.
__device__ __noinline__ Top2K resolve_units(Cand s, uint32_t slots, const int32_t* __restrict__ nbase,
                                            int t_cnt) {
    Top2K r{INT_MIN, INT_MIN, -1, -1};
    const bool second_first = s.c2 >= 0 && s.c2 < s.c1;
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        const bool take2 = (pass == 0) == second_first;
        const int c = take2 ? s.c2 : s.c1;
        if (c < 0) continue;
        const uint32_t src = slots + (take2 ? (s.s1 ^ 1) : s.s1) * kSlotStride;
#pragma unroll
        for (int g = 0; g < kUnit / 4; ++g) {
            const uint4 q = lds128(src + g * kSlotRow);
            const int4 nb = __ldg(reinterpret_cast<const int4*>(nbase + c) + g);
            const int vv[4] = {static_cast<int>(q.x), static_cast<int>(q.y), static_cast<int>(q.z),
                               static_cast<int>(q.w)};
            const int pp[4] = {nb.x & 1, nb.y & 1, nb.z & 1, nb.w & 1};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int col = c + 4 * g + e;
                if (col < t_cnt) top2k_insert(r, 2 * vv[e] - (pp[e] + 2 * kFoldC), col);
            }
        }
    }
    return r;
}

struct UnitVals {
    int v[kUnit];
};

// Rare path: element-wise exact update of one row with one unit (switching the row to exact
// mode first if it was still tracking units).
__device__ __noinline__ Cand exact_unit(Cand s, UnitVals in, int col, uint32_t slots,
                                        const int32_t* __restrict__ nbase, int t_cnt) {
    Top2K r;
    if (s.s1 < 2) {
        r = resolve_units(s, slots, nbase, t_cnt);
    } else {
        r = Top2K{s.m1, s.x, s.c1, s.c2};
    }
#pragma unroll
    for (int g = 0; g < kUnit / 4; ++g) {
        const int4 nb = __ldg(reinterpret_cast<const int4*>(nbase + col) + g);
        const int pp[4] = {nb.x & 1, nb.y & 1, nb.z & 1, nb.w & 1};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int c = col + 4 * g + e;
            if (c < t_cnt) top2k_insert(r, 2 * in.v[4 * g + e] - (pp[e] + 2 * kFoldC), c);
        }
    }
    s.m1 = r.k1;
    s.x = r.k2;
    s.c1 = r.i1;
    s.c2 = r.i2;
    s.s1 = 2;
    s.m2 = (r.k2 >> 1) + 1 + kFoldC;   // v can enter iff 2 (v - C) - p > k2 for some parity p
    return s;
}

// One unit whose maximum m reached the row's threshold.  `pub`: where this thread publishes
// its bound for the row's other thread (m2 in unit mode; threshold - 1 in exact mode, because an
// element one below it can still tie this thread's second key and win on the lower index).
__device__ __forceinline__ void unit_event(Cand& s, const int (&v)[kUnit], int m, int col, uint32_t slots,
                                           uint32_t pub, const int32_t* __restrict__ nbase, int t_cnt) {
    const bool new_best = m > s.m1;
    // a unit with maximum == m2 would stay unparked: the incoming one, or the evicted second
    const bool tie = m == s.m2 || (new_best && s.m2 == s.m1 && s.c2 >= 0);
    if (s.s1 < 2 && !tie) {
        const uint32_t dst = slots + (s.s1 ^ 1) * kSlotStride;
#pragma unroll
        for (int g = 0; g < kUnit / 4; ++g)
            sts128(dst + g * kSlotRow, v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
        if (new_best) {
            s.m2 = s.m1;
            s.c2 = s.c1;
            s.m1 = m;
            s.c1 = col;
            s.s1 ^= 1;
        } else {
            s.m2 = m;
            s.c2 = col;
        }
    } else {
        UnitVals in;
#pragma unroll
        for (int e = 0; e < kUnit; ++e) in.v[e] = v[e];
        s = exact_unit(s, in, col, slots, nbase, t_cnt);
    }
    s.thr = max(s.thr, s.m2);
    sts32(pub, static_cast<uint32_t>(s.s1 == 2 ? s.m2 - 1 : s.m2));
}

// 16 folded accumulators of one row = one unit.  MASKED: columns >= t_cnt (last, partial train
// tile) are forced to kNegV.  TREE_ONLY: timing probe (max tree, no candidate tracking).
template <bool MASKED, bool TREE_ONLY>
__device__ __forceinline__ void fold_chunk(Cand& s, const uint32_t (&acc)[16], int col0, int t_cnt,
                                           uint32_t slots, uint32_t pub, const int32_t* __restrict__ nbase) {
    int v[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        v[e] = static_cast<int>(acc[e]);
        if (MASKED) v[e] = (col0 + e < t_cnt) ? v[e] : kNegV;
    }
    const int a0 = __vimax3_s32(v[0], v[1], v[2]), a1 = __vimax3_s32(v[3], v[4], v[5]);
    const int a2 = __vimax3_s32(v[6], v[7], v[8]), a3 = __vimax3_s32(v[9], v[10], v[11]);
    const int a4 = __vimax3_s32(v[12], v[13], v[14]);
    const int m = max(__vimax3_s32(a0, a1, a2), __vimax3_s32(a3, a4, v[15]));
    if (TREE_ONLY) {
        s.m1 = max(s.m1, m);
        return;
    }
    if (m >= s.thr && (!MASKED || m > kNegV)) unit_event(s, v, m, col0, slots, pub, nbase, t_cnt);
}

struct SmemLayout {
    uint8_t* a;      // kAStages x (2 x 128 query rows x 128 B)
    uint8_t* b;      // kBStages x (256 train rows x 128 B | 256 x 32 B fold operand)
    uint8_t* ax;     // constant A operand of the fold step (128 x 32 B)
    uint4* slots;    // parked candidate units
    uint4* xchg;     // [2][256] per-row results handed from the upper to the lower column half
    int32_t* pub;    // [512] bound each epilogue thread publishes for its row's other thread
    uint64_t* a_full;
    uint64_t* a_empty;
    uint64_t* b_full;
    uint64_t* b_empty;
    uint64_t* t_full;    // [accumulator]
    uint64_t* t_empty;
    uint64_t* ax_full;
    uint32_t* tmem_slot;
    __device__ explicit SmemLayout(uint8_t* raw) {
        uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) &
                                                   ~static_cast<uintptr_t>(1023));
        a = base;
        b = a + size_t(kAStages) * kABlockBytes;
        ax = b + size_t(kBStages) * kStageBytes;
        slots = reinterpret_cast<uint4*>(ax + kAxBytes);
        xchg = reinterpret_cast<uint4*>(ax + kAxBytes + kSlotBytes);
        pub = reinterpret_cast<int32_t*>(ax + kAxBytes + kSlotBytes + 2 * SFM_QBLOCK * 16);
        uint64_t* bars = reinterpret_cast<uint64_t*>(base + kSmemData);
        a_full = bars;
        a_empty = a_full + kAStages;
        b_full = a_empty + kAStages;
        b_empty = b_full + kBStages;
        t_full = b_empty + kBStages;
        t_empty = t_full + kATiles;
        ax_full = t_empty + kATiles;
        tmem_slot = reinterpret_cast<uint32_t*>(ax_full + 1);
    }
};

// MODE 0: the kNN.  1: max tree only, 2: no epilogue (timing probes; output is not a kNN).
template <int MODE>
__global__ void __launch_bounds__(kThreads, 1)
knn2_wide_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_x,
                 const __grid_constant__ CUtensorMap tmap_ax, const WorkItem* __restrict__ work,
                 long long n_work, const int32_t* __restrict__ norms, const int32_t* __restrict__ norms_pad,
                 int32_t* __restrict__ flags, int force, uint32_t probe_flags, int4* __restrict__ out) {
    constexpr int kProducerWarp = kEpiWarps;
    constexpr int kMmaWarp = kEpiWarps + 1;
    // flags[0] != 0: some row's norm is beyond the fold range -> the generic kernel runs instead
    if (!force && *static_cast<volatile const int32_t*>(flags) != 0) return;

    extern __shared__ uint8_t smem_raw[];
    SmemLayout sm(smem_raw);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == kProducerWarp && lane == 0) {
        tma_prefetch_desc(&tmap);
        tma_prefetch_desc(&tmap_x);
        tma_prefetch_desc(&tmap_ax);
        for (int i = 0; i < kAStages; ++i) {
            mbar_init(&sm.a_full[i], 1);
            mbar_init(&sm.a_empty[i], 1);
        }
        for (int i = 0; i < kBStages; ++i) {
            mbar_init(&sm.b_full[i], 1);
            mbar_init(&sm.b_empty[i], 1);
        }
        for (int i = 0; i < kATiles; ++i) {
            mbar_init(&sm.t_full[i], 1);
            mbar_init(&sm.t_empty[i], kEpiWarps / kATiles);
        }
        mbar_init(sm.ax_full, 1);
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

    if (warp == kProducerWarp) {
        // warp-convergent loop, TMA issue predicated on one elected lane (see knn.cu)
        Pipe ap(kAStages), bp(kBStages);
        if (elect_one()) {
            mbar_arrive_expect_tx(sm.ax_full, kAxBytes);
            tma_load_2d(sm.ax, &tmap_ax, sm.ax_full, 0, 0);
        }
        __syncwarp();
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int q_row = w0.x, t_row = w0.z, t_cnt = w0.w;
            mbar_wait(&sm.a_empty[ap.stage], ap.phase ^ 1);
            if (elect_one()) {
                mbar_arrive_expect_tx(&sm.a_full[ap.stage], kABlockBytes);
                uint8_t* a_dst = sm.a + size_t(ap.stage) * kABlockBytes;
#pragma unroll
                for (int a = 0; a < kATiles; ++a)
                    tma_load_2d(a_dst + a * kATileBytes, &tmap, &sm.a_full[ap.stage], 0, q_row + a * kTileM);
            }
            __syncwarp();
            ap.advance();
            const int n_tiles = (t_cnt + kTileN - 1) / kTileN;
            for (int t = 0; t < n_tiles; ++t) {
                mbar_wait(&sm.b_empty[bp.stage], bp.phase ^ 1);
                if (elect_one()) {
                    uint8_t* dst = sm.b + size_t(bp.stage) * kStageBytes;
                    // rows beyond the train image are masked in the epilogue: a second half that
                    // lies wholly beyond it is not loaded (whatever the slot holds is multiplied)
                    const bool second = t * kTileN + kTileM < t_cnt;
                    mbar_arrive_expect_tx(&sm.b_full[bp.stage], second ? kStageBytes : kStageBytes - kATileBytes);
                    tma_load_2d(dst, &tmap, &sm.b_full[bp.stage], 0, t_row + t * kTileN);
                    if (second)
                        tma_load_2d(dst + kATileBytes, &tmap, &sm.b_full[bp.stage], 0, t_row + t * kTileN + kTileM);
                    // fold operand rows are parallel to the descriptor rows, after 128 constant rows
                    tma_load_2d(dst + kBTileBytes, &tmap_x, &sm.b_full[bp.stage], 0, kTileM + t_row + t * kTileN);
                }
                __syncwarp();
                bp.advance();
            }
        }
    } else if (warp == kMmaWarp) {
        Pipe ap(kAStages), bp(kBStages);
        uint32_t tphase = 0;
        mbar_wait(sm.ax_full, 0);
        const uint64_t axdesc = make_kmajor_sw32_desc(smem_u32(sm.ax));
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int n_tiles = (w0.w + kTileN - 1) / kTileN;
            mbar_wait(&sm.a_full[ap.stage], ap.phase);
            const uint64_t adesc0 = make_kmajor_sw128_desc(smem_u32(sm.a + size_t(ap.stage) * kABlockBytes));
            for (int t = 0; t < n_tiles; ++t) {
                mbar_wait(&sm.b_full[bp.stage], bp.phase);
                const uint32_t bsm = smem_u32(sm.b + size_t(bp.stage) * kStageBytes);
                const uint64_t bdesc0 = make_kmajor_sw128_desc(bsm);
                const uint64_t xdesc = make_kmajor_sw32_desc(bsm + kBTileBytes);
#pragma unroll
                for (int a = 0; a < kATiles; ++a) {
                    mbar_wait(&sm.t_empty[a], tphase ^ 1);
                    tc_fence_after();
                    const uint32_t d = tmem_base + a * kTileN;
                    const uint64_t adesc = adesc0 + uint64_t(a) * (kATileBytes >> 4);
                    if (elect_one()) {
                        if (!(probe_flags & 2u)) {
#pragma unroll
                            for (int k = 0; k < kKSteps; ++k)
                                mma_i8_ss(d, adesc + uint64_t(k) * (kUmmaK >> 4),
                                          bdesc0 + uint64_t(k) * (kUmmaK >> 4), kIdesc, k > 0);
                            mma_i8_ss(d, axdesc, xdesc, kIdesc, 1);   // + (1, 255 x 31) . digits(C - ||b||^2 / 2)
                        }
                        mma_commit(&sm.t_full[a]);
                        if (a == kATiles - 1) mma_commit(&sm.b_empty[bp.stage]);
                    }
                    __syncwarp();
                }
                bp.advance();
                tphase ^= 1;
            }
            if (elect_one()) mma_commit(&sm.a_empty[ap.stage]);
            __syncwarp();
            ap.advance();
        }
    } else {
        // ------------------------------------------------ epilogue warps
        const int quad = warp & 3;          // TMEM lane quadrant this warp may read
        const int acc = (warp >> 2) & 1;    // which 128-row accumulator
        const int half = warp >> 3;         // which 128-column half of every tile
        const int row = acc * kTileM + quad * 32 + lane;
        const int colbase = half * kColsPerThread;
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * kTileN + colbase;
        const uint32_t my_slots = smem_u32(sm.slots) + threadIdx.x * 16;
        const uint32_t my_pub = smem_u32(sm.pub) + threadIdx.x * 4;
        const uint32_t other_pub = smem_u32(sm.pub) + (threadIdx.x ^ (8 * 32)) * 4;
        sts32(my_pub, static_cast<uint32_t>(INT_MIN));
        asm volatile("bar.sync %0, 64;" ::"r"(1 + (warp & 7)) : "memory");
        uint64_t* full = &sm.t_full[acc];
        uint64_t* empty = &sm.t_empty[acc];
        uint32_t tphase = 0;
        int item = 0;
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x, ++item) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int4 w1 = __ldg(reinterpret_cast<const int4*>(work + w) + 1);
            const int q_row = w0.x, q_cnt = w0.y, t_cnt = w0.w, t_pad = w1.x;
            const long long out_row =
                static_cast<long long>(static_cast<uint32_t>(w1.z)) | (static_cast<long long>(w1.w) << 32);
            const bool valid = row < q_cnt;
            const int na = valid ? __ldg(norms + q_row + row) : 0;
            const int n_tiles = (t_cnt + kTileN - 1) / kTileN;
            const int32_t* nbase = norms_pad + t_pad;
            Cand sk{INT_MIN, INT_MIN, -1, -1, 0, 0, INT_MIN};
            for (int t = 0; t < n_tiles; ++t) {
                mbar_wait(full, tphase);
                tc_fence_after();
                tphase ^= 1;
                if (MODE == 2) {   // timing probe: TMA + MMA pipeline alone
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(empty);
                    continue;
                }
                const bool partial = (t + 1) * kTileN > t_cnt;
                sk.thr = max(sk.thr, static_cast<int>(lds32(other_pub)));
                // 16-column chunks, double-buffered: 32 live accumulator registers per thread
                uint32_t v[2][16];
                tmem_ld_32x16(taddr, v[0]);
#pragma unroll
                for (int c = 0; c < kColsPerThread / 16; ++c) {
                    const int col0 = t * kTileN + colbase + c * 16;
                    tmem_ld_wait();
                    if (c + 1 < kColsPerThread / 16) {
                        tmem_ld_32x16(taddr + (c + 1) * 16, v[(c + 1) & 1]);
                    } else {
                        // every column of this thread is in registers: hand the accumulator back
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(empty);
                    }
                    if (partial) {
                        fold_chunk<true, MODE == 1>(sk, v[c & 1], col0, t_cnt, my_slots, my_pub, nbase);
                    } else {
                        fold_chunk<false, MODE == 1>(sk, v[c & 1], col0, t_cnt, my_slots, my_pub, nbase);
                    }
                }
            }
            if (MODE != 0) {
                if (valid && half == 0) out[out_row + row] = make_int4(sk.m1, sk.m2, na, 0);   // keeps the probe live
                continue;
            }
            Top2K r;
            if (sk.s1 == 2) {
                r = Top2K{sk.m1, sk.x, sk.c1, sk.c2};
            } else {
                r = resolve_units(sk, my_slots, nbase, t_cnt);
            }
            const unsigned ex = __ballot_sync(0xffffffffu, valid && sk.s1 == 2);
            if (ex && lane == 0) atomicAdd(flags + 1, __popc(ex));   // statistics: rows in exact mode
            // the thread of the upper column half hands its top-2 to the lower half's thread,
            // which merges under the (key, lowest index) order and writes the row
            const uint32_t xaddr = smem_u32(sm.xchg) + ((item & 1) * SFM_QBLOCK + row) * 16;
            if (half == 1) sts128(xaddr, r.k1, r.k2, r.i1, r.i2);
            sts32(my_pub, static_cast<uint32_t>(INT_MIN));   // next work item starts unbounded
            asm volatile("bar.sync %0, 64;" ::"r"(1 + (warp & 7)) : "memory");
            if (half == 0) {
                const uint4 o = lds128(xaddr);
                Top2K b{static_cast<int>(o.x), static_cast<int>(o.y), static_cast<int>(o.z), static_cast<int>(o.w)};
                if (b.i1 < 0) b.i1 = INT_MAX;   // empty entries sort last
                if (b.i2 < 0) b.i2 = INT_MAX;
                if (r.i1 < 0) r.i1 = INT_MAX;
                if (r.i2 < 0) r.i2 = INT_MAX;
                Top2K mrg;
                if (key_before(r.k1, r.i1, b.k1, b.i1)) {
                    mrg.k1 = r.k1;
                    mrg.i1 = r.i1;
                    const bool mine = key_before(r.k2, r.i2, b.k1, b.i1);
                    mrg.k2 = mine ? r.k2 : b.k1;
                    mrg.i2 = mine ? r.i2 : b.i1;
                } else {
                    mrg.k1 = b.k1;
                    mrg.i1 = b.i1;
                    const bool theirs = key_before(b.k2, b.i2, r.k1, r.i1);
                    mrg.k2 = theirs ? b.k2 : r.k1;
                    mrg.i2 = theirs ? b.i2 : r.i1;
                }
                if (valid) {
                    const int i1 = mrg.i1 == INT_MAX ? -1 : mrg.i1;
                    const int i2 = mrg.i2 == INT_MAX ? -1 : mrg.i2;
                    const int d1 = i1 < 0 ? INT_MAX : na - mrg.k1;
                    const int d2 = i2 < 0 ? INT_MAX : na - mrg.k2;
                    out[out_row + row] = make_int4(i1, i2, d1, d2);
                }
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

// Fold-step operands, 32-byte rows: rows [0,128) are the constant A operand (1, 255 x 31);
// row 128 + r holds the u8 digits (weights 1, 255 x 31) of C - floor(||b_r||^2 / 2) for
// descriptor row r.  flags[0] is raised when a norm is beyond the encodable range.
__global__ void bias32_kernel(const int32_t* __restrict__ norms, long long total_rows,
                              uint8_t* __restrict__ bias, int32_t* __restrict__ flags) {
    const long long r = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    if (r >= total_rows + kTileM) return;
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = 0;
    if (r < kTileM) {
#pragma unroll
        for (int i = 0; i < 8; ++i) w[i] = 0xFFFFFFFFu;
        w[0] = 0xFFFFFF01u;
    } else {
        int h = norms[r - kTileM] >> 1;
        if (h > kFoldC) {
            h = kFoldC;
            atomicOr(flags, 1);
        }
        const int u = kFoldC - h;
        int q = u / 255;
        w[0] = static_cast<uint32_t>(u % 255);
#pragma unroll
        for (int k = 1; k < 32; ++k) {
            const int take = q < 255 ? q : 255;
            w[k >> 2] |= static_cast<uint32_t>(take) << (8 * (k & 3));
            q -= take;
        }
    }
    uint4* dst = reinterpret_cast<uint4*>(bias + r * kUmmaK);
    dst[0] = make_uint4(w[0], w[1], w[2], w[3]);
    dst[1] = make_uint4(w[4], w[5], w[6], w[7]);
}

// Debug: raw folded accumulators of one 128 x 256 tile through the same TMA / UMMA path.
__global__ void __launch_bounds__(128, 1)
debug_wide_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_x,
                  const __grid_constant__ CUtensorMap tmap_ax, int q_row, int t_row, int32_t* __restrict__ out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    uint8_t* sa = base;
    uint8_t* sb = sa + kATileBytes;
    uint8_t* sx = sb + kBTileBytes;
    uint8_t* sax = sx + kXTileBytes;
    uint64_t* bar_ld = reinterpret_cast<uint64_t*>(sax + kAxBytes);
    uint64_t* bar_mma = bar_ld + 1;
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar_mma + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(bar_ld, 1);
        mbar_init(bar_mma, 1);
        fence_barrier_init();
    }
    if (warp == 0) {
        tmem_alloc(slot, 256);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *slot;
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(bar_ld, kATileBytes + kBTileBytes + kXTileBytes + kAxBytes);
        tma_load_2d(sa, &tmap, bar_ld, 0, q_row);
        tma_load_2d(sb, &tmap, bar_ld, 0, t_row);
        tma_load_2d(sb + kATileBytes, &tmap, bar_ld, 0, t_row + kTileM);
        tma_load_2d(sx, &tmap_x, bar_ld, 0, kTileM + t_row);
        tma_load_2d(sax, &tmap_ax, bar_ld, 0, 0);
        mbar_wait(bar_ld, 0);
        tc_fence_after();
        const uint64_t ad = make_kmajor_sw128_desc(smem_u32(sa));
        const uint64_t bd = make_kmajor_sw128_desc(smem_u32(sb));
        for (int k = 0; k < kKSteps; ++k)
            mma_i8_ss(tmem_base, ad + uint64_t(k) * (kUmmaK >> 4), bd + uint64_t(k) * (kUmmaK >> 4), kIdesc, k > 0);
        mma_i8_ss(tmem_base, make_kmajor_sw32_desc(smem_u32(sax)), make_kmajor_sw32_desc(smem_u32(sx)), kIdesc, 1);
        mma_commit(bar_mma);
    }
    __syncwarp();
    mbar_wait(bar_mma, 0);
    tc_fence_after();
    const int row = warp * 32 + lane;
    for (int c = 0; c < kTileN / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(warp * 32) << 16) + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) out[row * kTileN + c * 32 + j] = static_cast<int32_t>(v[j]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 256);
    }
}

template <int MODE>
cudaError_t launch_variant(const CUtensorMap& tmap, const CUtensorMap& tmap_x, const CUtensorMap& tmap_ax,
                           const WorkItem* d_work, int64_t n_work, const int32_t* d_norms,
                           const int32_t* d_norms_pad, int32_t* d_flags, int force, uint32_t probe_flags,
                           int32_t* d_knn, unsigned grid, cudaStream_t st) {
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(knn2_wide_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(kSmemBytes));
        if (e != cudaSuccess) return e;
        configured = true;
    }
    knn2_wide_kernel<MODE><<<grid, kThreads, kSmemBytes, st>>>(tmap, tmap_x, tmap_ax, d_work, n_work, d_norms,
                                                              d_norms_pad, d_flags, force, probe_flags,
                                                              reinterpret_cast<int4*>(d_knn));
    return cudaGetLastError();
}

}  // namespace

void launch_bias32(const int32_t* d_norms, int64_t total_rows, uint8_t* d_bias32, int32_t* d_flags,
                   cudaStream_t st, int64_t* launches) {
    const long long n = total_rows + kTileM;
    bias32_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, st>>>(d_norms, total_rows, d_bias32, d_flags);
    ++*launches;
}

// variant 0: the kNN (exits at once when the fold-range flag is raised, unless force);
// 1: max tree only, 2: no epilogue (timing probes).
int launch_knn2_wide(const CUtensorMap& tmap, const CUtensorMap& tmap_x, const CUtensorMap& tmap_ax,
                     const WorkItem* d_work, int64_t n_work, const int32_t* d_norms, const int32_t* d_norms_pad,
                     int32_t* d_flags, int variant, int force, uint32_t probe_flags, int32_t* d_knn, int num_sms,
                     cudaStream_t st, int64_t* launches) {
    if (n_work <= 0) return 0;
    const unsigned grid = static_cast<unsigned>(n_work < num_sms ? n_work : num_sms);
    cudaError_t e;
    if (variant == 1)
        e = launch_variant<1>(tmap, tmap_x, tmap_ax, d_work, n_work, d_norms, d_norms_pad, d_flags, 1, probe_flags,
                              d_knn, grid, st);
    else if (variant == 2)
        e = launch_variant<2>(tmap, tmap_x, tmap_ax, d_work, n_work, d_norms, d_norms_pad, d_flags, 1, probe_flags,
                              d_knn, grid, st);
    else
        e = launch_variant<0>(tmap, tmap_x, tmap_ax, d_work, n_work, d_norms, d_norms_pad, d_flags, force,
                              probe_flags, d_knn, grid, st);
    ++*launches;
    return static_cast<int>(e);
}

int launch_debug_wide(const CUtensorMap& tmap, const CUtensorMap& tmap_x, const CUtensorMap& tmap_ax,
                      int32_t q_row, int32_t t_row, int32_t* d_out, cudaStream_t st, int64_t* launches) {
    const size_t smem = kATileBytes + kBTileBytes + kXTileBytes + kAxBytes + 64 + 1024;
    cudaError_t e = cudaFuncSetAttribute(debug_wide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    debug_wide_kernel<<<1, 128, smem, st>>>(tmap, tmap_x, tmap_ax, q_row, t_row, d_out);
    ++*launches;
    return static_cast<int>(cudaGetLastError());
}

}  // namespace sfm
