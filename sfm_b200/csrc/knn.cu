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
// K = 128 is so short that the epilogue (16384 accumulators per 270-cycle
// tile) is the bound, not the tensor pipe.  Two kernel variants:
//
//  kFold (default): the column-norm term is folded INTO the contraction.  A
//    fifth K-step multiplies a constant u8 row (1, 255 x 31) with a per-train-
//    row u8 row holding the digits of C - floor(||b||^2 / 2), C = 2,016,029, so
//    the accumulator already holds v = a.b + C - floor(||b||^2/2) and the exact
//    ranking key is 2(v - C) - (||b||^2 & 1) = 2 a.b - ||b||^2 (maximise;
//    d^2 = ||a||^2 - key).  All five K-steps share one instruction descriptor
//    (u8 x u8).  The epilogue is then a max3 tree (0.6 instr/element) compared
//    against a per-row threshold; only 4-column groups that can change the
//    top-2 are looked at exactly.  Needs ||b||^2 <= 4,032,059 (norm <= 2008;
//    SIFT is 512); sfm_set_images raises a device flag otherwise and the
//    generic variant runs instead.
//  kGeneric: keys computed per element from a padded column-norm array.
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
constexpr int kXStages = 2;                          // bias tiles (one per 4 train tiles)
constexpr int kXGroup = 4;                           // train tiles per bias tile
constexpr int kTmemStages = 2;
constexpr int kTmemColsPerStage = kATiles * kTileN;  // 256
constexpr int kTmemCols = kTmemStages * kTmemColsPerStage;  // 512 = all of TMEM
// epilogue warps: 8 for the generic variant (one thread per query row, 128 columns per tile);
// 16 for the folded variants (two threads per row, 64 columns each): four warps per scheduler
// hide the TMEM-load and compare latencies of the short per-chunk dependency chains.
constexpr int kEpiWarpsMax = 16;
__host__ __device__ constexpr int epi_warps(int mode) { return mode == 0 ? 8 : 16; }
__host__ __device__ constexpr int knn_threads(int mode) { return (epi_warps(mode) + 2) * 32; }
constexpr int kUmmaK = 32;                           // bytes of K per tcgen05.mma kind::i8
constexpr int kKSteps = kRowBytes / kUmmaK;          // 4
constexpr uint32_t kIdesc = make_idesc_u8(kTileM, kTileN);
constexpr int kPadNorm = 1 << 30;                    // column norm of a padding column
constexpr int kNegV = -(1 << 28);                    // folded accumulator of a masked column (2v - 2C must not wrap)
// fold constant C: the largest value the 32 u8 bias digits (weights 1, 255 x 31) can express
constexpr int kFoldC = 255 * 31 * 255 + 254;

enum KnnMode { kGeneric = 0, kFold = 1, kFoldFastOnly = 2, kFoldNoEpilogue = 3 };

constexpr size_t kSlotBytes = 2 * 2 * kEpiWarpsMax * 32 * 16;   // 2 parked 8-column units per epilogue thread (32 KiB)
constexpr size_t kXchgBytes = 2 * SFM_QBLOCK * 16 + kEpiWarpsMax * 32 * 4;   // per-row hand-over between the two column halves + published bounds
constexpr size_t kSmemData = size_t(kAStages) * kABlockBytes + size_t(kBStages) * kTileBytes +
                             size_t(kXStages + 1) * kTileBytes + kSlotBytes + kXchgBytes;
constexpr int kNumBars = 2 * kAStages + 2 * kBStages + 4 * kTmemStages + 2 * kXStages + 1;
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

// ------------------------------------------------------------------ folded epilogue
// The exact ranking key is K = 2 a.b - ||b||^2 = 2 (v - C) - p (maximised), v the folded
// accumulator and p the parity of ||b||^2.  Both nearest neighbours have v >= V2, the second
// largest v of the row (K >= 2 V2 - 1 forces v >= V2).  So per row the kernel only tracks the
// two 16-column UNITS with the largest maxima (m1 >= m2, earliest unit on equal maxima) and
// parks their 16 accumulators in shared memory; everything exact -- parity, indices, the
// lowest-index tie rule -- is done once per work item on those 32 values (resolve_units).
// The steady-state cost is a max3 tree plus two compares per 32 columns.
// The two units hold every element with v >= V2 unless a unit that is NOT parked has a
// maximum equal to m2 (an incoming unit with m == m2, or the old second unit evicted by a new
// best while m2 == m1).  At that moment the row switches to EXACT mode for the rest of the
// work item: an element-wise running top-2 of K (exact_unit), entered only when a unit
// maximum reaches the threshold kept in m2.  Exact mode is rare outside all-ties data
// (0.65 % of rows on the synthetic 8k set) and is always correct.
struct Cand {
    int m1, m2;   // units: largest / second largest unit maximum.   exact: k1 / entry threshold on v
    int c1, c2;   // units: first column of those units (-1: none).  exact: i1 / i2
    int s1;       // units: shared-memory slot (0/1) of unit c1.     exact: 2
    int x;        //                                                  exact: k2
    int thr;      // entry threshold on a unit maximum: max(m2, what the row's other thread published)
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

// (key, index) order of the tie rule: larger key first, then lower index.
__device__ __forceinline__ bool key_before(int ka, int ia, int kb, int ib) {
    return ka > kb || (ka == kb && ia < ib);
}

constexpr int kUnit = 8;
constexpr uint32_t kSlotRow = kEpiWarpsMax * 32 * 16;       // bytes between a thread's uint4s
constexpr uint32_t kSlotStride = (kUnit / 4) * kSlotRow;    // bytes per slot

// Exact top-2 of the (at most) two parked units, visited in ascending column order.
// slots: shared address of this thread's first uint4; layout [slot][4 x uint4][thread]
__device__ __forceinline__ Top2K resolve_units(const Cand& s, uint32_t slots,
                                               const int32_t* __restrict__ nbase, int t_cnt) {
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
// mode first if it was still tracking units).  Kept out of line: the hot loop only pays a call
// in an unlikely branch.
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

// `pub`: shared address where this thread publishes a bound for the row's other thread (the one
// scanning the other column half): unit maxima below it cannot matter for the row.  In unit
// mode that is m2 (two of this thread's units reach it); in exact mode the entry threshold
// minus one, because an element one below it can still TIE this thread's second key and win
// on the lower index.  Stale values are harmless: the bound only ever rises within a work item.
template <bool MASKED>
__device__ __forceinline__ void unit_event(Cand& s, const int (&v)[16], int u, int m, int col,
                                           uint32_t slots, uint32_t pub, const int32_t* __restrict__ nbase,
                                           int t_cnt) {
    if (m >= s.thr && (!MASKED || m > kNegV)) {
        const bool new_best = m > s.m1;
        // a unit with maximum == m2 would stay unparked: the incoming one, or the evicted second
        const bool tie = m == s.m2 || (new_best && s.m2 == s.m1 && s.c2 >= 0);
        if (s.s1 < 2 && !tie) {
            const uint32_t dst = slots + (s.s1 ^ 1) * kSlotStride;
#pragma unroll
            for (int g = 0; g < kUnit / 4; ++g)
                sts128(dst + g * kSlotRow, v[u * kUnit + 4 * g], v[u * kUnit + 4 * g + 1],
                       v[u * kUnit + 4 * g + 2], v[u * kUnit + 4 * g + 3]);
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
            for (int e = 0; e < kUnit; ++e) in.v[e] = v[u * kUnit + e];
            s = exact_unit(s, in, col, slots, nbase, t_cnt);
        }
        s.thr = max(s.thr, s.m2);
        sts32(pub, static_cast<uint32_t>(s.s1 == 2 ? s.m2 - 1 : s.m2));
    }
}

// 16 folded accumulators of one row = two units.  MASKED: columns >= t_cnt (last, partial
// train tile) are forced to kNegV.
template <bool MASKED, int MODE>
__device__ __forceinline__ void fold_chunk(Cand& s, const uint32_t (&acc)[16], int col0, int t_cnt,
                                           uint32_t slots, uint32_t pub, const int32_t* __restrict__ nbase) {
    int v[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        v[e] = static_cast<int>(acc[e]);
        if (MASKED) v[e] = (col0 + e < t_cnt) ? v[e] : kNegV;
    }
    int m8[2];
#pragma unroll
    for (int j = 0; j < 2; ++j)
        m8[j] = max(__vimax3_s32(__vimax3_s32(v[8 * j], v[8 * j + 1], v[8 * j + 2]),
                                 __vimax3_s32(v[8 * j + 3], v[8 * j + 4], v[8 * j + 5]), v[8 * j + 6]),
                    v[8 * j + 7]);
    if (MODE == kFoldFastOnly) {
        s.m1 = __vimax3_s32(s.m1, m8[0], m8[1]);   // timing probe: tree only (results are not a kNN)
        return;
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) unit_event<MASKED>(s, v, j, m8[j], col0 + j * kUnit, slots, pub, nbase, t_cnt);
}

struct SmemLayout {
    uint8_t* a;
    uint8_t* b;
    uint8_t* x;    // kXStages bias tiles
    uint8_t* ax;   // constant A operand of the fold step
    uint4* slots;  // parked candidate units of the folded epilogue
    uint4* xchg;   // [2][256] per-row results handed from the upper to the lower column half
    int32_t* pub;  // [512] bound each folded-epilogue thread publishes for its row's other thread
    uint64_t* a_full;
    uint64_t* a_empty;
    uint64_t* b_full;
    uint64_t* b_empty;
    uint64_t* t_full;
    uint64_t* t_empty;
    uint64_t* x_full;
    uint64_t* x_empty;
    uint64_t* ax_full;
    uint32_t* tmem_slot;
    __device__ explicit SmemLayout(uint8_t* raw) {
        uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) &
                                                   ~static_cast<uintptr_t>(1023));
        a = base;
        b = a + size_t(kAStages) * kABlockBytes;
        x = b + size_t(kBStages) * kTileBytes;
        ax = x + size_t(kXStages) * kTileBytes;
        slots = reinterpret_cast<uint4*>(ax + kTileBytes);
        xchg = reinterpret_cast<uint4*>(ax + kTileBytes + kSlotBytes);
        pub = reinterpret_cast<int32_t*>(ax + kTileBytes + kSlotBytes + 2 * SFM_QBLOCK * 16);
        uint64_t* bars = reinterpret_cast<uint64_t*>(base + kSmemData);
        a_full = bars;
        a_empty = a_full + kAStages;
        b_full = a_empty + kAStages;
        b_empty = b_full + kBStages;
        t_full = b_empty + kBStages;
        t_empty = t_full + 2 * kTmemStages;   // [stage][accumulator]
        x_full = t_empty + 2 * kTmemStages;
        x_empty = x_full + kXStages;
        ax_full = x_empty + kXStages;
        tmem_slot = reinterpret_cast<uint32_t*>(ax_full + 1);
    }
};

// `flags[0]` != 0: some registered row's norm is beyond what the fold step can
// encode -> only the generic variant may run.  Both variants are launched
// back to back (no host synchronisation); the one that does not apply exits.
template <int MODE>
__global__ void __launch_bounds__(knn_threads(MODE), 1)
knn2_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_x,
            const uint8_t* __restrict__ desc, const WorkItem* __restrict__ work, long long n_work,
            const int32_t* __restrict__ norms,
            const int32_t* __restrict__ norms_pad, int32_t* __restrict__ flags, int force,
            uint32_t wait_ns, uint32_t epi_wait_ns, uint32_t probe_flags, int4* __restrict__ out) {
    constexpr bool FOLD = MODE != kGeneric;
    constexpr int kEpiWarps = epi_warps(MODE);
    constexpr int kProducerWarp = kEpiWarps;
    constexpr int kMmaWarp = kEpiWarps + 1;
    constexpr int kHalves = kEpiWarps / 8;             // threads (column ranges) per query row
    constexpr int kColsPerThread = kTileN / kHalves;   // columns of each train tile per thread
    if (!force && ((*static_cast<volatile const int32_t*>(flags) != 0) == FOLD)) return;

    extern __shared__ uint8_t smem_raw[];
    SmemLayout sm(smem_raw);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == kProducerWarp && lane == 0) {
        tma_prefetch_desc(&tmap);
        if (FOLD) tma_prefetch_desc(&tmap_x);
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
        for (int i = 0; i < kXStages; ++i) {
            mbar_init(&sm.x_full[i], 1);
            mbar_init(&sm.x_empty[i], 1);
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

    // Producer and MMA warps run their loops warp-convergent (all 32 lanes wait and keep the
    // pipeline state) and only the TMA / tcgen05 issue is predicated on one elected lane: the
    // descriptors then live in uniform registers and each tcgen05.mma is a single predicated
    // instruction.  (Under `if (lane == 0)` ptxas wraps every UTCIMMA in an R2UR + elect loop
    // and the issuing thread, not the tensor pipe, paces the kernel.)
    if (warp == kProducerWarp) {
        Pipe ap(kAStages), bp(kBStages), xp(kXStages);
        if (FOLD && elect_one()) {   // constant A operand of the fold step: rows 0..127 of the bias array
            mbar_arrive_expect_tx(sm.ax_full, kTileBytes);
            tma_load_2d(sm.ax, &tmap_x, sm.ax_full, 0, 0);
        }
        __syncwarp();
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int4 w1 = __ldg(reinterpret_cast<const int4*>(work + w) + 1);
            const int q_row = w0.x, t_row = w0.z, t_cnt = w0.w, x_row = w1.y;
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
                if (FOLD && (t % kXGroup) == 0) {
                    mbar_wait_suspend(&sm.x_empty[xp.stage], xp.phase ^ 1, wait_ns);
                    if (elect_one()) {
                        mbar_arrive_expect_tx(&sm.x_full[xp.stage], kTileBytes);
                        tma_load_2d(sm.x + size_t(xp.stage) * kTileBytes, &tmap_x, &sm.x_full[xp.stage], 0,
                                    x_row + (t / kXGroup) * kTileM);
                    }
                    __syncwarp();
                    xp.advance();
                }
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
        Pipe ap(kAStages), bp(kBStages), tp(kTmemStages), xp(kXStages);
        uint64_t axdesc = 0;
        if (FOLD) {
            mbar_wait_suspend(sm.ax_full, 0, wait_ns);
            axdesc = make_kmajor_sw128_desc(smem_u32(sm.ax));
        }
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int n_tiles = (w0.w + kTileN - 1) / kTileN;
            mbar_wait_suspend(&sm.a_full[ap.stage], ap.phase, wait_ns);
            const uint64_t adesc0 = make_kmajor_sw128_desc(smem_u32(sm.a + size_t(ap.stage) * kABlockBytes));
            uint64_t xdesc0 = 0;
            for (int t = 0; t < n_tiles; ++t) {
                if (FOLD && (t % kXGroup) == 0) {
                    mbar_wait_suspend(&sm.x_full[xp.stage], xp.phase, wait_ns);
                    xdesc0 = make_kmajor_sw128_desc(smem_u32(sm.x + size_t(xp.stage) * kTileBytes));
                }
                mbar_wait_suspend(&sm.b_full[bp.stage], bp.phase, wait_ns);
                const uint64_t bdesc0 = make_kmajor_sw128_desc(smem_u32(sm.b + size_t(bp.stage) * kTileBytes));
                const uint64_t xdesc = xdesc0 + uint64_t(t % kXGroup) * (kUmmaK >> 4);
                const uint32_t d0 = tmem_base + tp.stage * kTmemColsPerStage;
                const bool x_done = FOLD && ((t % kXGroup) == kXGroup - 1 || t == n_tiles - 1);
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
                            if (FOLD)   // + (1, 255 x 31) . digits(C - floor(||b||^2 / 2))
                                mma_i8_ss(d, axdesc, xdesc, kIdesc, 1);
                        }
                        mma_commit(&sm.t_full[tp.stage * kATiles + a]);   // accumulator ready for the epilogue
                        if (a == kATiles - 1) {
                            mma_commit(&sm.b_empty[bp.stage]);   // B slot reusable once the MMAs retire
                            if (x_done) mma_commit(&sm.x_empty[xp.stage]);
                        }
                    }
                    __syncwarp();
                }
                __syncwarp();
                if (x_done) xp.advance();
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
        const int half = warp >> 3;         // which column range of every tile (folded variants: 0/1)
        const int row = acc * kTileM + quad * 32 + lane;
        const int colbase = half * kColsPerThread;
        const uint32_t lane_addr =
            tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * kTileN + colbase;
        const uint32_t my_slots = smem_u32(sm.slots) + threadIdx.x * 16;
        const uint32_t my_pub = smem_u32(sm.pub) + threadIdx.x * 4;
        const uint32_t other_pub = smem_u32(sm.pub) + (threadIdx.x ^ (8 * 32)) * 4;
        if (FOLD && kHalves == 2) {
            sts32(my_pub, static_cast<uint32_t>(INT_MIN));
            asm volatile("bar.sync %0, 64;" ::"r"(1 + (warp & 7)) : "memory");
        }
        Pipe tp(kTmemStages);
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
            Top2 s{INT_MAX, INT_MAX, -1, -1};
            Cand sk{INT_MIN, INT_MIN, -1, -1, 0, 0, INT_MIN};
            for (int t = 0; t < n_tiles; ++t) {
                uint64_t* full = &sm.t_full[tp.stage * kATiles + acc];
                uint64_t* empty = &sm.t_empty[tp.stage * kATiles + acc];
                mbar_wait_suspend(full, tp.phase, epi_wait_ns);
                tc_fence_after();
                if (MODE == kFoldNoEpilogue) {   // timing probe: TMA + MMA pipeline alone
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(empty);
                    tp.advance();
                    continue;
                }
                const uint32_t taddr = lane_addr + tp.stage * kTmemColsPerStage;
                const bool partial = FOLD && (t + 1) * kTileN > t_cnt;
                if (FOLD && kHalves == 2) sk.thr = max(sk.thr, static_cast<int>(lds32(other_pub)));
                if (!FOLD) {
                    uint32_t v[2][32];
                    tmem_ld_32x32(taddr, v[0]);
#pragma unroll
                    for (int c = 0; c < kColsPerThread / 32; ++c) {
                        const int col0 = t * kTileN + colbase + c * 32;
                        int4 nb[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            nb[j] = __ldg(reinterpret_cast<const int4*>(nbase) + col0 / 4 + j);
                        tmem_ld_wait();
                        if (c + 1 < kColsPerThread / 32) {
                            tmem_ld_32x32(taddr + (c + 1) * 32, v[(c + 1) & 1]);
                        } else {
                            // every column of this thread is in registers: hand the accumulator back
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(empty);
                        }
                        top2_chunk(s, v[c & 1], nb, col0);
                    }
                } else {
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
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(empty);
                        }
                        if (partial) {
                            fold_chunk<true, MODE>(sk, v[c & 1], col0, t_cnt, my_slots, my_pub, nbase);
                        } else {
                            fold_chunk<false, MODE>(sk, v[c & 1], col0, t_cnt, my_slots, my_pub, nbase);
                        }
                    }
                }
                tp.advance();
            }
            if (!FOLD) {
                if (valid) {
                    // padding columns (index >= t_cnt) can only surface when t_cnt < 2
                    if (s.i1 >= t_cnt) s.i1 = -1;
                    if (s.i2 >= t_cnt) s.i2 = -1;
                    const int d1 = s.i1 < 0 ? INT_MAX : s.d1 + na;
                    const int d2 = s.i2 < 0 ? INT_MAX : s.d2 + na;
                    out[out_row + row] = make_int4(s.i1, s.i2, d1, d2);
                }
            } else if (MODE != kFold) {
                if (valid && half == 0) out[out_row + row] = make_int4(sk.m1, sk.m2, na, 0);   // keeps the probe live
            } else {
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
                    Top2K b{static_cast<int>(o.x), static_cast<int>(o.y), static_cast<int>(o.z),
                            static_cast<int>(o.w)};
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
debug_dot_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_x,
                 int q_row, int t_row, int x_row, int x_chunk, int32_t* __restrict__ out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    uint8_t* sa = base;
    uint8_t* sb = base + kTileBytes;
    uint8_t* sx = base + 2 * kTileBytes;    // bias tile (fold step B operand)
    uint8_t* sax = base + 3 * kTileBytes;   // constant fold-step A operand
    uint64_t* bar_ld = reinterpret_cast<uint64_t*>(base + 4 * kTileBytes);
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
        const bool fold = x_row >= 0;
        mbar_arrive_expect_tx(bar_ld, (fold ? 4 : 2) * kTileBytes);
        tma_load_2d(sa, &tmap, bar_ld, 0, q_row);
        tma_load_2d(sb, &tmap, bar_ld, 0, t_row);
        if (fold) {
            tma_load_2d(sx, &tmap_x, bar_ld, 0, x_row);
            tma_load_2d(sax, &tmap_x, bar_ld, 0, 0);
        }
        mbar_wait(bar_ld, 0);
        tc_fence_after();
        const uint64_t ad = make_kmajor_sw128_desc(smem_u32(sa));
        const uint64_t bd = make_kmajor_sw128_desc(smem_u32(sb));
        for (int k = 0; k < kKSteps; ++k)
            mma_i8_ss(tmem_base, ad + uint64_t(k) * (kUmmaK >> 4), bd + uint64_t(k) * (kUmmaK >> 4),
                      kIdesc, k > 0);
        if (fold)
            mma_i8_ss(tmem_base, make_kmajor_sw128_desc(smem_u32(sax)),
                      make_kmajor_sw128_desc(smem_u32(sx)) + uint64_t(x_chunk) * (kUmmaK >> 4),
                      kIdesc, 1);
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

// Fold-step operands (see the header comment): row 0..127 of the bias array is
// the constant A operand (1, 255 x 31 in the first 32 bytes of each 128-byte
// row); from row x_off[img] on, 128-byte row (g * 128 + r) holds in its 32-byte
// chunk j the u8 digits of C - floor(||b||^2 / 2) for train row g * 512 + j * 128 + r
// of image img, i.e. one 128 x 128-byte SWIZZLE_128B TMA box serves 4 train tiles.
__global__ void bias_kernel(const int32_t* __restrict__ norms, const long long* __restrict__ img_off,
                            const long long* __restrict__ x_off, int n_images,
                            uint8_t* __restrict__ bias, int32_t* __restrict__ flags) {
    const int img = blockIdx.y;
    if (img >= n_images) return;
    const long long r0 = img_off[img], n = img_off[img + 1] - r0;
    const long long x0 = x_off[img], nx = (x_off[img + 1] - x0) * kXGroup;   // (row, chunk) slots
    for (long long sidx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; sidx < nx;
         sidx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long xr = sidx / kXGroup;          // bias row within the image
        const int j = static_cast<int>(sidx % kXGroup);
        const long long trow = (xr / kTileM) * (kXGroup * kTileM) + j * kTileM + (xr % kTileM);
        int h = 0;
        if (trow < n) {
            h = norms[r0 + trow] >> 1;
            if (h > kFoldC) {
                h = kFoldC;
                atomicOr(flags, 1);
            }
        }
        const int u = kFoldC - h;
        int q = u / 255;
        uint32_t w[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) w[i] = 0;
        w[0] = static_cast<uint32_t>(u % 255);          // digit 0, weight 1
#pragma unroll
        for (int k = 1; k < 32; ++k) {                  // digits 1..31, weight 255
            const int take = q < 255 ? q : 255;
            w[k >> 2] |= static_cast<uint32_t>(take) << (8 * (k & 3));
            q -= take;
        }
        uint4* dst = reinterpret_cast<uint4*>(bias + (x0 + xr) * kRowBytes + j * kUmmaK);
        dst[0] = make_uint4(w[0], w[1], w[2], w[3]);
        dst[1] = make_uint4(w[4], w[5], w[6], w[7]);
    }
}

__global__ void bias_const_kernel(uint8_t* __restrict__ bias) {
    // 128 rows x 128 bytes; bytes 0..31 = (1, 255 x 31), the rest 0
    for (int i = threadIdx.x; i < kTileM * kRowBytes; i += blockDim.x) {
        const int c = i % kRowBytes;
        bias[i] = c == 0 ? 1 : c < 32 ? 255 : 0;
    }
}

__global__ void build_work_kernel(const int32_t* __restrict__ pairs, int n_pairs,
                                  const long long* __restrict__ blk_prefix,
                                  const long long* __restrict__ row_off,
                                  const long long* __restrict__ img_off,
                                  const long long* __restrict__ pad_off,
                                  const long long* __restrict__ x_off, WorkItem* __restrict__ work) {
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
        it.x_row = static_cast<int32_t>(x_off[ti]);
        it.out_row = row_off[p] + b * SFM_QBLOCK;
        work[w0 + b] = it;
    }
}

}  // namespace

void launch_norms(const uint8_t* d_desc, int64_t total_rows, const int64_t* d_img_off,
                  const int64_t* d_pad_off, const int64_t* d_x_off, int32_t n_images,
                  int32_t* d_norms, int32_t* d_norms_pad, uint8_t* d_bias, int32_t* d_flags,
                  cudaStream_t st, int64_t* launches) {
    if (total_rows > 0) {
        const int threads = 256;
        const long long blocks = (total_rows * 32 + threads - 1) / threads;
        norms_kernel<<<static_cast<unsigned>(blocks), threads, 0, st>>>(d_desc, total_rows, d_norms);
        ++*launches;
    }
    cudaMemsetAsync(d_flags, 0, 4 * sizeof(int32_t), st);
    bias_const_kernel<<<1, 256, 0, st>>>(d_bias);
    ++*launches;
    // images are processed in slabs of 65535 (grid.y limit)
    for (int i0 = 0; i0 < n_images; i0 += 65535) {
        const int cnt = n_images - i0 < 65535 ? n_images - i0 : 65535;
        dim3 grid(8, cnt);
        pad_norms_kernel<<<grid, 256, 0, st>>>(d_norms, reinterpret_cast<const long long*>(d_img_off) + i0,
                                               reinterpret_cast<const long long*>(d_pad_off) + i0, cnt,
                                               d_norms_pad);
        bias_kernel<<<grid, 256, 0, st>>>(d_norms, reinterpret_cast<const long long*>(d_img_off) + i0,
                                          reinterpret_cast<const long long*>(d_x_off) + i0, cnt, d_bias,
                                          d_flags);
        *launches += 2;
    }
}

void launch_build_work(const int32_t* d_pairs, int32_t n_pairs, const int64_t* d_blk_prefix,
                       const int64_t* d_row_off, const int64_t* d_img_off, const int64_t* d_pad_off,
                       const int64_t* d_x_off, WorkItem* d_work, cudaStream_t st, int64_t* launches) {
    if (n_pairs <= 0) return;
    build_work_kernel<<<(n_pairs + 127) / 128, 128, 0, st>>>(
        d_pairs, n_pairs, reinterpret_cast<const long long*>(d_blk_prefix),
        reinterpret_cast<const long long*>(d_row_off), reinterpret_cast<const long long*>(d_img_off),
        reinterpret_cast<const long long*>(d_pad_off), reinterpret_cast<const long long*>(d_x_off), d_work);
    ++*launches;
}

namespace {
template <int MODE>
cudaError_t launch_knn2_variant(const CUtensorMap& tmap, const CUtensorMap& tmap_x, const uint8_t* d_desc,
                                const WorkItem* d_work,
                                int64_t n_work, const int32_t* d_norms, const int32_t* d_norms_pad,
                                int32_t* d_flags, int force, const KnnTuning& tn, int32_t* d_knn,
                                unsigned grid, cudaStream_t st) {
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(knn2_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(kSmemBytes));
        if (e != cudaSuccess) return e;
        configured = true;
    }
    knn2_kernel<MODE><<<grid, knn_threads(MODE), kSmemBytes, st>>>(tmap, tmap_x, d_desc, d_work, n_work, d_norms, d_norms_pad,
                                                          d_flags, force, tn.wait_ns, tn.epi_wait_ns, tn.probe_flags,
                                                          reinterpret_cast<int4*>(d_knn));
    return cudaGetLastError();
}
}  // namespace

// mode 0: automatic (fold variant unless the norm flag is raised, then generic);
// 1: generic only; 2: fold only (caller guarantees the norms fit);
// 3: fold fast path only, 4: fold without epilogue (timing probes, output is not a kNN).
int launch_knn2(const CUtensorMap& tmap, const CUtensorMap& tmap_x, const uint8_t* d_desc,
                const WorkItem* d_work, int64_t n_work,
                const int32_t* d_norms, const int32_t* d_norms_pad, int32_t* d_flags, int mode,
                const KnnTuning& tn, int32_t* d_knn, int num_sms, cudaStream_t st, int64_t* launches) {
    if (n_work <= 0) return 0;
    const unsigned grid = static_cast<unsigned>(n_work < num_sms ? n_work : num_sms);
    cudaError_t e = cudaSuccess;
    if (mode == 0 || mode == 2) {
        e = launch_knn2_variant<kFold>(tmap, tmap_x, d_desc, d_work, n_work, d_norms, d_norms_pad, d_flags, mode == 2,
                                       tn, d_knn, grid, st);
        ++*launches;
    }
    if (e == cudaSuccess && (mode == 0 || mode == 1 || mode == 9)) {   // 9: generic, only if the flag is raised
        e = launch_knn2_variant<kGeneric>(tmap, tmap_x, d_desc, d_work, n_work, d_norms, d_norms_pad, d_flags,
                                          mode == 1, tn, d_knn, grid, st);
        ++*launches;
    }
    if (e == cudaSuccess && mode == 3) {
        e = launch_knn2_variant<kFoldFastOnly>(tmap, tmap_x, d_desc, d_work, n_work, d_norms, d_norms_pad, d_flags, 1,
                                               tn, d_knn, grid, st);
        ++*launches;
    }
    if (e == cudaSuccess && mode == 4) {
        e = launch_knn2_variant<kFoldNoEpilogue>(tmap, tmap_x, d_desc, d_work, n_work, d_norms, d_norms_pad, d_flags,
                                                 1, tn, d_knn, grid, st);
        ++*launches;
    }
    return static_cast<int>(e);
}

int launch_debug_dot(const CUtensorMap& tmap, const CUtensorMap& tmap_x, int32_t q_row, int32_t t_row,
                     int32_t x_row, int32_t x_chunk, int32_t* d_out, cudaStream_t st, int64_t* launches) {
    const size_t smem = 4 * kTileBytes + 64 + 1024;
    cudaError_t e = cudaFuncSetAttribute(debug_dot_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    debug_dot_kernel<<<1, 128, smem, st>>>(tmap, tmap_x, q_row, t_row, x_row, x_chunk, d_out);
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
