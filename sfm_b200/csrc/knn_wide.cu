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
//  * candidate units are 16 columns and their tracking is branch-free (packed keys, a min/max
//    insertion network, predicated parking stores): the N = 128 kernel's divergent "unit event"
//    path was executed by half of all warp-chunks and cost more than its max tree;
//  * exactness at ties is restored at the end of a work item (warp-cooperative recompute of a
//    tied third unit, brute-force fix-up pass for double ties) instead of an exact mode inside
//    the loop.
// Measured on B200 (36 x 8192): 2133 TOP/s useful vs 1490 for knn.cu.
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
constexpr int kFoldC = 255 * 31 * 255 + 254;

constexpr int kUnit = 16;
constexpr uint32_t kSlotRow = kEpiWarps * 32 * 16;          // bytes between a thread's uint4s
constexpr uint32_t kSlotStride = (kUnit / 4) * kSlotRow;    // bytes per slot
constexpr size_t kSlotBytes = 2 * size_t(kSlotStride);      // 64 KiB
constexpr int kFlagged = -2;                                // out[row].x of a row left to knn_fix_kernel
constexpr size_t kXchgBytes = 2 * SFM_QBLOCK * 16 + kEpiWarps * 32 * 4;
constexpr size_t kSmemData = size_t(kAStages) * kABlockBytes + size_t(kBStages) * kStageBytes + kAxBytes +
                             kSlotBytes + kXchgBytes;
constexpr int kNumBars = 2 * kAStages + 2 * kBStages + 2 * kATiles + 1;
constexpr size_t kSmemBytes = kSmemData + kNumBars * 8 + 16;   // the dynamic window itself is 1024-aligned (checked)
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

// ---- folded epilogue: branch-free tracking of the best 16-column units per thread.
// The exact ranking key is K = 2 a.b - ||b||^2 = 2 (v - C) - p (maximised), v the folded
// accumulator (0 <= v < 2^23) and p the parity of ||b||^2.  Both nearest neighbours have v >= V2,
// the second largest v of the row, so they lie in units whose maximum m is >= m2, the second
// largest UNIT maximum (m2 <= V2).  Per unit a thread forms one packed u32 key
//     [ m >> sh | uidrev : ubits | slot : 1 ],   uidrev = (units of the work item - 1) - unit index
// (an earlier unit wins on equal m) and keeps the four largest keys k1 >= k2 >= dm >= dd with a
// min/max insertion network.  The units of k1 and k2 are PARKED: when a key enters the top two its
// 16 accumulators are stored to shared-memory slot `slot`, the slot of the key it evicts (k2).
// At the end of the work item the parked values give the exact top-2 (resolve_units).  A third
// unit can only matter if its (truncated) maximum equals k2's: then dm names it and its 16 keys
// are recomputed from the descriptors (recompute_units, ~2 % of rows); if dd ties as well the row
// is flagged for knn_fix_kernel (brute force).  ubits / sh depend on the train image size:
// 8192 rows need 8 bits and no truncation, 32768 rows 10 bits and sh = 2 (ties on the truncated
// maximum are merely more frequent; the logic above is exact for any monotone truncation).
struct Trk {
    uint32_t k1, k2, dm, dd;
    uint32_t thr;   // SKIP variant: maximum (key units) of the k2 unit -- lower units cannot matter any more
};

struct KeyFmt {
    uint32_t sh, mul, ubits;   // key = (m >> sh) * mul + (uidrev << 1) + slot, mul = 2^(ubits+1)
    int n_units;               // units per thread over the whole work item (8 per train tile)
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

// Insert under the full (key desc, index asc) order (columns arrive in any order).
__device__ __forceinline__ void top2k_insert_ordered(Top2K& s, int key, int col) {
    if (s.i1 < 0 || key_before(key, col, s.k1, s.i1)) {
        s.k2 = s.k1;
        s.i2 = s.i1;
        s.k1 = key;
        s.i1 = col;
    } else if (s.i2 < 0 || key_before(key, col, s.k2, s.i2)) {
        s.k2 = key;
        s.i2 = col;
    }
}

// 16 folded accumulators of one row = one unit, chunk c of its tile; ut = uidrev << 1 of chunk 0.
// MASKED: columns >= t_cnt (last, partial train tile) count as v = 0 (real accumulators are >= 0;
// resolve_units skips them by column).  TREE_ONLY: timing probe.
// SH0: train images up to 8192 rows (ubits = 8, no truncation): the key is m * 512 + ... with
// compile-time constants.
// SKIP: a unit whose maximum is below the maximum of the thread's k2 unit can neither enter the top two nor
// tie with them (k2 only grows), so when that holds for every lane of the warp the whole bookkeeping below
// is skipped (warp-uniform branch); lanes that are below while others are not take part with m = 0.
// Measured (16 x 8192, same box): 1950 TOP/s with the skip against 2050 without -- the vote and the
// reconvergence cost more than the 40 % of skipped updates save, so the default kernel is built without it
// (knn mode 15 runs it; bit-exact on every kNN test).
template <bool MASKED, bool TREE_ONLY, bool SH0, bool SKIP>
__device__ __forceinline__ void fold_chunk(Trk& s, const uint32_t (&acc)[16], int col0, int t_cnt,
                                           uint32_t slots, const KeyFmt& kf, uint32_t ut, uint32_t c) {
    int v[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        v[e] = static_cast<int>(acc[e]);
        if (MASKED) v[e] = (col0 + e < t_cnt) ? v[e] : 0;
    }
    const int a0 = __vimax3_s32(v[0], v[1], v[2]), a1 = __vimax3_s32(v[3], v[4], v[5]);
    const int a2 = __vimax3_s32(v[6], v[7], v[8]), a3 = __vimax3_s32(v[9], v[10], v[11]);
    const int a4 = __vimax3_s32(v[12], v[13], v[14]);
    const int m = max(__vimax3_s32(a0, a1, a2), __vimax3_s32(a3, a4, v[15]));
    if (TREE_ONLY) {
        s.k1 = max(s.k1, static_cast<uint32_t>(m));
        return;
    }
    uint32_t mk = SH0 ? static_cast<uint32_t>(m) : (static_cast<uint32_t>(m) >> kf.sh);
    if (SKIP) {
        const bool pass = mk >= s.thr;
        if (!__any_sync(0xffffffffu, pass)) return;
        mk = pass ? mk : 0u;
    }
    const uint32_t sb = s.k2 & 1u;
    // low key bits: (uidrev of chunk c) << 1 | slot.  The units of a tile count down from a value whose
    // low three bits are 111 (n_units is a multiple of 8), so ut - 2c == ut ^ 2c: a warp-uniform value
    // (uniform datapath) that one LOP3 merges with the slot bit; the rest is one IMAD on the FMA pipe.
    const uint32_t low = sb | (ut ^ (2u * c));
    const uint32_t key = SH0 ? mk * 512u + low : mk * kf.mul + low;
    if (key > s.k2) {
        const uint32_t dst = slots + sb * kSlotStride;
#pragma unroll
        for (int g = 0; g < kUnit / 4; ++g)
            sts128(dst + g * kSlotRow, v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
    }
    const uint32_t t1 = min(s.k1, key);
    s.k1 = max(s.k1, key);
    const uint32_t t2 = min(s.k2, t1);
    s.k2 = max(s.k2, t1);
    const uint32_t t3 = min(s.dm, t2);
    s.dm = max(s.dm, t2);
    s.dd = max(s.dd, t3);
    if (SKIP) s.thr = s.k2 >> (kf.ubits + 1);
}

// First train column of the unit a key names (colbase: first column of this thread's half).
__device__ __forceinline__ int key_col(uint32_t key, const KeyFmt& kf, int colbase) {
    const int uidrev = static_cast<int>((key >> 1) & ((1u << kf.ubits) - 1u));
    const int idx = kf.n_units - 1 - uidrev;
    return (idx >> 3) * kTileN + colbase + (idx & 7) * kUnit;
}

// Exact top-2 of the (at most) two parked units, branch-free.  The 32 parked accumulators become
// packed candidates  x = (2v + 2 - parity) << 5 | (31 - idx)  with idx = 0..15 for the unit with the
// lower columns and 16..31 for the other one: 2v - parity orders like the exact key (it is the key
// plus 2C), the tag makes the lowest column win among equal keys (OpenCV's rule), invalid candidates
// are 0 and every valid one is >= 32.  Two running maxima (3 min/max per candidate) replace the
// divergent insertion of one candidate at a time.
struct ParkedUnit {
    uint4 q[kUnit / 4];
    int4 nb[kUnit / 4];
};

// c < 0 (fewer than two units seen: tiny train images) reads unit 0 and masks everything.
__device__ __forceinline__ void resolve_load(ParkedUnit& u, int c, uint32_t src, const int32_t* __restrict__ nbase) {
    const int cc = max(c, 0);
#pragma unroll
    for (int g = 0; g < kUnit / 4; ++g) {
        u.nb[g] = __ldg(reinterpret_cast<const int4*>(nbase + cc) + g);
        u.q[g] = lds128(src + g * kSlotRow);
    }
}

__device__ __forceinline__ void resolve_pass(uint32_t& m1, uint32_t& m2, const ParkedUnit& u, int c, int t_cnt,
                                             int unit) {
    const int n_ok = c < 0 ? 0 : t_cnt - c;     // valid columns of the unit (>= 16 unless it is the last one)
#pragma unroll
    for (int g = 0; g < kUnit / 4; ++g) {
        const uint32_t vv[4] = {u.q[g].x, u.q[g].y, u.q[g].z, u.q[g].w};
        const int nn[4] = {u.nb[g].x, u.nb[g].y, u.nb[g].z, u.nb[g].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const uint32_t tag = 64u + 31u - static_cast<uint32_t>(unit * kUnit + 4 * g + e);
            uint32_t x = vv[e] * 64u + tag - 32u * static_cast<uint32_t>(nn[e] & 1);
            if (n_ok < kUnit && 4 * g + e >= n_ok) x = 0;
            m2 = max(m2, min(m1, x));
            m1 = max(m1, x);
        }
    }
}

__device__ __forceinline__ Top2K resolve_units(const Trk& s, int n_seen, const KeyFmt& kf, int colbase,
                                               uint32_t slots, const int32_t* __restrict__ nbase, int t_cnt) {
    const int s1 = s.k1 & 1, s2 = s.k2 & 1;
    const int c1 = n_seen >= 1 ? key_col(s.k1, kf, colbase) : -1;
    const int c2 = n_seen >= 2 ? key_col(s.k2, kf, colbase) : -1;
    const bool second_first = c2 >= 0 && c2 < c1;
    const int ca = second_first ? c2 : c1, cb = second_first ? c1 : c2;
    const uint32_t sa = slots + (second_first ? s2 : s1) * kSlotStride;
    const uint32_t sb = slots + (second_first ? s1 : s2) * kSlotStride;
    // all sixteen loads first: the norm rows come from L2, one latency for both units
    ParkedUnit ua, ub;
    resolve_load(ua, ca, sa, nbase);
    resolve_load(ub, cb, sb, nbase);
    uint32_t m1 = 0, m2 = 0;
    resolve_pass(m1, m2, ua, ca, t_cnt, 0);
    resolve_pass(m1, m2, ub, cb, t_cnt, 1);
    Top2K r{INT_MIN, INT_MIN, -1, -1};
    if (m1) {
        const int idx = 31 - static_cast<int>(m1 & 31u);
        r.k1 = static_cast<int>(m1 >> 5) - 2 - 2 * kFoldC;
        r.i1 = idx < kUnit ? ca + idx : cb + idx - kUnit;
    }
    if (m2) {
        const int idx = 31 - static_cast<int>(m2 & 31u);
        r.k2 = static_cast<int>(m2 >> 5) - 2 - 2 * kFoldC;
        r.i2 = idx < kUnit ? ca + idx : cb + idx - kUnit;
    }
    return r;
}

// Rare path (2 % of rows): the exact keys of one dropped unit, recomputed from the descriptors on
// the CUDA cores by the WHOLE warp for one row at a time -- lane e and lane 16 + e each take 64
// bytes of train row col + e (dp4a), the 16 keys are then fed to the row's lane in column order
// and merged under the (key desc, index asc) order.  All lanes must call it (uniform control flow).
__device__ __noinline__ Top2K recompute_units(Top2K r, bool need, int q_abs, int col, const uint8_t* __restrict__ desc_q,
                                              const uint8_t* __restrict__ desc, int t_row, int t_cnt,
                                              const int32_t* __restrict__ nbase) {
    const int lane = threadIdx.x & 31;
    unsigned todo = __ballot_sync(0xffffffffu, need);
    while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const int qa = __shfl_sync(0xffffffffu, q_abs, src);
        const int cc = __shfl_sync(0xffffffffu, col, src);
        const int c = cc + (lane & 15);
        const int part = lane >> 4;   // which 64-byte half of the 128-byte rows
        uint32_t dot = 0;
        if (c < t_cnt) {
            const uint4* ap = reinterpret_cast<const uint4*>(desc_q + static_cast<size_t>(qa) * kRowBytes) + 4 * part;
            const uint4* bp = reinterpret_cast<const uint4*>(desc + static_cast<size_t>(t_row + c) * kRowBytes) + 4 * part;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint4 a = __ldg(ap + i), b = __ldg(bp + i);
                dot = __dp4a(a.x, b.x, dot);
                dot = __dp4a(a.y, b.y, dot);
                dot = __dp4a(a.z, b.z, dot);
                dot = __dp4a(a.w, b.w, dot);
            }
        }
        dot += __shfl_xor_sync(0xffffffffu, dot, 16);
        const int key = c < t_cnt ? 2 * static_cast<int>(dot) - __ldg(nbase + c) : INT_MIN;
#pragma unroll 1
        for (int e = 0; e < kUnit; ++e) {
            const int ke = __shfl_sync(0xffffffffu, key, e);
            if (lane == src && cc + e < t_cnt) top2k_insert_ordered(r, ke, cc + e);
        }
    }
    return r;
}

struct SmemLayout {
    uint8_t* a;      // kAStages x (2 x 128 query rows x 128 B)
    uint8_t* b;      // kBStages x (256 train rows x 128 B | 256 x 32 B fold operand)
    uint8_t* ax;     // constant A operand of the fold step (128 x 32 B)
    uint4* slots;    // parked candidate units: [slot][4 x uint4][thread]
    uint4* xchg;     // [2][256] per-row results handed from the upper to the lower column half
    int32_t* pub;    // [2][256] tie flag handed over with them
    uint64_t* a_full;
    uint64_t* a_empty;
    uint64_t* b_full;
    uint64_t* b_empty;
    uint64_t* t_full;    // [accumulator]
    uint64_t* t_empty;
    uint64_t* ax_full;
    uint32_t* tmem_slot;
    __device__ explicit SmemLayout(uint8_t* raw) {
        uint8_t* base = raw;   // 1024-aligned (checked at kernel entry): no slack to spare for rounding up
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
template <int MODE, bool SKIP>
__global__ void __launch_bounds__(kThreads, 1)
knn2_wide_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap,
                 const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_ax,
                 const uint8_t* __restrict__ desc_q, const int32_t* __restrict__ norms_q,
                 const uint8_t* __restrict__ desc,
                 const WorkItem* __restrict__ work, long long n_work, const int32_t* __restrict__ norms_pad,
                 int32_t* __restrict__ flags, int force, uint32_t probe_flags, uint32_t wait_ns,
                 uint32_t epi_wait_ns, int2* __restrict__ fix_list, int fix_cap, int4* __restrict__ out) {
    constexpr int kProducerWarp = kEpiWarps;
    constexpr int kMmaWarp = kEpiWarps + 1;
    // flags[0] != 0: some row's norm is beyond the fold range -> the generic kernel runs instead
    if (!force && *static_cast<volatile const int32_t*>(flags) != 0) return;

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    if (smem_u32(smem_raw) & 1023u) __trap();   // SWIZZLE_128B tiles need 1024-byte alignment
    SmemLayout sm(smem_raw);
    // broadcast from lane 0: tells the compiler the warp index (and every TMEM address derived from it)
    // is warp-uniform, so it lives in uniform registers instead of being re-materialised per tcgen05.ld
    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;

    if (warp == kProducerWarp && lane == 0) {
        tma_prefetch_desc(&tmap_q);
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
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *sm.tmem_slot, 0);

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
                    tma_load_2d(a_dst + a * kATileBytes, &tmap_q, &sm.a_full[ap.stage], 0, q_row + a * kTileM);
            }
            __syncwarp();
            ap.advance();
            const int n_tiles = (t_cnt + kTileN - 1) / kTileN;
            for (int t = 0; t < n_tiles; ++t) {
                mbar_wait_suspend(&sm.b_empty[bp.stage], bp.phase ^ 1, wait_ns);
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
                mbar_wait_suspend(&sm.b_full[bp.stage], bp.phase, wait_ns);
                const uint32_t bsm = smem_u32(sm.b + size_t(bp.stage) * kStageBytes);
                const uint64_t bdesc0 = make_kmajor_sw128_desc(bsm);
                const uint64_t xdesc = make_kmajor_sw32_desc(bsm + kBTileBytes);
#pragma unroll
                for (int a = 0; a < kATiles; ++a) {
                    mbar_wait_suspend(&sm.t_empty[a], tphase ^ 1, wait_ns);
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
        uint64_t* full = &sm.t_full[acc];
        uint64_t* empty = &sm.t_empty[acc];
        uint32_t tphase = 0;
        int item = 0;
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x, ++item) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int4 w1 = __ldg(reinterpret_cast<const int4*>(work + w) + 1);
            const int q_row = w0.x, q_cnt = w0.y, t_row = w0.z, t_cnt = w0.w, t_pad = w1.x;
            const long long out_row =
                static_cast<long long>(static_cast<uint32_t>(w1.z)) | (static_cast<long long>(w1.w) << 32);
            const bool valid = row < q_cnt;
            const int na = valid ? __ldg(norms_q + q_row + row) : 0;
            const int n_tiles = (t_cnt + kTileN - 1) / kTileN;
            const int32_t* nbase = norms_pad + t_pad;
            KeyFmt kf;
            kf.n_units = n_tiles * (kColsPerThread / kUnit);
            kf.ubits = max(32 - __clz(kf.n_units - 1), 8);
            kf.sh = kf.ubits > 8 ? kf.ubits - 8 : 0;   // m < 2^23: (23 - sh) + ubits + 1 <= 32 bits
            kf.mul = 2u << kf.ubits;
            // units this thread will see: 8 per full tile, fewer in the last, partial one
            const int rem = t_cnt - (n_tiles - 1) * kTileN - colbase;
            const int n_seen = (n_tiles - 1) * (kColsPerThread / kUnit) +
                               min(max((rem + kUnit - 1) / kUnit, 0), kColsPerThread / kUnit);
            Trk sk{1, 0, 0, 0, 0};   // empty keys: k1 owns slot 1, k2 slot 0 (the first unit is parked in slot 0)
            for (int t = 0; t < n_tiles; ++t) {
                mbar_wait_suspend(full, tphase, epi_wait_ns);
                tc_fence_after();
                tphase ^= 1;
                if (MODE == 2) {   // timing probe: TMA + MMA pipeline alone
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(empty);
                    continue;
                }
                // 16-column chunks, double-buffered: 32 live accumulator registers per thread
                uint32_t v[2][16];
                tmem_ld_32x16(taddr, v[0]);
                const uint32_t ut = static_cast<uint32_t>(kf.n_units - 1 - t * (kColsPerThread / kUnit)) << 1;
                if ((t + 1) * kTileN <= t_cnt && kf.sh == 0) {
#pragma unroll
                    for (int c = 0; c < kColsPerThread / 16; ++c) {
                        tmem_ld_wait();
                        if (c + 1 < kColsPerThread / 16) {
                            tmem_ld_32x16(taddr + (c + 1) * 16, v[(c + 1) & 1]);
                        } else {
                            // every column of this thread is in registers: hand the accumulator back
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(empty);
                        }
                        fold_chunk<false, MODE == 1, true, SKIP>(sk, v[c & 1], t * kTileN + colbase + c * 16, t_cnt, my_slots,
                                                     kf, ut, c);
                    }
                } else if ((t + 1) * kTileN <= t_cnt) {
#pragma unroll
                    for (int c = 0; c < kColsPerThread / 16; ++c) {
                        tmem_ld_wait();
                        if (c + 1 < kColsPerThread / 16) {
                            tmem_ld_32x16(taddr + (c + 1) * 16, v[(c + 1) & 1]);
                        } else {
                            // every column of this thread is in registers: hand the accumulator back
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(empty);
                        }
                        fold_chunk<false, MODE == 1, false, SKIP>(sk, v[c & 1], t * kTileN + colbase + c * 16, t_cnt, my_slots,
                                                     kf, ut, c);
                    }
                } else {   // last, partial train tile: chunks wholly beyond the image are skipped
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
                        if (col0 < t_cnt) fold_chunk<true, MODE == 1, false, SKIP>(sk, v[c & 1], col0, t_cnt, my_slots, kf, ut, c);
                    }
                }
            }
            if (MODE != 0) {
                if (valid && half == 0) out[out_row + row] = make_int4(static_cast<int>(sk.k1), static_cast<int>(sk.k2), na, 0);   // keeps the probe live
                continue;
            }
            // SKIP: units below the threshold were never inserted; a key that was is >= 2 (the empty markers are 0, 1)
            const int n_in = SKIP ? (sk.k1 >= 2u) + (sk.k2 >= 2u) + (sk.dm >= 2u) + (sk.dd >= 2u) : n_seen;
            Top2K r = resolve_units(sk, n_in, kf, colbase, my_slots, nbase, t_cnt);
            // a dropped unit whose (truncated) maximum equals the second kept one may hold a candidate:
            // the best dropped one is recomputed here, a second one sends the row to knn_fix_kernel
            const uint32_t mf = kf.ubits + 1;
            r = recompute_units(r, valid && n_in >= 3 && (sk.dm >> mf) == (sk.k2 >> mf), q_row + row,
                                key_col(sk.dm, kf, colbase), desc_q, desc, t_row, t_cnt, nbase);
            const bool tie = n_in >= 4 && (sk.dd >> mf) == (sk.k2 >> mf);
            // the thread of the upper column half hands its top-2 to the lower half's thread,
            // which merges under the (key, lowest index) order and writes the row
            const uint32_t xaddr = smem_u32(sm.xchg) + ((item & 1) * SFM_QBLOCK + row) * 16;
            const uint32_t faddr = smem_u32(sm.pub) + ((item & 1) * SFM_QBLOCK + row) * 4;
            if (half == 1) {
                sts128(xaddr, r.k1, r.k2, r.i1, r.i2);
                sts32(faddr, tie ? 1u : 0u);
            }
            asm volatile("bar.sync %0, 64;" ::"r"(1 + (warp & 7)) : "memory");
            if (half == 0) {
                const uint4 o = lds128(xaddr);
                const bool flagged = tie || lds32(faddr) != 0;
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
                if (valid && flagged) {   // rare: left to knn_fix_kernel
                    atomicAdd(flags + 1, 1);   // statistics since sfm_set_images
                    const int e = atomicAdd(flags + 2, 1);
                    if (e < fix_cap) fix_list[e] = make_int2(static_cast<int>(w), row);
                }
                if (valid) {
                    const int i1 = mrg.i1 == INT_MAX ? -1 : mrg.i1;
                    const int i2 = mrg.i2 == INT_MAX ? -1 : mrg.i2;
                    const int d1 = i1 < 0 ? INT_MAX : na - mrg.k1;
                    const int d2 = i2 < 0 ? INT_MAX : na - mrg.k2;
                    out[out_row + row] = flagged ? make_int4(kFlagged, kFlagged, 0, 0) : make_int4(i1, i2, d1, d2);
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

// Rows the wide kernel flagged (two dropped units tie with the second kept one) are recomputed
// here exactly on the CUDA cores.  The main kernel marks them (out[row].x == kFlagged) and appends
// (work item, row) to a list -- flags[2] entries, at most fix_cap stored.  A block takes a row at a
// time: threads stride over the train columns (ascending per thread, strict compares), partial
// top-2s merged under the (key desc, index asc) order = OpenCV's lowest-index tie rule.  Rows
// beyond the list capacity (degenerate data: everything ties) are found by scanning the work items.
__device__ void fix_row(const WorkItem* __restrict__ work, long long w, int r, const uint8_t* __restrict__ desc_q,
                        const int32_t* __restrict__ norms_q, const uint8_t* __restrict__ desc,
                        const int32_t* __restrict__ norms, int4* __restrict__ out, Top2K* s_part) {
    const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
    const int4 w1 = __ldg(reinterpret_cast<const int4*>(work + w) + 1);
    const int q_row = w0.x, t_row = w0.z, t_cnt = w0.w;
    const long long out_row =
        static_cast<long long>(static_cast<uint32_t>(w1.z)) | (static_cast<long long>(w1.w) << 32);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint4 a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
        a[i] = __ldg(reinterpret_cast<const uint4*>(desc_q + static_cast<size_t>(q_row + r) * kRowBytes) + i);
    Top2K best{INT_MIN, INT_MIN, INT_MAX, INT_MAX};
#pragma unroll 2
    for (int c = threadIdx.x; c < t_cnt; c += SFM_QBLOCK) {
        const uint4* bp = reinterpret_cast<const uint4*>(desc + static_cast<size_t>(t_row + c) * kRowBytes);
        uint32_t dot = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint4 b = __ldg(bp + i);
            dot = __dp4a(a[i].x, b.x, dot);
            dot = __dp4a(a[i].y, b.y, dot);
            dot = __dp4a(a[i].z, b.z, dot);
            dot = __dp4a(a[i].w, b.w, dot);
        }
        top2k_insert(best, 2 * static_cast<int>(dot) - __ldg(norms + t_row + c), c);
    }
    auto merge = [](const Top2K& x, const Top2K& y) {
        Top2K m;
        if (key_before(x.k1, x.i1, y.k1, y.i1)) {
            m.k1 = x.k1;
            m.i1 = x.i1;
            const bool keep = key_before(x.k2, x.i2, y.k1, y.i1);
            m.k2 = keep ? x.k2 : y.k1;
            m.i2 = keep ? x.i2 : y.i1;
        } else {
            m.k1 = y.k1;
            m.i1 = y.i1;
            const bool keep = key_before(y.k2, y.i2, x.k1, x.i1);
            m.k2 = keep ? y.k2 : x.k1;
            m.i2 = keep ? y.i2 : x.i1;
        }
        return m;
    };
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Top2K b;
        b.k1 = __shfl_xor_sync(0xffffffffu, best.k1, o);
        b.k2 = __shfl_xor_sync(0xffffffffu, best.k2, o);
        b.i1 = __shfl_xor_sync(0xffffffffu, best.i1, o);
        b.i2 = __shfl_xor_sync(0xffffffffu, best.i2, o);
        best = merge(best, b);
    }
    if (lane == 0) s_part[warp] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
        Top2K m = s_part[0];
        for (int k = 1; k < SFM_QBLOCK / 32; ++k) m = merge(m, s_part[k]);
        const int na = __ldg(norms_q + q_row + r);
        const int i1 = m.i1 == INT_MAX ? -1 : m.i1;
        const int i2 = m.i2 == INT_MAX ? -1 : m.i2;
        out[out_row + r] = make_int4(i1, i2, i1 < 0 ? INT_MAX : na - m.k1, i2 < 0 ? INT_MAX : na - m.k2);
    }
    __syncthreads();
}

__global__ void __launch_bounds__(SFM_QBLOCK) knn_fix_kernel(const WorkItem* __restrict__ work, long long n_work,
                                                             const uint8_t* __restrict__ desc_q,
                                                             const int32_t* __restrict__ norms_q,
                                                             const uint8_t* __restrict__ desc,
                                                             const int32_t* __restrict__ norms,
                                                             const int32_t* __restrict__ flags,
                                                             const int2* __restrict__ fix_list, int fix_cap,
                                                             int4* __restrict__ out) {
    const int cnt = *(static_cast<volatile const int32_t*>(flags) + 2);
    if (cnt == 0) return;
    __shared__ Top2K s_part[SFM_QBLOCK / 32];
    __shared__ int s_list[SFM_QBLOCK];
    __shared__ int s_n;
    const int n_list = cnt < fix_cap ? cnt : fix_cap;
    for (int e = blockIdx.x; e < n_list; e += gridDim.x) {
        const int2 it = __ldg(fix_list + e);
        fix_row(work, it.x, it.y, desc_q, norms_q, desc, norms, out, s_part);
    }
    if (cnt <= fix_cap) return;
    for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {   // list overflow: scan for the markers
        const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
        const int4 w1 = __ldg(reinterpret_cast<const int4*>(work + w) + 1);
        const long long out_row =
            static_cast<long long>(static_cast<uint32_t>(w1.z)) | (static_cast<long long>(w1.w) << 32);
        if (threadIdx.x == 0) s_n = 0;
        __syncthreads();
        if (static_cast<int>(threadIdx.x) < w0.y && out[out_row + threadIdx.x].x == kFlagged)
            s_list[atomicAdd(&s_n, 1)] = threadIdx.x;
        __syncthreads();
        const int n = s_n;
        for (int k = 0; k < n; ++k) fix_row(work, w, s_list[k], desc_q, norms_q, desc, norms, out, s_part);
        __syncthreads();
    }
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

template <int MODE, bool SKIP>
cudaError_t launch_variant(const KnnQuery& q, const CUtensorMap& tmap, const CUtensorMap& tmap_x, const CUtensorMap& tmap_ax,
                           const uint8_t* d_desc, const WorkItem* d_work, int64_t n_work,
                           const int32_t* d_norms_pad, int32_t* d_flags, int force, const KnnTuning& tn,
                           int2* d_fix_list, int fix_cap, int32_t* d_knn, unsigned grid, cudaStream_t st) {
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(knn2_wide_kernel<MODE, SKIP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(kSmemBytes));
        if (e != cudaSuccess) return e;
        configured = true;
    }
    knn2_wide_kernel<MODE, SKIP><<<grid, kThreads, kSmemBytes, st>>>(*q.tmap, tmap, tmap_x, tmap_ax, q.desc, q.norms, d_desc,
                                                              d_work, n_work,
                                                              d_norms_pad, d_flags, force, tn.probe_flags, tn.wait_ns,
                                                              tn.epi_wait_ns, d_fix_list, fix_cap, reinterpret_cast<int4*>(d_knn));
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
                     const uint8_t* d_desc, const WorkItem* d_work, int64_t n_work, const int32_t* d_norms,
                     const int32_t* d_norms_pad,
                     int32_t* d_flags, int variant, int force, const KnnTuning& tn, int32_t* d_fix_list,
                     int32_t fix_cap, int32_t* d_knn, int num_sms, cudaStream_t st, int64_t* launches,
                     const KnnQuery* query) {
    if (n_work <= 0) return 0;
    // query rows: the registered descriptor array itself unless the caller brings its own (mutual check)
    const KnnQuery q = query ? *query : KnnQuery{&tmap, d_desc, d_norms};
    const unsigned grid = static_cast<unsigned>(n_work < num_sms ? n_work : num_sms);
    int2* fl = reinterpret_cast<int2*>(d_fix_list);
    cudaError_t e = cudaMemsetAsync(d_flags + 2, 0, sizeof(int32_t), st);   // length of this launch's fix list
    if (e != cudaSuccess) return static_cast<int>(e);
    if (variant == 1)
        e = launch_variant<1, false>(q, tmap, tmap_x, tmap_ax, d_desc, d_work, n_work, d_norms_pad, d_flags, 1,
                                     tn, fl, fix_cap, d_knn, grid, st);
    else if (variant == 2)
        e = launch_variant<2, false>(q, tmap, tmap_x, tmap_ax, d_desc, d_work, n_work, d_norms_pad, d_flags, 1,
                                     tn, fl, fix_cap, d_knn, grid, st);
    else if (variant == 3)   // the kNN WITH the warp-uniform skip of the unit bookkeeping (A/B: 5 % slower)
        e = launch_variant<0, true>(q, tmap, tmap_x, tmap_ax, d_desc, d_work, n_work, d_norms_pad, d_flags,
                                    force, tn, fl, fix_cap, d_knn, grid, st);
    else
        e = launch_variant<0, false>(q, tmap, tmap_x, tmap_ax, d_desc, d_work, n_work, d_norms_pad, d_flags,
                                     force, tn, fl, fix_cap, d_knn, grid, st);
    ++*launches;
    if (e == cudaSuccess && (variant == 0 || variant == 3) && !(tn.probe_flags & 4u)) {   // (probe bit 2: main kernel alone)
        knn_fix_kernel<<<2 * num_sms, SFM_QBLOCK, 0, st>>>(d_work, n_work, q.desc, q.norms, d_desc, d_norms, d_flags, fl,
                                                          fix_cap, reinterpret_cast<int4*>(d_knn));
        ++*launches;
        e = cudaGetLastError();
    }
    return static_cast<int>(e);
}

int launch_knn_fix(const WorkItem* d_work, int64_t n_work, const uint8_t* d_desc, const int32_t* d_norms,
                   const int32_t* d_flags, const int32_t* d_fix_list, int32_t fix_cap, int32_t* d_knn, int num_sms,
                   cudaStream_t st, int64_t* launches) {
    knn_fix_kernel<<<2 * num_sms, SFM_QBLOCK, 0, st>>>(d_work, n_work, d_desc, d_norms, d_desc, d_norms, d_flags,
                                                      reinterpret_cast<const int2*>(d_fix_list), fix_cap,
                                                      reinterpret_cast<int4*>(d_knn));
    ++*launches;
    return static_cast<int>(cudaGetLastError());
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
