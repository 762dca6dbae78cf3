// K1, CTA-pair variant (replaces cv2.BFMatcher(NORM_L2).knnMatch(k=2), reference
// utils/find_matches.py:144-147; same exact integer arithmetic and tie rule as knn_wide.cu).
//
// Why a pair.  The wide kernel (knn_wide.cu) is bound by two things that one SM cannot fix:
//  * hand-over: with two 256-column accumulators the chain MMA -> epilogue -> MMA of an accumulator is
//    serial, and its epilogue holds the accumulator for the whole reduction (there are not enough
//    registers to copy 128 columns per thread out first);
//  * ALU-pipe work: 16-column candidate units cost 18 ALU instructions per 16 accumulators; 32-column
//    units cost 13.5, but their parking slots (2 x 32 words per thread) do not fit beside the operand
//    rings of one SM.
// A cluster of two CTAs (the two SMs of a TPC) issues tcgen05.mma.cta_group::2 with M = 256: each CTA
// holds 128 query rows, their accumulators and HALF of every train tile.  Per CTA that halves the
// operand rings (A 2 x 16 KiB, B 4 x 10 KiB), which frees 128 KiB for the parking slots, and all 512
// TMEM columns serve 128 rows: FOUR 128-column accumulator stages.  An epilogue thread owns one row and
// one 32-column quarter of every group (= one candidate unit), copies it to registers in one
// tcgen05.ld and hands the stage back at once, so a stage is held for a TMEM round trip only and three
// other stages (960 tensor cycles) cover it; the loads are double buffered against the reduction.
//
// Work item = 256 query rows x one train image (as knn_wide.cu); cluster c takes items c, c + C, ...
#include <climits>

#include "kernels.h"
#include "ptx.cuh"

namespace sfm {

namespace {

constexpr int kRowBytes = SFM_DESC_DIM;
constexpr int kUmmaK = 32;
constexpr int kKSteps = kRowBytes / kUmmaK;          // 4
constexpr int kTileM = 128;                          // query rows per CTA (M = 256 over the pair)
constexpr int NT = 128;                              // UMMA N = train rows per accumulator group (= stage)
constexpr int kMmaN = NT;
constexpr int kBRows = kMmaN / 2;                    // train rows of a group staged by each CTA
constexpr int kAccStages = 4;
constexpr int kBStages = 4;
constexpr int kAStages = 2;
constexpr int kATileBytes = kTileM * kRowBytes;      // 16 KiB
constexpr int kBTileBytes = kBRows * kRowBytes;      // 8 KiB
constexpr int kXTileBytes = kBRows * kUmmaK;         // 2 KiB: fold operand rows of the half tile
constexpr int kAxBytes = kTileM * kUmmaK;            // 4 KiB: constant fold-step A operand
constexpr int kTmemCols = kAccStages * NT;           // 512
constexpr int kEpiWarps = 16;                        // 4 lane quadrants x 4 column quarters
constexpr int kMmaWarps = 2;                          // groups alternate between two issuing warps
constexpr int kThreads = (kEpiWarps + 1 + kMmaWarps) * 32;
constexpr int kUnit = 32;                            // columns per thread and group = one candidate unit
constexpr uint32_t kIdesc = make_idesc_u8(2 * kTileM, kMmaN);
constexpr int kFoldC = 255 * 31 * 255 + 254;
constexpr uint32_t kSlotRow = kEpiWarps * 32 * 16;          // bytes between a thread's uint4s
constexpr uint32_t kSlotStride = (kUnit / 4) * kSlotRow;    // bytes per slot (64 KiB)
constexpr size_t kSlotBytes = 2 * size_t(kSlotStride);      // 128 KiB
constexpr int kFlagged = -2;
constexpr size_t kXchgBytes = 2 * 3 * kTileM * 16;          // [item parity][quarter 1..3][row] partial top-2
constexpr size_t kPubBytes = 2 * 3 * kTileM * 4;            // tie flags handed over with them
constexpr size_t kOffA = 0;
constexpr size_t kOffB = kOffA + size_t(kAStages) * kATileBytes;
constexpr size_t kOffX = kOffB + size_t(kBStages) * kBTileBytes;
constexpr size_t kOffAx = kOffX + size_t(kBStages) * kXTileBytes;
constexpr size_t kOffSlots = kOffAx + kAxBytes;
constexpr size_t kOffXchg = kOffSlots + kSlotBytes;
constexpr size_t kOffPub = kOffXchg + kXchgBytes;
constexpr size_t kOffThr = kOffPub + kPubBytes;
constexpr size_t kThrBytes = 2 * 4 * kTileM * 4;            // [item parity][quarter][row] published thresholds
constexpr size_t kSmemData = kOffThr + kThrBytes;
constexpr int kNumBars = 1 + 2 * kAStages + 2 * kBStages + 2 * kAccStages;
constexpr size_t kSmemBytes = kSmemData + kNumBars * 8 + 16;
static_assert(kSmemBytes <= 232448, "shared memory budget");
static_assert(kOffB % 1024 == 0 && kOffX % 256 == 0 && kOffAx % 256 == 0, "swizzle atom alignment");

// K-major operand whose rows are 32 bytes (one UMMA K step), SWIZZLE_32B: 8-row groups of 256 bytes.
__device__ __forceinline__ uint64_t make_kmajor_sw32_desc_p(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
    d |= static_cast<uint64_t>(1) << 16;
    d |= static_cast<uint64_t>(256 >> 4) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(6) << 61;
    return d;
}

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

// ---- candidate tracking, exactly as knn_wide.cu but on 32-column units (one per thread and group):
// key = [ m >> sh | uidrev : ubits | slot : 1 ], the four largest keys k1 >= k2 >= dm >= dd are kept,
// the units of k1 and k2 are parked in shared memory.
struct Trk {
    uint32_t k1, k2, dm, dd;
};

struct KeyFmt {
    uint32_t sh, mul, ubits;   // key = (m >> sh) * mul + (uidrev << 1) + slot, mul = 2^(ubits+1)
    int n_units;               // units per thread over the whole work item (one per train tile)
};

struct Top2K {
    int k1, k2, i1, i2;
};

__device__ __forceinline__ bool key_before(int ka, int ia, int kb, int ib) {
    return ka > kb || (ka == kb && ia < ib);
}

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

// (key desc, index asc) merge of two partial top-2s; empty entries carry INT_MIN / INT_MAX.
__device__ __forceinline__ Top2K merge_top2(const Top2K& x, const Top2K& y) {
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
}

// Maximum of the 32 folded accumulators of a unit: a 3-input max tree, 16 instructions.
// MASKED: columns >= t_cnt count as 0 (in place; real accumulators are >= 0).
template <bool MASKED>
__device__ __forceinline__ int unit_max(uint32_t (&acc)[32], int col0, int t_cnt) {
    int v[32];
#pragma unroll
    for (int e = 0; e < 32; ++e) {
        if (MASKED) acc[e] = (col0 + e < t_cnt) ? acc[e] : 0u;
        v[e] = static_cast<int>(acc[e]);
    }
    int a[10];
#pragma unroll
    for (int i = 0; i < 10; ++i) a[i] = __vimax3_s32(v[3 * i], v[3 * i + 1], v[3 * i + 2]);
    const int b0 = __vimax3_s32(a[0], a[1], a[2]), b1 = __vimax3_s32(a[3], a[4], a[5]);
    const int b2 = __vimax3_s32(a[6], a[7], a[8]), b3 = __vimax3_s32(a[9], v[30], v[31]);
    return max(__vimax3_s32(b0, b1, b2), b3);
}

// mk: the unit's maximum in key units (m >> sh), or 0 for a unit below the row's threshold.
template <bool SH0>
__device__ __forceinline__ void unit_update(Trk& s, const uint32_t (&acc)[32], uint32_t mk, uint32_t slots,
                                            const KeyFmt& kf, uint32_t ulow) {
    const uint32_t sb = s.k2 & 1u;
    const uint32_t low = sb | ulow;
    const uint32_t key = SH0 ? mk * 512u + low : mk * kf.mul + low;
    if (key > s.k2) {
        const uint32_t dst = slots + sb * kSlotStride;
#pragma unroll
        for (int g = 0; g < kUnit / 4; ++g)
            sts128(dst + g * kSlotRow, acc[4 * g], acc[4 * g + 1], acc[4 * g + 2], acc[4 * g + 3]);
    }
    const uint32_t t1 = min(s.k1, key);
    s.k1 = max(s.k1, key);
    const uint32_t t2 = min(s.k2, t1);
    s.k2 = max(s.k2, t1);
    const uint32_t t3 = min(s.dm, t2);
    s.dm = max(s.dm, t2);
    s.dd = max(s.dd, t3);
}

// First train column of the unit a key names (colbase: first column of this thread's quarter).
__device__ __forceinline__ int key_col(uint32_t key, const KeyFmt& kf, int colbase) {
    const int uidrev = static_cast<int>((key >> 1) & ((1u << kf.ubits) - 1u));
    return (kf.n_units - 1 - uidrev) * NT + colbase;
}

// Exact top-2 of the (at most) two parked units, branch-free.  Candidate of unit u (0: lower columns),
// element e, accumulator v, norm parity p:   x = 128 v + 191 - (32 u + e) - 64 p.
// x >> 6 = 2 v + 2 - p orders like the exact key 2 a.b - ||b||^2 (= x >> 6 minus 2 + 2C), the low six
// bits make the lowest column win among equal keys (OpenCV's rule); invalid candidates are 0, valid >= 64.
__device__ __forceinline__ void resolve_pass(uint32_t& m1, uint32_t& m2, uint32_t src, uint32_t par, int c, int t_cnt,
                                             int unit) {
    const int n_ok = c < 0 ? 0 : t_cnt - c;     // valid columns of the unit (>= 32 unless it is the last one)
#pragma unroll
    for (int g = 0; g < kUnit / 4; ++g) {
        const uint4 q = lds128(src + g * kSlotRow);
        const uint32_t vv[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int idx = 4 * g + e;
            uint32_t x = vv[e] * 128u + (191u - static_cast<uint32_t>(unit * kUnit + idx)) - ((par >> idx) & 1u) * 64u;
            if (n_ok < kUnit && idx >= n_ok) x = 0;
            m2 = max(m2, min(m1, x));
            m1 = max(m1, x);
        }
    }
}

// A key that was inserted (and its unit parked) is >= 2; the empty markers are 0 and 1.
__device__ __forceinline__ Top2K resolve_units(const Trk& s, const KeyFmt& kf, int colbase, uint32_t slots,
                                               const uint32_t* __restrict__ parbase, int t_cnt) {
    const int s1 = s.k1 & 1, s2 = s.k2 & 1;
    const int c1 = s.k1 >= 2u ? key_col(s.k1, kf, colbase) : -1;
    const int c2 = s.k2 >= 2u ? key_col(s.k2, kf, colbase) : -1;
    const bool second_first = c2 >= 0 && c2 < c1;
    const int ca = second_first ? c2 : c1, cb = second_first ? c1 : c2;
    const uint32_t sa = slots + (second_first ? s2 : s1) * kSlotStride;
    const uint32_t sb = slots + (second_first ? s1 : s2) * kSlotStride;
    const uint32_t pa = __ldg(parbase + (max(ca, 0) >> 5)), pb = __ldg(parbase + (max(cb, 0) >> 5));
    uint32_t m1 = 0, m2 = 0;
    resolve_pass(m1, m2, sa, pa, ca, t_cnt, 0);
    resolve_pass(m1, m2, sb, pb, cb, t_cnt, 1);
    Top2K r{INT_MIN, INT_MIN, -1, -1};
    if (m1) {
        const int idx = 63 - static_cast<int>(m1 & 63u);
        r.k1 = static_cast<int>(m1 >> 6) - 2 - 2 * kFoldC;
        r.i1 = idx < kUnit ? ca + idx : cb + idx - kUnit;
    }
    if (m2) {
        const int idx = 63 - static_cast<int>(m2 & 63u);
        r.k2 = static_cast<int>(m2 >> 6) - 2 - 2 * kFoldC;
        r.i2 = idx < kUnit ? ca + idx : cb + idx - kUnit;
    }
    return r;
}

// Rare path: the exact keys of one dropped 32-column unit, recomputed from the descriptors by the
// WHOLE warp for one row at a time (lane l takes train row col + l: 128 bytes, dp4a); the 32 keys are
// then fed to the row's lane in column order.  All lanes must call it (uniform control flow).
__device__ __noinline__ Top2K recompute_unit(Top2K r, bool need, int q_abs, int col, const uint8_t* __restrict__ desc,
                                             int t_row, int t_cnt, const int32_t* __restrict__ nbase) {
    const int lane = threadIdx.x & 31;
    unsigned todo = __ballot_sync(0xffffffffu, need);
    while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const int qa = __shfl_sync(0xffffffffu, q_abs, src);
        const int cc = __shfl_sync(0xffffffffu, col, src);
        const int c = cc + lane;
        uint32_t dot = 0;
        if (c < t_cnt) {
            const uint4* ap = reinterpret_cast<const uint4*>(desc + static_cast<size_t>(qa) * kRowBytes);
            const uint4* bp = reinterpret_cast<const uint4*>(desc + static_cast<size_t>(t_row + c) * kRowBytes);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint4 a = __ldg(ap + i), b = __ldg(bp + i);
                dot = __dp4a(a.x, b.x, dot);
                dot = __dp4a(a.y, b.y, dot);
                dot = __dp4a(a.z, b.z, dot);
                dot = __dp4a(a.w, b.w, dot);
            }
        }
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
    uint8_t* a;      // kAStages x (128 query rows x 128 B)
    uint8_t* b;      // kBStages x (64 train rows x 128 B)
    uint8_t* x;      // kBStages x (64 x 32 B fold operand)
    uint8_t* ax;     // constant A operand of the fold step (128 x 32 B)
    uint4* slots;    // parked candidate units: [slot][8 x uint4][thread]
    uint4* xchg;     // [2][3][128] partial top-2 of column quarters 1..3
    int32_t* pub;    // [2][3][128] their tie flags
    uint32_t* thr;   // [2][4][128] second-largest unit maximum of every (quarter, row), shared per row
    uint64_t* ax_full;   // leader's: counts both CTAs' loads
    uint64_t* a_full;    // leader's
    uint64_t* a_empty;   // both (multicast commit)
    uint64_t* b_full;    // leader's
    uint64_t* b_empty;   // both
    uint64_t* t_full;    // both
    uint64_t* t_empty;   // leader's: 16 epilogue warps of each CTA
    uint32_t* tmem_slot;
    __device__ explicit SmemLayout(uint8_t* base) {
        a = base + kOffA;
        b = base + kOffB;
        x = base + kOffX;
        ax = base + kOffAx;
        slots = reinterpret_cast<uint4*>(base + kOffSlots);
        xchg = reinterpret_cast<uint4*>(base + kOffXchg);
        pub = reinterpret_cast<int32_t*>(base + kOffPub);
        thr = reinterpret_cast<uint32_t*>(base + kOffThr);
        uint64_t* bars = reinterpret_cast<uint64_t*>(base + kSmemData);
        ax_full = bars;
        a_full = ax_full + 1;
        a_empty = a_full + kAStages;
        b_full = a_empty + kAStages;
        b_empty = b_full + kBStages;
        t_full = b_empty + kBStages;
        t_empty = t_full + kAccStages;
        tmem_slot = reinterpret_cast<uint32_t*>(t_empty + kAccStages);
    }
};

// One candidate unit from registers.  MODE 1: max tree only (timing probe).
// thr: lower bound (key units) of the row's second-largest unit maximum over all four quarters.  A unit
// below it holds neither neighbour; when that is so for every lane the whole update is skipped.
template <bool MASKED, bool SH0, int MODE>
__device__ __forceinline__ void process_unit(Trk& sk, uint32_t (&acc)[32], int t, int t_cnt, int colbase,
                                             uint32_t my_slots, const KeyFmt& kf, uint32_t thr) {
    const int col0 = t * NT + colbase;
    if (MASKED && col0 >= t_cnt) return;
    const int m = unit_max<MASKED>(acc, col0, t_cnt);
    if (MODE == 1) {
        sk.k1 = max(sk.k1, static_cast<uint32_t>(m));
        return;
    }
    const uint32_t mk = SH0 ? static_cast<uint32_t>(m) : (static_cast<uint32_t>(m) >> kf.sh);
    const bool pass = mk >= thr;
    if (!__any_sync(0xffffffffu, pass)) return;
    unit_update<SH0>(sk, acc, pass ? mk : 0u, my_slots, kf, static_cast<uint32_t>(kf.n_units - 1 - t) << 1);
}

// MODE 0: the kNN.  1: max tree only, 2: no epilogue (timing probes; output is not a kNN).
template <int MODE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
knn2_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_ax,
                 const uint8_t* __restrict__ desc, const WorkItem* __restrict__ work, long long n_work,
                 const int32_t* __restrict__ norms, const int32_t* __restrict__ norms_pad,
                 const uint32_t* __restrict__ par_bits, int32_t* __restrict__ flags, int force, uint32_t probe_flags,
                 int2* __restrict__ fix_list, int fix_cap, int4* __restrict__ out) {
    constexpr int kProducerWarp = kEpiWarps;
    constexpr int kMmaWarp = kEpiWarps + 1;   // first of kMmaWarps issuing warps (leader CTA only)
    // flags[0] != 0: some row's norm is beyond the fold range -> the generic kernel runs instead
    if (!force && *static_cast<volatile const int32_t*>(flags) != 0) return;

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    if (smem_u32(smem_raw) & 1023u) __trap();   // SWIZZLE_128B tiles need 1024-byte alignment
    SmemLayout sm(smem_raw);
    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();          // 0 = leader: issues the MMAs, owns the "full" barriers
    const long long cid = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

    if (warp == kProducerWarp && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
        tma_prefetch_desc(&tmap_x);
        tma_prefetch_desc(&tmap_ax);
        mbar_init(sm.ax_full, 2);
        for (int i = 0; i < kAStages; ++i) {
            mbar_init(&sm.a_full[i], 2);
            mbar_init(&sm.a_empty[i], kMmaWarps);
        }
        for (int i = 0; i < kBStages; ++i) {
            mbar_init(&sm.b_full[i], 2);
            mbar_init(&sm.b_empty[i], 1);
        }
        for (int i = 0; i < kAccStages; ++i) {
            mbar_init(&sm.t_full[i], 1);
            mbar_init(&sm.t_empty[i], 2 * kEpiWarps);
        }
        fence_barrier_init();
    }
    if (warp < kEpiWarps) {   // published thresholds start at zero (both item parities)
        sm.thr[threadIdx.x] = 0u;
        sm.thr[kEpiWarps * 32 + threadIdx.x] = 0u;
    }
    if (warp == kMmaWarp) {   // the same warp of both CTAs
        tmem_alloc_2cta(sm.tmem_slot, kTmemCols);
        tmem_relinquish_2cta();
    }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *sm.tmem_slot, 0);

    if (warp == kProducerWarp) {
        // each CTA loads ITS query rows and ITS half of every train tile; all bytes are counted on the
        // leader's barriers (count 2: the leader's arrive.expect_tx for both halves + the peer's arrive)
        Pipe ap(kAStages), bp(kBStages);
        const uint32_t ax_full_l = mapa_shared(smem_u32(sm.ax_full), 0);
        const uint32_t a_full_l = mapa_shared(smem_u32(sm.a_full), 0);
        const uint32_t b_full_l = mapa_shared(smem_u32(sm.b_full), 0);
        if (elect_one()) {
            if (rank == 0)
                mbar_arrive_expect_tx(sm.ax_full, 2 * kAxBytes);
            else
                mbar_arrive_cluster(ax_full_l);
            tma_load_2d_2sm(sm.ax, &tmap_ax, ax_full_l, 0, 0);
        }
        __syncwarp();
        for (long long w = cid; w < n_work; w += n_clusters) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int q_row = w0.x, t_row = w0.z, t_cnt = w0.w;
            mbar_wait(&sm.a_empty[ap.stage], ap.phase ^ 1);
            if (elect_one()) {
                if (rank == 0)
                    mbar_arrive_expect_tx(&sm.a_full[ap.stage], 2 * kATileBytes);
                else
                    mbar_arrive_cluster(a_full_l + 8 * ap.stage);
                tma_load_2d_2sm(sm.a + size_t(ap.stage) * kATileBytes, &tmap_a, a_full_l + 8 * ap.stage, 0,
                                q_row + static_cast<int>(rank) * kTileM);
            }
            __syncwarp();
            ap.advance();
            const int n_tiles = (t_cnt + NT - 1) / NT;
            for (int t = 0; t < n_tiles; ++t) {
                mbar_wait(&sm.b_empty[bp.stage], bp.phase ^ 1);
                if (elect_one()) {
                    if (probe_flags & 1u) {   // timing probe: no train-tile traffic
                        mbar_arrive_cluster(b_full_l + 8 * bp.stage);
                    } else {
                        // rows beyond the train image are masked in the epilogue; beyond the array TMA fills zeros
                        if (rank == 0)
                            mbar_arrive_expect_tx(&sm.b_full[bp.stage], 2 * (kBTileBytes + kXTileBytes));
                        else
                            mbar_arrive_cluster(b_full_l + 8 * bp.stage);
                        const int r0 = t_row + t * kMmaN + static_cast<int>(rank) * kBRows;
                        tma_load_2d_2sm(sm.b + size_t(bp.stage) * kBTileBytes, &tmap_b, b_full_l + 8 * bp.stage, 0, r0);
                        // fold operand rows are parallel to the descriptor rows, after 128 constant rows
                        tma_load_2d_2sm(sm.x + size_t(bp.stage) * kXTileBytes, &tmap_x, b_full_l + 8 * bp.stage, 0,
                                        kTileM + r0);
                    }
                }
                __syncwarp();
                bp.advance();
            }
        }
    } else if (warp >= kMmaWarp) {
        // Two issuing warps take the groups in turn (even / odd): one warp's serial work per group (two
        // barrier waits, five MMAs, two commits: ~400 cycles measured) exceeds the 320 cycles a group takes.
        // Each tracks its own MMAs with its own commits; the accumulator and train-tile stages advance in step.
        // (N = 256 instructions over stage PAIRS were measured too: one issuing warp is then enough, but with
        // only two hand-over units the MMA pipeline alone drops from 3260 to 2120 TOP/s.)
        if (rank == 0) {
            const uint32_t mw = static_cast<uint32_t>(warp - kMmaWarp);
            Pipe ap(kAStages);
            uint32_t g = 0;   // group counter over all work items: stage g & 3, phase (g >> 2) & 1
            mbar_wait(sm.ax_full, 0);
            const uint64_t axdesc = make_kmajor_sw32_desc_p(smem_u32(sm.ax));
            const uint64_t bdesc_base = make_kmajor_sw128_desc(smem_u32(sm.b));
            const uint64_t xdesc_base = make_kmajor_sw32_desc_p(smem_u32(sm.x));
            // probe bit 8: cycle counters of issuing warp 0 (cluster 0 writes them over out[0..3])
            const bool prof = MODE == 2 && (probe_flags & 256u) && mw == 0;
            long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            const long long p_begin = clock64();
#define SFM_PC(i) do { if (prof) { const long long n__ = clock64(); pc[i] += n__ - p_t; p_t = n__; } } while (0)
            for (long long w = cid; w < n_work; w += n_clusters) {
                const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
                const int n_tiles = (w0.w + NT - 1) / NT;
                mbar_wait(&sm.a_full[ap.stage], ap.phase);
                const uint64_t adesc0 = make_kmajor_sw128_desc(smem_u32(sm.a + size_t(ap.stage) * kATileBytes));
                for (int t = 0; t < n_tiles; ++t, ++g) {
                    if ((g & 1u) != mw) continue;
                    const uint32_t st = g & 3u, ph = (g >> 2) & 1u;
                    long long p_t = prof ? clock64() : 0;
                    mbar_wait(&sm.b_full[st], ph);
                    SFM_PC(1);
                    mbar_wait(&sm.t_empty[st], ph ^ 1);
                    tc_fence_after();
                    SFM_PC(2);
                    const uint64_t bdesc0 = bdesc_base + uint64_t(st) * (kBTileBytes >> 4);
                    const uint64_t xdesc = xdesc_base + uint64_t(st) * (kXTileBytes >> 4);
                    const uint32_t d = tmem_base + st * NT;
                    if (elect_one()) {
                        if (!(probe_flags & 2u)) {
#pragma unroll
                            for (int k = 0; k < kKSteps; ++k)
                                mma_i8_ss_2cta(d, adesc0 + uint64_t(k) * (kUmmaK >> 4),
                                               bdesc0 + uint64_t(k) * (kUmmaK >> 4), kIdesc, k > 0);
                            mma_i8_ss_2cta(d, axdesc, xdesc, kIdesc, 1);   // + (1, 255 x 31) . digits(C - ||b||^2 / 2)
                        }
                        mma_commit_2cta(&sm.t_full[st]);
                        mma_commit_2cta(&sm.b_empty[st]);
                    }
                    __syncwarp();
                    SFM_PC(3);
                    pc[6] += 1;
                }
                if (elect_one()) mma_commit_2cta(&sm.a_empty[ap.stage]);   // this warp's MMAs of the item are done
                __syncwarp();
                ap.advance();
            }
#undef SFM_PC
            if (prof && cid == 0 && lane == 0) {
                pc[0] = clock64() - p_begin;
                long long* o = reinterpret_cast<long long*>(out);
                for (int i = 0; i < 8; ++i) o[i] = pc[i];
            }
        }
    } else {
        // ------------------------------------------------ epilogue warps (both CTAs)
        const int quad = warp & 3;          // TMEM lane quadrant this warp may read
        const int cq = warp >> 2;           // which 32-column quarter of every group
        const int row = quad * 32 + lane;   // row of this CTA's 128
        const int irow = static_cast<int>(rank) * kTileM + row;   // row of the work item
        const int colbase = cq * kUnit;
        const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + colbase;
        const uint32_t my_slots = smem_u32(sm.slots) + threadIdx.x * 16;
        const uint32_t t_empty_l = mapa_shared(smem_u32(sm.t_empty), 0);
        const uint32_t t_full_a = smem_u32(sm.t_full);
        uint32_t g = 0;   // group counter over all work items: stage g & 3, phase (g >> 2) & 1
        int item = 0;
        // probe bit 9: cycle counters of epilogue warp 0 (cluster 0 writes them after the kernel's rows)
        const bool eprof = (probe_flags & 512u) != 0;
        long long ec[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        long long e_t = eprof ? clock64() : 0;
#define SFM_EC(i) do { if (eprof) { const long long n__ = clock64(); ec[i] += n__ - e_t; e_t = n__; } } while (0)
        for (long long w = cid; w < n_work; w += n_clusters, ++item) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int4 w1 = __ldg(reinterpret_cast<const int4*>(work + w) + 1);
            const int q_row = w0.x, q_cnt = w0.y, t_row = w0.z, t_cnt = w0.w, t_pad = w1.x;
            const long long out_row =
                static_cast<long long>(static_cast<uint32_t>(w1.z)) | (static_cast<long long>(w1.w) << 32);
            const bool valid = irow < q_cnt;
            const int na = valid ? __ldg(norms + q_row + irow) : 0;
            const int n_tiles = (t_cnt + NT - 1) / NT;
            const int n_full = t_cnt / NT;
            const int32_t* nbase = norms_pad + t_pad;
            KeyFmt kf;
            kf.n_units = n_tiles;
            kf.ubits = max(32 - __clz(max(kf.n_units - 1, 1)), 8);
            kf.sh = kf.ubits > 8 ? kf.ubits - 8 : 0;   // m < 2^23: (23 - sh) + ubits + 1 <= 32 bits
            kf.mul = 2u << kf.ubits;
            Trk sk{1, 0, 0, 0};   // empty keys: k1 owns slot 1, k2 slot 0 (the first unit is parked in slot 0)
            // Row-level lower bound of the second-largest unit maximum (in key units: m >> sh).  A unit below it
            // cannot hold one of the row's two nearest neighbours whichever quarter it is in, so it is neither
            // tracked nor parked; the four threads of a row publish their own bound every fourth group.
            uint32_t thr = 0;
            const uint32_t thr_mine = smem_u32(sm.thr) + (((item & 1) * 4 + cq) * kTileM + row) * 4;
            const uint32_t thr_row = smem_u32(sm.thr) + ((item & 1) * 4 * kTileM + row) * 4;
            const uint32_t thr_next = smem_u32(sm.thr) + ((((item & 1) ^ 1) * 4 + cq) * kTileM + row) * 4;

            if (MODE == 2) {      // timing probe: TMA + MMA pipeline alone
                for (int t = 0; t < n_tiles; ++t, ++g) {
                    mbar_wait_a(t_full_a + 8 * (g & 3u), (g >> 2) & 1u);
                    tc_fence_after();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(t_empty_l + 8 * (g & 3u));
                }
            } else {
                // Double-buffered: the load of group t + 1 is in flight while group t is reduced, and a stage
                // goes back to the MMA warps as soon as its columns have landed in registers.
                uint32_t buf0[32], buf1[32];
                mbar_wait_a(t_full_a + 8 * (g & 3u), (g >> 2) & 1u);
                tc_fence_after();
                tmem_ld_32x32(lane_base + (g & 3u) * NT, buf0);
#define SFM_PAIR_STEP(CUR, NXT, T, MASKED, SH0)                                                   \
    do {                                                                                          \
        tmem_ld_wait();                                                                           \
        tc_fence_before();                                                                        \
        __syncwarp();                                                                             \
        if (lane == 0) mbar_arrive_cluster(t_empty_l + 8 * (g & 3u));                             \
        ++g;                                                                                      \
        if ((T) + 1 < n_tiles) {                                                                  \
            mbar_wait_a(t_full_a + 8 * (g & 3u), (g >> 2) & 1u);                                  \
            tc_fence_after();                                                                     \
            tmem_ld_32x32(lane_base + (g & 3u) * NT, NXT);                                        \
        }                                                                                         \
        process_unit<MASKED, SH0, MODE>(sk, CUR, (T), t_cnt, colbase, my_slots, kf, thr);         \
    } while (0)
                // every fourth group: publish this thread's bound, take the row's
#define SFM_PAIR_SHARE()                                                                          \
    do {                                                                                          \
        if (MODE == 0) {                                                                          \
            sts32(thr_mine, sk.k2 >> (kf.ubits + 1));                                             \
            thr = max(max(lds32(thr_row), lds32(thr_row + kTileM * 4)),                           \
                      max(lds32(thr_row + 2 * kTileM * 4), lds32(thr_row + 3 * kTileM * 4)));     \
        }                                                                                         \
    } while (0)
                int t = 0;
                if (kf.sh == 0) {
                    for (; t + 3 < n_full; t += 4) {
                        SFM_PAIR_STEP(buf0, buf1, t, false, true);
                        SFM_PAIR_STEP(buf1, buf0, t + 1, false, true);
                        SFM_PAIR_STEP(buf0, buf1, t + 2, false, true);
                        SFM_PAIR_STEP(buf1, buf0, t + 3, false, true);
                        SFM_PAIR_SHARE();
                    }
                } else {
                    for (; t + 3 < n_full; t += 4) {
                        SFM_PAIR_STEP(buf0, buf1, t, false, false);
                        SFM_PAIR_STEP(buf1, buf0, t + 1, false, false);
                        SFM_PAIR_STEP(buf0, buf1, t + 2, false, false);
                        SFM_PAIR_STEP(buf1, buf0, t + 3, false, false);
                        SFM_PAIR_SHARE();
                    }
                }
                // up to three more full tiles and the partial one (t is a multiple of four here)
                for (; t < n_tiles; t += 2) {
                    SFM_PAIR_STEP(buf0, buf1, t, true, false);
                    if (t + 1 < n_tiles) SFM_PAIR_STEP(buf1, buf0, t + 1, true, false);
                }
#undef SFM_PAIR_STEP
#undef SFM_PAIR_SHARE
                sts32(thr_next, 0u);   // the other parity's entry: the next item starts without a bound
            }
            SFM_EC(0);
            if (MODE != 0 || (probe_flags & 64u)) {   // (probe bit 6: no end-of-item work)
                if (valid && cq == 0 && !(probe_flags & 256u)) out[out_row + irow] = make_int4(static_cast<int>(sk.k1), static_cast<int>(sk.k2), na, 0);
                continue;
            }
            Top2K r = resolve_units(sk, kf, colbase, my_slots, par_bits + (t_pad >> 5), t_cnt);
            SFM_EC(1);
            // a dropped unit whose (truncated) maximum equals the second kept one may hold a candidate:
            // the best dropped one is recomputed here, a second one sends the row to knn_fix_kernel
            const uint32_t mf = kf.ubits + 1;
            r = recompute_unit(r, valid && sk.dm >= 2u && (sk.dm >> mf) == (sk.k2 >> mf), q_row + irow,
                               key_col(sk.dm, kf, colbase), desc, t_row, t_cnt, nbase);
            SFM_EC(2);
            ec[4] += __popc(__ballot_sync(0xffffffffu, valid && sk.dm >= 2u && (sk.dm >> mf) == (sk.k2 >> mf)));
            const bool tie = sk.dd >= 2u && (sk.dd >> mf) == (sk.k2 >> mf);
            if (r.i1 < 0) r.i1 = INT_MAX;   // empty entries sort last
            if (r.i2 < 0) r.i2 = INT_MAX;
            // quarters 1..3 hand their top-2 to quarter 0's thread of the same row, which merges under the
            // (key, lowest index) order and writes the row
            const uint32_t xbase = smem_u32(sm.xchg) + ((item & 1) * 3 * kTileM + row) * 16;
            const uint32_t fbase = smem_u32(sm.pub) + ((item & 1) * 3 * kTileM + row) * 4;
            if (cq > 0) {
                sts128(xbase + (cq - 1) * kTileM * 16, r.k1, r.k2, r.i1, r.i2);
                sts32(fbase + (cq - 1) * kTileM * 4, tie ? 1u : 0u);
            }
            asm volatile("bar.sync %0, 128;" ::"r"(1 + quad) : "memory");
            if (cq == 0) {
                bool flagged = tie;
                Top2K mrg = r;
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const uint4 o = lds128(xbase + k * kTileM * 16);
                    flagged = flagged || lds32(fbase + k * kTileM * 4) != 0;
                    const Top2K b{static_cast<int>(o.x), static_cast<int>(o.y), static_cast<int>(o.z), static_cast<int>(o.w)};
                    mrg = merge_top2(mrg, b);
                }
                if (valid && flagged) {   // rare: left to knn_fix_kernel
                    atomicAdd(flags + 1, 1);   // statistics since sfm_set_images
                    const int e = atomicAdd(flags + 2, 1);
                    if (e < fix_cap) fix_list[e] = make_int2(static_cast<int>(w), irow);
                }
                if (valid) {
                    const int i1 = mrg.i1 == INT_MAX ? -1 : mrg.i1;
                    const int i2 = mrg.i2 == INT_MAX ? -1 : mrg.i2;
                    const int d1 = i1 < 0 ? INT_MAX : na - mrg.k1;
                    const int d2 = i2 < 0 ? INT_MAX : na - mrg.k2;
                    out[out_row + irow] = flagged ? make_int4(kFlagged, kFlagged, 0, 0) : make_int4(i1, i2, d1, d2);
                }
            }
            SFM_EC(3);
            ec[5] += 1;
        }
#undef SFM_EC
        if (eprof && cid == 0 && warp == 0 && lane == 0 && rank == 0) {
            long long* o = reinterpret_cast<long long*>(fix_list + fix_cap / 2);
            for (int i = 0; i < 8; ++i) o[i] = ec[i];
        }
    }

    // both CTAs are done with each other's barriers and tensor memory before either leaves
    tc_fence_before();
    cluster_sync_all();
    if (warp == kMmaWarp) {
        tc_fence_after();
        tmem_dealloc_2cta(tmem_base, kTmemCols);
    }
}

// Debug: folded accumulators of one 256 x 128 tile through the pair path (out: [256][128]).
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
debug_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                  const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_ax, int q_row,
                  int t_row, int32_t* __restrict__ out) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    if (smem_u32(smem_raw) & 1023u) __trap();
    uint8_t* sa = smem_raw;
    uint8_t* sb = sa + kATileBytes;
    uint8_t* sx = sb + kBTileBytes;
    uint8_t* sax = sx + kXTileBytes;
    uint64_t* bar_ld = reinterpret_cast<uint64_t*>(sax + kAxBytes);
    uint64_t* bar_mma = bar_ld + 1;
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar_mma + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    if (threadIdx.x == 0) {
        mbar_init(bar_ld, 2);
        mbar_init(bar_mma, 1);
        fence_barrier_init();
    }
    if (warp == 0) {
        tmem_alloc_2cta(slot, kMmaN);
        tmem_relinquish_2cta();
    }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *slot;
    if (threadIdx.x == 0) {
        const uint32_t bar_l = mapa_shared(smem_u32(bar_ld), 0);
        constexpr uint32_t kBytes = kATileBytes + kBTileBytes + kXTileBytes + kAxBytes;
        if (rank == 0)
            mbar_arrive_expect_tx(bar_ld, 2 * kBytes);
        else
            mbar_arrive_cluster(bar_l);
        tma_load_2d_2sm(sa, &tmap_a, bar_l, 0, q_row + static_cast<int>(rank) * kTileM);
        tma_load_2d_2sm(sb, &tmap_b, bar_l, 0, t_row + static_cast<int>(rank) * kBRows);
        tma_load_2d_2sm(sx, &tmap_x, bar_l, 0, kTileM + t_row + static_cast<int>(rank) * kBRows);
        tma_load_2d_2sm(sax, &tmap_ax, bar_l, 0, 0);
        if (rank == 0) {
            mbar_wait(bar_ld, 0);
            tc_fence_after();
            const uint64_t ad = make_kmajor_sw128_desc(smem_u32(sa));
            const uint64_t bd = make_kmajor_sw128_desc(smem_u32(sb));
            for (int k = 0; k < kKSteps; ++k)
                mma_i8_ss_2cta(tmem_base, ad + uint64_t(k) * (kUmmaK >> 4), bd + uint64_t(k) * (kUmmaK >> 4), kIdesc, k > 0);
            mma_i8_ss_2cta(tmem_base, make_kmajor_sw32_desc_p(smem_u32(sax)), make_kmajor_sw32_desc_p(smem_u32(sx)), kIdesc, 1);
            mma_commit_2cta(bar_mma);
        }
    }
    __syncwarp();
    mbar_wait(bar_mma, 0);
    tc_fence_after();
    const int row = static_cast<int>(rank) * kTileM + warp * 32 + lane;
    for (int c = 0; c < kMmaN / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(warp * 32) << 16) + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) out[row * kMmaN + c * 32 + j] = static_cast<int32_t>(v[j]);
    }
    tc_fence_before();
    cluster_sync_all();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc_2cta(tmem_base, kMmaN);
    }
}

// Norm parities of the padded column space, one bit per column, 32 columns per word.
__global__ void parity_bits_kernel(const int32_t* __restrict__ norms_pad, long long pad_total,
                                   uint32_t* __restrict__ par_bits) {
    const long long c = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    const bool odd = c < pad_total && (norms_pad[c] & 1);
    const unsigned bal = __ballot_sync(0xffffffffu, odd);
    if ((threadIdx.x & 31) == 0 && c < pad_total) par_bits[c >> 5] = bal;
}

template <int MODE>
cudaError_t launch_pair_variant(const CUtensorMap& tmap_a, const CUtensorMap& tmap_b, const CUtensorMap& tmap_x,
                                const CUtensorMap& tmap_ax, const uint8_t* d_desc, const WorkItem* d_work,
                                int64_t n_work, const int32_t* d_norms, const int32_t* d_norms_pad,
                                const uint32_t* d_par_bits, int32_t* d_flags, int force, const KnnTuning& tn,
                                int2* d_fix_list, int fix_cap, int32_t* d_knn, unsigned grid, cudaStream_t st) {
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(knn2_pair_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(kSmemBytes));
        if (e != cudaSuccess) return e;
        configured = true;
    }
    knn2_pair_kernel<MODE><<<grid, kThreads, kSmemBytes, st>>>(
        tmap_a, tmap_b, tmap_x, tmap_ax, d_desc, d_work, n_work, d_norms, d_norms_pad, d_par_bits, d_flags, force,
        tn.probe_flags, d_fix_list, fix_cap, reinterpret_cast<int4*>(d_knn));
    return cudaGetLastError();
}

}  // namespace

void launch_parity_bits(const int32_t* d_norms_pad, int64_t pad_total, uint32_t* d_par_bits, cudaStream_t st,
                        int64_t* launches) {
    if (pad_total <= 0) return;
    parity_bits_kernel<<<static_cast<unsigned>((pad_total + 255) / 256), 256, 0, st>>>(d_norms_pad, pad_total,
                                                                                      d_par_bits);
    ++*launches;
}

// variant 0: the kNN (exits at once when the fold-range flag is raised, unless force), followed by the
// fix-up pass of knn_wide.cu; 1: max tree only, 2: no epilogue (timing probes).
int launch_knn2_pair(const CUtensorMap& tmap_a, const CUtensorMap& tmap_b, const CUtensorMap& tmap_x,
                     const CUtensorMap& tmap_ax, const uint8_t* d_desc, const WorkItem* d_work, int64_t n_work,
                     const int32_t* d_norms, const int32_t* d_norms_pad, const uint32_t* d_par_bits,
                     int32_t* d_flags, int variant, int force, const KnnTuning& tn, int32_t* d_fix_list,
                     int32_t fix_cap, int32_t* d_knn, int num_sms, cudaStream_t st, int64_t* launches) {
    if (n_work <= 0) return 0;
    const int64_t max_clusters = num_sms / 2;
    const unsigned grid = 2u * static_cast<unsigned>(n_work < max_clusters ? n_work : max_clusters);
    int2* fl = reinterpret_cast<int2*>(d_fix_list);
    cudaError_t e = cudaMemsetAsync(d_flags + 2, 0, sizeof(int32_t), st);   // length of this launch's fix list
    if (e != cudaSuccess) return static_cast<int>(e);
#define SFM_PAIR_LAUNCH(M, F)                                                                                \
    launch_pair_variant<M>(tmap_a, tmap_b, tmap_x, tmap_ax, d_desc, d_work, n_work, d_norms, d_norms_pad,    \
                           d_par_bits, d_flags, F, tn, fl, fix_cap, d_knn, grid, st)
    if (variant == 1)
        e = SFM_PAIR_LAUNCH(1, 1);
    else if (variant == 2)
        e = SFM_PAIR_LAUNCH(2, 1);
    else
        e = SFM_PAIR_LAUNCH(0, force);
#undef SFM_PAIR_LAUNCH
    ++*launches;
    if (e == cudaSuccess && variant == 0 && !(tn.probe_flags & 4u))   // (probe bit 2: main kernel alone)
        return launch_knn_fix(d_work, n_work, d_desc, d_norms, d_flags, d_fix_list, fix_cap, d_knn, num_sms, st,
                              launches);
    return static_cast<int>(e);
}

int launch_debug_pair(const CUtensorMap& tmap_a, const CUtensorMap& tmap_b, const CUtensorMap& tmap_x,
                      const CUtensorMap& tmap_ax, int32_t q_row, int32_t t_row, int32_t* d_out, cudaStream_t st,
                      int64_t* launches) {
    const size_t smem = kATileBytes + kBTileBytes + kXTileBytes + kAxBytes + 64;
    cudaError_t e = cudaFuncSetAttribute(debug_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    debug_pair_kernel<<<2, 128, smem, st>>>(tmap_a, tmap_b, tmap_x, tmap_ax, q_row, t_row, d_out);
    ++*launches;
    return static_cast<int>(cudaGetLastError());
}

}  // namespace sfm
