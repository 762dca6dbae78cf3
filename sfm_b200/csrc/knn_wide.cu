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
// 16 accumulators are stored to shared-memory slot `slot`, the slot of the key it evicts (k2).  (The lean
// tracker variant, see fold_chunk, keeps the slot as an address instead of a key bit.)
// At the end of the work item the parked values give the exact top-2 (resolve_units).  A third
// unit can only matter if its (truncated) maximum equals k2's: then dm names it and its 16 keys
// are recomputed from the descriptors (recompute_units, ~2 % of rows); if dd ties as well the row
// is flagged for knn_fix_kernel (brute force).  ubits / sh depend on the train image size:
// 8192 rows need 8 bits and no truncation, 32768 rows 10 bits and sh = 2 (ties on the truncated
// maximum are merely more frequent; the logic above is exact for any monotone truncation).
struct Trk {
    uint32_t k1, k2, dm, dd;
    uint32_t thr;   // SKIP variant: maximum (key units) of the k2 unit -- lower units cannot matter any more
    uint32_t dst;   // shared-memory address of the slot that holds k2's unit (the next one to be overwritten)
    uint32_t ev;    // evicted key of the even chunk, waiting for the odd one's (dm / dd take two at a time)
};

// Per-thread constants of the tracker variants (fold_chunk).
struct TrkConst {
    uint32_t slots;     // address of this thread's slot 0
    uint32_t flipc;     // slot 0 address + slot 1 address
    bool park;          // false: timing probe without the parking stores (probe bit 5; results meaningless)
};

struct KeyFmt {
    uint32_t sh, mul, ubits;   // key = (m >> sh) * mul + (uidrev << 1) + slot, mul = 2^(ubits+1)
    int n_units;               // units per thread over the whole work item
    int ushift;                // log2(units per thread and train tile): 3 (wide kernel) or 2 (ping-pong kernel)
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

// (key desc, index asc) merge of two per-thread top-2 results of one row; empty entries carry index INT_MAX.
__device__ __forceinline__ Top2K merge_top2(const Top2K& r, const Top2K& b) {
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
    return mrg;
}

// 16 folded accumulators of one row = one unit, chunk c of its tile; ut = uidrev << 1 of chunk 0.
// MASKED: columns >= t_cnt (last, partial train tile) count as v = 0 (real accumulators are >= 0;
// resolve_units skips them by column).  TREE_ONLY: timing probe.
// SH0: train images up to 8192 rows (ubits = 8, no truncation of the maximum).
// TRK selects the bookkeeping (all three give the same kNN bit for bit; knn modes 16 / 15 / 5):
//   0  lean tracker (knn mode 16 and the ping-pong kernel).  The slot is not a key bit (bit 0 of every key is 0) but one address per thread,
//      `dst` = the slot of k2's unit, the next to be overwritten.  It changes exactly when a key becomes k1 (the
//      old k1 turns into k2 and keeps its slot): one compare + a predicated IMAD on the FMA pipe instead of two
//      LOP3 on the ALU pipe.  dm / dd take the evicted keys of TWO chunks per update (5 instead of 6 min/max).
//      16.75 ALU-pipe instructions per chunk against 18 -- and the same speed (2177 against 2190 TOP/s on one box):
//      the kernel's rate is not set by its ALU instruction count ("where the wide kernel waits").
//   1  lean tracker + warp-uniform skip: a unit whose maximum is below the maximum of the thread's k2 unit can
//      neither enter the top two nor tie with them (k2 only grows), so when that holds for every lane of the warp
//      the bookkeeping is skipped; lanes that are below while others are not take part with m = 0.  Measured
//      (16 x 8192, same box): 1950 TOP/s with the skip against 2050 without -- the vote and the reconvergence
//      cost more than the 40 % of skipped updates save (bit-exact on every kNN test).
//   2  round-1 tracker (default): slot bit in bit 0 of the key (k2 & 1 names the slot to overwrite), dm / dd per chunk.
//      (A third variant kept the slot bit in the key but merged it and formed the slot address on the FMA pipe --
//      x = k2 << 31 as an IMAD, key and address as IMAD.HI of x, 15.5 ALU-pipe instructions per chunk: 2150 TOP/s
//      against 2190 / 2195 for variants 0 / 2 on the same box; removed.)
// batch (a constant once the chunk loop is unrolled): 0 = dm / dd updated at once, 1 = even chunk (evicted key
// stashed), 2 = odd chunk (both merged).
template <bool MASKED, bool TREE_ONLY, bool SH0, int TRK>
__device__ __forceinline__ void fold_chunk(Trk& s, const uint32_t (&acc)[16], int col0, int t_cnt,
                                           const TrkConst& tc, const KeyFmt& kf, uint32_t ut, uint32_t c, int batch) {
    constexpr bool SKIP = TRK == 1;
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
    // low key bits: (uidrev of chunk c) << 1 [| slot].  The units of a tile count down from a value whose bits
    // 1..3 are 111 (n_units is a multiple of 8), so ut - 2c == ut ^ 2c: a warp-uniform value (uniform datapath).
    const uint32_t low = ut ^ (2u * c);
    uint32_t key, dst;
    if (TRK == 2) {
        const uint32_t sb = s.k2 & 1u;
        key = SH0 ? mk * 512u + (sb | low) : mk * kf.mul + (sb | low);
        dst = tc.slots + sb * kSlotStride;
    } else {
        key = mk * kf.mul + low;   // mul: a register even for SH0 (512), or the compiler picks LEA (ALU pipe)
        dst = s.dst;
    }
    // top two (keys are unique, every compare is strict)
    const bool enters = key > s.k2 && tc.park;
    const bool first = (TRK == 0 || TRK == 1) && key > s.k1;
    const uint32_t t1 = min(s.k1, key);
    s.k1 = max(s.k1, key);
    const uint32_t t2 = min(s.k2, t1);
    s.k2 = max(s.k2, t1);
    if (enters) {   // park the unit where k2's unit was
#pragma unroll
        for (int g = 0; g < kUnit / 4; ++g)
            sts128(dst + g * kSlotRow, v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
    }
    if (TRK == 0 || TRK == 1) {
        // it became k1: the old k1 is k2 now and its slot the next to go (a predicated IMAD on the FMA pipe)
        asm("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %1, 0;\n\t@p mad.lo.u32 %0, %0, 0xffffffff, %2;\n\t}"
            : "+r"(s.dst)
            : "r"(static_cast<uint32_t>(first)), "r"(tc.flipc));
    }
    if (batch == 0) {
        const uint32_t t3 = min(s.dm, t2);
        s.dm = max(s.dm, t2);
        s.dd = max(s.dd, t3);
    } else if (batch == 1) {
        s.ev = t2;
    } else {
        // the two largest of {dm, dd, ev, t2}
        const uint32_t u = max(s.ev, t2), l = min(s.ev, t2);
        const uint32_t x = min(s.dm, u);
        s.dm = max(s.dm, u);
        s.dd = __vimax3_u32(s.dd, x, l);
    }
    if (SKIP) s.thr = s.k2 >> (kf.ubits + 1);
}

// First train column of the unit a key names (colbase: first column of this thread's half).
__device__ __forceinline__ int key_col(uint32_t key, const KeyFmt& kf, int colbase) {
    const int uidrev = static_cast<int>((key >> 1) & ((1u << kf.ubits) - 1u));
    const int idx = kf.n_units - 1 - uidrev;
    return (idx >> kf.ushift) * kTileN + colbase + (idx & ((1 << kf.ushift) - 1)) * kUnit;
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

template <int TRK>
__device__ __forceinline__ Top2K resolve_units(const Trk& s, int n_seen, const KeyFmt& kf, int colbase,
                                               uint32_t slots, const int32_t* __restrict__ nbase, int t_cnt) {
    // lean tracker: k2's unit is in the slot dst names, k1's in the other; else the slot is bit 0 of the key
    const int s2 = (TRK == 0 || TRK == 1) ? s.dst != slots : s.k2 & 1, s1 = (TRK == 0 || TRK == 1) ? s2 ^ 1 : s.k1 & 1;
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

template <int MODE, int TRK, bool PROF>
__device__ __forceinline__ void knn2_wide_body(const CUtensorMap& tmap_q, const CUtensorMap& tmap, const CUtensorMap& tmap_x,
                                               const CUtensorMap& tmap_ax, const uint8_t* __restrict__ desc_q,
                                               const int32_t* __restrict__ norms_q, const uint8_t* __restrict__ desc,
                                               const WorkItem* __restrict__ work, long long n_work,
                                               const int32_t* __restrict__ norms_pad, int32_t* __restrict__ flags, int force,
                                               uint32_t probe_flags, uint32_t wait_ns, uint32_t epi_wait_ns,
                                               int2* __restrict__ fix_list, int fix_cap, int4* __restrict__ tie_list,
                                               int tie_cap, int4* __restrict__ out) {
    constexpr int kProducerWarp = kEpiWarps;
    constexpr int kMmaWarp = kEpiWarps + 1;
    constexpr bool kBatch = TRK == 0;   // dm / dd take two chunks per update
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
        // probe bit 4: cycle counters of CTA 0 (read back with sfm_debug_pair_counters): [4] this warp's loop,
        // [5] of it waiting for train tiles, [6] waiting for an accumulator to be handed back
        const bool prof = PROF && (probe_flags & 16u) && blockIdx.x == 0;
        long long pc_b = 0, pc_t = 0;
        const long long pc_begin = clock64();
        mbar_wait(sm.ax_full, 0);
        const uint64_t axdesc = make_kmajor_sw32_desc(smem_u32(sm.ax));
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int n_tiles = (w0.w + kTileN - 1) / kTileN;
            mbar_wait(&sm.a_full[ap.stage], ap.phase);
            const uint64_t adesc0 = make_kmajor_sw128_desc(smem_u32(sm.a + size_t(ap.stage) * kABlockBytes));
            for (int t = 0; t < n_tiles; ++t) {
                long long p0 = prof ? clock64() : 0;
                mbar_wait_suspend(&sm.b_full[bp.stage], bp.phase, wait_ns);
                if (prof) pc_b += clock64() - p0;
                const uint32_t bsm = smem_u32(sm.b + size_t(bp.stage) * kStageBytes);
                const uint64_t bdesc0 = make_kmajor_sw128_desc(bsm);
                const uint64_t xdesc = make_kmajor_sw32_desc(bsm + kBTileBytes);
#pragma unroll
                for (int a = 0; a < kATiles; ++a) {
                    p0 = prof ? clock64() : 0;
                    mbar_wait_suspend(&sm.t_empty[a], tphase ^ 1, wait_ns);
                    if (prof) pc_t += clock64() - p0;
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
        if (prof && lane == 0) {
            long long* pc = reinterpret_cast<long long*>(fix_list + (1 << 17));   // api.cu: d_fix_list + kFixCap ints
            pc[4] = clock64() - pc_begin;
            pc[5] = pc_b;
            pc[6] = pc_t;
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
        const TrkConst tc{my_slots, 2u * my_slots + kSlotStride, !(PROF && (probe_flags & 32u))};
        uint64_t* full = &sm.t_full[acc];
        uint64_t* empty = &sm.t_empty[acc];
        uint32_t tphase = 0;
        int item = 0;
        // probe bit 4: cycle counters of CTA 0's first epilogue warp: [0] its loop, [1] of it waiting for a full
        // accumulator, [2] waiting for tcgen05.ld results, [3] at the end of the work items, [7] tiles
        const bool prof = PROF && (probe_flags & 16u) && blockIdx.x == 0 && warp == 0;
        long long pc_f = 0, pc_l = 0, pc_e = 0, pc_n = 0;
        const long long pc_begin = clock64();
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
            kf.ushift = 3;
            kf.ubits = max(32 - __clz(kf.n_units - 1), 8);
            kf.sh = kf.ubits > 8 ? kf.ubits - 8 : 0;   // m < 2^23: (23 - sh) + ubits + 1 <= 32 bits
            kf.mul = 2u << kf.ubits;
            // units this thread will see: 8 per full tile, fewer in the last, partial one
            const int rem = t_cnt - (n_tiles - 1) * kTileN - colbase;
            const int n_seen = (n_tiles - 1) * (kColsPerThread / kUnit) +
                               min(max((rem + kUnit - 1) / kUnit, 0), kColsPerThread / kUnit);
            // empty keys: k1 owns slot 1, k2 slot 0 (the first unit is parked in slot 0)
            Trk sk{1, 0, 0, 0, 0, my_slots, 0};
            for (int t = 0; t < n_tiles; ++t) {
                long long p0 = prof ? clock64() : 0;
                mbar_wait_suspend(full, tphase, epi_wait_ns);
                if (prof) {
                    pc_f += clock64() - p0;
                    ++pc_n;
                }
                tc_fence_after();
                tphase ^= 1;
                // MODE 2: timing probe, TMA + MMA pipeline alone.  Probe bit 8: the second row block's warps hand
                // their accumulator back at once, so the first one's eight warps have the SM to themselves.
                if (MODE == 2 || (PROF && (probe_flags & 256u) && acc == 1)) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(empty);
                    continue;
                }
                const uint32_t ut = static_cast<uint32_t>(kf.n_units - 1 - t * (kColsPerThread / kUnit)) << 1;
                // 16-column chunks, double-buffered: 32 live accumulator registers per thread
                uint32_t v[2][16];
                tmem_ld_32x16(taddr, v[0]);
                if ((t + 1) * kTileN <= t_cnt && kf.sh == 0) {
                    // probe bit 10: only the first of a tile's eight TMEM loads is issued (the other chunks reuse
                    // whatever the registers hold): the epilogue's own dependency chains without the loads
                    const bool no_ld = PROF && (probe_flags & 1024u);
#pragma unroll
                    for (int c = 0; c < kColsPerThread / 16; ++c) {
                        p0 = prof ? clock64() : 0;
                        if (!no_ld || c == 0) tmem_ld_wait();
                        if (prof) pc_l += clock64() - p0;
                        if (c + 1 < kColsPerThread / 16) {
                            if (!no_ld) tmem_ld_32x16(taddr + (c + 1) * 16, v[(c + 1) & 1]);
                        } else {
                            // every column of this thread is in registers: hand the accumulator back
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(empty);
                        }
                        fold_chunk<false, MODE == 1, true, TRK>(sk, v[c & 1], t * kTileN + colbase + c * 16, t_cnt, tc, kf,
                                                                ut, c, kBatch ? 1 + (c & 1) : 0);
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
                        fold_chunk<false, MODE == 1, false, TRK>(sk, v[c & 1], t * kTileN + colbase + c * 16, t_cnt, tc, kf,
                                                                 ut, c, kBatch ? 1 + (c & 1) : 0);
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
                        if (col0 < t_cnt) fold_chunk<true, MODE == 1, false, TRK>(sk, v[c & 1], col0, t_cnt, tc, kf, ut, c, 0);
                    }
                }
            }
            const long long pe0 = prof ? clock64() : 0;
            if (MODE != 0 || (PROF && (((probe_flags & 256u) && acc == 1) || (probe_flags & 1024u)))) {
                if (valid && half == 0) out[out_row + row] = make_int4(static_cast<int>(sk.k1), static_cast<int>(sk.k2), na, 0);   // keeps the probe live
                continue;
            }
            // skip variant: units below the threshold were never inserted; a key that was is >= 2 (the empty markers are 0, 1)
            const int n_in = TRK == 1 ? (sk.k1 >= 2u) + (sk.k2 >= 2u) + (sk.dm >= 2u) + (sk.dd >= 2u) : n_seen;
            Top2K r = resolve_units<TRK>(sk, n_in, kf, colbase, my_slots, nbase, t_cnt);
            // a dropped unit whose (truncated) maximum equals the second kept one may hold a candidate:
            // the best dropped one is recomputed here, a second one sends the row to knn_fix_kernel
            const uint32_t mf = kf.ubits + 1;
            // With a tie list the unit is not recomputed here: the row goes out as it is and (row, unit) is appended
            // for knn_tie_kernel, which merges the unit's 16 keys into the row afterwards (see knn2_pp_kernel).
            const bool need = valid && n_in >= 3 && (sk.dm >> mf) == (sk.k2 >> mf);
            int ncol = -1;
            if (tie_list)
                ncol = need ? key_col(sk.dm, kf, colbase) : -1;
            else
                r = recompute_units(r, need, q_row + row, key_col(sk.dm, kf, colbase), desc_q, desc, t_row, t_cnt, nbase);
            const bool tie = n_in >= 4 && (sk.dd >> mf) == (sk.k2 >> mf);
            // the thread of the upper column half hands its top-2 to the lower half's thread,
            // which merges under the (key, lowest index) order and writes the row
            const uint32_t xaddr = smem_u32(sm.xchg) + ((item & 1) * SFM_QBLOCK + row) * 16;
            const uint32_t faddr = smem_u32(sm.pub) + ((item & 1) * SFM_QBLOCK + row) * 4;
            if (half == 1) {
                sts128(xaddr, r.k1, r.k2, r.i1, r.i2);
                sts32(faddr, (static_cast<uint32_t>(ncol + 1) << 1) | (tie ? 1u : 0u));
            }
            asm volatile("bar.sync %0, 64;" ::"r"(1 + (warp & 7)) : "memory");
            if (half == 0) {
                const uint4 o = lds128(xaddr);
                const uint32_t pw = lds32(faddr);
                bool flagged = tie || (pw & 1u) != 0;
                const int ncol_b = static_cast<int>(pw >> 1) - 1;
                Top2K b{static_cast<int>(o.x), static_cast<int>(o.y), static_cast<int>(o.z), static_cast<int>(o.w)};
                if (b.i1 < 0) b.i1 = INT_MAX;   // empty entries sort last
                if (b.i2 < 0) b.i2 = INT_MAX;
                if (r.i1 < 0) r.i1 = INT_MAX;
                if (r.i2 < 0) r.i2 = INT_MAX;
                const Top2K mrg = merge_top2(r, b);
                if (valid && !flagged && (ncol & ncol_b) >= 0) {   // some unit to patch in
                    const int e = atomicAdd(flags + 3, 1);
                    if (e < tie_cap) {
                        const long long o_idx = out_row + row;
                        tie_list[2 * e] = make_int4(static_cast<int>(static_cast<uint32_t>(o_idx)),
                                                    static_cast<int>(o_idx >> 32), q_row + row, static_cast<int>(w));
                        tie_list[2 * e + 1] = make_int4(ncol, ncol_b, -1, -1);
                    } else {
                        flagged = true;   // list full: brute force instead
                    }
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
            if (prof) pc_e += clock64() - pe0;
        }
        if (prof && lane == 0) {
            long long* pc = reinterpret_cast<long long*>(fix_list + (1 << 17));
            pc[0] = clock64() - pc_begin;
            pc[1] = pc_f;
            pc[2] = pc_l;
            pc[3] = pc_e;
            pc[7] = pc_n;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// MODE 0: the kNN.  1: max tree only, 2: no epilogue (timing probes; output is not a kNN).
#define SFM_WIDE_PARAMS                                                                                                  \
    const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap,                                \
        const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_ax,                         \
        const uint8_t *__restrict__ desc_q, const int32_t *__restrict__ norms_q, const uint8_t *__restrict__ desc,       \
        const WorkItem *__restrict__ work, long long n_work, const int32_t *__restrict__ norms_pad,                      \
        int32_t *__restrict__ flags, int force, uint32_t probe_flags, uint32_t wait_ns, uint32_t epi_wait_ns,            \
        int2 *__restrict__ fix_list, int fix_cap, int4 *__restrict__ tie_list, int tie_cap, int4 *__restrict__ out
#define SFM_WIDE_FWD                                                                                                     \
    tmap_q, tmap, tmap_x, tmap_ax, desc_q, norms_q, desc, work, n_work, norms_pad, flags, force, probe_flags, wait_ns,   \
        epi_wait_ns, fix_list, fix_cap, tie_list, tie_cap, out
template <int MODE, int TRK, bool PROF>
__global__ void __launch_bounds__(kThreads, 1) knn2_wide_kernel(SFM_WIDE_PARAMS) {
    knn2_wide_body<MODE, TRK, PROF>(SFM_WIDE_FWD);
}
#undef SFM_WIDE_PARAMS
#undef SFM_WIDE_FWD

// ---------------------------------------------------------------------------------------------------------
// Ping-pong variant (knn mode 17).  Cycle counters of the wide kernel (probe bit 4, profiles/
// r02_probe_wide_counters.json) show where its time goes: per pair of train tiles the epilogue warps WAIT 880
// cycles for a full accumulator and work 1290, the MMA warp waits 1310 for accumulators to come back -- the two
// 128-row accumulators run in phase, so the tensor pipe idles while both are drained and the ALU pipe idles while
// both are filled (the chain release -> issue -> 5 MMAs -> commit -> visible is ~1000 cycles long).  Here the two
// accumulators alternate over TRAIN TILES of one 128-row query block instead, and ALL sixteen epilogue warps drain
// one accumulator (a 64-column quarter of it per thread) while the tensor pipe fills the other one: the same
// tracker per thread (one row, four units per tile), the same parking slots, the same fix-up pass.  The price is
// that a train image is streamed through shared memory once per 128 query rows instead of once per 256 (three B
// stages instead of two; the query-row ring shrinks to 2 x 16 KiB).  A work item (256 query rows) is processed as
// two such passes.  The four threads of a row merge at the end of a pass through their own parking slots.
constexpr int kPpBStages = 3;
constexpr int kPpColsPerThread = kTileN / 4;   // 64
constexpr size_t kPpSmemData = size_t(kAStages) * kATileBytes + size_t(kPpBStages) * kStageBytes + kAxBytes + kSlotBytes;
constexpr int kPpNumBars = 2 * kAStages + 2 * kPpBStages + 2 * 2 + 1;
constexpr size_t kPpSmemBytes = kPpSmemData + kPpNumBars * 8 + 16;
static_assert(kPpSmemBytes <= 232448, "shared memory budget");

template <int MODE, int TRK, bool PROF>
__global__ void __launch_bounds__(kThreads, 1)
knn2_pp_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap,
               const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_ax,
               const uint8_t* __restrict__ desc_q, const int32_t* __restrict__ norms_q,
               const uint8_t* __restrict__ desc,
               const WorkItem* __restrict__ work, long long n_work, const int32_t* __restrict__ norms_pad,
               int32_t* __restrict__ flags, int force, uint32_t probe_flags, uint32_t wait_ns,
               uint32_t epi_wait_ns, int2* __restrict__ fix_list, int fix_cap, int4* __restrict__ tie_list, int tie_cap,
               int4* __restrict__ out) {
    constexpr int kProducerWarp = kEpiWarps;
    constexpr int kMmaWarp = kEpiWarps + 1;
    constexpr bool kBatch = TRK == 0;
    if (!force && *static_cast<volatile const int32_t*>(flags) != 0) return;

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    if (smem_u32(smem_raw) & 1023u) __trap();
    uint8_t* const s_a = smem_raw;                                         // kAStages x 128 query rows x 128 B
    uint8_t* const s_b = s_a + size_t(kAStages) * kATileBytes;             // kPpBStages x (B tile | fold operand)
    uint8_t* const s_ax = s_b + size_t(kPpBStages) * kStageBytes;
    uint4* const s_slots = reinterpret_cast<uint4*>(s_ax + kAxBytes);
    uint64_t* const a_full = reinterpret_cast<uint64_t*>(smem_raw + kPpSmemData);
    uint64_t* const a_empty = a_full + kAStages;
    uint64_t* const b_full = a_empty + kAStages;
    uint64_t* const b_empty = b_full + kPpBStages;
    uint64_t* const t_full = b_empty + kPpBStages;
    uint64_t* const t_empty = t_full + 2;
    uint64_t* const ax_full = t_empty + 2;
    uint32_t* const tmem_slot = reinterpret_cast<uint32_t*>(ax_full + 1);
    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;

    if (warp == kProducerWarp && lane == 0) {
        tma_prefetch_desc(&tmap_q);
        tma_prefetch_desc(&tmap);
        tma_prefetch_desc(&tmap_x);
        tma_prefetch_desc(&tmap_ax);
        for (int i = 0; i < kAStages; ++i) {
            mbar_init(&a_full[i], 1);
            mbar_init(&a_empty[i], 1);
        }
        for (int i = 0; i < kPpBStages; ++i) {
            mbar_init(&b_full[i], 1);
            mbar_init(&b_empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&t_full[i], 1);
            mbar_init(&t_empty[i], kEpiWarps);
        }
        mbar_init(ax_full, 1);
        fence_barrier_init();
    }
    if (warp == kMmaWarp) {
        tmem_alloc(tmem_slot, kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

    if (warp == kProducerWarp) {
        Pipe ap(kAStages), bp(kPpBStages);
        if (elect_one()) {
            mbar_arrive_expect_tx(ax_full, kAxBytes);
            tma_load_2d(s_ax, &tmap_ax, ax_full, 0, 0);
        }
        __syncwarp();
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int q_row = w0.x, q_cnt = w0.y, t_row = w0.z, t_cnt = w0.w;
            const int n_tiles = (t_cnt + kTileN - 1) / kTileN;
            const int n_sub = (q_cnt + kTileM - 1) / kTileM;
            for (int sub = 0; sub < n_sub; ++sub) {
                mbar_wait(&a_empty[ap.stage], ap.phase ^ 1);
                if (elect_one()) {
                    mbar_arrive_expect_tx(&a_full[ap.stage], kATileBytes);
                    tma_load_2d(s_a + size_t(ap.stage) * kATileBytes, &tmap_q, &a_full[ap.stage], 0, q_row + sub * kTileM);
                }
                __syncwarp();
                ap.advance();
                for (int t = 0; t < n_tiles; ++t) {
                    mbar_wait_suspend(&b_empty[bp.stage], bp.phase ^ 1, wait_ns);
                    if (elect_one()) {
                        uint8_t* dst = s_b + size_t(bp.stage) * kStageBytes;
                        const bool second = t * kTileN + kTileM < t_cnt;
                        mbar_arrive_expect_tx(&b_full[bp.stage], second ? kStageBytes : kStageBytes - kATileBytes);
                        tma_load_2d(dst, &tmap, &b_full[bp.stage], 0, t_row + t * kTileN);
                        if (second)
                            tma_load_2d(dst + kATileBytes, &tmap, &b_full[bp.stage], 0, t_row + t * kTileN + kTileM);
                        tma_load_2d(dst + kBTileBytes, &tmap_x, &b_full[bp.stage], 0, kTileM + t_row + t * kTileN);
                    }
                    __syncwarp();
                    bp.advance();
                }
            }
        }
    } else if (warp == kMmaWarp) {
        Pipe ap(kAStages), bp(kPpBStages);
        uint32_t g = 0;   // train tiles issued so far: accumulator g & 1, its barrier phase (g >> 1) & 1
        const bool prof = PROF && (probe_flags & 16u) && blockIdx.x == 0;
        long long pc_b = 0, pc_t = 0;
        const long long pc_begin = clock64();
        mbar_wait(ax_full, 0);
        const uint64_t axdesc = make_kmajor_sw32_desc(smem_u32(s_ax));
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int n_tiles = (w0.w + kTileN - 1) / kTileN;
            const int n_sub = (w0.y + kTileM - 1) / kTileM;
            for (int sub = 0; sub < n_sub; ++sub) {
                mbar_wait(&a_full[ap.stage], ap.phase);
                const uint64_t adesc = make_kmajor_sw128_desc(smem_u32(s_a + size_t(ap.stage) * kATileBytes));
                for (int t = 0; t < n_tiles; ++t, ++g) {
                    long long p0 = prof ? clock64() : 0;
                    mbar_wait_suspend(&b_full[bp.stage], bp.phase, wait_ns);
                    if (prof) pc_b += clock64() - p0;
                    const uint32_t bsm = smem_u32(s_b + size_t(bp.stage) * kStageBytes);
                    const uint64_t bdesc0 = make_kmajor_sw128_desc(bsm);
                    const uint64_t xdesc = make_kmajor_sw32_desc(bsm + kBTileBytes);
                    const uint32_t a = g & 1u;
                    p0 = prof ? clock64() : 0;
                    mbar_wait_suspend(&t_empty[a], ((g >> 1) & 1u) ^ 1u, wait_ns);
                    if (prof) pc_t += clock64() - p0;
                    tc_fence_after();
                    const uint32_t d = tmem_base + a * kTileN;
                    if (elect_one()) {
                        if (!(probe_flags & 2u)) {
#pragma unroll
                            for (int k = 0; k < kKSteps; ++k)
                                mma_i8_ss(d, adesc + uint64_t(k) * (kUmmaK >> 4), bdesc0 + uint64_t(k) * (kUmmaK >> 4),
                                          kIdesc, k > 0);
                            mma_i8_ss(d, axdesc, xdesc, kIdesc, 1);
                        }
                        mma_commit(&t_full[a]);
                        mma_commit(&b_empty[bp.stage]);
                    }
                    __syncwarp();
                    bp.advance();
                }
                if (elect_one()) mma_commit(&a_empty[ap.stage]);
                __syncwarp();
                ap.advance();
            }
        }
        if (prof && lane == 0) {
            long long* pc = reinterpret_cast<long long*>(fix_list + (1 << 17));
            pc[4] = clock64() - pc_begin;
            pc[5] = pc_b;
            pc[6] = pc_t;
        }
    } else {
        // ------------------------------------------------ epilogue warps: every warp reads BOTH accumulators
        const int quad = warp & 3;     // TMEM lane quadrant = rows 32 quad .. 32 quad + 31 of the query block
        const int colq = warp >> 2;    // which 64-column quarter of every tile
        const int row_l = quad * 32 + lane;
        const int colbase = colq * kPpColsPerThread;
        const uint32_t taddr0 = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + colbase;
        const uint32_t slots0 = smem_u32(s_slots);
        const uint32_t my_slots = slots0 + threadIdx.x * 16;
        const TrkConst tc{my_slots, 2u * my_slots + kSlotStride, !(PROF && (probe_flags & 32u))};
        uint32_t g = 0;
        const bool prof = PROF && (probe_flags & 16u) && blockIdx.x == 0 && warp == 0;
        long long pc_f = 0, pc_l = 0, pc_e = 0, pc_n = 0;
        const long long pc_begin = clock64();
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int4 w1 = __ldg(reinterpret_cast<const int4*>(work + w) + 1);
            const int q_row = w0.x, q_cnt = w0.y, t_row = w0.z, t_cnt = w0.w, t_pad = w1.x;
            const long long out_row =
                static_cast<long long>(static_cast<uint32_t>(w1.z)) | (static_cast<long long>(w1.w) << 32);
            const int n_tiles = (t_cnt + kTileN - 1) / kTileN;
            const int n_sub = (q_cnt + kTileM - 1) / kTileM;
            const int32_t* nbase = norms_pad + t_pad;
            KeyFmt kf;
            kf.n_units = n_tiles * (kPpColsPerThread / kUnit);
            kf.ushift = 2;
            kf.ubits = max(32 - __clz(kf.n_units - 1), 8);
            kf.sh = kf.ubits > 8 ? kf.ubits - 8 : 0;
            kf.mul = 2u << kf.ubits;
            const int rem = t_cnt - (n_tiles - 1) * kTileN - colbase;
            const int n_seen = (n_tiles - 1) * (kPpColsPerThread / kUnit) +
                               min(max((rem + kUnit - 1) / kUnit, 0), kPpColsPerThread / kUnit);
            for (int sub = 0; sub < n_sub; ++sub) {
                const int row = sub * kTileM + row_l;   // row of the work item
                const bool valid = row < q_cnt;
                const int na = valid ? __ldg(norms_q + q_row + row) : 0;
                Trk sk{1, 0, 0, 0, 0, my_slots, 0};
                for (int t = 0; t < n_tiles; ++t, ++g) {
                    const uint32_t a = g & 1u;
                    long long p0 = prof ? clock64() : 0;
                    mbar_wait_suspend(&t_full[a], (g >> 1) & 1u, epi_wait_ns);
                    if (prof) {
                        pc_f += clock64() - p0;
                        ++pc_n;
                    }
                    tc_fence_after();
                    if (MODE == 2) {
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&t_empty[a]);
                        continue;
                    }
                    const uint32_t taddr = taddr0 + a * kTileN;
                    uint32_t v[2][16];
                    tmem_ld_32x16(taddr, v[0]);
                    const uint32_t ut = static_cast<uint32_t>(kf.n_units - 1 - t * (kPpColsPerThread / kUnit)) << 1;
                    const bool full_tile = (t + 1) * kTileN <= t_cnt;
                    if (full_tile && kf.sh == 0) {
#pragma unroll
                        for (int c = 0; c < kPpColsPerThread / 16; ++c) {
                            p0 = prof ? clock64() : 0;
                            tmem_ld_wait();
                            if (prof) pc_l += clock64() - p0;
                            if (c + 1 < kPpColsPerThread / 16) {
                                tmem_ld_32x16(taddr + (c + 1) * 16, v[(c + 1) & 1]);
                            } else {
                                tc_fence_before();
                                __syncwarp();
                                if (lane == 0) mbar_arrive(&t_empty[a]);
                            }
                            fold_chunk<false, MODE == 1, true, TRK>(sk, v[c & 1], t * kTileN + colbase + c * 16, t_cnt, tc, kf,
                                                                    ut, c, kBatch ? 1 + (c & 1) : 0);
                        }
                    } else if (full_tile) {
#pragma unroll
                        for (int c = 0; c < kPpColsPerThread / 16; ++c) {
                            tmem_ld_wait();
                            if (c + 1 < kPpColsPerThread / 16) {
                                tmem_ld_32x16(taddr + (c + 1) * 16, v[(c + 1) & 1]);
                            } else {
                                tc_fence_before();
                                __syncwarp();
                                if (lane == 0) mbar_arrive(&t_empty[a]);
                            }
                            fold_chunk<false, MODE == 1, false, TRK>(sk, v[c & 1], t * kTileN + colbase + c * 16, t_cnt, tc, kf,
                                                                     ut, c, kBatch ? 1 + (c & 1) : 0);
                        }
                    } else {
#pragma unroll
                        for (int c = 0; c < kPpColsPerThread / 16; ++c) {
                            const int col0 = t * kTileN + colbase + c * 16;
                            tmem_ld_wait();
                            if (c + 1 < kPpColsPerThread / 16) {
                                tmem_ld_32x16(taddr + (c + 1) * 16, v[(c + 1) & 1]);
                            } else {
                                tc_fence_before();
                                __syncwarp();
                                if (lane == 0) mbar_arrive(&t_empty[a]);
                            }
                            if (col0 < t_cnt) fold_chunk<true, MODE == 1, false, TRK>(sk, v[c & 1], col0, t_cnt, tc, kf, ut, c, 0);
                        }
                    }
                }
                const long long pe0 = prof ? clock64() : 0;
                if (MODE != 0 || (PROF && (probe_flags & 64u))) {   // probe bit 6: no end-of-pass work (results meaningless)
                    if (valid && colq == 0) out[out_row + row] = make_int4(static_cast<int>(sk.k1), static_cast<int>(sk.k2), na, 0);
                    continue;
                }
                const int n_in = TRK == 1 ? (sk.k1 >= 2u) + (sk.k2 >= 2u) + (sk.dm >= 2u) + (sk.dd >= 2u) : n_seen;
                Top2K r = resolve_units<TRK>(sk, n_in, kf, colbase, my_slots, nbase, t_cnt);
                const uint32_t mf = kf.ubits + 1;
                // A dropped unit whose (truncated) maximum equals the second kept one may hold a candidate.  With a
                // tie list the row goes out as it is and (row, unit) is appended for knn_tie_kernel, which merges the
                // unit's 16 keys into the row afterwards (the merge order does not matter): the warp-cooperative
                // recompute in here made all sixteen warps -- and with them the tensor pipe -- wait for the slowest
                // warp's L2 round trips at the end of every pass.
                const bool need = valid && n_in >= 3 && (sk.dm >> mf) == (sk.k2 >> mf);
                int ncol = -1;
                if (tie_list)
                    ncol = need ? key_col(sk.dm, kf, colbase) : -1;
                else
                    r = recompute_units(r, need, q_row + row, key_col(sk.dm, kf, colbase), desc_q, desc, t_row, t_cnt, nbase);
                const bool tie = n_in >= 4 && (sk.dd >> mf) == (sk.k2 >> mf);
                // the threads of column quarters 1..3 hand their top-2 to the quarter-0 thread of the row through
                // their own (now free) parking slots; the second barrier keeps the next pass's parking stores
                // behind the reads
                if (colq != 0) {
                    sts128(my_slots, r.k1, r.k2, r.i1, r.i2);
                    sts128(my_slots + kSlotRow, tie ? 1u : 0u, static_cast<uint32_t>(ncol), 0u, 0u);
                }
                asm volatile("bar.sync %0, 128;" ::"r"(1 + quad) : "memory");
                if (colq == 0) {
                    bool flagged = tie;
                    int cols[4] = {ncol, -1, -1, -1};
                    if (r.i1 < 0) r.i1 = INT_MAX;   // empty entries sort last
                    if (r.i2 < 0) r.i2 = INT_MAX;
                    Top2K mrg = r;
#pragma unroll
                    for (int j = 1; j < 4; ++j) {
                        const uint32_t paddr = my_slots + j * (kTileM * 16);
                        const uint4 o = lds128(paddr);
                        const uint4 f = lds128(paddr + kSlotRow);
                        flagged = flagged || f.x != 0;
                        cols[j] = static_cast<int>(f.y);
                        Top2K b{static_cast<int>(o.x), static_cast<int>(o.y), static_cast<int>(o.z), static_cast<int>(o.w)};
                        if (b.i1 < 0) b.i1 = INT_MAX;
                        if (b.i2 < 0) b.i2 = INT_MAX;
                        mrg = merge_top2(mrg, b);
                    }
                    if (valid && !flagged && (cols[0] & cols[1] & cols[2] & cols[3]) >= 0) {   // some unit to patch in
                        const int e = atomicAdd(flags + 3, 1);
                        if (e < tie_cap) {
                            const long long o_idx = out_row + row;
                            tie_list[2 * e] = make_int4(static_cast<int>(static_cast<uint32_t>(o_idx)),
                                                        static_cast<int>(o_idx >> 32), q_row + row, static_cast<int>(w));
                            tie_list[2 * e + 1] = make_int4(cols[0], cols[1], cols[2], cols[3]);
                        } else {
                            flagged = true;   // list full: brute force instead
                        }
                    }
                    if (valid && flagged) {   // rare: left to knn_fix_kernel
                        atomicAdd(flags + 1, 1);
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
                asm volatile("bar.sync %0, 128;" ::"r"(5 + quad) : "memory");
                if (prof) pc_e += clock64() - pe0;
            }
        }
        if (prof && lane == 0) {
            long long* pc = reinterpret_cast<long long*>(fix_list + (1 << 17));
            pc[0] = clock64() - pc_begin;
            pc[1] = pc_f;
            pc[2] = pc_l;
            pc[3] = pc_e;
            pc[7] = pc_n;
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

// Patches of the ping-pong kernel's tie list: entry e = {output row (i64), query row, work item} {first column of
// up to four dropped units, -1 = none}.  One warp per entry: the 16 keys of each unit are recomputed from the
// descriptors (lane e and lane 16 + e take 64 bytes each of train row col + e) and merged into the row under the
// (key desc, index asc) order.  Rows left to knn_fix_kernel are skipped.
__global__ void __launch_bounds__(256) knn_tie_kernel(const WorkItem* __restrict__ work, const uint8_t* __restrict__ desc_q,
                                                      const int32_t* __restrict__ norms_q, const uint8_t* __restrict__ desc,
                                                      const int32_t* __restrict__ norms, const int32_t* __restrict__ flags,
                                                      const int4* __restrict__ tie_list, int tie_cap, int4* __restrict__ out) {
    const int cnt_all = *(static_cast<volatile const int32_t*>(flags) + 3);
    const int cnt = cnt_all < tie_cap ? cnt_all : tie_cap;
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    for (int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; e < cnt; e += warps) {
        const int4 ea = __ldg(tie_list + 2 * e), eb = __ldg(tie_list + 2 * e + 1);
        const long long o_idx = static_cast<long long>(static_cast<uint32_t>(ea.x)) | (static_cast<long long>(ea.y) << 32);
        const int4 o = out[o_idx];
        if (o.x == kFlagged) continue;
        const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + ea.w));
        const int t_row = w0.z, t_cnt = w0.w;
        const int na = __ldg(norms_q + ea.z);
        Top2K r{o.x < 0 ? INT_MIN : na - o.z, o.y < 0 ? INT_MIN : na - o.w, o.x, o.y};
        const int cols[4] = {eb.x, eb.y, eb.z, eb.w};
        const int part = lane >> 4;
        uint4 a[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
            a[i] = __ldg(reinterpret_cast<const uint4*>(desc_q + static_cast<size_t>(ea.z) * kRowBytes) + 4 * part + i);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int cc = cols[u];
            if (cc < 0) continue;   // warp-uniform
            const int c = cc + (lane & 15);
            uint32_t dot = 0;
            if (c < t_cnt) {
                const uint4* bp = reinterpret_cast<const uint4*>(desc + static_cast<size_t>(t_row + c) * kRowBytes) + 4 * part;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint4 b = __ldg(bp + i);
                    dot = __dp4a(a[i].x, b.x, dot);
                    dot = __dp4a(a[i].y, b.y, dot);
                    dot = __dp4a(a[i].z, b.z, dot);
                    dot = __dp4a(a[i].w, b.w, dot);
                }
            }
            dot += __shfl_xor_sync(0xffffffffu, dot, 16);
            const int key = c < t_cnt ? 2 * static_cast<int>(dot) - __ldg(norms + t_row + c) : INT_MIN;
#pragma unroll 1
            for (int k = 0; k < kUnit; ++k) {
                const int ke = __shfl_sync(0xffffffffu, key, k);
                if (cc + k < t_cnt) top2k_insert_ordered(r, ke, cc + k);   // every lane keeps the same r
            }
        }
        if (lane == 0)
            out[o_idx] = make_int4(r.i1, r.i2, r.i1 < 0 ? INT_MAX : na - r.k1, r.i2 < 0 ? INT_MAX : na - r.k2);
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

// PROF: the instantiation with the cycle counters and probe switches of probe bits 4..6 (they cost the default
// kernel 15 % when compiled in: two predicated clock reads per chunk).
template <int MODE, int TRK, bool PROF = false>
cudaError_t launch_variant(const KnnQuery& q, const CUtensorMap& tmap, const CUtensorMap& tmap_x, const CUtensorMap& tmap_ax,
                           const uint8_t* d_desc, const WorkItem* d_work, int64_t n_work,
                           const int32_t* d_norms_pad, int32_t* d_flags, int force, const KnnTuning& tn,
                           int2* d_fix_list, int fix_cap, int4* d_tie_list, int tie_cap, int32_t* d_knn, unsigned grid,
                           cudaStream_t st) {
    auto kern = knn2_wide_kernel<MODE, TRK, PROF>;
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemBytes));
        if (e != cudaSuccess) return e;
        configured = true;
    }
    kern<<<grid, kThreads, kSmemBytes, st>>>(*q.tmap, tmap, tmap_x, tmap_ax, q.desc, q.norms, d_desc, d_work, n_work,
                                             d_norms_pad, d_flags, force, tn.probe_flags, tn.wait_ns, tn.epi_wait_ns,
                                             d_fix_list, fix_cap, d_tie_list, tie_cap, reinterpret_cast<int4*>(d_knn));
    return cudaGetLastError();
}

template <int MODE, int TRK, bool PROF = false>
cudaError_t launch_pp(const KnnQuery& q, const CUtensorMap& tmap, const CUtensorMap& tmap_x, const CUtensorMap& tmap_ax,
                      const uint8_t* d_desc, const WorkItem* d_work, int64_t n_work, const int32_t* d_norms_pad,
                      int32_t* d_flags, int force, const KnnTuning& tn, int2* d_fix_list, int fix_cap, int4* d_tie_list,
                      int tie_cap, int32_t* d_knn, unsigned grid, cudaStream_t st) {
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(knn2_pp_kernel<MODE, TRK, PROF>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(kPpSmemBytes));
        if (e != cudaSuccess) return e;
        configured = true;
    }
    knn2_pp_kernel<MODE, TRK, PROF><<<grid, kThreads, kPpSmemBytes, st>>>(*q.tmap, tmap, tmap_x, tmap_ax, q.desc, q.norms, d_desc,
                                                                d_work, n_work, d_norms_pad, d_flags, force,
                                                                tn.probe_flags, tn.wait_ns, tn.epi_wait_ns, d_fix_list,
                                                                fix_cap, d_tie_list, tie_cap, reinterpret_cast<int4*>(d_knn));
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
// 1: max tree only, 2: no epilogue (timing probes); 3: with the warp-uniform skip, 4: with the lean tracker;
// 5: ping-pong kernel (d_tie_list: tie_cap entries of 8 ints for its deferred unit recomputes), 6 / 7: its probes.
int launch_knn2_wide(const CUtensorMap& tmap, const CUtensorMap& tmap_x, const CUtensorMap& tmap_ax,
                     const uint8_t* d_desc, const WorkItem* d_work, int64_t n_work, const int32_t* d_norms,
                     const int32_t* d_norms_pad,
                     int32_t* d_flags, int variant, int force, const KnnTuning& tn, int32_t* d_fix_list,
                     int32_t fix_cap, int32_t* d_knn, int num_sms, cudaStream_t st, int64_t* launches,
                     const KnnQuery* query, int32_t* d_tie_list, int32_t tie_cap) {
    if (n_work <= 0) return 0;
    int4* tl = (tn.probe_flags & 128u) ? nullptr : reinterpret_cast<int4*>(d_tie_list);   // probe bit 7: recompute in the kernel
    // query rows: the registered descriptor array itself unless the caller brings its own (mutual check)
    const KnnQuery q = query ? *query : KnnQuery{&tmap, d_desc, d_norms};
    const unsigned grid = static_cast<unsigned>(n_work < num_sms ? n_work : num_sms);
    int2* fl = reinterpret_cast<int2*>(d_fix_list);
    cudaError_t e = cudaMemsetAsync(d_flags + 2, 0, 2 * sizeof(int32_t), st);   // lengths of this launch's fix and tie lists
    if (e != cudaSuccess) return static_cast<int>(e);
#define SFM_WIDE_ARGS q, tmap, tmap_x, tmap_ax, d_desc, d_work, n_work, d_norms_pad, d_flags
    const bool prof = (tn.probe_flags & (16u | 32u | 64u | 256u | 1024u)) != 0;   // cycle counters / probe switches compiled in
    if (variant == 1)
        e = prof ? launch_variant<1, 0, true>(SFM_WIDE_ARGS, 1, tn, fl, fix_cap, tl, tie_cap, d_knn, grid, st)
                 : launch_variant<1, 0>(SFM_WIDE_ARGS, 1, tn, fl, fix_cap, tl, tie_cap, d_knn, grid, st);
    else if (variant == 2)
        e = prof ? launch_variant<2, 0, true>(SFM_WIDE_ARGS, 1, tn, fl, fix_cap, tl, tie_cap, d_knn, grid, st)
                 : launch_variant<2, 0>(SFM_WIDE_ARGS, 1, tn, fl, fix_cap, tl, tie_cap, d_knn, grid, st);
    else if (variant == 3)   // the kNN WITH the warp-uniform skip of the unit bookkeeping (A/B: 5 % slower)
        e = launch_variant<0, 1>(SFM_WIDE_ARGS, force, tn, fl, fix_cap, tl, tie_cap, d_knn, grid, st);
    else if (variant == 4)   // lean tracker (A/B: 7 % fewer ALU-pipe instructions, same speed)
        e = launch_variant<0, 0>(SFM_WIDE_ARGS, force, tn, fl, fix_cap, tl, tie_cap, d_knn, grid, st);
    else if (variant == 5)   // ping-pong kernel
        e = prof ? launch_pp<0, 0, true>(SFM_WIDE_ARGS, force, tn, fl, fix_cap, tl, tie_cap, d_knn, grid, st)
                 : launch_pp<0, 0>(SFM_WIDE_ARGS, force, tn, fl, fix_cap, tl, tie_cap, d_knn, grid, st);
    else if (variant == 6)   // ... its max-tree-only probe
        e = prof ? launch_pp<1, 0, true>(SFM_WIDE_ARGS, 1, tn, fl, fix_cap, tl, tie_cap, d_knn, grid, st)
                 : launch_pp<1, 0>(SFM_WIDE_ARGS, 1, tn, fl, fix_cap, tl, tie_cap, d_knn, grid, st);
    else if (variant == 7)   // ... its TMA + MMA pipeline alone
        e = prof ? launch_pp<2, 0, true>(SFM_WIDE_ARGS, 1, tn, fl, fix_cap, tl, tie_cap, d_knn, grid, st)
                 : launch_pp<2, 0>(SFM_WIDE_ARGS, 1, tn, fl, fix_cap, tl, tie_cap, d_knn, grid, st);
    else
        e = prof ? launch_variant<0, 2, true>(SFM_WIDE_ARGS, force, tn, fl, fix_cap, tl, tie_cap, d_knn, grid, st)
                 : launch_variant<0, 2>(SFM_WIDE_ARGS, force, tn, fl, fix_cap, tl, tie_cap, d_knn, grid, st);
#undef SFM_WIDE_ARGS
    ++*launches;
    if (e == cudaSuccess && (variant == 0 || (variant >= 3 && variant <= 5)) && tl && !(tn.probe_flags & 4u)) {
        knn_tie_kernel<<<4 * num_sms, 256, 0, st>>>(d_work, q.desc, q.norms, d_desc, d_norms, d_flags, tl, tie_cap,
                                                    reinterpret_cast<int4*>(d_knn));
        ++*launches;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess && (variant == 0 || (variant >= 3 && variant <= 5)) && !(tn.probe_flags & 4u)) {   // (probe bit 2: main kernel alone)
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
