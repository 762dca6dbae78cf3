// K1, TS variant -- probes and debug tile (the kernel itself follows in this file once the
// numbers below justify it).
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


// ============================================================================================
// The kernel.  One persistent CTA per SM, warp-specialised as knn_wide.cu: 16 epilogue warps, a
// TMA producer warp, an MMA warp.  A work item is 256 query rows (two 128-row blocks, resident in
// TMEM as the A operands) against one whole train image streamed as 128-row tiles through a
// 4-deep shared-memory ring; every tile feeds two MMA groups (one per query block) of five
// 128x128x32 instructions (four K steps + the norm-fold step) into a ring of three 128-column
// accumulator stages.  Epilogue warp e = (half, block, quadrant): the eight warps of a block take
// that block's groups, each thread one row x 64 columns (a column half) per group.
//
// Epilogue arithmetic (exactness argument as in knn_wide.cu): candidate UNITS are 32 columns --
// one packed key [max >> sh | reversed unit id | slot] and one 7-instruction insertion network
// per 32 accumulators instead of per 16; the four largest keys are kept, the units of the best two
// are parked in shared memory (8 predicated STS.128), resolved exactly at the end of the work item
// with the norm parities read from a packed bitmap (one word per unit), a tied third unit is
// recomputed from the descriptors by the warp, a tied fourth sends the row to knn_fix_kernel.
constexpr int NT = 128;                              // train rows per tile = UMMA N
constexpr int kAccStages = 3;
constexpr int kBStages = 4;
constexpr int kBTileBytes = NT * kRowBytes;          // 16 KiB
constexpr int kXTileBytes = NT * kUmmaK;             // 4 KiB
constexpr int kStageBytes = kBTileBytes + kXTileBytes;
constexpr int kEpiWarps = 16;
constexpr int kThreads = (kEpiWarps + 2) * 32;
constexpr int kColsPerThread = NT / 2;               // 64 columns of every group per thread
constexpr int kUnit = 32;
constexpr int kUnitsPerTile = kColsPerThread / kUnit;   // 2
constexpr uint32_t kIdesc = make_idesc_u8(kTileM, NT);
constexpr int kFoldC = 255 * 31 * 255 + 254;
constexpr uint32_t kSlotRow = kEpiWarps * 32 * 16;          // bytes between a thread's uint4s
constexpr uint32_t kSlotStride = (kUnit / 4) * kSlotRow;    // bytes per slot (64 KiB)
constexpr size_t kSlotBytes = 2 * size_t(kSlotStride);      // 128 KiB
constexpr int kFlagged = -2;
constexpr size_t kXchgBytes = 2 * SFM_QBLOCK * 16 + 2 * SFM_QBLOCK * 4;
constexpr size_t kSmemData = size_t(kBStages) * kStageBytes + kSlotBytes + kXchgBytes;
constexpr int kNumBars = 2 + 2 * kBStages + 2 * kAccStages;
constexpr size_t kSmemBytes = kSmemData + kNumBars * 8 + 16;
static_assert(kSmemBytes <= 232448, "shared memory budget");
static_assert(kAccStages * NT == kAccCols, "accumulator ring fills the columns below the A operands");

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

struct Trk {
    uint32_t k1, k2, dm, dd;
};

struct KeyFmt {
    uint32_t sh, mul, ubits;   // key = (m >> sh) * mul + (uidrev << 1) + slot, mul = 2^(ubits+1)
    int n_units;               // units per thread over the whole work item (2 per train tile)
    bool no_park;              // timing probe: skip the parking stores (results become meaningless)
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

// Maximum of the 16 folded accumulators acc[OFF..OFF+16) (in place; MASKED: columns >= t_cnt count as 0).
template <bool MASKED, int OFF>
__device__ __forceinline__ int chunk_max(uint32_t (&acc)[64], int col0, int t_cnt) {
    int v[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        if (MASKED) acc[OFF + e] = (col0 + OFF + e < t_cnt) ? acc[OFF + e] : 0u;
        v[e] = static_cast<int>(acc[OFF + e]);
    }
    const int a0 = __vimax3_s32(v[0], v[1], v[2]), a1 = __vimax3_s32(v[3], v[4], v[5]);
    const int a2 = __vimax3_s32(v[6], v[7], v[8]), a3 = __vimax3_s32(v[9], v[10], v[11]);
    const int a4 = __vimax3_s32(v[12], v[13], v[14]);
    return max(__vimax3_s32(a0, a1, a2), __vimax3_s32(a3, a4, v[15]));
}

// One 32-column unit acc[OFF..OFF+32): m its maximum, ulow = uidrev << 1.
template <bool SH0, int OFF>
__device__ __forceinline__ void unit_update(Trk& s, const uint32_t (&acc)[64], int m, uint32_t slots,
                                            const KeyFmt& kf, uint32_t ulow) {
    const uint32_t sb = s.k2 & 1u;
    const uint32_t low = sb | ulow;
    const uint32_t key = SH0 ? static_cast<uint32_t>(m) * 512u + low
                             : (static_cast<uint32_t>(m) >> kf.sh) * kf.mul + low;
    if (key > s.k2 && !kf.no_park) {
        const uint32_t dst = slots + sb * kSlotStride;
#pragma unroll
        for (int g = 0; g < 8; ++g)
            sts128(dst + g * kSlotRow, acc[OFF + 4 * g], acc[OFF + 4 * g + 1], acc[OFF + 4 * g + 2], acc[OFF + 4 * g + 3]);
    }
    const uint32_t t1 = min(s.k1, key);
    s.k1 = max(s.k1, key);
    const uint32_t t2 = min(s.k2, t1);
    s.k2 = max(s.k2, t1);
    const uint32_t t3 = min(s.dm, t2);
    s.dm = max(s.dm, t2);
    s.dd = max(s.dd, t3);
}

// First train column of the unit a key names (colbase: first column of this thread's half).
__device__ __forceinline__ int key_col(uint32_t key, const KeyFmt& kf, int colbase) {
    const int uidrev = static_cast<int>((key >> 1) & ((1u << kf.ubits) - 1u));
    const int idx = kf.n_units - 1 - uidrev;
    return (idx >> 1) * NT + colbase + (idx & 1) * kUnit;
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

__device__ __forceinline__ Top2K resolve_units(const Trk& s, int n_seen, const KeyFmt& kf, int colbase, uint32_t slots,
                                               const uint32_t* __restrict__ parbase, int t_cnt) {
    const int s1 = s.k1 & 1, s2 = s.k2 & 1;
    const int c1 = n_seen >= 1 ? key_col(s.k1, kf, colbase) : -1;
    const int c2 = n_seen >= 2 ? key_col(s.k2, kf, colbase) : -1;
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
    uint8_t* b;      // kBStages x (128 train rows x 128 B | 128 x 32 B fold operand)
    uint4* slots;    // parked candidate units: [slot][8 x uint4][thread]
    uint4* xchg;     // [2][256] per-row results handed from the upper to the lower column half
    int32_t* pub;    // [2][256] tie flag handed over with them
    uint64_t* a_full;    // [block]: the block's query rows of the next work item are in TMEM
    uint64_t* b_full;
    uint64_t* b_empty;
    uint64_t* t_full;    // [accumulator stage]
    uint64_t* t_empty;
    uint32_t* tmem_slot;
    __device__ explicit SmemLayout(uint8_t* base) {
        b = base;
        slots = reinterpret_cast<uint4*>(b + size_t(kBStages) * kStageBytes);
        xchg = reinterpret_cast<uint4*>(b + size_t(kBStages) * kStageBytes + kSlotBytes);
        pub = reinterpret_cast<int32_t*>(b + size_t(kBStages) * kStageBytes + kSlotBytes + 2 * SFM_QBLOCK * 16);
        uint64_t* bars = reinterpret_cast<uint64_t*>(base + kSmemData);
        a_full = bars;
        b_full = a_full + 2;
        b_empty = b_full + kBStages;
        t_full = b_empty + kBStages;
        t_empty = t_full + kAccStages;
        tmem_slot = reinterpret_cast<uint32_t*>(t_empty + kAccStages);
    }
};

// This thread's 64 bytes (column half `half`) of query row `row` of a work item, clamped into the block.
__device__ __forceinline__ void load_a_half(uint32_t (&a)[16], const uint8_t* __restrict__ desc, int q_row, int q_cnt,
                                            int row, int half) {
    const int r = q_row + min(row, q_cnt - 1);
    const uint4* src = reinterpret_cast<const uint4*>(desc + static_cast<size_t>(r) * kRowBytes + 64 * half);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint4 w = __ldg(src + i);
        a[4 * i] = w.x; a[4 * i + 1] = w.y; a[4 * i + 2] = w.z; a[4 * i + 3] = w.w;
    }
}

// One accumulator group of this thread: its 64 columns come out of TMEM in ONE load and the stage goes
// back to the MMA warp as soon as they have landed -- the MMA -> epilogue -> MMA chain of a stage then
// holds one TMEM round trip instead of the whole reduction (measured: the stage was held 560 cycles with
// four 16-column loads interleaved with the work, and the hand-over round trip itself is ~800).  The two
// 32-column units are reduced from registers afterwards, overlapping the next groups' MMAs.
template <bool MASKED, bool SH0, int MODE>
__device__ __forceinline__ void process_group(Trk& sk, uint32_t taddr, int t, int t_cnt, int colbase, uint32_t my_slots,
                                              const KeyFmt& kf, uint64_t* empty, int lane, long long* ev3 = nullptr,
                                              long long ev_base = 0) {
    uint32_t b[64];
    const uint32_t ut = static_cast<uint32_t>(kf.n_units - 1 - t * kUnitsPerTile) << 1;
    const int col0 = t * NT + colbase;
    tmem_ld_32x64(taddr, b);
    tmem_ld_wait();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(empty);
    if (ev3 && lane == 0) *ev3 = clock64() - ev_base;
    {
        const int ma = chunk_max<MASKED, 0>(b, col0, t_cnt), mb = chunk_max<MASKED, 16>(b, col0, t_cnt);
        if (MODE == 1) {
            sk.k1 = max(sk.k1, static_cast<uint32_t>(max(ma, mb)));
        } else if (!MASKED || col0 < t_cnt) {
            unit_update<SH0, 0>(sk, b, max(ma, mb), my_slots, kf, ut);
        }
    }
    {
        const int ma = chunk_max<MASKED, 32>(b, col0, t_cnt), mb = chunk_max<MASKED, 48>(b, col0, t_cnt);
        if (MODE == 1) {
            sk.k1 = max(sk.k1, static_cast<uint32_t>(max(ma, mb)));
        } else if (!MASKED || col0 + kUnit < t_cnt) {
            unit_update<SH0, 32>(sk, b, max(ma, mb), my_slots, kf, ut ^ 2u);
        }
    }
}

// Cycle counters of the profiling build (PROF): per CTA, 16 words.
//  [0] MMA warp: total loop cycles   [1] waiting b_full   [2] waiting t_empty   [3] waiting a_full
//  [4] producer: waiting b_empty     [5] groups issued
//  [8] epilogue warp 0: total        [9] waiting t_full   [10] in process_group [11] end of work item (resolve..write)
//  [12] epilogue warp 0: A refill (st + wait + arrive)
// plus, for CTA 0, time stamps (cycles since the CTA's start) of the first kProfEvents groups after the
// counters: [k][g], k = 0 MMA warp passed t_empty, 1 commit issued, 2 epilogue (quadrant 0, lower half of
// the group's block) passed t_full, 3 it handed the stage back, 4 it finished the group.
constexpr int kProfEvents = 4096;
#define SFM_PROF_EV(kind, g) do { if (PROF && blockIdx.x == 0 && lane == 0 && (g) < kProfEvents) \
    prof_out[gridDim.x * 16 + (kind) * kProfEvents + (g)] = clock64() - prof_base; } while (0)
#define SFM_PROF_T0() long long prof_t0 = PROF ? clock64() : 0
#define SFM_PROF_ADD(slot) do { if (PROF) { const long long t1__ = clock64(); prof[slot] += t1__ - prof_t0; prof_t0 = t1__; } } while (0)

// MODE 0: the kNN.  1: max tree only, 2: no epilogue (timing probes; output is not a kNN).
template <int MODE, bool PROF>
__global__ void __launch_bounds__(kThreads, 1)
knn2_ts_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_x,
               const uint8_t* __restrict__ desc, const uint8_t* __restrict__ bias32,
               const WorkItem* __restrict__ work, long long n_work, const int32_t* __restrict__ norms,
               const int32_t* __restrict__ norms_pad, const uint32_t* __restrict__ par_bits,
               int32_t* __restrict__ flags, int force, uint32_t probe_flags, uint32_t wait_ns, uint32_t epi_wait_ns,
               int2* __restrict__ fix_list, int fix_cap, int4* __restrict__ out, long long* __restrict__ prof_out) {
    long long prof[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) prof[i] = 0;
    constexpr int kProducerWarp = kEpiWarps;
    constexpr int kMmaWarp = kEpiWarps + 1;
    // flags[0] != 0: some row's norm is beyond the fold range -> the generic kernel runs instead
    if (!force && *static_cast<volatile const int32_t*>(flags) != 0) return;

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    if (smem_u32(smem_raw) & 1023u) __trap();   // SWIZZLE_128B tiles need 1024-byte alignment
    SmemLayout sm(smem_raw);
    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;

    if (warp == kProducerWarp && lane == 0) {
        tma_prefetch_desc(&tmap);
        tma_prefetch_desc(&tmap_x);
        for (int i = 0; i < 2; ++i) mbar_init(&sm.a_full[i], kEpiWarps / 2);
        for (int i = 0; i < kBStages; ++i) {
            mbar_init(&sm.b_full[i], 1);
            mbar_init(&sm.b_empty[i], 1);
        }
        for (int i = 0; i < kAccStages; ++i) {
            mbar_init(&sm.t_full[i], 1);
            mbar_init(&sm.t_empty[i], kEpiWarps / 2);
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
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *sm.tmem_slot, 0);
    const long long prof_base = PROF ? clock64() : 0;
    // probe bits 8..10: use only this many stages of the train-tile ring (0 = all; a latency probe)
    const int b_depth = ((probe_flags >> 8) & 7u) ? min(static_cast<int>((probe_flags >> 8) & 7u), kBStages) : kBStages;

    if (warp == kProducerWarp) {
        Pipe bp(b_depth);
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int t_row = w0.z, t_cnt = w0.w;
            const int n_tiles = (t_cnt + NT - 1) / NT;
            for (int t = 0; t < n_tiles; ++t) {
                SFM_PROF_T0();
                mbar_wait_suspend(&sm.b_empty[bp.stage], bp.phase ^ 1, wait_ns);
                SFM_PROF_ADD(4);
                if (elect_one()) {
                    uint8_t* dst = sm.b + size_t(bp.stage) * kStageBytes;
                    // rows beyond the train image are masked in the epilogue; beyond the array TMA fills zeros
                    mbar_arrive_expect_tx(&sm.b_full[bp.stage], kStageBytes);
                    tma_load_2d(dst, &tmap, &sm.b_full[bp.stage], 0, t_row + t * NT);
                    // fold operand rows are parallel to the descriptor rows, after 128 constant rows
                    tma_load_2d(dst + kBTileBytes, &tmap_x, &sm.b_full[bp.stage], 0, kTileM + t_row + t * NT);
                }
                __syncwarp();
                bp.advance();
            }
        }
        if (PROF && lane == 0) prof_out[blockIdx.x * 16 + 4] = prof[4];
    } else if (warp == kMmaWarp) {
        Pipe bp(b_depth);
        uint32_t s = 0, sph = 0;
        int item = 0;
        const long long prof_begin = PROF ? clock64() : 0;
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x, ++item) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int n_tiles = (w0.w + NT - 1) / NT;
            for (int t = 0; t < n_tiles; ++t) {
                SFM_PROF_T0();
                mbar_wait(&sm.b_full[bp.stage], bp.phase);
                SFM_PROF_ADD(1);
                const uint32_t bsm = smem_u32(sm.b + size_t(bp.stage) * kStageBytes);
                const uint64_t bdesc0 = make_kmajor_sw128_desc(bsm);
                const uint64_t xdesc = make_kmajor_sw32_desc_ts(bsm + kBTileBytes);
#pragma unroll
                for (int a = 0; a < 2; ++a) {
                    if (PROF) prof_t0 = clock64();
                    if (t == 0) mbar_wait(&sm.a_full[a], item & 1);   // this block's query rows are in TMEM
                    SFM_PROF_ADD(3);
                    mbar_wait(&sm.t_empty[s], sph ^ 1);
                    SFM_PROF_ADD(2);
                    SFM_PROF_EV(0, static_cast<int>(prof[5]));
                    tc_fence_after();
                    const uint32_t d = tmem_base + s * NT;
                    const uint32_t ta = tmem_base + kAColBase + 32 * a;
                    if (elect_one()) {
                        if (!(probe_flags & 2u)) {
#pragma unroll
                            for (int k = 0; k < kKSteps; ++k)
                                mma_i8_ts(d, ta + 8 * k, bdesc0 + uint64_t(k) * (kUmmaK >> 4), kIdesc, k > 0);
                            mma_i8_ts(d, tmem_base + kAxCol, xdesc, kIdesc, 1);   // + (1, 255 x 31) . digits(C - ||b||^2 / 2)
                        }
                        mma_commit(&sm.t_full[s]);
                        if (a == 1) mma_commit(&sm.b_empty[bp.stage]);
                    }
                    __syncwarp();
                    SFM_PROF_EV(1, static_cast<int>(prof[5]));
                    if (PROF) prof[5] += 1;
                    if (PROF && (probe_flags & 16u)) {   // diagnostic: how long until THIS warp sees the group complete
                        const long long tq = clock64();
                        mbar_wait(&sm.t_full[s], sph);
                        prof[6] += clock64() - tq;
                    }
                    if (++s == kAccStages) {
                        s = 0;
                        sph ^= 1;
                    }
                }
                bp.advance();
            }
        }
        if (PROF && lane == 0) {
            prof_out[blockIdx.x * 16 + 0] = clock64() - prof_begin;
            prof_out[blockIdx.x * 16 + 1] = prof[1];
            prof_out[blockIdx.x * 16 + 2] = prof[2];
            prof_out[blockIdx.x * 16 + 3] = prof[3];
            prof_out[blockIdx.x * 16 + 5] = prof[5];
            prof_out[blockIdx.x * 16 + 6] = prof[6];
        }
    } else {
        // ------------------------------------------------ epilogue warps
        const int quad = warp & 3;          // TMEM lane quadrant this warp may access
        const int blk = (warp >> 2) & 1;    // which 128-row query block
        const int half = warp >> 3;         // which 64-column half of every group
        const int row = blk * kTileM + quad * 32 + lane;
        const int colbase = half * kColsPerThread;
        const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
        const uint32_t a_taddr = lane_base + kAColBase + 32 * blk + 16 * half;
        const uint32_t my_slots = smem_u32(sm.slots) + threadIdx.x * 16;
        uint32_t s = blk, sph = 0;          // accumulator stage / phase of this block's next group
        uint32_t a_next[16];
        // first work item: query rows (and, once, the constant fold-step operand) into TMEM
        if (static_cast<long long>(blockIdx.x) < n_work) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + blockIdx.x));
            load_a_half(a_next, desc, w0.x, w0.y, row, half);
            tmem_st_32x16(a_taddr, a_next);
            if (blk == 0 && half == 0) {
                uint32_t c[8];
                const uint4 c0 = __ldg(reinterpret_cast<const uint4*>(bias32));
                const uint4 c1 = __ldg(reinterpret_cast<const uint4*>(bias32) + 1);
                c[0] = c0.x; c[1] = c0.y; c[2] = c0.z; c[3] = c0.w; c[4] = c1.x; c[5] = c1.y; c[6] = c1.z; c[7] = c1.w;
                tmem_st_32x8(lane_base + kAxCol, c);
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.a_full[blk]);
        }
        int item = 0;
        const long long prof_begin = PROF ? clock64() : 0;
        int prof_g = blk;                       // global group index of this block's next group
        const bool prof_me = PROF && quad == 0 && half == 0;
        for (long long w = blockIdx.x; w < n_work; w += gridDim.x, ++item) {
            const int4 w0 = __ldg(reinterpret_cast<const int4*>(work + w));
            const int4 w1 = __ldg(reinterpret_cast<const int4*>(work + w) + 1);
            const int q_row = w0.x, q_cnt = w0.y, t_row = w0.z, t_cnt = w0.w, t_pad = w1.x;
            const long long out_row =
                static_cast<long long>(static_cast<uint32_t>(w1.z)) | (static_cast<long long>(w1.w) << 32);
            const bool valid = row < q_cnt;
            const int na = valid ? __ldg(norms + q_row + row) : 0;
            const int n_tiles = (t_cnt + NT - 1) / NT;
            const int32_t* nbase = norms_pad + t_pad;
            const bool has_next = w + gridDim.x < n_work;
            int nq_row = 0, nq_cnt = 1;
            if (has_next) {
                const int4 n0 = __ldg(reinterpret_cast<const int4*>(work + w + gridDim.x));
                nq_row = n0.x;
                nq_cnt = n0.y;
            }
            const int pre_t = n_tiles - 1;              // the next work item's rows are fetched behind the last wait
            KeyFmt kf;
            kf.n_units = n_tiles * kUnitsPerTile;
            kf.ubits = max(32 - __clz(kf.n_units - 1), 8);
            kf.sh = kf.ubits > 8 ? kf.ubits - 8 : 0;   // m < 2^23: (23 - sh) + ubits + 1 <= 32 bits
            kf.mul = 2u << kf.ubits;
            kf.no_park = (probe_flags & 32u) != 0;
            // units this thread will see: 2 per full tile, fewer in the last, partial one
            const int rem = t_cnt - (n_tiles - 1) * NT - colbase;
            const int n_seen = (n_tiles - 1) * kUnitsPerTile + min(max((rem + kUnit - 1) / kUnit, 0), kUnitsPerTile);
            Trk sk{1, 0, 0, 0};   // empty keys: k1 owns slot 1, k2 slot 0 (the first unit is parked in slot 0)
            for (int t = 0; t < n_tiles; ++t) {
                if (has_next && t == pre_t) load_a_half(a_next, desc, nq_row, nq_cnt, row, half);
                SFM_PROF_T0();
                mbar_wait_suspend(&sm.t_full[s], sph, epi_wait_ns);
                SFM_PROF_ADD(9);
                if (prof_me) SFM_PROF_EV(2, prof_g);
                tc_fence_after();
                if (has_next && t == n_tiles - 1) {
                    // the last group that reads this block's A operand has completed: next rows in
                    tmem_st_32x16(a_taddr, a_next);
                    tmem_st_wait();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&sm.a_full[blk]);
                    SFM_PROF_ADD(12);
                }
                uint64_t* empty = &sm.t_empty[s];
                const uint32_t taddr = lane_base + s * NT + colbase;
                s += 2;
                if (s >= kAccStages) {
                    s -= kAccStages;
                    sph ^= 1;
                }
                if (MODE == 2) {   // timing probe: TMA + MMA pipeline alone
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(empty);
                    continue;
                }
                long long* ev3 = (prof_me && blockIdx.x == 0 && prof_g < kProfEvents)
                                     ? prof_out + gridDim.x * 16 + 3 * kProfEvents + prof_g : nullptr;
                if ((t + 1) * NT <= t_cnt) {
                    if (kf.sh == 0)
                        process_group<false, true, MODE>(sk, taddr, t, t_cnt, colbase, my_slots, kf, empty, lane, ev3, prof_base);
                    else
                        process_group<false, false, MODE>(sk, taddr, t, t_cnt, colbase, my_slots, kf, empty, lane, ev3, prof_base);
                } else {
                    process_group<true, false, MODE>(sk, taddr, t, t_cnt, colbase, my_slots, kf, empty, lane, ev3, prof_base);
                }
                SFM_PROF_ADD(10);
                if (prof_me) SFM_PROF_EV(4, prof_g);
                prof_g += 2;
            }
            SFM_PROF_T0();
            if (MODE != 0 || (probe_flags & 64u)) {
                if (valid && half == 0) out[out_row + row] = make_int4(static_cast<int>(sk.k1), static_cast<int>(sk.k2), na, 0);
                continue;
            }
            Top2K r = resolve_units(sk, n_seen, kf, colbase, my_slots, par_bits + (t_pad >> 5), t_cnt);
            // a dropped unit whose (truncated) maximum equals the second kept one may hold a candidate:
            // the best dropped one is recomputed here, a second one sends the row to knn_fix_kernel
            const uint32_t mf = kf.ubits + 1;
            r = recompute_unit(r, valid && n_seen >= 3 && (sk.dm >> mf) == (sk.k2 >> mf), q_row + row,
                               key_col(sk.dm, kf, colbase), desc, t_row, t_cnt, nbase);
            const bool tie = n_seen >= 4 && (sk.dd >> mf) == (sk.k2 >> mf);
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
            SFM_PROF_ADD(11);
        }
        if (PROF && warp == 0 && lane == 0) {
            prof_out[blockIdx.x * 16 + 8] = clock64() - prof_begin;
            prof_out[blockIdx.x * 16 + 9] = prof[9];
            prof_out[blockIdx.x * 16 + 10] = prof[10];
            prof_out[blockIdx.x * 16 + 11] = prof[11];
            prof_out[blockIdx.x * 16 + 12] = prof[12];
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
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

template <int MODE, bool PROF>
cudaError_t launch_ts_variant(const CUtensorMap& tmap, const CUtensorMap& tmap_x, const uint8_t* d_desc,
                              const uint8_t* d_bias32, const WorkItem* d_work, int64_t n_work, const int32_t* d_norms,
                              const int32_t* d_norms_pad, const uint32_t* d_par_bits, int32_t* d_flags, int force,
                              const KnnTuning& tn, int2* d_fix_list, int fix_cap, int32_t* d_knn, unsigned grid,
                              long long* d_prof, cudaStream_t st) {
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(knn2_ts_kernel<MODE, PROF>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(kSmemBytes));
        if (e != cudaSuccess) return e;
        configured = true;
    }
    knn2_ts_kernel<MODE, PROF><<<grid, kThreads, kSmemBytes, st>>>(
        tmap, tmap_x, d_desc, d_bias32, d_work, n_work, d_norms, d_norms_pad, d_par_bits, d_flags, force,
        tn.probe_flags, tn.wait_ns, tn.epi_wait_ns, d_fix_list, fix_cap, reinterpret_cast<int4*>(d_knn), d_prof);
    return cudaGetLastError();
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

void launch_parity_bits(const int32_t* d_norms_pad, int64_t pad_total, uint32_t* d_par_bits, cudaStream_t st,
                        int64_t* launches) {
    if (pad_total <= 0) return;
    parity_bits_kernel<<<static_cast<unsigned>((pad_total + 255) / 256), 256, 0, st>>>(d_norms_pad, pad_total,
                                                                                      d_par_bits);
    ++*launches;
}

// variant 0: the kNN (exits at once when the fold-range flag is raised, unless force), followed by the
// fix-up pass of knn_wide.cu; 1: max tree only, 2: no epilogue (timing probes).
int launch_knn2_ts(const CUtensorMap& tmap, const CUtensorMap& tmap_x, const uint8_t* d_desc, const uint8_t* d_bias32,
                   const WorkItem* d_work, int64_t n_work, const int32_t* d_norms, const int32_t* d_norms_pad,
                   const uint32_t* d_par_bits, int32_t* d_flags, int variant, int force, const KnnTuning& tn,
                   int32_t* d_fix_list, int32_t fix_cap, int32_t* d_knn, int num_sms, cudaStream_t st,
                   int64_t* launches, long long* d_prof) {
    if (n_work <= 0) return 0;
    const unsigned grid = static_cast<unsigned>(n_work < num_sms ? n_work : num_sms);
    int2* fl = reinterpret_cast<int2*>(d_fix_list);
    cudaError_t e = cudaMemsetAsync(d_flags + 2, 0, sizeof(int32_t), st);   // length of this launch's fix list
    if (e != cudaSuccess) return static_cast<int>(e);
#define SFM_TS_LAUNCH(M, P, F)                                                                                   \
    launch_ts_variant<M, P>(tmap, tmap_x, d_desc, d_bias32, d_work, n_work, d_norms, d_norms_pad, d_par_bits,    \
                            d_flags, F, tn, fl, fix_cap, d_knn, grid, d_prof, st)
    if (variant == 1)
        e = d_prof ? SFM_TS_LAUNCH(1, true, 1) : SFM_TS_LAUNCH(1, false, 1);
    else if (variant == 2)
        e = d_prof ? SFM_TS_LAUNCH(2, true, 1) : SFM_TS_LAUNCH(2, false, 1);
    else
        e = d_prof ? SFM_TS_LAUNCH(0, true, force) : SFM_TS_LAUNCH(0, false, force);
#undef SFM_TS_LAUNCH
    ++*launches;
    if (e == cudaSuccess && variant == 0 && !(tn.probe_flags & 4u))   // (probe bit 2: main kernel alone)
        return launch_knn_fix(d_work, n_work, d_desc, d_norms, d_flags, d_fix_list, fix_cap, d_knn, num_sms, st,
                              launches);
    return static_cast<int>(e);
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
