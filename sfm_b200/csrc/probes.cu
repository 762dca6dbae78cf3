// Micro-probes used to size the kNN epilogue (measurement helpers, not on the product path).
#include "kernels.h"
#include "ptx.cuh"

namespace sfm {
namespace {

__device__ __forceinline__ void tmem_ld_x64(uint32_t taddr) {
    uint32_t v[64];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
        "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
        "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]),
          "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]),
          "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]),
          "=r"(v[29]), "=r"(v[30]), "=r"(v[31]), "=r"(v[32]), "=r"(v[33]), "=r"(v[34]), "=r"(v[35]),
          "=r"(v[36]), "=r"(v[37]), "=r"(v[38]), "=r"(v[39]), "=r"(v[40]), "=r"(v[41]), "=r"(v[42]),
          "=r"(v[43]), "=r"(v[44]), "=r"(v[45]), "=r"(v[46]), "=r"(v[47]), "=r"(v[48]), "=r"(v[49]),
          "=r"(v[50]), "=r"(v[51]), "=r"(v[52]), "=r"(v[53]), "=r"(v[54]), "=r"(v[55]), "=r"(v[56]),
          "=r"(v[57]), "=r"(v[58]), "=r"(v[59]), "=r"(v[60]), "=r"(v[61]), "=r"(v[62]), "=r"(v[63])
        : "r"(taddr)
        : "memory");
}

__device__ __forceinline__ void tmem_ld_x32_discard(uint32_t taddr) {
    uint32_t v[32];
    tmem_ld_32x32(taddr, v);
}

// 16 lanes x 256 bits, x8: 32 registers per thread, covers 16 lanes x 64 columns.
__device__ __forceinline__ void tmem_ld_16x256_x8(uint32_t taddr) {
    uint32_t v[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x8.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
          "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
          "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]),
          "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]),
          "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}

// variant 0: 32x32b.x32, wait after each load        (4 KiB per warp-instruction)
//         1: 32x32b.x32, wait after every 4 loads
//         2: 32x32b.x64, wait after each load         (8 KiB)
//         3: 16x256b.x8, wait after every 4 loads     (4 KiB... 16 lanes x 64 columns)
//         4 / 5 / 6: 32x32b.x16 (2 KiB), wait after each / every second / every eighth load
__global__ void tmem_ld_probe_kernel(int iters, int variant, long long* out) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        tmem_alloc(&slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t base = slot + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        const uint32_t col = (i * 64) & 511;
        if (variant == 0) {
            tmem_ld_x32_discard(base + (col & 480));
            tmem_ld_wait();
        } else if (variant == 1) {
            tmem_ld_x32_discard(base + (col & 480));
            if ((i & 3) == 3) tmem_ld_wait();
        } else if (variant == 2) {
            tmem_ld_x64(base + (col & 448));
            tmem_ld_wait();
        } else if (variant == 3) {
            tmem_ld_16x256_x8(base + (col & 448));
            if ((i & 3) == 3) tmem_ld_wait();
        } else if (variant == 4) {   // the kNN epilogue's own shape: 32x32b.x16, wait after each
            uint32_t v[16];
            tmem_ld_32x16(base + (col & 496), v);
            tmem_ld_wait();
        } else if (variant == 5) {   // ... wait after every second load
            uint32_t v[16];
            tmem_ld_32x16(base + (col & 496), v);
            if (i & 1) tmem_ld_wait();
        } else {                     // ... wait after every eighth load
            uint32_t v[16];
            tmem_ld_32x16(base + (col & 496), v);
            if ((i & 7) == 7) tmem_ld_wait();
        }
    }
    tmem_ld_wait();
    const long long t1 = clock64();
    if ((threadIdx.x & 31) == 0) out[blockIdx.x * (blockDim.x >> 5) + warp] = t1 - t0;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(slot, 512);
    }
}

// MMA issue pattern of the kNN kernel without any data movement: `steps` steps of
// 2 accumulators x 5 tcgen05.mma (128 x N x 32, u8), two TMEM stages, issued warp-convergent
// by one elected lane (as the kNN kernel does).
//   flags bit 0: tcgen05.commit to a "full" barrier after every step
//         bit 1: consumer warps wait "full" and arrive "empty"; the issuer waits "empty"
//                before it reuses a stage (the accumulator hand-shake of the real kernel)
//         bit 2: one extra commit per step to a barrier nobody waits on (the B-slot release)
//         bit 3: no tcgen05.fence::after_thread_sync per step
//         bit 4: same A and B tile every time (else A differs per accumulator, B rotates)
//         bit 5: consumer warps spin on a barrier that never completes (shared-memory polling load)
//         bit 6: N = 256 (one accumulator of 256 columns per stage... two MMAs groups share A)
//         bit 7: N = 160, one accumulator per step in a ring of THREE stages (480 TMEM columns)
// blockDim = 32 * (1 + consumers); warp 0 issues.
template <int N>
__global__ void __launch_bounds__(320, 1) mma_handshake_probe_kernel(int steps, int flags) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    constexpr int kTile = 128 * 128;
    uint8_t* sa = base;              // 2 A tiles
    uint8_t* sb = base + 2 * kTile;  // 4 B tiles of 128 rows (2 of 256)
    uint64_t* full = reinterpret_cast<uint64_t*>(base + 6 * kTile);
    uint64_t* empty = full + 3;
    uint64_t* other = empty + 3;
    uint64_t* fin = other + 1;
    uint64_t* never = fin + 1;
    uint32_t* slot = reinterpret_cast<uint32_t*>(never + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int consumers = (blockDim.x >> 5) - 1;
    for (int i = threadIdx.x; i < 6 * kTile / 4; i += blockDim.x)
        reinterpret_cast<uint32_t*>(base)[i] = 0x01010101u * (i & 3);
    fence_proxy_async();
    if (threadIdx.x == 0) {
        for (int i = 0; i < 3; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], consumers > 0 ? consumers : 1);
        }
        mbar_init(other, 1);
        mbar_init(fin, 1);
        mbar_init(never, 1);
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
    constexpr uint32_t idesc = make_idesc_u8(128, N);
    constexpr int kAcc = N == 160 ? 1 : 256 / N;   // accumulators per stage
    constexpr int kStages = N == 160 ? 3 : 2;
    constexpr int kStageCols = N == 160 ? 160 : 256;
    if (warp == 0) {
        uint32_t phase = 0;
        int stage = 0;
        for (int st = 0; st < steps; ++st) {
            if (flags & 2) mbar_wait(&empty[stage], phase ^ 1);
            if (!(flags & 8)) tc_fence_after();
            const int bsel = (flags & 16) ? 0 : (N == 128 ? (st & 3) : (st & 1) * 2);   // N = 160: 160 rows of a 256-row region
            const uint64_t bd = make_kmajor_sw128_desc(smem_u32(sb + bsel * kTile));
            const uint64_t ad0 = make_kmajor_sw128_desc(smem_u32(sa));
            const uint64_t astep = (flags & 16) ? 0 : (kTile >> 4);
            const uint32_t d0 = tmem_base + stage * kStageCols;
            if (elect_one()) {
#pragma unroll
                for (int a = 0; a < kAcc; ++a) {
#pragma unroll
                    for (int k = 0; k < 5; ++k)
                        mma_i8_ss(d0 + a * N, ad0 + a * astep + uint64_t(k & 3) * 2, bd + uint64_t(k & 3) * 2, idesc,
                                  k > 0);
                }
                if (flags & 4) mma_commit(other);
                if (flags & 1) mma_commit(&full[stage]);
            }
            __syncwarp();
            if (++stage == kStages) {
                stage = 0;
                phase ^= 1;
            }
        }
        if (elect_one()) mma_commit(fin);
        __syncwarp();
        mbar_wait(fin, 0);
        if ((flags & 32) && elect_one()) mbar_arrive(never);   // release the spinners
    } else if (flags & 2) {
        uint32_t phase = 0;
        int stage = 0;
        for (int st = 0; st < steps; ++st) {
            mbar_wait(&full[stage], phase);
            tc_fence_after();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[stage]);
            if (++stage == kStages) {
                stage = 0;
                phase ^= 1;
            }
        }
    } else if (flags & 32) {
        mbar_wait(never, 0);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace

int launch_mma_handshake_probe(int steps, int k_per_acc, int flags, int consumers, int num_sms,
                               cudaStream_t st, int64_t* launches) {
    (void)k_per_acc;   // fixed at 5 (the folded kNN K-loop)
    const size_t smem = 6 * 128 * 128 + 128 + 1024;
    cudaError_t e = cudaFuncSetAttribute(mma_handshake_probe_kernel<128>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(mma_handshake_probe_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 static_cast<int>(smem));
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(mma_handshake_probe_kernel<160>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    if (flags & 128)
        mma_handshake_probe_kernel<160><<<num_sms, 32 * (1 + consumers), smem, st>>>(steps, flags);
    else if (flags & 64)
        mma_handshake_probe_kernel<256><<<num_sms, 32 * (1 + consumers), smem, st>>>(steps, flags);
    else
        mma_handshake_probe_kernel<128><<<num_sms, 32 * (1 + consumers), smem, st>>>(steps, flags);
    ++*launches;
    return static_cast<int>(cudaGetLastError());
}

int launch_tmem_ld_probe(int warps, int variant, int iters, int num_sms, long long* d_out, cudaStream_t st,
                         int64_t* launches) {
    tmem_ld_probe_kernel<<<num_sms, warps * 32, 0, st>>>(iters, variant, d_out);
    ++*launches;
    return static_cast<int>(cudaGetLastError());
}

namespace {
// FP64 instruction-rate probe: eight independent DFMA chains per thread, 1024 threads per SM x 2 CTAs.
// The radius matcher (radius.cu) is bound by the FP64 pipe's ISSUE rate (its DADD / DMUL count one each).
__global__ void __launch_bounds__(512) fp64_probe_kernel(int iters, double seed, double* __restrict__ out) {
    double x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = seed + i + threadIdx.x;
    const double a = 1.0000001, b = 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = fma(x[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += x[i];
    if (s == 12345.678) out[0] = s;   // keeps the chains live
}
}  // namespace

int launch_fp64_probe(int iters, int num_sms, double* d_out, cudaStream_t st, int64_t* launches) {
    fp64_probe_kernel<<<num_sms * 4, 512, 0, st>>>(iters, 0.5, d_out);
    ++*launches;
    return static_cast<int>(cudaGetLastError());
}

}  // namespace sfm
