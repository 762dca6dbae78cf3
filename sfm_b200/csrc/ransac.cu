// K3 + K4: batched RANSAC fundamental matrix and geometric verification.
//
// Replaces, per image pair (reference utils/find_matches.py):
//   :278-279  point gather from the keypoints
//   :282      cv2.findFundamentalMat(pts1, pts2, cv2.FM_RANSAC, 3.0)
//   :157-201  ImageMatcher.geometric_verification
//   :203-214  ImageMatcher.verify_match_quality
//
// One CTA per pair.  Correspondences are staged in shared memory (SoA float4
// per match, 2048 at a time); every warp owns a batch of 32 hypotheses, one
// per lane: the lane draws its 8-sample with a counter-based RNG, solves the
// normalised 8-point system in registers and then walks the staged points
// (shared-memory broadcast reads) counting Sampson inliers.  After each round
// of 8 x 32 hypotheses the block keeps the best model and re-evaluates the
// adaptive stopping rule.  The best model is polished by least-squares refits
// on its inliers (float64 normal equations, block-reduced), and the final F is
// pushed through the reference's verification arithmetic: epilines in float64
// rounded to float32 exactly as cv2.computeCorrespondEpilines does, then the
// reference's float32 NumPy expressions without FMA contraction, so that for
// an identical F the inlier mask is bit-identical to the reference's.
#include <cfloat>

#include "kernels.h"
#include "ransac_core.h"

namespace sfm {

namespace {

// 128 threads, FOUR CTAs per SM (launch bound: 128 registers, 268 bytes of spills): the per-pair phases are
// latency chains and residency hides them -- measured on the 200 x 8k workload: 3 CTAs (166 registers, no
// spill) 15.6 ms, 4 CTAs 14.1 ms, 5 CTAs (96 registers, 840 bytes of spills) 15.7 ms.
constexpr int kThreads = 128;
constexpr int kWarps = kThreads / 32;
constexpr int kChunk = 2048;   // staged correspondences per pass (32 KiB)
constexpr int kMinSample = 8;

struct Shared {
    float4 pts[kChunk];
    double red[kWarps][48];
    double res[48];
    double jA[9][9], jV[9][9];   // warp-cooperative Jacobi of the refit (warp_refit_from_normal)
    double Fd[9];
    float bestF[9];
    float warpF[kWarps][9];
    int warp_cnt[kWarps];
    float warp_cost[kWarps];
    int warp_hyp[kWarps];
    int best_cnt;
    float best_cost;
    int need;
    int flag;
    float np_res[6];   // NumPy-exact float32 gate quantities of a borderline pair: std x1, y1, x2, y2, mean inlier error, mean error
};

template <int N>
__device__ void block_sum(double (&v)[N], Shared& sh) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < N; ++i) {
        double x = v[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0) sh.red[warp][i] = x;
    }
    __syncthreads();
    if (threadIdx.x < N) {
        double s = 0;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) s += sh.red[w][threadIdx.x];
        sh.res[threadIdx.x] = s;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = sh.res[i];
    __syncthreads();
}

// ---- the reference's NumPy float32 arithmetic, restated instruction for instruction -------------------------
// verify_match_quality's gates read np.std(pts[inlier_mask], axis=0) and np.mean(symmetric_errors[inlier_mask])
// (utils/find_matches.py:176-186): float32 arrays, so NumPy sums in float32 -- SEQUENTIALLY along axis 0 of the
// (I, 2) point arrays (the reduction axis is the outer one) and PAIRWISE along the contiguous 1-D error array
// (eight accumulators per block of <= 128, halves split at multiples of 8).  A pair whose spread or error sits
// within float32 rounding of a gate would otherwise flip against the reference.  One thread per quantity walks the
// pair's matches in order; no FMA contraction (__f*_rn).
// np.std of coordinate `k` (0..3 = x1, y1, x2, y2) over the inliers.
// np.std of two coordinates (src[stride * i], src[stride * i + 1] of match i) over the inliers; err[i] is the
// match's symmetric error (inlier <=> err[i] < thr, the comparison that made the mask).  Two independent chains per
// loop.  Shared or global memory (generic pointers).
__device__ __noinline__ float2 np_std2_f32(const float* src, int stride, const float* err, float thr, int m, int n_in) {
    // (loads outside the condition: they are independent of the running sums and pipeline)
    float s0 = 0.f, s1 = 0.f;
#pragma unroll 8
    for (int i = 0; i < m; ++i) {
        const float e = err[i], x0 = src[stride * i], x1 = src[stride * i + 1];
        if (e < thr) {
            s0 = __fadd_rn(s0, x0);
            s1 = __fadd_rn(s1, x1);
        }
    }
    const float fn = static_cast<float>(n_in);
    const float m0 = __fdiv_rn(s0, fn), m1 = __fdiv_rn(s1, fn);
    float q0 = 0.f, q1 = 0.f;
#pragma unroll 8
    for (int i = 0; i < m; ++i) {
        const float e = err[i];
        const float d0 = __fsub_rn(src[stride * i], m0), d1 = __fsub_rn(src[stride * i + 1], m1);
        const float p0 = __fmul_rn(d0, d0), p1 = __fmul_rn(d1, d1);
        if (e < thr) {
            q0 = __fadd_rn(q0, p0);
            q1 = __fadd_rn(q1, p1);
        }
    }
    return make_float2(__fsqrt_rn(__fdiv_rn(q0, fn)), __fsqrt_rn(__fdiv_rn(q1, fn)));
}

// Cursor over the (masked) error array: the next selected value, in index order.
struct ErrCursor {
    const float* v;
    float thr;
    bool masked;   // only the values below thr (the inliers); else every element
    int i;
    __device__ __forceinline__ float next() {
        if (masked)
            while (!(v[i] < thr)) ++i;
        return v[i++];
    }
};

// NumPy's pairwise_sum over the next n values of the cursor (numpy/core/src/umath/loops_utils.h.src): a block of
// at most 128 values.
__device__ __forceinline__ float np_pairwise_leaf_f32(ErrCursor& c, int n) {
    if (n < 8) {
        float res = 0.f;
        for (int i = 0; i < n; ++i) res = __fadd_rn(res, c.next());
        return res;
    }
    float r[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) r[k] = c.next();
    int i = 8;
    for (; i < n - (n % 8); i += 8) {
#pragma unroll
        for (int k = 0; k < 8; ++k) r[k] = __fadd_rn(r[k], c.next());
    }
    float res = __fadd_rn(__fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])),
                          __fadd_rn(__fadd_rn(r[4], r[5]), __fadd_rn(r[6], r[7])));
    for (; i < n; ++i) res = __fadd_rn(res, c.next());
    return res;
}

// ... and its recursion (n > 128: the halves split at a multiple of 8, left half first), with an explicit stack:
// device threads have 1 KiB of call stack by default.
__device__ __forceinline__ float np_pairwise_f32(ErrCursor& c, int n) {   // inlined: the cursor stays in registers
    constexpr int kDepth = 28;   // n halves every level
    int ns[kDepth];
    float left[kDepth];
    unsigned char stage[kDepth];
    int sp = 0;
    ns[0] = n;
    stage[0] = 0;
    float ret = 0.f;
    while (true) {
        if (stage[sp] == 0) {
            if (ns[sp] <= 128) {
                ret = np_pairwise_leaf_f32(c, ns[sp]);
                if (--sp < 0) break;
                continue;
            }
            int n2 = ns[sp] / 2;
            n2 -= n2 % 8;
            stage[sp] = 1;
            ns[sp + 1] = n2;
            stage[sp + 1] = 0;
            ++sp;
        } else if (stage[sp] == 1) {
            left[sp] = ret;
            int n2 = ns[sp] / 2;
            n2 -= n2 % 8;
            stage[sp] = 2;
            ns[sp + 1] = ns[sp] - n2;
            stage[sp + 1] = 0;
            ++sp;
        } else {
            ret = __fadd_rn(left[sp], ret);
            if (--sp < 0) break;
        }
    }
    return ret;
}

// np.mean of the selected errors (float32 sum / count).
__device__ __noinline__ float np_mean_f32(const float* v, float thr, bool masked, int n_sel) {
    ErrCursor c{v, thr, masked, 0};
    return __fdiv_rn(np_pairwise_f32(c, n_sel), static_cast<float>(n_sel));
}

__device__ __forceinline__ float4 load_match(const float* __restrict__ p1,
                                             const float* __restrict__ p2, long long i) {
    const float2 a = __ldg(reinterpret_cast<const float2*>(p1) + i);
    const float2 b = __ldg(reinterpret_cast<const float2*>(p2) + i);
    return make_float4(a.x, a.y, b.x, b.y);
}

// The reference's symmetric epipolar error of one correspondence, bit for bit
// (oracle/fundamental.py::_epilines + _verification_from_lines).
__device__ __forceinline__ float symmetric_error(const double (&F)[9], float x1, float y1, float x2,
                                                 float y2) {
    auto line = [](double f0, double f1, double f2, double f3, double f4, double f5, double f6,
                   double f7, double f8, float x, float y, float& la, float& lb, float& lc) {
        const double xd = x, yd = y;
        const double a = __dadd_rn(__dadd_rn(__dmul_rn(f0, xd), __dmul_rn(f1, yd)), f2);
        const double b = __dadd_rn(__dadd_rn(__dmul_rn(f3, xd), __dmul_rn(f4, yd)), f5);
        const double c = __dadd_rn(__dadd_rn(__dmul_rn(f6, xd), __dmul_rn(f7, yd)), f8);
        double nu = __dadd_rn(__dmul_rn(a, a), __dmul_rn(b, b));
        nu = nu != 0.0 ? __ddiv_rn(1.0, __dsqrt_rn(nu)) : 1.0;
        la = static_cast<float>(__dmul_rn(a, nu));
        lb = static_cast<float>(__dmul_rn(b, nu));
        lc = static_cast<float>(__dmul_rn(c, nu));
    };
    float a1, b1, c1, a2, b2, c2;
    // lines1: epilines in image 1 of the image-2 points, F^T x2   (:160)
    line(F[0], F[3], F[6], F[1], F[4], F[7], F[2], F[5], F[8], x2, y2, a1, b1, c1);
    // lines2: epilines in image 2 of the image-1 points, F x1     (:162)
    line(F[0], F[1], F[2], F[3], F[4], F[5], F[6], F[7], F[8], x1, y1, a2, b2, c2);
    const float e1 = __fdiv_rn(fabsf(__fadd_rn(__fadd_rn(__fmul_rn(x1, a1), __fmul_rn(y1, b1)), c1)),
                               __fsqrt_rn(__fadd_rn(__fmul_rn(a1, a1), __fmul_rn(b1, b1))));
    const float e2 = __fdiv_rn(fabsf(__fadd_rn(__fadd_rn(__fmul_rn(x2, a2), __fmul_rn(y2, b2)), c2)),
                               __fsqrt_rn(__fadd_rn(__fmul_rn(a2, a2), __fmul_rn(b2, b2))));
    return __fmul_rn(__fadd_rn(e1, e2), 0.5f);
}

// (c, s) of the Jacobi rotation that annihilates a_pq, t = sgn(theta) / (|theta| + sqrt(theta^2 + 1)) with
// theta = (a_qq - a_pp) / (2 a_pq), written with one square root, one division and one reciprocal square
// root: t = +-2 a_pq / (|d| + sqrt(d^2 + 4 a_pq^2)), c = rsqrt(t^2 + 1).  (The textbook form costs three
// divisions and two square roots, each a dependent chain of ~200-300 cycles in double precision; the
// eigen-solves of the refits are pure latency.)
__device__ __forceinline__ void jacobi_cs(double app, double aqq, double apq, double& c, double& s) {
    c = 1.0;
    s = 0.0;
    if (apq == 0.0) return;
    const double d = aqq - app;
    const double h = fabs(d) + sqrt(d * d + 4.0 * apq * apq);
    if (!(h > 0.0)) return;                      // NaN, or d = 0 with a_pq^2 underflowing: nothing to rotate
    double t = (apq + apq) / h;
    if (d < 0.0) t = -t;
    c = rsqrt(t * t + 1.0);
    s = t * c;
}

// rc::smallest_eigvec3<double> with the rotation above and a convergence test (the generic version always
// runs 8 sweeps = 24 rotations).
__device__ void smallest_eigvec3_d(double a00, double a01, double a02, double a11, double a12, double a22,
                                   double (&v)[3]) {
    double A[3][3] = {{a00, a01, a02}, {a01, a11, a12}, {a02, a12, a22}};
    double V[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
#pragma unroll 1
    for (int sweep = 0; sweep < 10; ++sweep) {
        const double off = A[0][1] * A[0][1] + A[0][2] * A[0][2] + A[1][2] * A[1][2];
        const double diag = A[0][0] * A[0][0] + A[1][1] * A[1][1] + A[2][2] * A[2][2];
        if (off <= 1e-32 * diag) break;
#pragma unroll
        for (int pq = 0; pq < 3; ++pq) {
            const int p = pq == 2 ? 1 : 0;
            const int q = pq == 0 ? 1 : 2;
            double c, s;
            jacobi_cs(A[p][p], A[q][q], A[p][q], c, s);
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const double akp = A[k][p], akq = A[k][q];
                A[k][p] = c * akp - s * akq;
                A[k][q] = s * akp + c * akq;
            }
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const double apk = A[p][k], aqk = A[q][k];
                A[p][k] = c * apk - s * aqk;
                A[q][k] = s * apk + c * aqk;
            }
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const double vkp = V[k][p], vkq = V[k][q];
                V[k][p] = c * vkp - s * vkq;
                V[k][q] = s * vkp + c * vkq;
            }
        }
    }
    int m = 0;
    if (A[1][1] < A[m][m]) m = 1;
    if (A[2][2] < A[m][m]) m = 2;
    v[0] = m == 0 ? V[0][0] : (m == 1 ? V[0][1] : V[0][2]);
    v[1] = m == 0 ? V[1][0] : (m == 1 ? V[1][1] : V[1][2]);
    v[2] = m == 0 ? V[2][0] : (m == 1 ? V[2][1] : V[2][2]);
}

// rc::enforce_rank2<double> on the eigen-solver above: F - (F v3) v3^T.
__device__ void enforce_rank2_d(double (&F)[9]) {
    double g[6];
    g[0] = F[0] * F[0] + F[3] * F[3] + F[6] * F[6];
    g[1] = F[0] * F[1] + F[3] * F[4] + F[6] * F[7];
    g[2] = F[0] * F[2] + F[3] * F[5] + F[6] * F[8];
    g[3] = F[1] * F[1] + F[4] * F[4] + F[7] * F[7];
    g[4] = F[1] * F[2] + F[4] * F[5] + F[7] * F[8];
    g[5] = F[2] * F[2] + F[5] * F[5] + F[8] * F[8];
    double v[3];
    smallest_eigvec3_d(g[0], g[1], g[2], g[3], g[4], g[5], v);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const double fv = F[3 * r] * v[0] + F[3 * r + 1] * v[1] + F[3 * r + 2] * v[2];
        F[3 * r] -= fv * v[0];
        F[3 * r + 1] -= fv * v[1];
        F[3 * r + 2] -= fv * v[2];
    }
}

// refit_from_normal (ransac_core.h) with the 9x9 Jacobi eigen-solve run by one warp in PARALLEL ORDER:
// a sweep is nine rounds of four rotations in disjoint planes (round-robin pairing of the nine indices,
// index r sits out in round r), the four (c, s) pairs are computed side by side by the lanes' groups
// (lane & 3) and applied together: first the column pairs of A and V (36 + 36 element pairs over the
// lanes), then the row pairs of A.  Rotations in disjoint planes commute, so this is the cyclic Jacobi
// method with another visiting order: the scalar chain (sqrt, division, rsqrt: ~700 cycles) is paid 9
// times per sweep instead of 36.  Round 1 ran the 36 rotations of a sweep one after the other: with two
// refits per pair that serial eigen-solve was half of the kernel (12.4 ms with, 5.8 ms without refits
// on the 19,900 pairs of the 200 x 8k set).  Called by warp 0 only; result in F (all lanes).
__device__ bool warp_refit_from_normal(const double (&ata)[45], double c1x, double c1y, double s1, double c2x,
                                       double c2y, double s2, Shared& sh, double (&F)[9]) {
    const int lane = threadIdx.x & 31;
    for (int idx = lane; idx < 81; idx += 32) {
        const int i = idx / 9, j = idx % 9;
        const int r = i < j ? i : j, c = i < j ? j : i;
        sh.jA[i][j] = ata[r * 9 - r * (r - 1) / 2 + (c - r)];
        sh.jV[i][j] = i == j ? 1.0 : 0.0;
    }
    __syncwarp();
    const int g = lane & 3, k0 = lane >> 2;
    for (int sweep = 0; sweep < 12; ++sweep) {
        // squared norms of the strict upper triangle and of the diagonal: row `lane`, then a butterfly
        double off = 0, diag = 0;
        if (lane < 9) {
            diag = sh.jA[lane][lane] * sh.jA[lane][lane];
            for (int j = lane + 1; j < 9; ++j) off += sh.jA[lane][j] * sh.jA[lane][j];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            off += __shfl_xor_sync(0xffffffffu, off, o);
            diag += __shfl_xor_sync(0xffffffffu, diag, o);
        }
        if (off <= 1e-30 * diag) break;
        for (int r = 0; r < 9; ++r) {
            int a = r + 1 + g, b = r + 8 - g;          // (r + i) mod 9 and (r - i) mod 9, i = g + 1
            a = a >= 9 ? a - 9 : a;
            b = b >= 9 ? b - 9 : b;
            const int p = a < b ? a : b, q = a < b ? b : a;
            double c, s;
            jacobi_cs(sh.jA[p][p], sh.jA[q][q], sh.jA[p][q], c, s);
            __syncwarp();
            for (int k = k0; k < 9; k += 8) {           // A <- A J, V <- V J: columns p, q of row k
                const double akp = sh.jA[k][p], akq = sh.jA[k][q];
                sh.jA[k][p] = c * akp - s * akq;
                sh.jA[k][q] = s * akp + c * akq;
                const double vkp = sh.jV[k][p], vkq = sh.jV[k][q];
                sh.jV[k][p] = c * vkp - s * vkq;
                sh.jV[k][q] = s * vkp + c * vkq;
            }
            __syncwarp();
            for (int k = k0; k < 9; k += 8) {           // A <- J^T A: rows p, q of column k
                const double apk = sh.jA[p][k], aqk = sh.jA[q][k];
                sh.jA[p][k] = c * apk - s * aqk;
                sh.jA[q][k] = s * apk + c * aqk;
            }
            __syncwarp();
        }
    }
    int m = 0;
    for (int i = 1; i < 9; ++i)
        if (sh.jA[i][i] < sh.jA[m][m]) m = i;
    for (int i = 0; i < 9; ++i) F[i] = sh.jV[i][m];
    enforce_rank2_d(F);
    rc::denormalise<double>(F, c1x, c1y, s1, c2x, c2y, s2);
    rc::scale_unit<double>(F);
    bool ok = true;
    for (int i = 0; i < 9; ++i) ok = ok && (fabs(F[i]) <= 1.0);
    return ok;
}

// A skipped pair leaves defined outputs behind (F = 0, no inlier, errors 0): what a caller packs with
// only_ok = 0 is then a pure function of the pair, whichever rank or chunk computed it.
__device__ void write_skip(int p, int m, int status, long long base, uint8_t* mask, float* sym, double* F_out,
                           sfm_pair_metrics* metrics, int32_t* status_out) {
    if (threadIdx.x < 9 && F_out) F_out[9ll * p + threadIdx.x] = 0.0;
    for (int i = threadIdx.x; i < m; i += kThreads) {
        mask[base + i] = 0;
        sym[base + i] = 0.f;
    }
    if (threadIdx.x == 0) {
        sfm_pair_metrics mt;
        mt.total_matches = m;
        mt.inliers = 0;
        mt.inlier_ratio = 0.f;
        mt.reprojection_error = __int_as_float(0x7f800000);
        mt.symmetric_error = 0.f;
        mt.well_distributed = 0;
        mt.ransac_inliers = 0;
        mt.ransac_iters = 0;
        metrics[p] = mt;
        status_out[p] = status;
    }
}

__global__ void __launch_bounds__(kThreads, 4)
ransac_verify_kernel(int n_pairs, const long long* __restrict__ off, const int32_t* __restrict__ cnt,
                     const int32_t* __restrict__ pair_ids, const float* __restrict__ pts1,
                     const float* __restrict__ pts2, const double* __restrict__ F_in,
                     sfm_verify_params prm, double* __restrict__ F_out, uint8_t* __restrict__ mask,
                     float* __restrict__ sym, sfm_pair_metrics* __restrict__ metrics,
                     int32_t* __restrict__ status_out) {
    __shared__ Shared sh;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float thr2 = static_cast<float>(prm.ransac_thresh * prm.ransac_thresh);
    const int max_iters = (prm.max_iters + 31) & ~31;
    const bool hyp_sym = (prm.score_mode & 1) != 0, lo_sym = (prm.score_mode & 2) != 0;

    for (int p = blockIdx.x; p < n_pairs; p += gridDim.x) {
        __syncthreads();
        const long long base = off[p];
        const int m = cnt ? cnt[p] : static_cast<int>(off[p + 1] - base);
        const float* p1 = pts1 + 2 * base;
        const float* p2 = pts2 + 2 * base;
        int ransac_inl = 0, ransac_it = 0;

        if (F_in) {
            if (threadIdx.x < 9) sh.Fd[threadIdx.x] = F_in[9ll * p + threadIdx.x];
            __syncthreads();
        } else {
            if (m < prm.min_matches) {
                write_skip(p, m, SFM_PAIR_TOO_FEW_MATCHES, base, mask, sym, F_out, metrics, status_out);
                continue;
            }
            if (m < kMinSample) {
                write_skip(p, m, SFM_PAIR_NO_MODEL, base, mask, sym, F_out, metrics, status_out);
                continue;
            }
            const uint64_t key = pair_ids ? rc::pair_key(prm.seed, pair_ids[2 * p], pair_ids[2 * p + 1])
                                          : rc::pair_key(prm.seed, static_cast<uint32_t>(p), 0xffffffffu);
            // Small pairs are cheap and their first rounds easily miss a good model: never stop before
            // hypotheses x matches >= min_work.  The floor is for bunny-scale pairs (m <= 256, where the 148
            // golden pairs need it to stay inside the SURVEY 8c envelope: with min_work = 2^16 no bound or
            // Jaccard failure and at most 2 boundary gate flips over three seeds, the same as with 2^18, while
            // without it up to 6 flips and a Jaccard failure -- tests/ransac_tune.py); it tapers to nothing at m = 384:
            // on 147 reference-passing synthetic 8k-config pairs (median 412 matches) the envelope against
            // cv2.findFundamentalMat is the same with and without it (Jaccard median 0.970, p10 0.908, no
            // bound failure, no gate flip; tests/ransac_synth_envelope.py), at 0.65 instead of 1.55 us per pair.
            // (Fewer matches than the inlier gate asks for cannot pass it whatever F is: no floor either.)
            int floor_h = 0;
            if (prm.min_work > 0 && m < 384 && m >= prm.min_inliers) {
                const long long w = m <= 256 ? prm.min_work : static_cast<long long>(prm.min_work) * (384 - m) / 128;
                floor_h = static_cast<int>(((w + m - 1) / m + 255) & ~255ll);
            }
            floor_h = floor_h > 4096 ? 4096 : floor_h;
            const int limit = max_iters > floor_h ? max_iters : floor_h;
            if (threadIdx.x == 0) {
                sh.best_cnt = -1;
                sh.best_cost = FLT_MAX;
                sh.need = limit;
            }
            const bool resident = m <= kChunk;
            if (resident)
                for (int i = threadIdx.x; i < m; i += kThreads) sh.pts[i] = load_match(p1, p2, i);
            __syncthreads();

            // ------------------------------------------------ hypothesis rounds
            int done = 0;
            while (done < sh.need) {
                const int hyp = done + warp * 32 + lane;
                float F[9];
                bool ok = false;
                if (hyp < limit) {
                    uint32_t idx[8];
                    rc::sample8(key, hyp, m, idx);
                    float s[8][4];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float4 q = load_match(p1, p2, idx[j]);
                        s[j][0] = q.x; s[j][1] = q.y; s[j][2] = q.z; s[j][3] = q.w;
                    }
                    ok = rc::solve8(s, F);
                }
                int c = 0;
                float cost = 0.f;
                for (int c0 = 0; c0 < m; c0 += kChunk) {
                    const int n = min(kChunk, m - c0);
                    if (!resident) {
                        __syncthreads();
                        for (int i = threadIdx.x; i < n; i += kThreads)
                            sh.pts[i] = load_match(p1, p2, c0 + i);
                        __syncthreads();
                    }
                    if (ok) {
#pragma unroll 4
                        for (int i = 0; i < n; ++i) {
                            const float4 q = sh.pts[i];
                            const float e = rc::residual_sq<float>(F, q.x, q.y, q.z, q.w, hyp_sym);
                            c += e < thr2;
                            cost += fminf(e, thr2);
                        }
                    }
                }
                if (!ok) { c = -1; cost = FLT_MAX; }
                // warp arg-best: more inliers, then lower truncated cost, then lower id
                int bc = c, bh = hyp;
                float bcost = cost;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
                    const float ocost = __shfl_xor_sync(0xffffffffu, bcost, o);
                    const int oh = __shfl_xor_sync(0xffffffffu, bh, o);
                    const bool better = oc > bc || (oc == bc && (ocost < bcost || (ocost == bcost && oh < bh)));
                    if (better) { bc = oc; bcost = ocost; bh = oh; }
                }
                if (bh == hyp) {
                    sh.warp_cnt[warp] = bc;
                    sh.warp_cost[warp] = bcost;
                    sh.warp_hyp[warp] = bh;
#pragma unroll
                    for (int i = 0; i < 9; ++i) sh.warpF[warp][i] = ok ? F[i] : 0.f;
                }
                __syncthreads();
                if (threadIdx.x == 0) {
                    int bw = -1;
                    int cc = sh.best_cnt;
                    float cs = sh.best_cost;
                    for (int w = 0; w < kWarps; ++w) {
                        if (sh.warp_cnt[w] > cc || (sh.warp_cnt[w] == cc && sh.warp_cost[w] < cs)) {
                            cc = sh.warp_cnt[w];
                            cs = sh.warp_cost[w];
                            bw = w;
                        }
                    }
                    if (bw >= 0) {
                        sh.best_cnt = cc;
                        sh.best_cost = cs;
                        for (int i = 0; i < 9; ++i) sh.bestF[i] = sh.warpF[bw][i];
                    }
                    const int need = rc::iters_needed(sh.best_cnt, m, prm.confidence, max_iters);
                    sh.need = need > floor_h ? need : floor_h;
                }
                done += kWarps * 32;
                __syncthreads();
            }
            ransac_it = done < limit ? done : limit;
            if (sh.best_cnt < kMinSample) {
                write_skip(p, m, SFM_PAIR_NO_MODEL, base, mask, sym, F_out, metrics, status_out);
                continue;
            }
            ransac_inl = sh.best_cnt;
            if (threadIdx.x < 9) sh.Fd[threadIdx.x] = static_cast<double>(sh.bestF[threadIdx.x]);
            __syncthreads();

            // ------------------------------------------------ inlier refits
            for (int it = 0; it < prm.lo_iters; ++it) {
                float Ff[9];
#pragma unroll
                for (int i = 0; i < 9; ++i) Ff[i] = static_cast<float>(sh.Fd[i]);
                double mom[5] = {0, 0, 0, 0, 0};
                for (int i = threadIdx.x; i < m; i += kThreads) {
                    const float4 q = load_match(p1, p2, i);
                    if (rc::residual_sq<float>(Ff, q.x, q.y, q.z, q.w, lo_sym) < thr2) {
                        mom[0] += 1.0; mom[1] += q.x; mom[2] += q.y; mom[3] += q.z; mom[4] += q.w;
                    }
                }
                block_sum<5>(mom, sh);
                const double n_in = mom[0];
                if (n_in < 8.0) break;
                const double c1x = mom[1] / n_in, c1y = mom[2] / n_in;
                const double c2x = mom[3] / n_in, c2y = mom[4] / n_in;
                double acc[47];
#pragma unroll
                for (int i = 0; i < 47; ++i) acc[i] = 0.0;
                for (int i = threadIdx.x; i < m; i += kThreads) {
                    const float4 q = load_match(p1, p2, i);
                    if (rc::residual_sq<float>(Ff, q.x, q.y, q.z, q.w, lo_sym) < thr2) {
                        const double x1 = q.x - c1x, y1 = q.y - c1y, x2 = q.z - c2x, y2 = q.w - c2y;
                        acc[45] += sqrt(x1 * x1 + y1 * y1);
                        acc[46] += sqrt(x2 * x2 + y2 * y2);
                        double a[9];
                        rc::design_row(x1, y1, x2, y2, a);
                        int k = 0;
#pragma unroll
                        for (int r = 0; r < 9; ++r)
#pragma unroll
                            for (int c = r; c < 9; ++c) acc[k++] += a[r] * a[c];
                    }
                }
                block_sum<47>(acc, sh);
                if (threadIdx.x == 0) sh.flag = 0;
                if (warp == 0 && acc[45] > 1e-9 && acc[46] > 1e-9) {
                    const double s1 = 1.4142135623730951 * n_in / acc[45];
                    const double s2 = 1.4142135623730951 * n_in / acc[46];
                    const double wgt[9] = {s1 * s2, s1 * s2, s2, s1 * s2, s1 * s2, s2, s1, s1, 1.0};
                    double ata[45];
                    int k = 0;
#pragma unroll
                    for (int r = 0; r < 9; ++r)
#pragma unroll
                        for (int c = r; c < 9; ++c) {
                            ata[k] = acc[k] * wgt[r] * wgt[c];
                            ++k;
                        }
                    double Fn[9];
                    const bool ok = warp_refit_from_normal(ata, c1x, c1y, s1, c2x, c2y, s2, sh, Fn);
                    if (lane == 0 && ok) {
                        for (int i = 0; i < 9; ++i) sh.res[i] = Fn[i];
                        sh.flag = 1;
                    }
                }
                __syncthreads();
                if (!sh.flag) break;
                float Fn[9];
#pragma unroll
                for (int i = 0; i < 9; ++i) Fn[i] = static_cast<float>(sh.res[i]);
                __syncthreads();
                double sc[2] = {0, 0};
                for (int i = threadIdx.x; i < m; i += kThreads) {
                    const float4 q = load_match(p1, p2, i);
                    const float e = rc::residual_sq<float>(Fn, q.x, q.y, q.z, q.w, lo_sym);
                    sc[0] += e < thr2 ? 1.0 : 0.0;
                    sc[1] += fminf(e, thr2);
                }
                block_sum<2>(sc, sh);
                const int new_cnt = static_cast<int>(sc[0]);
                const bool accept = new_cnt > sh.best_cnt ||
                                    (new_cnt == sh.best_cnt && static_cast<float>(sc[1]) < sh.best_cost);
                __syncthreads();
                if (!accept) break;
                if (threadIdx.x < 9) sh.Fd[threadIdx.x] = static_cast<double>(Fn[threadIdx.x]);
                if (threadIdx.x == 0) {
                    sh.best_cnt = new_cnt;
                    sh.best_cost = static_cast<float>(sc[1]);
                }
                __syncthreads();
            }
            ransac_inl = sh.best_cnt;
            if (threadIdx.x == 0) {
                double Fn[9];
                for (int i = 0; i < 9; ++i) Fn[i] = sh.Fd[i];
                enforce_rank2_d(Fn);               // exact rank 2 in float64
                rc::normalise_f22(Fn);
                for (int i = 0; i < 9; ++i) sh.Fd[i] = Fn[i];
            }
            __syncthreads();
            if (threadIdx.x < 9) F_out[9ll * p + threadIdx.x] = sh.Fd[threadIdx.x];
        }

        // ---------------------------------------------------- K4 verification
        double Fv[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) Fv[i] = sh.Fd[i];
        const float vthr = static_cast<float>(prm.verify_thresh);
        double st[11];
#pragma unroll
        for (int i = 0; i < 11; ++i) st[i] = 0.0;
        for (int i = threadIdx.x; i < m; i += kThreads) {
            const float4 q = load_match(p1, p2, i);
            const float e = symmetric_error(Fv, q.x, q.y, q.z, q.w);
            const bool in = e < vthr;
            sym[base + i] = e;
            mask[base + i] = in ? 1 : 0;
            st[1] += e;
            if (in) {
                st[0] += 1.0;
                st[2] += e;
                st[3] += q.x; st[4] += q.y; st[5] += q.z; st[6] += q.w;
                st[7] += static_cast<double>(q.x) * q.x; st[8] += static_cast<double>(q.y) * q.y;
                st[9] += static_cast<double>(q.z) * q.z; st[10] += static_cast<double>(q.w) * q.w;
            }
        }
        block_sum<11>(st, sh);   // (its barriers also publish the error array to the whole block)
        // The gates in float64 first.  NumPy's float32 results (sequential sums of <= n terms, pairwise for the means)
        // lie within gate * (n + 5) * 2^-24 of these, so a pair is decided unless a quantity sits within
        // gate * (n + 16) * 2^-22 of its gate; only such a BORDERLINE pair (rare) pays for the exact float32 walks.
        const double n_in = st[0];
        const int n_inl = static_cast<int>(n_in);
        const double tol = (n_in + 16.0) * (1.0 / 4194304.0);
        double sd64[4] = {0, 0, 0, 0};
        bool border = false;
        if (n_inl > 0) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const double mean = st[3 + k] / n_in;
                double var = st[7 + k] / n_in - mean * mean;
                var = var > 0 ? var : 0;
                sd64[k] = sqrt(var);
                border = border || fabs(sd64[k] - prm.min_spread) <= prm.min_spread * tol;
            }
            border = border || fabs(st[2] / n_in - prm.max_error) <= prm.max_error * tol;
        }
        if (border) {   // block-uniform: every thread holds the same sums
            const int task = (threadIdx.x & 31) == 0 ? static_cast<int>(threadIdx.x >> 5) : -1;
            const float* ev = sym + base;
            if (task == 0 || task == 1) {
                const float2 sd = np_std2_f32(task == 0 ? p1 : p2, 2, ev, vthr, m, n_inl);
                sh.np_res[2 * task] = sd.x;
                sh.np_res[2 * task + 1] = sd.y;
            } else if (task == 2) {
                sh.np_res[4] = np_mean_f32(ev, vthr, true, n_inl);
            } else if (task == 3) {
                sh.np_res[5] = np_mean_f32(ev, vthr, false, m);
            }
            __syncthreads();
        }
        if (threadIdx.x == 0) {
            sfm_pair_metrics mt;
            mt.total_matches = m;
            mt.inliers = n_inl;
            const double ratio = m > 0 ? n_in / m : 0.0;
            mt.inlier_ratio = static_cast<float>(ratio);
            mt.reprojection_error = n_inl > 0 ? (border ? sh.np_res[4] : static_cast<float>(st[2] / n_in))
                                              : __int_as_float(0x7f800000);
            mt.symmetric_error = m > 0 ? (border ? sh.np_res[5] : static_cast<float>(st[1] / m)) : 0.f;
            bool spread = n_inl > 0;
            if (spread) {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    spread = spread && ((border ? sh.np_res[k] : static_cast<float>(sd64[k])) > static_cast<float>(prm.min_spread));
            }
            mt.well_distributed = spread ? 1 : 0;
            mt.ransac_inliers = ransac_inl;
            mt.ransac_iters = ransac_it;
            metrics[p] = mt;
            const bool pass = mt.inliers >= prm.min_inliers && ratio >= prm.min_ratio &&
                              mt.reprojection_error <= static_cast<float>(prm.max_error) && spread;
            status_out[p] = pass ? SFM_PAIR_OK : SFM_PAIR_QUALITY_FAIL;
        }
    }
}

__global__ void gather_points_kernel(const int32_t* __restrict__ pairs, int n_pairs,
                                     const long long* __restrict__ img_off,
                                     const long long* __restrict__ row_off,
                                     const int32_t* __restrict__ mq, const int32_t* __restrict__ mt,
                                     const int32_t* __restrict__ count, const float2* __restrict__ kps,
                                     float2* __restrict__ pts1, float2* __restrict__ pts2) {
    for (int p = blockIdx.x; p < n_pairs; p += gridDim.x) {
        const long long q0 = img_off[pairs[2 * p]], t0 = img_off[pairs[2 * p + 1]];
        const long long base = row_off[p];
        const int m = count[p];
        for (int i = threadIdx.x; i < m; i += blockDim.x) {
            pts1[base + i] = __ldg(kps + q0 + mq[base + i]);
            pts2[base + i] = __ldg(kps + t0 + mt[base + i]);
        }
    }
}

}  // namespace

void launch_gather_points(const int32_t* d_pairs, int32_t n_pairs, const int64_t* d_img_off,
                          const int64_t* d_row_off, const int32_t* d_mq, const int32_t* d_mt,
                          const int32_t* d_count, const float* d_kps, float* d_pts1, float* d_pts2,
                          cudaStream_t st, int64_t* launches) {
    if (n_pairs <= 0) return;
    const int grid = n_pairs < 148 * 8 ? n_pairs : 148 * 8;
    gather_points_kernel<<<grid, 256, 0, st>>>(
        d_pairs, n_pairs, reinterpret_cast<const long long*>(d_img_off),
        reinterpret_cast<const long long*>(d_row_off), d_mq, d_mt, d_count,
        reinterpret_cast<const float2*>(d_kps), reinterpret_cast<float2*>(d_pts1),
        reinterpret_cast<float2*>(d_pts2));
    ++*launches;
}

int launch_ransac_verify(int32_t n_pairs, const int64_t* d_off, const int32_t* d_cnt,
                             const int32_t* d_pair_ids, const float* d_pts1, const float* d_pts2,
                             const double* d_F_in, const sfm_verify_params& prm, double* d_F,
                             uint8_t* d_mask, float* d_sym, sfm_pair_metrics* d_metrics,
                             int32_t* d_status, cudaStream_t st, int64_t* launches) {
    if (n_pairs <= 0) return 0;
    const int grid = n_pairs;   // one CTA per pair: the hardware scheduler balances the ragged pair costs
    ransac_verify_kernel<<<grid, kThreads, 0, st>>>(
        n_pairs, reinterpret_cast<const long long*>(d_off), d_cnt, d_pair_ids, d_pts1, d_pts2, d_F_in,
        prm, d_F, d_mask, d_sym, d_metrics, d_status);
    ++*launches;
    return static_cast<int>(cudaGetLastError());
}

}  // namespace sfm
