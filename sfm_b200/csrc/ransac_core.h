// Per-hypothesis mathematics of the fundamental-matrix RANSAC (device code of
// ransac.cu; SFM_HD so that the same functions compile for the host).  No
// reference code: OpenCV's findFundamentalMat (reference call site
// utils/find_matches.py:282) is a 7-point RANSAC with its own RNG; this is the
// normalised 8-point solve with a counter-based RNG, so that the result is a
// pure function of (seed, pair key, correspondences).  Hypotheses are scored
// with the symmetric epipolar error (OpenCV's own measure) or, by flag, with
// the Sampson error the north star names (sfm_verify_params::score_mode: the
// Sampson variant leaves the SURVEY 8c envelope on the golden pairs).
#pragma once
#include <cmath>
#include <cstdint>

#if defined(__CUDACC__)
#define SFM_HD __host__ __device__ __forceinline__
#else
#define SFM_HD inline
#endif

namespace sfm {
namespace rc {

// ------------------------------------------------------------------ RNG
SFM_HD uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

SFM_HD uint64_t pair_key(uint64_t seed, uint32_t a, uint32_t b) {
    return mix64(seed ^ mix64((static_cast<uint64_t>(a) << 32) | b));
}

// 8 distinct indices in [0, m), m >= 8, for hypothesis `hyp` of the pair `key`.
SFM_HD void sample8(uint64_t key, uint32_t hyp, uint32_t m, uint32_t (&idx)[8]) {
    uint64_t ctr = (static_cast<uint64_t>(hyp) << 20);
    int n = 0;
    while (n < 8) {
        const uint64_t r = mix64(key + ctr++);
        const uint32_t cand = static_cast<uint32_t>(((r >> 32) * static_cast<uint64_t>(m)) >> 32);
        bool dup = false;
        for (int j = 0; j < 8; ++j) dup = dup || (j < n && idx[j] == cand);
        if (!dup) idx[n++] = cand;
    }
}

// ------------------------------------------------------------ 3x3 helpers
// Jacobi eigen-decomposition of a symmetric 3x3 (a00 a01 a02 a11 a12 a22);
// returns the unit eigenvector of the smallest eigenvalue.
template <typename T>
SFM_HD void smallest_eigvec3(T a00, T a01, T a02, T a11, T a12, T a22, T (&v)[3]) {
    T A[3][3] = {{a00, a01, a02}, {a01, a11, a12}, {a02, a12, a22}};
    T V[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
#pragma unroll 1
    for (int sweep = 0; sweep < 8; ++sweep) {
#pragma unroll
        for (int pq = 0; pq < 3; ++pq) {
            const int p = pq == 2 ? 1 : 0;
            const int q = pq == 0 ? 1 : 2;
            const T apq = A[p][q];
            if (apq != T(0)) {
                const T theta = (A[q][q] - A[p][p]) / (T(2) * apq);
                const T t = (theta >= T(0) ? T(1) : T(-1)) /
                            (fabs(theta) + sqrt(theta * theta + T(1)));
                const T c = T(1) / sqrt(t * t + T(1));
                const T s = t * c;
#pragma unroll
                for (int k = 0; k < 3; ++k) {   // A <- A J
                    const T akp = A[k][p], akq = A[k][q];
                    A[k][p] = c * akp - s * akq;
                    A[k][q] = s * akp + c * akq;
                }
#pragma unroll
                for (int k = 0; k < 3; ++k) {   // A <- J^T A
                    const T apk = A[p][k], aqk = A[q][k];
                    A[p][k] = c * apk - s * aqk;
                    A[q][k] = s * apk + c * aqk;
                }
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const T vkp = V[k][p], vkq = V[k][q];
                    V[k][p] = c * vkp - s * vkq;
                    V[k][q] = s * vkp + c * vkq;
                }
            }
        }
    }
    int m = 0;
    if (A[1][1] < A[m][m]) m = 1;
    if (A[2][2] < A[m][m]) m = 2;
    // static selection keeps V in registers on the device
    v[0] = m == 0 ? V[0][0] : (m == 1 ? V[0][1] : V[0][2]);
    v[1] = m == 0 ? V[1][0] : (m == 1 ? V[1][1] : V[1][2]);
    v[2] = m == 0 ? V[2][0] : (m == 1 ? V[2][1] : V[2][2]);
}

// Closest rank-2 matrix in Frobenius norm: F - (F v3) v3^T, v3 = right singular
// vector of the smallest singular value (= smallest eigenvector of F^T F).
template <typename T>
SFM_HD void enforce_rank2(T (&F)[9]) {
    T g[6];  // F^T F, upper triangle
    g[0] = F[0] * F[0] + F[3] * F[3] + F[6] * F[6];
    g[1] = F[0] * F[1] + F[3] * F[4] + F[6] * F[7];
    g[2] = F[0] * F[2] + F[3] * F[5] + F[6] * F[8];
    g[3] = F[1] * F[1] + F[4] * F[4] + F[7] * F[7];
    g[4] = F[1] * F[2] + F[4] * F[5] + F[7] * F[8];
    g[5] = F[2] * F[2] + F[5] * F[5] + F[8] * F[8];
    T v[3];
    smallest_eigvec3<T>(g[0], g[1], g[2], g[3], g[4], g[5], v);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const T fv = F[3 * r] * v[0] + F[3 * r + 1] * v[1] + F[3 * r + 2] * v[2];
        F[3 * r] -= fv * v[0];
        F[3 * r + 1] -= fv * v[1];
        F[3 * r + 2] -= fv * v[2];
    }
}

// Hartley normalisation x' = s (x - c): F = T2^T F' T1.
template <typename T>
SFM_HD void denormalise(T (&F)[9], T c1x, T c1y, T s1, T c2x, T c2y, T s2) {
    // F' T1 : columns scaled by s1, third column gets -s1*(c1 . row)
    T G[9];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        G[3 * r] = F[3 * r] * s1;
        G[3 * r + 1] = F[3 * r + 1] * s1;
        G[3 * r + 2] = F[3 * r + 2] - s1 * (F[3 * r] * c1x + F[3 * r + 1] * c1y);
    }
    // T2^T G : rows 0,1 scaled by s2, row 2 gets -s2*(c2 . col)
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        F[c] = s2 * G[c];
        F[3 + c] = s2 * G[3 + c];
        F[6 + c] = G[6 + c] - s2 * (c2x * G[c] + c2y * G[3 + c]);
    }
}

template <typename T>
SFM_HD void scale_unit(T (&F)[9]) {
    T n = 0;
#pragma unroll
    for (int i = 0; i < 9; ++i) n += F[i] * F[i];
    const T inv = n > T(0) ? T(1) / sqrt(n) : T(0);
#pragma unroll
    for (int i = 0; i < 9; ++i) F[i] *= inv;
}

// ------------------------------------------------------- minimal solver
// Normalised 8-point algorithm on exactly 8 correspondences p[i] = (x1,y1,x2,y2).
// Row-major F with x2^T F x1 = 0.  Returns false for degenerate samples.
SFM_HD bool solve8(const float (&p)[8][4], float (&F)[9]) {
    float c1x = 0, c1y = 0, c2x = 0, c2y = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        c1x += p[i][0];
        c1y += p[i][1];
        c2x += p[i][2];
        c2y += p[i][3];
    }
    c1x *= 0.125f; c1y *= 0.125f; c2x *= 0.125f; c2y *= 0.125f;
    float d1 = 0, d2 = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        d1 += sqrtf((p[i][0] - c1x) * (p[i][0] - c1x) + (p[i][1] - c1y) * (p[i][1] - c1y));
        d2 += sqrtf((p[i][2] - c2x) * (p[i][2] - c2x) + (p[i][3] - c2y) * (p[i][3] - c2y));
    }
    if (!(d1 > 1e-6f) || !(d2 > 1e-6f)) return false;
    const float s1 = 8.0f * 1.41421356f / d1, s2 = 8.0f * 1.41421356f / d2;

    float A[8][9];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float x1 = (p[i][0] - c1x) * s1, y1 = (p[i][1] - c1y) * s1;
        const float x2 = (p[i][2] - c2x) * s2, y2 = (p[i][3] - c2y) * s2;
        A[i][0] = x2 * x1; A[i][1] = x2 * y1; A[i][2] = x2;
        A[i][3] = y2 * x1; A[i][4] = y2 * y1; A[i][5] = y2;
        A[i][6] = x1;      A[i][7] = y1;      A[i][8] = 1.0f;
    }
    // Gauss-Jordan with partial (row) pivoting on the leading 8 columns; all
    // indices are compile-time so the matrix lives in registers on the device.
    float minpiv = 1e30f;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        int piv = c;
        float best = fabsf(A[c][c]);
#pragma unroll
        for (int r = c + 1; r < 8; ++r) {
            const float v = fabsf(A[r][c]);
            if (v > best) { best = v; piv = r; }
        }
        minpiv = fminf(minpiv, best);
#pragma unroll
        for (int r = c + 1; r < 8; ++r) {
            if (r == piv) {
#pragma unroll
                for (int k = c; k < 9; ++k) { const float t = A[c][k]; A[c][k] = A[r][k]; A[r][k] = t; }
            }
        }
        const float inv = 1.0f / A[c][c];
#pragma unroll
        for (int k = c; k < 9; ++k) A[c][k] *= inv;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            if (r != c) {
                const float f = A[r][c];
#pragma unroll
                for (int k = c; k < 9; ++k) A[r][k] -= f * A[c][k];
            }
        }
    }
    if (!(minpiv > 1e-7f)) return false;
#pragma unroll
    for (int i = 0; i < 8; ++i) F[i] = -A[i][8];
    F[8] = 1.0f;
    scale_unit<float>(F);
    enforce_rank2<float>(F);
    denormalise<float>(F, c1x, c1y, s1, c2x, c2y, s2);
    scale_unit<float>(F);
    bool ok = true;
#pragma unroll
    for (int i = 0; i < 9; ++i) ok = ok && (fabsf(F[i]) <= 1.0f);   // rejects NaN/inf
    return ok;
}

// Sampson (first-order geometric) squared error of x2^T F x1 = 0.
template <typename T>
SFM_HD T sampson_sq(const T (&F)[9], T x1, T y1, T x2, T y2) {
    const T l2x = F[0] * x1 + F[1] * y1 + F[2];       // F x1   (line in image 2)
    const T l2y = F[3] * x1 + F[4] * y1 + F[5];
    const T l2z = F[6] * x1 + F[7] * y1 + F[8];
    const T l1x = F[0] * x2 + F[3] * y2 + F[6];       // F^T x2 (line in image 1)
    const T l1y = F[1] * x2 + F[4] * y2 + F[7];
    const T r = x2 * l2x + y2 * l2y + l2z;
    const T den = l2x * l2x + l2y * l2y + l1x * l1x + l1y * l1y;
    return den > T(0) ? (r * r) / den : T(1e30);
}

// Symmetric epipolar distance (mean of the two point-to-epiline distances),
// the quantity thresholded by the reference's geometric_verification
// (utils/find_matches.py:165-172); fast single-precision form for scoring.
template <typename T>
SFM_HD T symmetric_dist(const T (&F)[9], T x1, T y1, T x2, T y2) {
    const T l2x = F[0] * x1 + F[1] * y1 + F[2];
    const T l2y = F[3] * x1 + F[4] * y1 + F[5];
    const T l2z = F[6] * x1 + F[7] * y1 + F[8];
    const T l1x = F[0] * x2 + F[3] * y2 + F[6];
    const T l1y = F[1] * x2 + F[4] * y2 + F[7];
    const T r = fabs(x2 * l2x + y2 * l2y + l2z);
    const T n1 = l1x * l1x + l1y * l1y, n2 = l2x * l2x + l2y * l2y;
    if (!(n1 > T(0)) || !(n2 > T(0))) return T(1e15);
    return T(0.5) * r * (T(1) / sqrt(n1) + T(1) / sqrt(n2));
}

#ifdef __CUDA_ARCH__
// Hypothesis scoring form (device, float): two MUFU.RSQ instead of two sqrt + two divisions;
// the hot loop of the RANSAC rounds (hypotheses x matches evaluations).
template <>
__device__ __forceinline__ float symmetric_dist<float>(const float (&F)[9], float x1, float y1, float x2,
                                                       float y2) {
    const float l2x = fmaf(F[0], x1, fmaf(F[1], y1, F[2]));
    const float l2y = fmaf(F[3], x1, fmaf(F[4], y1, F[5]));
    const float l2z = fmaf(F[6], x1, fmaf(F[7], y1, F[8]));
    const float l1x = fmaf(F[0], x2, fmaf(F[3], y2, F[6]));
    const float l1y = fmaf(F[1], x2, fmaf(F[4], y2, F[7]));
    const float r = fabsf(fmaf(x2, l2x, fmaf(y2, l2y, l2z)));
    const float n1 = fmaf(l1x, l1x, l1y * l1y), n2 = fmaf(l2x, l2x, l2y * l2y);
    const float d = 0.5f * r * (rsqrtf(n1) + rsqrtf(n2));
    return (n1 > 0.f && n2 > 0.f) ? d : 1e15f;
}
#endif

// Residual used for inlier decisions: squared, compared against thr^2.
template <typename T>
SFM_HD T residual_sq(const T (&F)[9], T x1, T y1, T x2, T y2, bool symmetric) {
    if (symmetric) {
        const T d = symmetric_dist<T>(F, x1, y1, x2, y2);
        return d * d;
    }
    return sampson_sq<T>(F, x1, y1, x2, y2);
}

// Standard adaptive stopping rule: hypotheses needed to draw one all-inlier
// 8-sample with probability `conf` at inlier ratio w.
SFM_HD int iters_needed(int inliers, int m, double conf, int max_iters) {
    if (inliers <= 0 || m <= 0) return max_iters;
    const double w = static_cast<double>(inliers) / static_cast<double>(m);
    double p8 = w * w;
    p8 = p8 * p8;
    p8 = p8 * p8;
    if (p8 >= 1.0 - 1e-12) return 1;
    if (p8 < 1e-12) return max_iters;
    const double n = log(1.0 - conf) / log(1.0 - p8);
    if (!(n < static_cast<double>(max_iters))) return max_iters;
    return n < 1.0 ? 1 : static_cast<int>(ceil(n));
}

// ------------------------------------------------ least-squares refit
// Smallest eigenvector of a symmetric 9x9 (full storage) by cyclic Jacobi.
SFM_HD void smallest_eigvec9(double (&A)[9][9], double (&v)[9]) {
    double V[9][9];
    for (int i = 0; i < 9; ++i)
        for (int j = 0; j < 9; ++j) V[i][j] = i == j ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 12; ++sweep) {
        double off = 0, diag = 0;
        for (int i = 0; i < 9; ++i) {
            diag += A[i][i] * A[i][i];
            for (int j = i + 1; j < 9; ++j) off += A[i][j] * A[i][j];
        }
        if (off <= 1e-30 * diag) break;
        for (int p = 0; p < 8; ++p)
            for (int q = p + 1; q < 9; ++q) {
                const double apq = A[p][q];
                if (apq == 0.0) continue;
                const double theta = (A[q][q] - A[p][p]) / (2.0 * apq);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < 9; ++k) {
                    const double akp = A[k][p], akq = A[k][q];
                    A[k][p] = c * akp - s * akq;
                    A[k][q] = s * akp + c * akq;
                }
                for (int k = 0; k < 9; ++k) {
                    const double apk = A[p][k], aqk = A[q][k];
                    A[p][k] = c * apk - s * aqk;
                    A[q][k] = s * apk + c * aqk;
                }
                for (int k = 0; k < 9; ++k) {
                    const double vkp = V[k][p], vkq = V[k][q];
                    V[k][p] = c * vkp - s * vkq;
                    V[k][q] = s * vkp + c * vkq;
                }
            }
    }
    int m = 0;
    for (int i = 1; i < 9; ++i)
        if (A[i][i] < A[m][m]) m = i;
    for (int i = 0; i < 9; ++i) v[i] = V[i][m];
}

// Moments gathered over an inlier set for the normalised least-squares refit.
struct Moments {
    double n, sx1, sy1, sx2, sy2;   // pass 1: count and coordinate sums
    double d1, d2;                  // pass 2a: sum of distances to the centroid
};

// F from the 45 unique entries of A^T A (normalised coordinates) and the
// normalisation parameters.  ata is the upper triangle, row-major.
SFM_HD bool refit_from_normal(const double (&ata)[45], double c1x, double c1y, double s1,
                              double c2x, double c2y, double s2, double (&F)[9]) {
    double A[9][9];
    int k = 0;
    for (int i = 0; i < 9; ++i)
        for (int j = i; j < 9; ++j) {
            A[i][j] = ata[k];
            A[j][i] = ata[k];
            ++k;
        }
    double v[9];
    smallest_eigvec9(A, v);
    for (int i = 0; i < 9; ++i) F[i] = v[i];
    enforce_rank2<double>(F);
    denormalise<double>(F, c1x, c1y, s1, c2x, c2y, s2);
    scale_unit<double>(F);
    bool ok = true;
    for (int i = 0; i < 9; ++i) ok = ok && (fabs(F[i]) <= 1.0);
    return ok;
}

// One row of the 8-point design matrix in normalised coordinates.
SFM_HD void design_row(double x1, double y1, double x2, double y2, double (&a)[9]) {
    a[0] = x2 * x1; a[1] = x2 * y1; a[2] = x2;
    a[3] = y2 * x1; a[4] = y2 * y1; a[5] = y2;
    a[6] = x1;      a[7] = y1;      a[8] = 1.0;
}

// cv2 convention of the returned matrix: F[2][2] = 1 unless it is tiny.
SFM_HD void normalise_f22(double (&F)[9]) {
    scale_unit<double>(F);
    if (fabs(F[8]) > 1.1920929e-07) {
        const double inv = 1.0 / F[8];
        for (int i = 0; i < 9; ++i) F[i] *= inv;
        F[8] = 1.0;
    }
}

}  // namespace rc
}  // namespace sfm
