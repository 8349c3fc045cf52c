// essential.cu — essential-matrix estimation of matched pairs (second half): batched replacement of
//     E, mask = cv2.findEssentialMat(pts1_norm, pts2_norm, focal=1.0, pp=(0, 0), method=cv2.RANSAC, prob, threshold, maxIters)
// as called per image pair by /root/reference/CameraPoseEstimation/essential_estimation.py:118-126 on points normalised by
// cv2.undistortPoints(pts, K, None) (:98-108).  Input is what the matcher's hand-off already holds in HBM: the CSR `ptr`
// over pairs and the float32 [K, 4] pixel correspondences of b200m_gather_correspondences, plus the two cameras'
// (fx, fy, cx, cy) per pair.  One CTA per image pair.
//
// cv2's RANSAC is a deterministic sequential program (its RNG restarts from the same state on every call); the kernel
// keeps that sequence and parallelises inside it, as ransac.cu does for the homography:
//   * one thread draws HYP 5-tuples per round (distinct indices by rejection; there is no checkSubset for E);
//   * HYP threads run Nister's 5-point solver side by side, each on its own workspace in shared memory: the null-space
//     basis exactly as OpenCV's SVD completes it (fixed pseudo-random vectors, Gram-Schmidt), the ten cubic constraints
//     over 20 monomials, elimination, the 3 x 3 polynomial matrix B(z), its determinant (degree 10), the roots by
//     cv::solvePoly's Durand-Kerner sweeps -- kept operation for operation, because the ORDER in which the real roots
//     come out is the order of the candidate models and decides ties -- and up to 10 models per sample;
//   * all threads count the inliers of all models of the round (Sampson error in float64 rounded to float32, compared
//     with float32(threshold^2), no fused multiply-add);
//   * one thread replays the sequential decisions in order: "strictly more inliers wins", the shrinking iteration
//     limit, the stop.  Hypotheses drawn past the stop are discarded, exactly as if they had never been drawn.
// There is no refit: E is the winning model, the mask its inliers.  Compiled with -fmad=false (build.py).
#include <float.h>
#include <math.h>
#include <string.h>
#include <math_constants.h>
#include "common.cuh"

namespace {

constexpr int ET = 128;          // threads per CTA
constexpr int HYP = 96;          // hypotheses (5-tuples) per round = solver threads (3 of the 4 warps)
// per-thread workspace, element-major in shared memory: element e of thread t lives at ws[e * HYP + t].  256 doubles per
// thread = 192 KiB for 96 threads: rows 0..4 of V (needed only while the basis is built) overlay A, which is zeroed later.
constexpr int OFF_N = 0;         // 36  the basis X, Y, Z, W (rows 5..8 of V)
constexpr int OFF_A = 36;        // 200 A[10][20]; before it is built: V rows 0..4; after the elimination: B, polynomial, roots, models
constexpr int OFF_T = 236;       // 10  tr(E E^T)
constexpr int OFF_Q = 246;       // 10  quadratic temporary
constexpr int SCR = 256;
constexpr int OFF_B = OFF_A;             // 39  B[3][13]          (rows 0, 1 of A)
constexpr int OFF_C = OFF_A + 40;        // 11  c[0..10]          (row 2)
constexpr int OFF_RE = OFF_A + 60;       // 10  roots, real parts  (row 3)
constexpr int OFF_IM = OFF_A + 70;       // 10  roots, imaginary parts
constexpr int OFF_M = OFF_A + 80;        // 90  models[10][9]     (rows 4..8, free once B is built)
__device__ __forceinline__ int vidx(int i, int k) { return (i < 5 ? OFF_A + 9 * i : OFF_N + 9 * (i - 5)) + k; }   // V[i][k]

__device__ const int8_t LLT[4][4] = {{0, 3, 4, 6}, {3, 1, 5, 7}, {4, 5, 2, 8}, {6, 7, 8, 9}};          // linear x linear -> quadratic
__device__ const int8_t QLT[10][4] = {{0, 2, 4, 5}, {3, 1, 6, 7}, {10, 13, 16, 17}, {2, 3, 8, 9}, {4, 8, 10, 11},
                                      {8, 6, 13, 14}, {5, 9, 11, 12}, {9, 7, 14, 15}, {11, 14, 17, 18}, {12, 15, 18, 19}};   // quadratic x linear -> cubic (Nister's column order)

struct WS {                      // one solver thread's workspace
    double* p;
    __device__ __forceinline__ double& operator()(int e) const { return p[e * HYP]; }
};

__device__ __forceinline__ u32 rng_next(u64& s) {
    s = (u64)(u32)s * 4164903690ull + (s >> 32);
    return (u32)s;
}

// rows 5..8 of Vt of cv::SVD(Q [5 x 9], FULL_UV); Q's rows are already in V rows 0..4.  false: the sample is degenerate.
__device__ bool null_basis(const WS& w) {
    for (int i = 0; i < 5; ++i) {
        for (int it = 0; it < 2; ++it)
            for (int j = 0; j < i; ++j) {
                double sd = 0;
                for (int k = 0; k < 9; ++k) sd += w(vidx(i, k)) * w(vidx(j, k));
                for (int k = 0; k < 9; ++k) w(vidx(i, k)) -= sd * w(vidx(j, k));
            }
        double nn = 0;
        for (int k = 0; k < 9; ++k) nn += w(vidx(i, k)) * w(vidx(i, k));
        if (!(nn > 1e-300)) return false;
        nn = 1.0 / sqrt(nn);
        for (int k = 0; k < 9; ++k) w(vidx(i, k)) *= nn;
    }
    u64 st = 0x12345678ull;
    for (int i = 5; i < 9; ++i) {
        for (int k = 0; k < 9; ++k) w(vidx(i, k)) = (rng_next(st) & 256u) != 0 ? 1.0 / 9 : -1.0 / 9;
        for (int it = 0; it < 2; ++it)
            for (int j = 0; j < i; ++j) {
                double sd = 0, asum = 0;
                for (int k = 0; k < 9; ++k) sd += w(vidx(i, k)) * w(vidx(j, k));
                for (int k = 0; k < 9; ++k) {
                    const double t = w(vidx(i, k)) - sd * w(vidx(j, k));
                    w(vidx(i, k)) = t;
                    asum += fabs(t);
                }
                asum = asum > DBL_EPSILON * 100 ? 1 / asum : 0;
                for (int k = 0; k < 9; ++k) w(vidx(i, k)) *= asum;
            }
        double sd = 0;
        for (int k = 0; k < 9; ++k) sd += w(vidx(i, k)) * w(vidx(i, k));
        sd = sqrt(sd);
        const double s = sd > DBL_MIN ? 1 / sd : 0;
        for (int k = 0; k < 9; ++k) w(vidx(i, k)) *= s;
    }
    return true;
}

// linear polynomial of entry (r, c) of E = x X + y Y + z Z + W: coefficient k is basis vector k's element 3 r + c
__device__ __forceinline__ double Elin(const WS& w, int r, int c, int k) { return w(OFF_N + 9 * k + 3 * r + c); }

// q (10 quadratic coefficients at offset oq) += s * E[r1][c1] * E[r2][c2]
__device__ void mul_ll(const WS& w, int oq, int r1, int c1, int r2, int c2, double s) {
    for (int i = 0; i < 4; ++i) {
        const double a = Elin(w, r1, c1, i);
        for (int j = 0; j < 4; ++j) w(oq + LLT[i][j]) += s * a * Elin(w, r2, c2, j);
    }
}
// row (20 cubic coefficients at offset orow) += s * q * E[r][c]
__device__ void mul_ql(const WS& w, int orow, int oq, int r, int c, double s) {
    for (int i = 0; i < 10; ++i) {
        const double a = w(oq + i);
        for (int j = 0; j < 4; ++j) w(orow + QLT[i][j]) += s * a * Elin(w, r, c, j);
    }
}

// A[10][20]: det E = 0 (row 0) and 2 E E^T E - tr(E E^T) E = 0 (rows 1 + 3 r + c)
__device__ void coeff_matrix(const WS& w) {
    for (int i = 0; i < 200; ++i) w(OFF_A + i) = 0;
    for (int r = 0; r < 3; ++r) {                                                 // tr = ((EEt00 + EEt11) + EEt22), each summed on its own
        for (int q = 0; q < 10; ++q) w(OFF_Q + q) = 0;
        for (int k = 0; k < 3; ++k) mul_ll(w, OFF_Q, r, k, r, k, 1.0);
        for (int q = 0; q < 10; ++q) w(OFF_T + q) = r == 0 ? w(OFF_Q + q) : w(OFF_T + q) + w(OFF_Q + q);
    }
    const int P[6][3] = {{0, 1, 2}, {1, 2, 0}, {2, 0, 1}, {2, 1, 0}, {1, 0, 2}, {0, 2, 1}};
    for (int i = 0; i < 6; ++i) {
        for (int q = 0; q < 10; ++q) w(OFF_Q + q) = 0;
        mul_ll(w, OFF_Q, 0, P[i][0], 1, P[i][1], 1.0);
        mul_ql(w, OFF_A, OFF_Q, 2, P[i][2], i < 3 ? 1.0 : -1.0);
    }
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) {
            const int orow = OFF_A + 20 * (1 + 3 * r + c);
            for (int k = 0; k < 3; ++k) {
                for (int q = 0; q < 10; ++q) w(OFF_Q + q) = 0;
                for (int m = 0; m < 3; ++m) mul_ll(w, OFF_Q, r, m, k, m, 1.0);    // (E E^T)[r][k]
                mul_ql(w, orow, OFF_Q, k, c, 2.0);
            }
            mul_ql(w, orow, OFF_T, r, c, -1.0);
        }
}

// A[:, 10:] <- A[:, :10]^-1 A[:, 10:] (Gauss-Jordan with partial pivoting); false when singular
__device__ bool reduce(const WS& w) {
    for (int k = 0; k < 10; ++k) {
        int p = k;
        double best = fabs(w(OFF_A + 20 * k + k));
        for (int i = k + 1; i < 10; ++i) {
            const double v = fabs(w(OFF_A + 20 * i + k));
            if (v > best) { best = v; p = i; }
        }
        if (best < DBL_MIN) return false;
        if (p != k)
            for (int j = 0; j < 20; ++j) {
                const double t = w(OFF_A + 20 * k + j);
                w(OFF_A + 20 * k + j) = w(OFF_A + 20 * p + j);
                w(OFF_A + 20 * p + j) = t;
            }
        const double inv = 1.0 / w(OFF_A + 20 * k + k);
        for (int j = k; j < 20; ++j) w(OFF_A + 20 * k + j) *= inv;
        for (int i = 0; i < 10; ++i) {
            if (i == k) continue;
            const double f = w(OFF_A + 20 * i + k);
            if (f != 0)
                for (int j = k; j < 20; ++j) w(OFF_A + 20 * i + j) -= f * w(OFF_A + 20 * k + j);
        }
    }
    return true;
}

// The same iteration for the common case n == 10 with the polynomial and the roots in REGISTERS (both inner loops fully
// unrolled): in practice the roots never reach a fixed point -- they keep flipping in their last bits -- so all 300 sweeps run,
// 300 x 10 x 20 complex multiplications per sample; with every operand in shared memory that was ~6 ms per round of 64
// samples.  Same operations in the same order as solve_poly() below, hence the same bits.
__device__ void solve_poly10(const WS& w) {
    double c[11], re[10], im[10];
#pragma unroll
    for (int k = 0; k < 11; ++k) c[k] = w(OFF_C + k);
    {
        double pr = 1, pi = 0;
#pragma unroll
        for (int i = 0; i < 10; ++i) {
            re[i] = pr; im[i] = pi;
            const double t = pr - pi; pi = pr + pi; pr = t;
        }
    }
#pragma unroll 1
    for (int iter = 0; iter < 300; ++iter) {
        bool changed = false;        // (cv's maxDiff <= 0 test implies "nothing changed", which ends the iteration here anyway)
#pragma unroll
        for (int i = 0; i < 10; ++i) {
            const double xr = re[i], xi = im[i];
            double nr = c[10], ni = 0, dr = c[10], di = 0;
            int same = 1;
#pragma unroll
            for (int j = 0; j < 10; ++j) {
                const double t = nr * xr - ni * xi;
                ni = nr * xi + ni * xr;
                nr = t + c[9 - j];
                if (j != i) {
                    const double er = xr - re[j], ei = xi - im[j];
                    if (er != 0 || ei != 0) { const double u = dr * er - di * ei; di = dr * ei + di * er; dr = u; }
                    else ++same;
                }
            }
            const double t = 1.0 / (dr * dr + di * di);
            double qr = (nr * dr + ni * di) * t, qi = (-nr * di + ni * dr) * t;
            if (same > 1) {
                const double mag = pow(sqrt(qr * qr + qi * qi), 1.0 / same), ang = atan2(qi, qr) / same;
                qr = mag * cos(ang); qi = mag * sin(ang);
            }
            const double yr = xr - qr, yi = xi - qi;
            changed |= yr != xr || yi != xi;
            re[i] = yr; im[i] = yi;
        }
        if (!changed) break;
    }
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        w(OFF_RE + i) = re[i];
        w(OFF_IM + i) = fabs(im[i]) < 1e-100 ? 0.0 : im[i];
    }
}

// cv::solvePoly (Durand-Kerner, Gauss-Seidel updates, start values (1 + i)^k, 300 sweeps) on c[0..n] at OFF_C; roots to
// OFF_RE / OFF_IM; returns their number.  A sweep that leaves every root unchanged ends the iteration: every later sweep
// would recompute the same values.
__device__ int solve_poly(const WS& w) {
    int n = 10;
    for (; n > 1; --n)
        if (fabs(w(OFF_C + n)) > DBL_EPSILON) break;
    if (n == 10) { solve_poly10(w); return 10; }
    double pr = 1, pi = 0;
    for (int i = 0; i < n; ++i) {
        w(OFF_RE + i) = pr; w(OFF_IM + i) = pi;
        const double t = pr - pi; pi = pr + pi; pr = t;
    }
    const double lead = w(OFF_C + n);
    for (int iter = 0; iter < 300; ++iter) {
        double maxdiff = 0;
        bool changed = false;
        for (int i = 0; i < n; ++i) {
            const double xr = w(OFF_RE + i), xi = w(OFF_IM + i);
            double nr = lead, ni = 0, dr = lead, di = 0;
            int same = 1;
            for (int j = 0; j < n; ++j) {
                const double t = nr * xr - ni * xi;
                ni = nr * xi + ni * xr;
                nr = t + w(OFF_C + n - j - 1);
                if (j != i) {
                    const double er = xr - w(OFF_RE + j), ei = xi - w(OFF_IM + j);
                    if (er != 0 || ei != 0) { const double u = dr * er - di * ei; di = dr * ei + di * er; dr = u; }
                    else ++same;
                }
            }
            const double t = 1.0 / (dr * dr + di * di);
            double qr = (nr * dr + ni * di) * t, qi = (-nr * di + ni * dr) * t;
            if (same > 1) {                      // coinciding estimates: the same-th root of the correction (principal branch)
                const double mag = pow(sqrt(qr * qr + qi * qi), 1.0 / same), ang = atan2(qi, qr) / same;
                qr = mag * cos(ang); qi = mag * sin(ang);
            }
            const double yr = xr - qr, yi = xi - qi;
            changed |= yr != xr || yi != xi;
            w(OFF_RE + i) = yr; w(OFF_IM + i) = yi;
            maxdiff = fmax(maxdiff, sqrt(qr * qr + qi * qi));
        }
        if (maxdiff <= 0 || !changed) break;
    }
    for (int i = 0; i < n; ++i)
        if (fabs(w(OFF_IM + i)) < 1e-100) w(OFF_IM + i) = 0;
    return n;
}

// EMEstimatorCallback::runKernel for one sample: the 5 normalised correspondences are in m[5][4] = (x1, y1, x2, y2).
// Returns the number of models; model k (3 x 3 row-major, unit Frobenius norm) at OFF_M + 9 k.
__device__ int five_point(const WS& w, const double (*m)[4]) {
    for (int i = 0; i < 5; ++i) {
        const double x1 = m[i][0], y1 = m[i][1], x2 = m[i][2], y2 = m[i][3];
        const double q[9] = {x1 * x2, y1 * x2, x2, x1 * y2, y1 * y2, y2, x1, y1, 1.0};     // = E row by row for x2^T E x1
        for (int k = 0; k < 9; ++k) w(vidx(i, k)) = q[k];
    }
    if (!null_basis(w)) return 0;
    coeff_matrix(w);
    if (!reduce(w)) return 0;
    // B[i] = row_{2i+4} - z row_{2i+5} over [x: z^3..1 | y: z^3..1 | 1: z^4..1]; B lives in rows 0, 1 of A (no longer needed)
    for (int i = 0; i < 39; ++i) w(OFF_B + i) = 0;
    for (int i = 0; i < 3; ++i) {
        const int r1 = OFF_A + 20 * (2 * i + 4) + 10, r2 = OFF_A + 20 * (2 * i + 5) + 10, b = OFF_B + 13 * i;
        for (int k = 0; k < 3; ++k) {
            w(b + 1 + k) += w(r1 + k); w(b + 5 + k) += w(r1 + 3 + k);
            w(b + k) -= w(r2 + k); w(b + 4 + k) -= w(r2 + 3 + k);
        }
        for (int k = 0; k < 4; ++k) { w(b + 9 + k) += w(r1 + 6 + k); w(b + 8 + k) -= w(r2 + 6 + k); }
    }
    {   // c[k] = coefficient of z^k of det B(z); entry polynomials lowest degree first: P(i, 0), P(i, 1) cubic, P(i, 2) quartic
        auto P = [&](int i, int col, int k) -> double {
            if (col < 2) return k < 4 ? w(OFF_B + 13 * i + 4 * col + 3 - k) : 0.0;
            return w(OFF_B + 13 * i + 12 - k);
        };
        for (int k = 0; k < 11; ++k) w(OFF_C + k) = 0;
        const int T[6][3] = {{0, 1, 2}, {1, 2, 0}, {2, 0, 1}, {2, 1, 0}, {1, 0, 2}, {0, 2, 1}};
        for (int t = 0; t < 6; ++t) {
            double ab[9];
#pragma unroll
            for (int a = 0; a < 9; ++a) ab[a] = 0;
#pragma unroll
            for (int a = 0; a < 5; ++a)
#pragma unroll
                for (int b = 0; b < 5; ++b) ab[a + b] += P(0, T[t][0], a) * P(1, T[t][1], b);
            const double sg = t < 3 ? 1.0 : -1.0;
#pragma unroll
            for (int a = 0; a < 9; ++a)
#pragma unroll
                for (int b = 0; b < 5; ++b)
                    if (a + b <= 10) w(OFF_C + a + b) += sg * ab[a] * P(2, T[t][2], b);
        }
    }
    const int n = solve_poly(w);
    int count = 0;
    for (int i = 0; i < n; ++i) {
        if (fabs(w(OFF_IM + i)) > 1e-10) continue;
        const double z = w(OFF_RE + i), z2 = z * z, z3 = z2 * z, z4 = z3 * z;
        double bz[3][3], v[3] = {0, 0, 0}, best = -1;
        for (int k = 0; k < 3; ++k) {
            const int b = OFF_B + 13 * k;
            bz[k][0] = w(b + 0) * z3 + w(b + 1) * z2 + w(b + 2) * z + w(b + 3);
            bz[k][1] = w(b + 4) * z3 + w(b + 5) * z2 + w(b + 6) * z + w(b + 7);
            bz[k][2] = w(b + 8) * z4 + w(b + 9) * z3 + w(b + 10) * z2 + w(b + 11) * z + w(b + 12);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) {             // null vector of the (rank 2) matrix: the largest cross product of two rows
            const double* a = bz[k];
            const double* b = bz[(k + 1) % 3];
            const double c0 = a[1] * b[2] - a[2] * b[1], c1 = a[2] * b[0] - a[0] * b[2], c2 = a[0] * b[1] - a[1] * b[0];
            const double nn = c0 * c0 + c1 * c1 + c2 * c2;
            if (nn > best) { best = nn; v[0] = c0; v[1] = c1; v[2] = c2; }
        }
        if (!(best > 0)) continue;
        const double s = 1.0 / sqrt(best);
        if (fabs(v[2] * s) < 1e-10) continue;
        const double x = v[0] / v[2], y = v[1] / v[2];
        double nn = 0;
        for (int k = 0; k < 9; ++k) {
            const double e = w(OFF_N + k) * x + w(OFF_N + 9 + k) * y + w(OFF_N + 18 + k) * z + w(OFF_N + 27 + k);
            w(OFF_M + 9 * count + k) = e;
            nn += e * e;
        }
        nn = 1.0 / sqrt(nn);
        for (int k = 0; k < 9; ++k) w(OFF_M + 9 * count + k) *= nn;
        ++count;
    }
    return count;
}

// EMEstimatorCallback::computeError + the threshold test for one correspondence: cv stores the Sampson distance
// s^2 / d (float64) as float32 and keeps the point iff that float <= float32(threshold^2).  (float)q <= t  <=>  q <= M for the
// double M the host derives from t (the midpoint to the next float, or the double below it, by the parity of t: round to
// nearest even), and fl(s^2 / d) <= M is decided by one multiplication unless s^2 lies within 2^-50 of M d: only then is the
// division carried out.  Same decisions, bit for bit; the division was a third of the counting loop.
__device__ __forceinline__ bool sampson_inlier(const double* E, double x1, double y1, double x2, double y2, float thr2, double M) {
    const double a0 = E[0] * x1 + E[1] * y1 + E[2], a1 = E[3] * x1 + E[4] * y1 + E[5], a2 = E[6] * x1 + E[7] * y1 + E[8];
    const double b0 = E[0] * x2 + E[3] * y2 + E[6], b1 = E[1] * x2 + E[4] * y2 + E[7];
    const double s = x2 * a0 + y2 * a1 + a2;
    const double num = s * s, den = a0 * a0 + a1 * a1 + b0 * b0 + b1 * b1;
    const double p = M * den;
    if (num < p * (1.0 - 0x1p-50)) return true;
    if (num > p * (1.0 + 0x1p-50)) return false;
    return (float)(num / den) <= thr2;            // the boundary, a zero or a non-finite denominator: exactly as cv computes it
}

__device__ int update_num_iters5(double p, double ep, int max_iters) {      // RANSACUpdateNumIters, modelPoints = 5
    p = fmin(fmax(p, 0.0), 1.0);
    ep = fmin(fmax(ep, 0.0), 1.0);
    double num = fmax(1.0 - p, DBL_MIN);
    const double g = 1.0 - ep;
    double denom = 1.0 - g * g * g * g * g;
    if (denom < DBL_MIN) return 0;
    num = log(num);
    denom = log(denom);
    return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : (int)rint(num / denom);
}

struct ESmem {
    int tup[HYP][5];
    int nmodels[HYP];
    int cnt[HYP][10];
    double bestE[9];
    double cam[8];               // fx1, fy1, cx1, cy1, fx2, fy2, cx2, cy2
    int flag, best_cnt, iters_run;
    u64 rng;
};

constexpr int SUB = 16;          // hypotheses whose inliers are counted between two replays of cv's decisions

__global__ void __launch_bounds__(ET, 1)
ransac_essential_kernel(const float4* __restrict__ corr, const long long* __restrict__ ptr, int num_pairs,
                        const double* __restrict__ cam, float thr2, double thrM, double prob, int max_iters, double* __restrict__ E_out,
                        uint8_t* __restrict__ mask_out, int32_t* __restrict__ inliers_out, int32_t* __restrict__ iters_out) {
    __shared__ ESmem sm;
    extern __shared__ double scratch[];          // [SCR][HYP]
    const int pair = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    const WS w{scratch + (tid < HYP ? tid : 0)};
    const long long beg = ptr[pair];
    const int count = (int)(ptr[pair + 1] - beg);
    const float4* pts = corr + beg;
    uint8_t* mask = mask_out + beg;
    double* Eg = E_out + (size_t)pair * 9;
    if (tid < 8) {
        // cv2.undistortPoints without distortion: x' = (x - cx) * (1 / fx); the reciprocal is formed once, as OpenCV does
        const double v = cam ? cam[(size_t)pair * 8 + tid] : ((tid & 3) < 2 ? 1.0 : 0.0);
        sm.cam[tid] = (tid & 3) < 2 ? 1.0 / v : v;
    }
    __syncthreads();
    const double ifx1 = sm.cam[0], ify1 = sm.cam[1], cx1 = sm.cam[2], cy1 = sm.cam[3];
    const double ifx2 = sm.cam[4], ify2 = sm.cam[5], cx2 = sm.cam[6], cy2 = sm.cam[7];
    auto norm_pt = [&](float4 p, double& x1, double& y1, double& x2, double& y2) {
        x1 = ((double)p.x - cx1) * ifx1; y1 = ((double)p.y - cy1) * ify1;
        x2 = ((double)p.z - cx2) * ifx2; y2 = ((double)p.w - cy2) * ify2;
    };
    auto fail = [&]() {          // E = None
        for (int i = tid; i < count; i += ET) mask[i] = 0;
        if (tid < 9) Eg[tid] = CUDART_NAN;
        if (tid == 0) { inliers_out[pair] = -1; if (iters_out) iters_out[pair] = 0; }
    };
    if (count < 5) { fail(); return; }
    auto solve_tuple = [&](const int* idx) -> int {
        double m[5][4];
#pragma unroll
        for (int i = 0; i < 5; ++i) norm_pt(__ldg(pts + idx[i]), m[i][0], m[i][1], m[i][2], m[i][3]);
        return five_point(w, m);
    };

    if (count == 5) {            // exactly modelPoints: one kernel run, its first model, every point an inlier (ptsetreg.cpp)
        if (tid == 0) {
            const int idx[5] = {0, 1, 2, 3, 4};
            sm.flag = solve_tuple(idx);
            if (sm.flag > 0)
                for (int k = 0; k < 9; ++k) sm.bestE[k] = w(OFF_M + k);
        }
        __syncthreads();
        if (sm.flag <= 0) { fail(); return; }
        if (tid < 9) Eg[tid] = sm.bestE[tid];
        if (tid < 5) mask[tid] = 1;
        if (tid == 0) { inliers_out[pair] = 5; if (iters_out) iters_out[pair] = 0; }
        return;
    }

    int niters = max(max_iters, 1), iter = 0, max_good = 0;      // thread 0 only
    if (tid == 0) { sm.rng = 0xFFFFFFFFFFFFFFFFull; sm.flag = 0; sm.best_cnt = 0; }
    const u32 recip = 0xFFFFFFFFu / (u32)count;
    __syncthreads();
    for (;;) {
        if (tid == 0) {                         // the next HYP samples of cv's random sequence
            u64 s = sm.rng;
#pragma unroll 1
            for (int t = 0; t < HYP; ++t) {
                int id[5];
#pragma unroll
                for (int i = 0; i < 5; ++i) {
                    int v;
                    bool dup;
                    do {
                        const u32 x = rng_next(s);
                        u32 r = x - __umulhi(x, recip) * (u32)count;       // x % count: the estimate is low by at most 2
                        while (r >= (u32)count) r -= (u32)count;
                        v = (int)r;
                        dup = false;
#pragma unroll
                        for (int j = 0; j < 5; ++j) dup |= j < i && id[j] == v;
                    } while (dup);
                    id[i] = v;
                }
#pragma unroll
                for (int i = 0; i < 5; ++i) sm.tup[t][i] = id[i];
            }
            sm.rng = s;
        }
        __syncthreads();
        if (tid < HYP) sm.nmodels[tid] = solve_tuple(sm.tup[tid]);
        __syncthreads();
        for (int h0 = 0; h0 < HYP; h0 += SUB) {
            for (int k = tid; k < SUB * 10; k += ET) sm.cnt[h0 + k / 10][k % 10] = 0;
            __syncthreads();
            for (int i0 = 0; i0 < count; i0 += ET) {
                const int i = i0 + tid;
                const bool in = i < count;
                double x1 = 0, y1 = 0, x2 = 0, y2 = 0;
                if (in) norm_pt(__ldg(pts + i), x1, y1, x2, y2);
                for (int h = h0; h < h0 + SUB; ++h) {
                    const int nm = sm.nmodels[h];
                    for (int m = 0; m < nm; ++m) {
                        double E[9];
#pragma unroll
                        for (int k = 0; k < 9; ++k) E[k] = scratch[(OFF_M + 9 * m + k) * HYP + h];
                        const bool good = in && sampson_inlier(E, x1, y1, x2, y2, thr2, thrM);
                        const int c = __popc(__ballot_sync(0xffffffffu, good));
                        if (lane == 0 && c) atomicAdd(&sm.cnt[h][m], c);
                    }
                }
            }
            __syncthreads();
            if (tid == 0) {                     // cv's decisions, in cv's order
                int done = 0;
                for (int h = h0; h < h0 + SUB; ++h) {
                    if (iter >= niters) { done = 1; break; }
                    for (int m = 0; m < sm.nmodels[h]; ++m) {
                        const int good = sm.cnt[h][m];
                        if (good > max(max_good, 4)) {
                            max_good = good;
                            for (int k = 0; k < 9; ++k) sm.bestE[k] = scratch[(OFF_M + 9 * m + k) * HYP + h];
                            niters = update_num_iters5(prob, (double)(count - good) / count, niters);
                        }
                    }
                    ++iter;
                }
                if (!done && iter >= niters) done = 1;
                sm.flag = done;
                sm.best_cnt = max_good;
                sm.iters_run = iter;
            }
            __syncthreads();
            if (sm.flag) break;
        }
        if (sm.flag) break;
    }
    if (sm.best_cnt == 0) { fail(); return; }
    double E[9];
    for (int k = 0; k < 9; ++k) E[k] = sm.bestE[k];
    for (int i = tid; i < count; i += ET) {
        double x1, y1, x2, y2;
        norm_pt(__ldg(pts + i), x1, y1, x2, y2);
        mask[i] = sampson_inlier(E, x1, y1, x2, y2, thr2, thrM);
    }
    if (tid < 9) Eg[tid] = E[tid];
    if (tid == 0) { inliers_out[pair] = sm.best_cnt; if (iters_out) iters_out[pair] = sm.iters_run; }
}

}  // namespace

extern "C" {

int b200m_ransac_essential(const float* corr_xyxy, const int64_t* ptr, int num_pairs, const double* cam, double threshold,
                           double prob, int max_iters, double* E_out, uint8_t* mask_out, int32_t* inliers_out,
                           int32_t* iters_out, void* stream) {
    if (num_pairs < 0) return B200M_EINVAL;
    if (num_pairs == 0) return B200M_OK;
    if (!corr_xyxy || !ptr || !E_out || !mask_out || !inliers_out) return B200M_EINVAL;
    if (!(prob > 0.0 && prob < 1.0) || !(threshold > 0.0)) return B200M_EINVAL;      // CV_Assert in RANSACPointSetRegistrator::run
    if ((uintptr_t)corr_xyxy & 15) return B200M_EINVAL;
    constexpr size_t dyn = (size_t)SCR * HYP * sizeof(double);
    static B200mPerDevice once;                   // the shared-memory opt-in is per device
    int dev = 0;
    if (!b200m_current_device(&dev, nullptr)) return B200M_ENODEVICE;
    if (!once.done[dev]) {
        B200M_CUDA_CHECK(cudaFuncSetAttribute(ransac_essential_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn),
                         "ransac_essential_kernel smem attribute");
        once.done[dev] = 1;
    }
    // (float)q <= t  <=>  q <= M: M = the midpoint between t and the next float if t's mantissa is even (a tie rounds to t),
    // else the double just below that midpoint
    const float t = (float)(threshold * threshold);
    if (!(t > 0.0f) || !isfinite(t)) return B200M_EINVAL;
    const float tn = nextafterf(t, INFINITY);
    uint32_t tbits;
    memcpy(&tbits, &t, 4);
    double M = 0.5 * ((double)t + (double)tn);
    if (tbits & 1u) M = nextafter(M, 0.0);
    ransac_essential_kernel<<<(unsigned)num_pairs, ET, dyn, (cudaStream_t)stream>>>(
        (const float4*)corr_xyxy, (const long long*)ptr, num_pairs, cam, t, M, prob, max_iters, E_out,
        mask_out, inliers_out, iters_out);
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "ransac_essential_kernel launch");
    return B200M_OK;
}

}  // extern "C"
