// ransac.cu — geometric verification of matched pairs: batched replacement of
//     H, mask = cv2.findHomography(pts1, pts2, cv2.RANSAC, thr, confidence=conf, maxIters=max_iters)
// as called per image pair by /root/reference/FeatureMatchingExtraction/utils.py:141 (enhanced_filter_matches_with_homography).
// Input is what the matcher's hand-off already holds in HBM: the CSR `ptr` over pairs and the float32 [K, 4]
// correspondences (x1, y1, x2, y2) of b200m_gather_correspondences.  One CTA per image pair.
//
// cv2's RANSAC is a deterministic sequential program (its RNG restarts from the same state on every call), so the
// result is reproducible and parity is testable; the kernel keeps that sequence and parallelises inside it:
//   * the random stream is a geometry-independent sequence of 4-tuples (distinct indices by rejection); tuples that
//     fail checkSubset are skipped, the others are hypotheses 0, 1, 2 ... .  One thread draws TUPLES tuples per
//     round, TUPLES threads test them and fit their models side by side (normalised DLT, 9x9 Jacobi in fp64, the
//     rotation order of OpenCV's solver), all threads count inliers of all models of the round (float32 error, no
//     fused multiply-add: the comparison against thr^2 must see the bits cv2 sees), then one thread replays the
//     sequential decisions in order: "strictly more inliers wins", the shrinking iteration limit, the stop.
//     Models drawn past the stop are discarded, exactly as if they had never been drawn.
//   * after the loop: DLT over the winner's inliers and <= 10 Levenberg-Marquardt iterations on 8 parameters,
//     normal equations by block reduction, the 8x8 eigen-solve on one thread; the returned mask is recomputed from
//     the refined H (what cv2 4.13 returns).
// fp64 sums run in a different order than OpenCV's serial loops: H agrees to ~1e-9 relative, the mask bit for bit
// unless a point's error sits within rounding of the threshold (tests count such cases).
// Compiled with -fmad=false (build.py): a*b+c must round twice, as in OpenCV's SSE2 baseline build.
#include <float.h>
#include <math_constants.h>
#include "common.cuh"

namespace {

#ifndef RANSAC_TUPLES
#define RANSAC_TUPLES 64
#endif
#ifndef RANSAC_MINB
#define RANSAC_MINB 2
#endif
constexpr int RT = 128;          // threads per CTA
constexpr int NW = RT / 32;
constexpr int TUPLES = RANSAC_TUPLES;   // 4-tuples drawn per round
constexpr int GMAX = 512;        // tuples drawn per batch
constexpr int TS = TUPLES;       // stride between the elements of one thread's scratch matrix
constexpr int NRED = 45;         // sums per block reduction: upper triangle of 9x9, or 36 + 8 + 1 for the LM normal equations

struct LmState {                 // thread 0's state across the refinement (shared memory)
    double A[64], v[8], D[8], d[8], S, lam, lc, rmax;
    int iter;
};

struct Smem {
    double red[NW][NRED + 1];    // per-warp partial sums (+1: max slot)
    double tot[NRED + 1];
    double Hd[TUPLES][9];        // models of the round
    float Hf[TUPLES][8];
    alignas(16) int tup[GMAX][4];   // tuples of the batch
    short hlist[GMAX];           // the ones that passed checkSubset, in order
    uint8_t pass[GMAX];
    int npass;
    int cnt[TUPLES];
    int state[TUPLES];           // 1 = hypothesis without a model (no spread), 2 = hypothesis with a model
    double bestH[9];
    double x[8], xd[8];
    int flag, best_cnt, iters_run, n_inl;
    u64 rng;
    LmState lm;
};

__device__ __forceinline__ u32 rng_next(u64& s) {
    s = (u64)(u32)s * 4164903690ull + (s >> 32);
    return (u32)s;
}

__device__ __forceinline__ double hypot2(double a, double b) {      // scaled form, as the solver being matched
    a = fabs(a); b = fabs(b);
    if (a > b) { b /= a; return a * sqrt(1.0 + b * b); }
    if (b > 0) { a /= b; return b * sqrt(1.0 + a * a); }
    return 0.0;
}

// Per-thread scratch matrices live in SHARED memory, element-major: element e of thread t's matrix is at
// base[e * TS + t] (TS = TUPLES threads side by side), so lanes with different pivots still hit different banks.
// (A first version kept them in local memory: 2.6 KB per thread thrashed L1, ncu showed 2.5 GB of DRAM writes.)
template <int N> struct Scratch {
    double* m;                                    // this thread's column
    __device__ __forceinline__ double& A(int i, int j) const { return m[(i * N + j) * TS]; }
    __device__ __forceinline__ double& V(int i, int j) const { return m[(N * N + i * N + j) * TS]; }
    __device__ __forceinline__ double& W(int i) const { return m[(2 * N * N + i) * TS]; }
};
constexpr int SCRATCH_DOUBLES = 2 * 81 + 9;       // per thread, N = 9

__device__ __forceinline__ int nib_get(u64 w, int i) { return (int)((w >> (4 * i)) & 15u); }
__device__ __forceinline__ u64 nib_set(u64 w, int i, int v) { return (w & ~(15ull << (4 * i))) | ((u64)v << (4 * i)); }

// Symmetric eigen-decomposition, Jacobi with the pivot = largest off-diagonal element tracked through per-row and
// per-column argmax indices: the rotation order of OpenCV's solver (core/src/lapack.cpp), so that the eigenvector of
// a (numerically) rank-deficient L^T L comes out as cv2's does.  Upper triangle of A used and destroyed, eigenvectors
// in the ROWS of V, eigenvalues in W (unsorted; sorted_index() gives the order OpenCV's selection sort would).
// All loops have fixed trip counts and predicated bodies: lanes of a warp work on different matrices with
// different pivots.  The index arrays are nibbles of two registers.
template <int N>
__device__ void jacobi_eig(const Scratch<N>& S) {
    u64 indR = 0, indC = 0;
    auto row_arg = [&](int k) {                    // first maximum of |A[k][i]|, i > k
        int m = k + 1;
        double mv = -1.0;
#pragma unroll
        for (int i = 1; i < N; ++i)
            if (i > k) { const double v = fabs(S.A(k, i)); if (mv < v) { mv = v; m = i; } }
        return m;
    };
    auto col_arg = [&](int k) {                    // first maximum of |A[i][k]|, i < k
        int m = 0;
        double mv = -1.0;
#pragma unroll
        for (int i = 0; i < N - 1; ++i)
            if (i < k) { const double v = fabs(S.A(i, k)); if (mv < v) { mv = v; m = i; } }
        return m;
    };
#pragma unroll 1
    for (int i = 0; i < N; ++i) {
#pragma unroll
        for (int j = 0; j < N; ++j) S.V(i, j) = i == j ? 1.0 : 0.0;
        S.W(i) = S.A(i, i);
        if (i < N - 1) indR = nib_set(indR, i, row_arg(i));
        if (i > 0) indC = nib_set(indC, i, col_arg(i));
    }
#pragma unroll 1
    for (int it = 0; it < N * N * 30; ++it) {
        int k = 0;
        double mv = fabs(S.A(0, nib_get(indR, 0)));
#pragma unroll
        for (int i = 1; i < N - 1; ++i) { const double v = fabs(S.A(i, nib_get(indR, i))); if (mv < v) { mv = v; k = i; } }
        int l = nib_get(indR, k);
#pragma unroll
        for (int i = 1; i < N; ++i) {
            const int ci = nib_get(indC, i);
            const double v = fabs(S.A(ci, i));
            if (mv < v) { mv = v; k = ci; l = i; }
        }
        const double p = S.A(k, l);
        if (fabs(p) <= DBL_EPSILON) break;
        const double y = (S.W(l) - S.W(k)) * 0.5;
        double t = fabs(y) + hypot2(p, y);
        double s = hypot2(p, t);
        const double c = t / s;
        s = p / s;
        t = (p / t) * p;
        if (y < 0) { s = -s; t = -t; }
        S.A(k, l) = 0.0;
        S.W(k) -= t;
        S.W(l) += t;
#pragma unroll
        for (int i = 0; i < N; ++i) {
            {   // eigenvector rows k and l
                double& u = S.V(k, i); double& w = S.V(l, i);
                const double a0 = u, b0 = w;
                u = a0 * c - b0 * s; w = a0 * s + b0 * c;
            }
            if (i != k && i != l) {                // the upper-triangle elements that pair i with k and with l
                double& u = S.A(min(i, k), max(i, k)); double& w = S.A(min(i, l), max(i, l));
                const double a0 = u, b0 = w;
                u = a0 * c - b0 * s; w = a0 * s + b0 * c;
            }
        }
        if (k < N - 1) indR = nib_set(indR, k, row_arg(k));
        if (k > 0) indC = nib_set(indC, k, col_arg(k));
        if (l < N - 1) indR = nib_set(indR, l, row_arg(l));
        if (l > 0) indC = nib_set(indC, l, col_arg(l));
    }
}

// position (row of V) that OpenCV's descending selection sort leaves LAST: the eigenvector DLT keeps
template <int N>
__device__ int last_after_sort(const Scratch<N>& S) {
    u64 perm = 0;
    for (int i = 0; i < N; ++i) perm = nib_set(perm, i, i);
    for (int k = 0; k < N - 1; ++k) {
        int m = k;
        for (int i = k + 1; i < N; ++i) if (S.W(m) < S.W(i)) m = i;
        if (k != m) {
            const double tw = S.W(m); S.W(m) = S.W(k); S.W(k) = tw;
            const int pm = nib_get(perm, m), pk = nib_get(perm, k);
            perm = nib_set(nib_set(perm, m, pk), k, pm);
        }
    }
    return nib_get(perm, N - 1);
}

// rows of the DLT system for one normalised correspondence, accumulated into the upper triangle (45 sums, row-major j <= k)
__device__ __forceinline__ void ltl_accumulate(double* acc, double X, double Y, double x, double y) {
    const double Lx[9] = {X, Y, 1.0, 0.0, 0.0, 0.0, -x * X, -x * Y, -x};
    const double Ly[9] = {0.0, 0.0, 0.0, X, Y, 1.0, -y * X, -y * Y, -y};
    int o = 0;
#pragma unroll
    for (int j = 0; j < 9; ++j)
#pragma unroll
        for (int k = j; k < 9; ++k) acc[o++] += Lx[j] * Lx[k] + Ly[j] * Ly[k];
}

// A (upper triangle filled) -> eigenvector of the smallest eigenvalue, de-normalised: H = inv(Hnorm_dst) * H0 * Hnorm_src, h33 = 1
__device__ void dlt_finish(const Scratch<9>& S, const double* nrm, double* H) {
    jacobi_eig<9>(S);
    const int r8 = last_after_sort<9>(S);
    double h0[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) h0[i] = S.V(r8, i);
    const double cMx = nrm[0], cMy = nrm[1], cmx = nrm[2], cmy = nrm[3], sMx = nrm[4], sMy = nrm[5], smx = nrm[6], smy = nrm[7];
    const double inv[9] = {1.0 / smx, 0.0, cmx, 0.0, 1.0 / smy, cmy, 0.0, 0.0, 1.0};
    const double n2[9] = {sMx, 0.0, -cMx * sMx, 0.0, sMy, -cMy * sMy, 0.0, 0.0, 1.0};
    double T[9], R[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) T[r * 3 + c] = inv[r * 3] * h0[c] + inv[r * 3 + 1] * h0[3 + c] + inv[r * 3 + 2] * h0[6 + c];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) R[r * 3 + c] = T[r * 3] * n2[c] + T[r * 3 + 1] * n2[3 + c] + T[r * 3 + 2] * n2[6 + c];
    const double sc = 1.0 / R[8];
#pragma unroll
    for (int i = 0; i < 9; ++i) H[i] = R[i] * sc;
}

// normalised DLT of exactly four correspondences on one thread; false when a coordinate has no spread
__device__ bool dlt_four(const Scratch<9>& S, const float4* p, double* H) {
    double cMx = 0, cMy = 0, cmx = 0, cmy = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) { cmx += p[i].z; cmy += p[i].w; cMx += p[i].x; cMy += p[i].y; }
    cmx /= 4; cmy /= 4; cMx /= 4; cMy /= 4;
    double smx = 0, smy = 0, sMx = 0, sMy = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        smx += fabs(p[i].z - cmx); smy += fabs(p[i].w - cmy);
        sMx += fabs(p[i].x - cMx); sMy += fabs(p[i].y - cMy);
    }
    if (fabs(smx) < DBL_EPSILON || fabs(smy) < DBL_EPSILON || fabs(sMx) < DBL_EPSILON || fabs(sMy) < DBL_EPSILON) return false;
    smx = 4 / smx; smy = 4 / smy; sMx = 4 / sMx; sMy = 4 / sMy;
    double tri[NRED];
#pragma unroll
    for (int i = 0; i < NRED; ++i) tri[i] = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i)
        ltl_accumulate(tri, (p[i].x - cMx) * sMx, (p[i].y - cMy) * sMy, (p[i].z - cmx) * smx, (p[i].w - cmy) * smy);
    int o = 0;
#pragma unroll
    for (int j = 0; j < 9; ++j)
#pragma unroll
        for (int k = j; k < 9; ++k) S.A(j, k) = tri[o++];
    const double nrm[8] = {cMx, cMy, cmx, cmy, sMx, sMy, smx, smy};
    dlt_finish(S, nrm, H);
    return true;
}

__device__ __forceinline__ bool collinear_last(const double* x, const double* y) {
    // only the LAST of four points is tested against the pairs before it
    for (int j = 0; j < 3; ++j) {
        const double dx1 = x[j] - x[3], dy1 = y[j] - y[3];
        for (int k = 0; k < j; ++k) {
            const double dx2 = x[k] - x[3], dy2 = y[k] - y[3];
            if (fabs(dx2 * dy1 - dy2 * dx1) <= (double)FLT_EPSILON * (fabs(dx1) + fabs(dy1) + fabs(dx2) + fabs(dy2))) return true;
        }
    }
    return false;
}

__device__ __forceinline__ double det_tri(const double* x, const double* y, int a, int b, int c) {
    // determinant of [[xa ya 1] [xb yb 1] [xc yc 1]], cofactor expansion along the first row
    return x[a] * (y[b] - y[c]) - y[a] * (x[b] - x[c]) + (x[b] * y[c] - y[b] * x[c]);
}

__device__ bool subset_ok(const float4* p) {
    double sx[4], sy[4], dx[4], dy[4];
    for (int i = 0; i < 4; ++i) { sx[i] = p[i].x; sy[i] = p[i].y; dx[i] = p[i].z; dy[i] = p[i].w; }
    if (collinear_last(sx, sy) || collinear_last(dx, dy)) return false;
    const int tt[4][3] = {{0, 1, 2}, {1, 2, 3}, {0, 2, 3}, {0, 1, 3}};
    int negative = 0;
    for (int i = 0; i < 4; ++i)
        negative += det_tri(sx, sy, tt[i][0], tt[i][1], tt[i][2]) * det_tri(dx, dy, tt[i][0], tt[i][1], tt[i][2]) < 0;
    return negative == 0 || negative == 4;
}

// float32 reprojection error, every operation rounded on its own
__device__ __forceinline__ float reproj_err(const float* h, float4 p) {
    const float ww = __fdiv_rn(1.0f, __fadd_rn(__fadd_rn(__fmul_rn(h[6], p.x), __fmul_rn(h[7], p.y)), 1.0f));
    const float dx = __fsub_rn(__fmul_rn(__fadd_rn(__fadd_rn(__fmul_rn(h[0], p.x), __fmul_rn(h[1], p.y)), h[2]), ww), p.z);
    const float dy = __fsub_rn(__fmul_rn(__fadd_rn(__fadd_rn(__fmul_rn(h[3], p.x), __fmul_rn(h[4], p.y)), h[5]), ww), p.w);
    return __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
}

__device__ int update_num_iters(double p, double ep, int max_iters) {
    p = fmin(fmax(p, 0.0), 1.0);
    ep = fmin(fmax(ep, 0.0), 1.0);
    double num = fmax(1.0 - p, DBL_MIN);
    double denom = 1.0 - pow(1.0 - ep, 4.0);
    if (denom < DBL_MIN) return 0;
    num = log(num);
    denom = log(denom);
    return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : (int)rint(num / denom);
}

// block-wide sums of nv <= NRED doubles per thread (fixed tree: run-to-run deterministic) and one max; result in sm.tot
template <int NV>
__device__ void block_reduce(Smem& sm, double* acc, double mx) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        double v = acc[i];
#pragma unroll
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) sm.red[warp][i] = v;
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) sm.red[warp][NRED] = mx;
    __syncthreads();
    if (threadIdx.x < NV) {
        double v = 0;
        for (int w = 0; w < NW; ++w) v += sm.red[w][threadIdx.x];
        sm.tot[threadIdx.x] = v;
    }
    if (threadIdx.x == 32) {
        double v = 0;
        for (int w = 0; w < NW; ++w) v = fmax(v, sm.red[w][NRED]);
        sm.tot[NRED] = v;
    }
    __syncthreads();
}

// residuals and normal equations of the refinement at parameters h over the inliers: acc[0..35] upper triangle of
// J^T J, acc[36..43] J^T r, acc[44] |r|^2; returns max |r|
__device__ __forceinline__ double lm_accumulate(double* acc, const double* h, float4 p) {
    const double Mx = p.x, My = p.y;
    double ww = h[6] * Mx + h[7] * My + 1.0;
    ww = fabs(ww) > DBL_EPSILON ? 1.0 / ww : 0.0;
    const double xi = (h[0] * Mx + h[1] * My + h[2]) * ww, yi = (h[3] * Mx + h[4] * My + h[5]) * ww;
    const double rx = xi - p.z, ry = yi - p.w;
    const double Jx[8] = {Mx * ww, My * ww, ww, 0.0, 0.0, 0.0, -Mx * ww * xi, -My * ww * xi};
    const double Jy[8] = {0.0, 0.0, 0.0, Mx * ww, My * ww, ww, -Mx * ww * yi, -My * ww * yi};
    int o = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
        for (int k = j; k < 8; ++k) acc[o++] += Jx[j] * Jx[k] + Jy[j] * Jy[k];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[36 + j] += Jx[j] * rx + Jy[j] * ry;
    acc[44] += rx * rx + ry * ry;
    return fmax(fabs(rx), fabs(ry));
}


__device__ void unpack_normal(const double* tot, double* A, double* v) {
    int o = 0;
    for (int j = 0; j < 8; ++j)
        for (int k = j; k < 8; ++k) { A[j * 8 + k] = tot[o]; A[k * 8 + j] = tot[o]; ++o; }
    for (int j = 0; j < 8; ++j) v[j] = tot[36 + j];
}

// x = (A + lam diag(D))^-1 b by Cholesky, fully unrolled in registers; false if a pivot is not safely positive.
// cv2 solves the same system through an eigen-decomposition (DECOMP_EIG); for the well-conditioned normal equations
// of the refinement both are accurate to a few ulps of the solution, and the LM iteration converges to the same
// minimiser.  The 8 x 8 Jacobi on one thread (eig_solve8) cost ~0.2 ms per LM step with every other thread of the
// CTA waiting: ncu showed 29 % of all stall samples at that barrier.  It stays as the fallback and for the diagonal
// of the inverse.
__device__ bool chol_solve8(const double* A, double lam, const double* D, const double* b, double* x) {
    double Lm[8][8], y[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        double d = A[j * 8 + j] + lam * D[j];
#pragma unroll
        for (int k = 0; k < 8; ++k) if (k < j) d -= Lm[j][k] * Lm[j][k];
        if (!(d > 1e-300)) return false;
        const double r = sqrt(d);
        Lm[j][j] = r;
#pragma unroll
        for (int i = 0; i < 8; ++i)
            if (i > j) {
                double t = A[i * 8 + j];
#pragma unroll
                for (int k = 0; k < 8; ++k) if (k < j) t -= Lm[i][k] * Lm[j][k];
                Lm[i][j] = t / r;
            }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        double t = b[i];
#pragma unroll
        for (int k = 0; k < 8; ++k) if (k < i) t -= Lm[i][k] * y[k];
        y[i] = t / Lm[i][i];
    }
#pragma unroll
    for (int i = 7; i >= 0; --i) {
        double t = y[i];
#pragma unroll
        for (int k = 0; k < 8; ++k) if (k > i) t -= Lm[k][i] * x[k];
        x[i] = t / Lm[i][i];
    }
    bool ok = true;
#pragma unroll
    for (int i = 0; i < 8; ++i) ok &= isfinite(x[i]);
    return ok;
}

// x = (A + lam diag(D))^-1 b through the eigen-decomposition (cv::solve(..., DECOMP_EIG)); optionally the diagonal of the inverse
__device__ void eig_solve8(const Scratch<8>& S, const double* A, double lam, const double* D, const double* b, double* x,
                           double* inv_diag) {
    for (int i = 0; i < 8; ++i)
        for (int j = i; j < 8; ++j) S.A(i, j) = A[i * 8 + j] + (i == j ? lam * D[i] : 0.0);
    jacobi_eig<8>(S);
    double y[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        double t = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) t += S.V(i, j) * b[j];
        const double w = S.W(i);
        y[i] = fabs(w) > DBL_EPSILON ? t / w : 0.0;
    }
    if (x)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            double t = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) t += S.V(i, j) * y[i];
            x[j] = t;
        }
    if (inv_diag)
        for (int j = 0; j < 8; ++j) {
            double t = 0;
            for (int i = 0; i < 8; ++i) { const double w = S.W(i), e = S.V(i, j); t += fabs(w) > DBL_EPSILON ? e * e / w : 0.0; }
            inv_diag[j] = t;
        }
}

__global__ void __launch_bounds__(RT, RANSAC_MINB)
ransac_homography_kernel(const float4* __restrict__ corr, const long long* __restrict__ ptr, int num_pairs, float thr2,
                         double confidence, int max_iters, double* __restrict__ H_out, uint8_t* __restrict__ mask_out,
                         int32_t* __restrict__ inliers_out, int32_t* __restrict__ iters_out) {
    __shared__ Smem sm;
    extern __shared__ double scratch[];          // [SCRATCH_DOUBLES][TUPLES]: one Jacobi workspace per hypothesis thread
    const Scratch<9> S9{scratch + (threadIdx.x < TUPLES ? threadIdx.x : 0)};
    const Scratch<8> S8{scratch};
    const int pair = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    const long long beg = ptr[pair];
    const int count = (int)(ptr[pair + 1] - beg);
    const float4* pts = corr + beg;
    uint8_t* mask = mask_out + beg;
    double* Hg = H_out + (size_t)pair * 9;

    auto fail = [&]() {          // H = None
        for (int i = tid; i < count; i += RT) mask[i] = 0;
        if (tid < 9) Hg[tid] = CUDART_NAN;
        if (tid == 0) { inliers_out[pair] = -1; if (iters_out) iters_out[pair] = 0; }
    };
    if (count < 4) { fail(); return; }       // the reference returns ([], None) before calling cv2 (utils.py:135-136)

    if (count == 4) {
        if (tid == 0) {
            float4 p[4] = {pts[0], pts[1], pts[2], pts[3]};
            double H[9];
            sm.flag = dlt_four(S9, p, H);
            for (int i = 0; i < 9; ++i) sm.bestH[i] = H[i];
        }
        __syncthreads();
        if (!sm.flag) { fail(); return; }
        if (tid < 9) Hg[tid] = sm.bestH[tid];
        if (tid < 4) mask[tid] = 1;
        if (tid == 0) { inliers_out[pair] = 4; if (iters_out) iters_out[pair] = 0; }
        return;
    }

    // ---------------------------------------------------------------- hypothesis loop
    // Batches of tuples (128 in the first, GMAX afterwards: easy pairs stop within the first few hypotheses); only the
    // tuples that pass checkSubset (often ~10 %) become hypotheses, compacted in order into sub-rounds of TUPLES models.
    int niters = max(max_iters, 1), iter = 0, max_good = 0, fails = 0;      // thread 0 only
    if (tid == 0) { sm.rng = 0xFFFFFFFFFFFFFFFFull; sm.flag = 0; sm.best_cnt = 0; }
    const u32 recip = 0xFFFFFFFFu / (u32)count;
    int G = 128;
    __syncthreads();
    for (;;) {
        if (tid == 0) {
            u64 s = sm.rng;
#pragma unroll 1
            for (int t = 0; t < G; ++t) {
                int id[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    int v;
                    bool dup;
                    do {
                        const u32 x = rng_next(s);
                        u32 r = x - __umulhi(x, recip) * (u32)count;       // x % count: the estimate is low by at most 2
                        while (r >= (u32)count) r -= (u32)count;
                        v = (int)r;
                        dup = false;
#pragma unroll
                        for (int j = 0; j < 4; ++j) dup |= j < i && id[j] == v;
                    } while (dup);
                    id[i] = v;
                }
                *(int4*)sm.tup[t] = make_int4(id[0], id[1], id[2], id[3]);
            }
            sm.rng = s;
        }
        __syncthreads();
        for (int t = tid; t < G; t += RT) {
            float4 p[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) p[i] = __ldg(pts + sm.tup[t][i]);
            sm.pass[t] = subset_ok(p);
        }
        __syncthreads();
        if (tid < 32) {                         // ordered compaction: hlist[k] = k-th passing tuple of the batch
            int base = 0;
            for (int c = 0; c < G; c += 32) {
                const bool f = sm.pass[c + lane];
                const u32 bal = __ballot_sync(0xffffffffu, f);
                if (f) sm.hlist[base + __popc(bal & ((1u << lane) - 1u))] = (short)(c + lane);
                base += __popc(bal);
            }
            if (lane == 0) sm.npass = base;
        }
        __syncthreads();
        const int npass = sm.npass;
        int prev_tuple = -1;                    // thread 0: last tuple of the batch already replayed
        for (int h0 = 0; h0 < npass; h0 += TUPLES) {
            const int nh = min(TUPLES, npass - h0);
            if (tid < TUPLES) sm.cnt[tid] = 0;
            if (tid < nh) {
                const int t = sm.hlist[h0 + tid];
                float4 p[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) p[i] = __ldg(pts + sm.tup[t][i]);
                double H[9];
                int st = 1;
                if (dlt_four(S9, p, H)) {
                    st = 2;
#pragma unroll
                    for (int i = 0; i < 9; ++i) sm.Hd[tid][i] = H[i];
#pragma unroll
                    for (int i = 0; i < 8; ++i) sm.Hf[tid][i] = (float)H[i];
                }
                sm.state[tid] = st;
            }
            __syncthreads();
            for (int i0 = 0; i0 < count; i0 += RT) {
                const int i = i0 + tid;
                const bool in = i < count;
                const float4 p = in ? __ldg(pts + i) : make_float4(0, 0, 0, 0);
                for (int h = 0; h < nh; ++h) {
                    if (sm.state[h] != 2) continue;
                    const bool good = in && reproj_err(sm.Hf[h], p) <= thr2;
                    const int c = __popc(__ballot_sync(0xffffffffu, good));
                    if (lane == 0 && c) atomicAdd(&sm.cnt[h], c);
                }
            }
            __syncthreads();
            if (tid == 0) {
                int done = 0;
                for (int h = 0; h < nh && !done; ++h) {
                    if (iter >= niters) { done = 1; break; }
                    const int t = sm.hlist[h0 + h];
                    fails += t - prev_tuple - 1;                 // tuples that failed checkSubset before this one
                    prev_tuple = t;
                    if (fails >= 10000) { done = iter == 0 ? 2 : 1; break; }     // getSubset gave up
                    fails = 0;
                    if (sm.state[h] == 2) {
                        const int good = sm.cnt[h];
                        if (good > max(max_good, 3)) {
                            max_good = good;
                            for (int i = 0; i < 9; ++i) sm.bestH[i] = sm.Hd[h][i];
                            niters = update_num_iters(confidence, (double)(count - good) / count, niters);
                        }
                    }
                    ++iter;
                }
                if (!done && iter >= niters) done = 1;
                sm.flag = done;
            }
            __syncthreads();
            if (sm.flag) break;
        }
        if (tid == 0 && !sm.flag) {             // the batch's trailing failures
            fails += G - 1 - prev_tuple;
            if (fails >= 10000) sm.flag = iter == 0 ? 2 : 1;
        }
        if (tid == 0) { sm.best_cnt = max_good; sm.iters_run = iter; }
        __syncthreads();
        if (sm.flag) break;
        G = GMAX;
    }
    if (sm.flag == 2 || sm.best_cnt == 0) { fail(); return; }

    // ---------------------------------------------------------------- refit on the winner's inliers
    float hf[8];
    for (int i = 0; i < 8; ++i) hf[i] = (float)sm.bestH[i];
    double acc[NRED];
    for (int i = 0; i < NRED; ++i) acc[i] = 0.0;
    for (int i = tid; i < count; i += RT) {
        const float4 p = __ldg(pts + i);
        const bool good = reproj_err(hf, p) <= thr2;
        mask[i] = good;
        if (good) { acc[0] += p.x; acc[1] += p.y; acc[2] += p.z; acc[3] += p.w; acc[4] += 1.0; }
    }
    block_reduce<5>(sm, acc, 0.0);
    const int n_inl = (int)sm.tot[4];
    const double cMx = sm.tot[0] / n_inl, cMy = sm.tot[1] / n_inl, cmx = sm.tot[2] / n_inl, cmy = sm.tot[3] / n_inl;
    __syncthreads();
    for (int i = 0; i < 4; ++i) acc[i] = 0.0;
    for (int i = tid; i < count; i += RT)
        if (mask[i]) {
            const float4 p = __ldg(pts + i);
            acc[0] += fabs(p.x - cMx); acc[1] += fabs(p.y - cMy); acc[2] += fabs(p.z - cmx); acc[3] += fabs(p.w - cmy);
        }
    block_reduce<4>(sm, acc, 0.0);
    const bool spread = !(fabs(sm.tot[0]) < DBL_EPSILON || fabs(sm.tot[1]) < DBL_EPSILON || fabs(sm.tot[2]) < DBL_EPSILON ||
                          fabs(sm.tot[3]) < DBL_EPSILON);
    const double sMx = n_inl / sm.tot[0], sMy = n_inl / sm.tot[1], smx = n_inl / sm.tot[2], smy = n_inl / sm.tot[3];
    __syncthreads();
    if (spread) {                                  // block-uniform
        for (int i = 0; i < NRED; ++i) acc[i] = 0.0;
        for (int i = tid; i < count; i += RT)
            if (mask[i]) {
                const float4 p = __ldg(pts + i);
                ltl_accumulate(acc, (p.x - cMx) * sMx, (p.y - cMy) * sMy, (p.z - cmx) * smx, (p.w - cmy) * smy);
            }
        block_reduce<NRED>(sm, acc, 0.0);
        if (tid == 0) {
            const double nrm[8] = {cMx, cMy, cmx, cmy, sMx, sMy, smx, smy};
            double H[9];
            int o = 0;
            for (int j = 0; j < 9; ++j)
                for (int k = j; k < 9; ++k) S9.A(j, k) = sm.tot[o++];
            dlt_finish(S9, nrm, H);
            for (int i = 0; i < 9; ++i) sm.bestH[i] = H[i];
        }
        __syncthreads();
    }

    // ---------------------------------------------------------------- Levenberg-Marquardt, 8 parameters, <= 10 iterations
    LmState& lm = sm.lm;
    if (tid < 8) sm.xd[tid] = sm.bestH[tid];
    __syncthreads();
    for (int pass = 0;; ++pass) {
        // normal equations at xd (pass 0: the starting point)
        double h[8];
        for (int i = 0; i < 8; ++i) h[i] = sm.xd[i];
        for (int i = 0; i < NRED; ++i) acc[i] = 0.0;
        double rmax = 0.0;
        for (int i = tid; i < count; i += RT)
            if (mask[i]) rmax = fmax(rmax, lm_accumulate(acc, h, __ldg(pts + i)));
        block_reduce<NRED>(sm, acc, rmax);
        if (tid == 0) {
            const double eps = (double)FLT_EPSILON;
            bool proceed = true;
            if (pass == 0) {
                unpack_normal(sm.tot, lm.A, lm.v);
                for (int i = 0; i < 8; ++i) { lm.D[i] = lm.A[i * 9]; sm.x[i] = sm.xd[i]; }
                lm.S = sm.tot[44]; lm.rmax = sm.tot[NRED];
                lm.lam = 1.0; lm.lc = 0.75; lm.iter = 0;
            } else {
                const double* d = lm.d;
                const double Sd = sm.tot[44];
                double dS = 0, dv = 0, dmax = 0;
                for (int i = 0; i < 8; ++i) {
                    double t = 0;
                    for (int j = 0; j < 8; ++j) t += lm.A[i * 8 + j] * d[j];
                    dS += d[i] * (2.0 * lm.v[i] - t);
                    dv += d[i] * lm.v[i];
                    dmax = fmax(dmax, fabs(d[i]));
                }
                const double R = (lm.S - Sd) / (fabs(dS) > DBL_EPSILON ? dS : 1.0);
                if (R > 0.75) {
                    lm.lam *= 0.5;
                    if (lm.lam < lm.lc) lm.lam = 0.0;
                } else if (R < 0.25) {
                    double nu = (Sd - lm.S) / (fabs(dv) > DBL_EPSILON ? dv : 1.0) + 2.0;
                    nu = fmin(fmax(nu, 2.0), 10.0);
                    if (lm.lam == 0.0) {
                        double idg[8], mxv = DBL_EPSILON;
                        eig_solve8(S8, lm.A, 0.0, lm.D, lm.v, nullptr, idg);
                        for (int i = 0; i < 8; ++i) mxv = fmax(mxv, fabs(idg[i]));
                        lm.lam = lm.lc = 1.0 / mxv;
                        nu *= 0.5;
                    }
                    lm.lam *= nu;
                }
                if (Sd < lm.S) {
                    lm.S = Sd;
                    for (int i = 0; i < 8; ++i) sm.x[i] = sm.xd[i];
                    unpack_normal(sm.tot, lm.A, lm.v);
                    lm.rmax = sm.tot[NRED];
                }
                ++lm.iter;
                proceed = lm.iter < 10 && dmax >= eps && lm.rmax >= eps;
            }
            if (proceed) {
                double dd[8];
                if (chol_solve8(lm.A, lm.lam, lm.D, lm.v, dd)) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) lm.d[i] = dd[i];
                } else {
                    eig_solve8(S8, lm.A, lm.lam, lm.D, lm.v, lm.d, nullptr);
                }
                for (int i = 0; i < 8; ++i) sm.xd[i] = sm.x[i] - lm.d[i];
            }
            sm.flag = proceed;
        }
        __syncthreads();
        if (!sm.flag) break;
    }

    // ---------------------------------------------------------------- mask of the refined model
    for (int i = 0; i < 8; ++i) hf[i] = (float)sm.x[i];
    int c = 0;
    for (int i = tid; i < count; i += RT) {
        const bool good = reproj_err(hf, __ldg(pts + i)) <= thr2;
        mask[i] = good;
        c += good;
    }
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) sm.cnt[tid >> 5] = c;
    __syncthreads();
    if (tid == 0) {
        int tot = 0;
        for (int w = 0; w < NW; ++w) tot += sm.cnt[w];
        inliers_out[pair] = tot;
        if (iters_out) iters_out[pair] = sm.iters_run;
    }
    if (tid < 8) Hg[tid] = sm.x[tid];
    if (tid == 8) Hg[8] = 1.0;
}

}  // namespace

extern "C" {

int b200m_ransac_homography(const float* corr_xyxy, const int64_t* ptr, int num_pairs, double thr, double confidence,
                            int max_iters, double* H_out, uint8_t* mask_out, int32_t* inliers_out, int32_t* iters_out,
                            void* stream) {
    if (num_pairs < 0) return B200M_EINVAL;
    if (num_pairs == 0) return B200M_OK;
    if (!corr_xyxy || !ptr || !H_out || !mask_out || !inliers_out) return B200M_EINVAL;
    if (!(confidence > 0.0 && confidence < 1.0)) return B200M_EINVAL;            // CV_Assert in RANSACPointSetRegistrator::run
    if ((uintptr_t)corr_xyxy & 15) return B200M_EINVAL;
    constexpr size_t dyn = (size_t)SCRATCH_DOUBLES * TUPLES * sizeof(double);
    static B200mPerDevice once;                   // the shared-memory opt-in is per device
    int dev = 0;
    if (!b200m_current_device(&dev, nullptr)) return B200M_ENODEVICE;
    if (!once.done[dev]) {
        B200M_CUDA_CHECK(cudaFuncSetAttribute(ransac_homography_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn),
                         "ransac_homography_kernel smem attribute");
        once.done[dev] = 1;
    }
    ransac_homography_kernel<<<(unsigned)num_pairs, RT, dyn, (cudaStream_t)stream>>>(
        (const float4*)corr_xyxy, (const long long*)ptr, num_pairs, (float)(thr * thr), confidence, max_iters, H_out, mask_out,
        inliers_out, iters_out);
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "ransac_homography_kernel launch");
    return B200M_OK;
}

}  // extern "C"
