/* oracle/ransac_oracle.c — TEST INFRASTRUCTURE ONLY (the checker; never linked into the product path).
 *
 * Plain-C restatement of cv2.findHomography(pts1, pts2, cv2.RANSAC, thr, confidence, maxIters) as the reference calls it
 * at /root/reference/FeatureMatchingExtraction/utils.py:141 (enhanced_filter_matches_with_homography, :118-153).
 * Same algorithm as oracle/ransac_oracle.py (see its header for the OpenCV files restated and the parity status); this
 * file exists so that the iteration-capped golden cases, which cost the pure-Python oracle 10 s each, are checked on the
 * CPU in milliseconds.  Built by oracle/ransac_oracle.py:build_c() with -ffp-contract=off: a*b+c must round twice, as in
 * OpenCV's SSE2 baseline build.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define API __attribute__((visibility("default")))

static uint32_t rng_next(uint64_t* s) {
    *s = (uint64_t)(uint32_t)(*s) * 4164903690ull + (*s >> 32);
    return (uint32_t)(*s);
}

static double hyp(double a, double b) { /* lapack.cpp's own scaled hypot */
    a = fabs(a); b = fabs(b);
    if (a > b) { b /= a; return a * sqrt(1 + b * b); }
    if (b > 0) { a /= b; return b * sqrt(1 + a * a); }
    return 0;
}

/* JacobiImpl_<double>: A n x n (upper triangle used, destroyed), W eigenvalues descending, V eigenvectors as rows */
static void jacobi(double* A, int n, double* W, double* V) {
    int indR[9], indC[9], i, j, k, l, m, it;
    double mv, val;
    for (i = 0; i < n; i++) for (j = 0; j < n; j++) V[i * n + j] = i == j;
    for (k = 0; k < n; k++) {
        W[k] = A[k * n + k];
        if (k < n - 1) {
            for (m = k + 1, mv = fabs(A[k * n + m]), i = k + 2; i < n; i++) { val = fabs(A[k * n + i]); if (mv < val) mv = val, m = i; }
            indR[k] = m;
        }
        if (k > 0) {
            for (m = 0, mv = fabs(A[k]), i = 1; i < k; i++) { val = fabs(A[i * n + k]); if (mv < val) mv = val, m = i; }
            indC[k] = m;
        }
    }
    for (it = 0; it < n * n * 30; it++) {
        for (k = 0, mv = fabs(A[indR[0]]), i = 1; i < n - 1; i++) { val = fabs(A[i * n + indR[i]]); if (mv < val) mv = val, k = i; }
        l = indR[k];
        for (i = 1; i < n; i++) { val = fabs(A[indC[i] * n + i]); if (mv < val) mv = val, k = indC[i], l = i; }
        double p = A[k * n + l];
        if (fabs(p) <= DBL_EPSILON) break;
        double y = (W[l] - W[k]) * 0.5, t = fabs(y) + hyp(p, y), s = hyp(p, t), c = t / s, a0, b0;
        s = p / s; t = (p / t) * p;
        if (y < 0) s = -s, t = -t;
        A[k * n + l] = 0;
        W[k] -= t; W[l] += t;
#define ROT(v0, v1) a0 = v0, b0 = v1, v0 = a0 * c - b0 * s, v1 = a0 * s + b0 * c
        for (i = 0; i < k; i++) ROT(A[i * n + k], A[i * n + l]);
        for (i = k + 1; i < l; i++) ROT(A[k * n + i], A[i * n + l]);
        for (i = l + 1; i < n; i++) ROT(A[k * n + i], A[l * n + i]);
        for (i = 0; i < n; i++) ROT(V[k * n + i], V[l * n + i]);
#undef ROT
        for (j = 0; j < 2; j++) {
            int idx = j == 0 ? k : l;
            if (idx < n - 1) {
                for (m = idx + 1, mv = fabs(A[idx * n + m]), i = idx + 2; i < n; i++) { val = fabs(A[idx * n + i]); if (mv < val) mv = val, m = i; }
                indR[idx] = m;
            }
            if (idx > 0) {
                for (m = 0, mv = fabs(A[idx]), i = 1; i < idx; i++) { val = fabs(A[i * n + idx]); if (mv < val) mv = val, m = i; }
                indC[idx] = m;
            }
        }
    }
    for (k = 0; k < n - 1; k++) {
        m = k;
        for (i = k + 1; i < n; i++) if (W[m] < W[i]) m = i;
        if (k != m) {
            double tw = W[m]; W[m] = W[k]; W[k] = tw;
            for (i = 0; i < n; i++) { double tv = V[m * n + i]; V[m * n + i] = V[k * n + i]; V[k * n + i] = tv; }
        }
    }
}

/* HomographyEstimatorCallback::runKernel over the points listed in idx (count of them); 0 if a coordinate has no spread */
static int run_kernel(const float* p1, const float* p2, const int* idx, int count, double* H) {
    double cMx = 0, cMy = 0, cmx = 0, cmy = 0, sMx = 0, sMy = 0, smx = 0, smy = 0, LtL[81], W[9], V[81];
    int i, j, k;
    for (i = 0; i < count; i++) { int q = idx[i]; cmx += p2[2 * q]; cmy += p2[2 * q + 1]; cMx += p1[2 * q]; cMy += p1[2 * q + 1]; }
    cmx /= count; cmy /= count; cMx /= count; cMy /= count;
    for (i = 0; i < count; i++) {
        int q = idx[i];
        smx += fabs(p2[2 * q] - cmx); smy += fabs(p2[2 * q + 1] - cmy);
        sMx += fabs(p1[2 * q] - cMx); sMy += fabs(p1[2 * q + 1] - cMy);
    }
    if (fabs(smx) < DBL_EPSILON || fabs(smy) < DBL_EPSILON || fabs(sMx) < DBL_EPSILON || fabs(sMy) < DBL_EPSILON) return 0;
    smx = count / smx; smy = count / smy; sMx = count / sMx; sMy = count / sMy;
    memset(LtL, 0, sizeof LtL);
    for (i = 0; i < count; i++) {
        int q = idx[i];
        double x = (p2[2 * q] - cmx) * smx, y = (p2[2 * q + 1] - cmy) * smy;
        double X = (p1[2 * q] - cMx) * sMx, Y = (p1[2 * q + 1] - cMy) * sMy;
        double Lx[9] = {X, Y, 1, 0, 0, 0, -x * X, -x * Y, -x}, Ly[9] = {0, 0, 0, X, Y, 1, -y * X, -y * Y, -y};
        for (j = 0; j < 9; j++) for (k = j; k < 9; k++) LtL[j * 9 + k] += Lx[j] * Lx[k] + Ly[j] * Ly[k];
    }
    for (j = 0; j < 9; j++) for (k = 0; k < j; k++) LtL[j * 9 + k] = LtL[k * 9 + j];
    jacobi(LtL, 9, W, V);
    const double* h0 = V + 72;
    double inv[9] = {1. / smx, 0, cmx, 0, 1. / smy, cmy, 0, 0, 1}, n2[9] = {sMx, 0, -cMx * sMx, 0, sMy, -cMy * sMy, 0, 0, 1}, T[9], R[9];
    for (i = 0; i < 3; i++) for (j = 0; j < 3; j++) T[i * 3 + j] = inv[i * 3] * h0[j] + inv[i * 3 + 1] * h0[3 + j] + inv[i * 3 + 2] * h0[6 + j];
    for (i = 0; i < 3; i++) for (j = 0; j < 3; j++) R[i * 3 + j] = T[i * 3] * n2[j] + T[i * 3 + 1] * n2[3 + j] + T[i * 3 + 2] * n2[6 + j];
    for (i = 0; i < 9; i++) H[i] = R[i] * (1. / R[8]);
    return 1;
}

static int collinear_last(const float* p, const int* id) { /* haveCollinearPoints(m, 4): the last point against earlier pairs */
    int j, k;
    double xi = p[2 * id[3]], yi = p[2 * id[3] + 1];
    for (j = 0; j < 3; j++) {
        double dx1 = p[2 * id[j]] - xi, dy1 = p[2 * id[j] + 1] - yi;
        for (k = 0; k < j; k++) {
            double dx2 = p[2 * id[k]] - xi, dy2 = p[2 * id[k] + 1] - yi;
            if (fabs(dx2 * dy1 - dy2 * dx1) <= FLT_EPSILON * (fabs(dx1) + fabs(dy1) + fabs(dx2) + fabs(dy2))) return 1;
        }
    }
    return 0;
}

static double det3(const float* p, int a, int b, int c) {
    double m[9] = {p[2 * a], p[2 * a + 1], 1, p[2 * b], p[2 * b + 1], 1, p[2 * c], p[2 * c + 1], 1};
    return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) + m[2] * (m[3] * m[7] - m[4] * m[6]);
}

static int check_subset(const float* p1, const float* p2, const int* id) {
    static const int tt[4][3] = {{0, 1, 2}, {1, 2, 3}, {0, 2, 3}, {0, 1, 3}};
    int i, negative = 0;
    if (collinear_last(p1, id) || collinear_last(p2, id)) return 0;
    for (i = 0; i < 4; i++)
        negative += det3(p1, id[tt[i][0]], id[tt[i][1]], id[tt[i][2]]) * det3(p2, id[tt[i][0]], id[tt[i][1]], id[tt[i][2]]) < 0;
    return negative == 0 || negative == 4;
}

/* computeError + findInliers: float32 throughout; returns the inlier count */
static int find_inliers(const float* p1, const float* p2, int n, const double* H, float t, uint8_t* mask) {
    float h[8];
    int i, nz = 0;
    for (i = 0; i < 8; i++) h[i] = (float)H[i];
    for (i = 0; i < n; i++) {
        float Mx = p1[2 * i], My = p1[2 * i + 1];
        float ww = 1.f / (h[6] * Mx + h[7] * My + 1.f);
        float dx = (h[0] * Mx + h[1] * My + h[2]) * ww - p2[2 * i];
        float dy = (h[3] * Mx + h[4] * My + h[5]) * ww - p2[2 * i + 1];
        float e = dx * dx + dy * dy;
        mask[i] = e <= t;
        nz += mask[i];
    }
    return nz;
}

static int update_num_iters(double p, double ep, int max_iters) {
    double num, denom;
    p = p < 0 ? 0 : p > 1 ? 1 : p;
    ep = ep < 0 ? 0 : ep > 1 ? 1 : ep;
    num = 1. - p > DBL_MIN ? 1. - p : DBL_MIN;
    denom = 1. - pow(1. - ep, 4);
    if (denom < DBL_MIN) return 0;
    num = log(num); denom = log(denom);
    return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : (int)nearbyint(num / denom);
}

/* x = A^-1 b for symmetric 8 x 8 A through the Jacobi decomposition (cv::solve DECOMP_EIG); inv_diag optional */
static void eig_solve8(const double* A, const double* b, double* x, double* inv_diag) {
    double T[64], W[8], V[64], y[8];
    int i, j;
    memcpy(T, A, sizeof T);
    jacobi(T, 8, W, V);
    for (i = 0; i < 8; i++) {
        double s = 0;
        for (j = 0; j < 8; j++) s += V[i * 8 + j] * b[j];
        y[i] = fabs(W[i]) > DBL_EPSILON ? s / W[i] : 0;
    }
    for (j = 0; j < 8; j++) { double s = 0; for (i = 0; i < 8; i++) s += V[i * 8 + j] * y[i]; x[j] = s; }
    if (inv_diag)
        for (j = 0; j < 8; j++) {
            double s = 0;
            for (i = 0; i < 8; i++) s += fabs(W[i]) > DBL_EPSILON ? V[i * 8 + j] * V[i * 8 + j] / W[i] : 0;
            inv_diag[j] = s;
        }
}

/* HomographyRefineCallback::compute over the listed points: A = J^T J, v = J^T r, returns |r|^2, *rmax = max |r| */
static double refine_eval(const float* p1, const float* p2, const int* idx, int cnt, const double* h, double* A, double* v, double* rmax) {
    double S = 0;
    int i, j, k;
    if (A) { memset(A, 0, 64 * sizeof(double)); memset(v, 0, 8 * sizeof(double)); }
    *rmax = 0;
    for (i = 0; i < cnt; i++) {
        int q = idx[i];
        double Mx = p1[2 * q], My = p1[2 * q + 1], ww = h[6] * Mx + h[7] * My + 1.;
        ww = fabs(ww) > DBL_EPSILON ? 1. / ww : 0;
        double xi = (h[0] * Mx + h[1] * My + h[2]) * ww, yi = (h[3] * Mx + h[4] * My + h[5]) * ww;
        double rx = xi - p2[2 * q], ry = yi - p2[2 * q + 1];
        S += rx * rx + ry * ry;
        if (fabs(rx) > *rmax) *rmax = fabs(rx);
        if (fabs(ry) > *rmax) *rmax = fabs(ry);
        if (A) {
            double Jx[8] = {Mx * ww, My * ww, ww, 0, 0, 0, -Mx * ww * xi, -My * ww * xi};
            double Jy[8] = {0, 0, 0, Mx * ww, My * ww, ww, -Mx * ww * yi, -My * ww * yi};
            for (j = 0; j < 8; j++) {
                for (k = 0; k < 8; k++) A[j * 8 + k] += Jx[j] * Jx[k] + Jy[j] * Jy[k];
                v[j] += Jx[j] * rx + Jy[j] * ry;
            }
        }
    }
    return S;
}

static void lm_refine(const float* p1, const float* p2, const int* idx, int cnt, double* H) {
    double x[8], xd[8], A[64], v[8], D[8], d[8], Ap[64], An[64], vn[8], rmax, rmaxd, lam = 1, lc = 0.75;
    const double eps = FLT_EPSILON;
    int i, j, iter = 0;
    memcpy(x, H, sizeof x);
    double S = refine_eval(p1, p2, idx, cnt, x, A, v, &rmax);
    for (i = 0; i < 8; i++) D[i] = A[i * 9];
    for (;;) {
        memcpy(Ap, A, sizeof Ap);
        for (i = 0; i < 8; i++) Ap[i * 9] += lam * D[i];
        eig_solve8(Ap, v, d, 0);
        for (i = 0; i < 8; i++) xd[i] = x[i] - d[i];
        double Sd = refine_eval(p1, p2, idx, cnt, xd, An, vn, &rmaxd), dS = 0, dv = 0, dmax = 0;
        for (i = 0; i < 8; i++) {
            double t = 0;
            for (j = 0; j < 8; j++) t += A[i * 8 + j] * d[j];
            dS += d[i] * (2 * v[i] - t); dv += d[i] * v[i];
            if (fabs(d[i]) > dmax) dmax = fabs(d[i]);
        }
        double R = (S - Sd) / (fabs(dS) > DBL_EPSILON ? dS : 1);
        if (R > 0.75) { lam *= 0.5; if (lam < lc) lam = 0; }
        else if (R < 0.25) {
            double nu = (Sd - S) / (fabs(dv) > DBL_EPSILON ? dv : 1) + 2;
            nu = nu < 2 ? 2 : nu > 10 ? 10 : nu;
            if (lam == 0) {
                double dummy[8], idg[8], mx = DBL_EPSILON;
                eig_solve8(A, v, dummy, idg);
                for (i = 0; i < 8; i++) if (fabs(idg[i]) > mx) mx = fabs(idg[i]);
                lam = lc = 1. / mx; nu *= 0.5;
            }
            lam *= nu;
        }
        if (Sd < S) { S = Sd; memcpy(x, xd, sizeof x); memcpy(A, An, sizeof A); memcpy(v, vn, sizeof v); rmax = rmaxd; }
        iter++;
        if (!(iter < 10 && dmax >= eps && rmax >= eps)) break;
    }
    memcpy(H, x, sizeof x);
    H[8] = 1;
}

/* returns 1 and fills H (9 doubles), mask (n bytes), *iters when cv2 returns a model; 0 when it returns None */
API int ransac_find_homography(const float* p1, const float* p2, int n, double thr, double conf, int max_iters, double* H,
                               uint8_t* mask, int* iters) {
    int i, id[4], niters = max_iters > 1 ? max_iters : 1, max_good = 0, it = 0, found_any = 0;
    float t = (float)(thr * thr);
    double best[9], Hc[9];
    uint64_t s = 0xFFFFFFFFFFFFFFFFull;
    *iters = 0;
    memset(mask, 0, (size_t)(n > 0 ? n : 0));
    if (n < 4) return 0;
    if (n == 4) {
        int all[4] = {0, 1, 2, 3};
        if (!run_kernel(p1, p2, all, 4, H)) return 0;
        memset(mask, 1, 4);
        return 1;
    }
    uint8_t* m = (uint8_t*)malloc((size_t)n);
    int* inl = (int*)malloc(sizeof(int) * (size_t)n);
    for (it = 0; it < niters; it++) {
        int attempt, ok = 0;
        for (attempt = 0; attempt < 10000 && !ok; attempt++) {
            for (i = 0; i < 4; i++) {
                int v, j, dup;
                do { v = (int)(rng_next(&s) % (uint32_t)n); dup = 0; for (j = 0; j < i; j++) dup |= id[j] == v; } while (dup);
                id[i] = v;
            }
            ok = check_subset(p1, p2, id);
        }
        if (!ok) { if (it == 0) { free(m); free(inl); return 0; } break; }
        if (!run_kernel(p1, p2, id, 4, Hc)) continue;
        int good = find_inliers(p1, p2, n, Hc, t, m);
        if (good > (max_good > 3 ? max_good : 3)) {
            max_good = good; found_any = 1;
            memcpy(best, Hc, sizeof best);
            niters = update_num_iters(conf, (double)(n - good) / n, niters);
        }
    }
    *iters = it;
    if (!found_any) { free(m); free(inl); return 0; }
    int cnt = 0;
    find_inliers(p1, p2, n, best, t, m);
    for (i = 0; i < n; i++) if (m[i]) inl[cnt++] = i;
    memcpy(H, best, sizeof best);
    if (run_kernel(p1, p2, inl, cnt, Hc)) memcpy(H, Hc, sizeof Hc);
    lm_refine(p1, p2, inl, cnt, H);
    find_inliers(p1, p2, n, H, t, mask);          /* 4.13: the returned mask follows the refined H */
    free(m); free(inl);
    return 1;
}
