/* oracle/essential_oracle.c — TEST INFRASTRUCTURE ONLY (the checker; never linked into the product path).
 *
 * Plain-C restatement of cv2.findEssentialMat(p1, p2, focal=1, pp=(0,0), method=cv2.RANSAC, prob, threshold, maxIters) as
 * the reference calls it at /root/reference/CameraPoseEstimation/essential_estimation.py:118-126.  Same algorithm as
 * oracle/essential_oracle.py (see its header for the OpenCV files restated and the parity status); this file exists so
 * that iteration-capped cases (5000 iterations x up to 10 models), which cost the pure-Python oracle minutes, are checked
 * on the CPU in a second.  Built by oracle/essential_oracle.py:build_c() with -ffp-contract=off.
 *
 * One deliberate shortcut, exactly equivalent: solvePoly's 300 Durand-Kerner sweeps stop as soon as a whole sweep leaves
 * every root unchanged -- every later sweep would recompute the same values.
 */
#include <complex.h>
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

#define API __attribute__((visibility("default")))

static uint32_t rng_next(uint64_t* s) {
    *s = (uint64_t)(uint32_t)(*s) * 4164903690ull + (*s >> 32);
    return (uint32_t)(*s);
}

/* monomials (exponents of x, y, z): linear, quadratic, and the 20 cubic columns in Nister's order */
static const int LIN[4][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {0, 0, 0}};
static const int QUAD[10][3] = {{2, 0, 0}, {0, 2, 0}, {0, 0, 2}, {1, 1, 0}, {1, 0, 1}, {0, 1, 1}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {0, 0, 0}};
static const int CUB[20][3] = {{3, 0, 0}, {0, 3, 0}, {2, 1, 0}, {1, 2, 0}, {2, 0, 1}, {2, 0, 0}, {0, 2, 1}, {0, 2, 0}, {1, 1, 1}, {1, 1, 0},
                               {1, 0, 2}, {1, 0, 1}, {1, 0, 0}, {0, 1, 2}, {0, 1, 1}, {0, 1, 0}, {0, 0, 3}, {0, 0, 2}, {0, 0, 1}, {0, 0, 0}};
static int LL[4][4], QL[10][4], tables_ready = 0;

static int find(const int (*tab)[3], int n, int a, int b, int c) {
    for (int i = 0; i < n; i++) if (tab[i][0] == a && tab[i][1] == b && tab[i][2] == c) return i;
    return -1;
}
static void init_tables(void) {
    if (tables_ready) return;
    for (int a = 0; a < 4; a++) for (int b = 0; b < 4; b++)
        LL[a][b] = find(QUAD, 10, LIN[a][0] + LIN[b][0], LIN[a][1] + LIN[b][1], LIN[a][2] + LIN[b][2]);
    for (int q = 0; q < 10; q++) for (int l = 0; l < 4; l++)
        QL[q][l] = find(CUB, 20, QUAD[q][0] + LIN[l][0], QUAD[q][1] + LIN[l][1], QUAD[q][2] + LIN[l][2]);
    tables_ready = 1;
}
static void mul_ll(const double* a, const double* b, double s, double* q) {      /* q += s a b */
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) q[LL[i][j]] += s * a[i] * b[j];
}
static void mul_ql(const double* q, const double* l, double s, double* c) {      /* c += s q l */
    for (int i = 0; i < 10; i++) for (int j = 0; j < 4; j++) c[QL[i][j]] += s * q[i] * l[j];
}

/* rows 5..8 of Vt of cv::SVD(Q[5x9], FULL_UV): see null_space_basis() in essential_oracle.py */
static int null_basis(const double Q[5][9], double N[4][9]) {
    double V[9][9];
    int i, j, k, it;
    for (i = 0; i < 5; i++) {                       /* an orthonormal basis of the row space (modified Gram-Schmidt, twice) */
        for (k = 0; k < 9; k++) V[i][k] = Q[i][k];
        for (it = 0; it < 2; it++)
            for (j = 0; j < i; j++) {
                double sd = 0;
                for (k = 0; k < 9; k++) sd += V[i][k] * V[j][k];
                for (k = 0; k < 9; k++) V[i][k] -= sd * V[j][k];
            }
        double nn = 0;
        for (k = 0; k < 9; k++) nn += V[i][k] * V[i][k];
        if (!(nn > 1e-300)) return 0;
        nn = 1.0 / sqrt(nn);
        for (k = 0; k < 9; k++) V[i][k] *= nn;
    }
    uint64_t st = 0x12345678;
    for (i = 5; i < 9; i++) {
        for (k = 0; k < 9; k++) V[i][k] = (rng_next(&st) & 256) != 0 ? 1.0 / 9 : -1.0 / 9;
        for (it = 0; it < 2; it++)
            for (j = 0; j < i; j++) {
                double sd = 0, asum = 0;
                for (k = 0; k < 9; k++) sd += V[i][k] * V[j][k];
                for (k = 0; k < 9; k++) { double t = V[i][k] - sd * V[j][k]; V[i][k] = t; asum += fabs(t); }
                asum = asum > DBL_EPSILON * 100 ? 1 / asum : 0;
                for (k = 0; k < 9; k++) V[i][k] *= asum;
            }
        double sd = 0;
        for (k = 0; k < 9; k++) sd += V[i][k] * V[i][k];
        sd = sqrt(sd);
        double s = sd > DBL_MIN ? 1 / sd : 0;
        for (k = 0; k < 9; k++) V[i][k] *= s;
    }
    memcpy(N, &V[5][0], sizeof(double) * 36);
    return 1;
}

/* A[10][20]: det E = 0 and 2 E E^T E - tr(E E^T) E = 0 for E = x X + y Y + z Z + W, entries of the 9-vectors read row-major */
static void coeff_matrix(const double N[4][9], double A[10][20]) {
    double E[3][3][4], EEt[3][3][10], tr[10];
    int r, c, k, i;
    for (r = 0; r < 3; r++) for (c = 0; c < 3; c++) for (k = 0; k < 4; k++) E[r][c][k] = N[k][3 * r + c];
    memset(EEt, 0, sizeof(EEt));
    for (r = 0; r < 3; r++) for (c = 0; c < 3; c++) for (k = 0; k < 3; k++) mul_ll(E[r][k], E[c][k], 1.0, EEt[r][c]);
    for (i = 0; i < 10; i++) tr[i] = EEt[0][0][i] + EEt[1][1][i] + EEt[2][2][i];
    memset(A, 0, sizeof(double) * 200);
    {   /* determinant */
        static const int P[6][3] = {{0, 1, 2}, {1, 2, 0}, {2, 0, 1}, {2, 1, 0}, {1, 0, 2}, {0, 2, 1}};
        static const double S[6] = {1, 1, 1, -1, -1, -1};
        for (i = 0; i < 6; i++) {
            double q[10];
            memset(q, 0, sizeof(q));
            mul_ll(E[0][P[i][0]], E[1][P[i][1]], 1.0, q);
            mul_ql(q, E[2][P[i][2]], S[i], A[0]);
        }
    }
    for (r = 0; r < 3; r++) for (c = 0; c < 3; c++) {
        double* row = A[1 + 3 * r + c];
        for (k = 0; k < 3; k++) mul_ql(EEt[r][k], E[k][c], 2.0, row);
        mul_ql(tr, E[r][c], -1.0, row);
    }
}

/* X = A[:, :10]^-1 A[:, 10:] by Gaussian elimination with partial pivoting; returns 0 when singular */
static int reduce(double A[10][20]) {
    for (int k = 0; k < 10; k++) {
        int p = k;
        for (int i = k + 1; i < 10; i++) if (fabs(A[i][k]) > fabs(A[p][k])) p = i;
        if (fabs(A[p][k]) < DBL_MIN) return 0;
        if (p != k) for (int j = 0; j < 20; j++) { double t = A[k][j]; A[k][j] = A[p][j]; A[p][j] = t; }
        double inv = 1.0 / A[k][k];
        for (int j = k; j < 20; j++) A[k][j] *= inv;
        for (int i = 0; i < 10; i++) if (i != k) {
            double f = A[i][k];
            if (f != 0) for (int j = k; j < 20; j++) A[i][j] -= f * A[k][j];
        }
    }
    return 1;
}

/* cv::solvePoly, real coefficients c[0..n]: returns the number of roots written to (re, im) */
static int solve_poly(const double* c, int n, double* re, double* im) {
    for (; n > 1; n--) if (fabs(c[n]) > DBL_EPSILON) break;
    double pr = 1, pi = 0;
    for (int i = 0; i < n; i++) { re[i] = pr; im[i] = pi; double t = pr - pi; pi = pr + pi; pr = t; }   /* p *= (1 + i) */
    for (int iter = 0; iter < 300; iter++) {
        double maxdiff = 0;
        int changed = 0;
        for (int i = 0; i < n; i++) {
            double xr = re[i], xi = im[i], nr = c[n], ni = 0, dr = c[n], di = 0;
            int same = 1;
            for (int j = 0; j < n; j++) {
                double t = nr * xr - ni * xi; ni = nr * xi + ni * xr; nr = t + c[n - j - 1];
                if (j != i) {
                    double er = xr - re[j], ei = xi - im[j];
                    if (er != 0 || ei != 0) { double u = dr * er - di * ei; di = dr * ei + di * er; dr = u; }
                    else same++;
                }
            }
            double t = 1.0 / (dr * dr + di * di);
            double qr = (nr * dr + ni * di) * t, qi = (-nr * di + ni * dr) * t;
            if (same > 1) { double complex z = cpow(qr + qi * I, 1.0 / same); qr = creal(z); qi = cimag(z); }
            double yr = xr - qr, yi = xi - qi;
            if (yr != xr || yi != xi) changed = 1;
            re[i] = yr; im[i] = yi;
            double a = sqrt(qr * qr + qi * qi);
            if (a > maxdiff) maxdiff = a;
        }
        if (maxdiff <= 0 || !changed) break;
    }
    for (int i = 0; i < n; i++) if (fabs(im[i]) < 1e-100) im[i] = 0;
    return n;
}

/* EMEstimatorCallback::runKernel: 5 correspondences -> up to 10 models (3x3 row-major, unit Frobenius norm) in cv's order */
static int run_kernel(const double* m1, const double* m2, const int* idx, double models[10][9]) {
    double Q[5][9], N[4][9], A[10][20], B[3][13], c[11], re[10], im[10];
    int i, k;
    for (i = 0; i < 5; i++) {
        double x1 = m1[2 * idx[i]], y1 = m1[2 * idx[i] + 1], x2 = m2[2 * idx[i]], y2 = m2[2 * idx[i] + 1];
        double q[9] = {x1 * x2, y1 * x2, x2, x1 * y2, y1 * y2, y2, x1, y1, 1.0};     /* = E row by row for x2^T E x1 */
        memcpy(Q[i], q, sizeof(q));
    }
    if (!null_basis(Q, N)) return 0;
    coeff_matrix(N, A);
    if (!reduce(A)) return 0;
    memset(B, 0, sizeof(B));
    for (i = 0; i < 3; i++) {
        const double *r1 = &A[2 * i + 4][10], *r2 = &A[2 * i + 5][10];
        for (k = 0; k < 3; k++) { B[i][1 + k] += r1[k]; B[i][5 + k] += r1[3 + k]; B[i][k] -= r2[k]; B[i][4 + k] -= r2[3 + k]; }
        for (k = 0; k < 4; k++) { B[i][9 + k] += r1[6 + k]; B[i][8 + k] -= r2[6 + k]; }
    }
    {   /* c[k] = coefficient of z^k of det B(z); entries lowest degree first: P[i][0], P[i][1] cubic, P[i][2] quartic */
        double P[3][3][5];
        memset(P, 0, sizeof(P));
        for (i = 0; i < 3; i++) {
            for (k = 0; k < 4; k++) { P[i][0][k] = B[i][3 - k]; P[i][1][k] = B[i][7 - k]; }
            for (k = 0; k < 5; k++) P[i][2][k] = B[i][12 - k];
        }
        memset(c, 0, sizeof(c));
        static const int T[6][3] = {{0, 1, 2}, {1, 2, 0}, {2, 0, 1}, {2, 1, 0}, {1, 0, 2}, {0, 2, 1}};
        static const double S[6] = {1, 1, 1, -1, -1, -1};
        for (int t = 0; t < 6; t++) {
            double ab[9];
            memset(ab, 0, sizeof(ab));
            for (int a = 0; a < 5; a++) for (int b = 0; b < 5; b++) ab[a + b] += P[0][T[t][0]][a] * P[1][T[t][1]][b];
            for (int a = 0; a < 9; a++) for (int b = 0; b < 5; b++) if (a + b <= 10) c[a + b] += S[t] * ab[a] * P[2][T[t][2]][b];
        }
    }
    int n = solve_poly(c, 10, re, im), count = 0;
    for (i = 0; i < n; i++) {
        if (fabs(im[i]) > 1e-10) continue;
        double z = re[i], z2 = z * z, z3 = z2 * z, z4 = z3 * z, bz[3][3], v[3], best = -1;
        for (k = 0; k < 3; k++) {
            bz[k][0] = B[k][0] * z3 + B[k][1] * z2 + B[k][2] * z + B[k][3];
            bz[k][1] = B[k][4] * z3 + B[k][5] * z2 + B[k][6] * z + B[k][7];
            bz[k][2] = B[k][8] * z4 + B[k][9] * z3 + B[k][10] * z2 + B[k][11] * z + B[k][12];
        }
        for (k = 0; k < 3; k++) {                   /* null vector of the (rank 2) matrix: the largest cross product of two rows */
            const double *a = bz[k], *b = bz[(k + 1) % 3];
            double w[3] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
            double nn = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
            if (nn > best) { best = nn; memcpy(v, w, sizeof(w)); }
        }
        if (!(best > 0)) continue;
        double s = 1.0 / sqrt(best);
        if (fabs(v[2] * s) < 1e-10) continue;
        double x = v[0] / v[2], y = v[1] / v[2], e[9], nn = 0;
        for (k = 0; k < 9; k++) { e[k] = N[0][k] * x + N[1][k] * y + N[2][k] * z + N[3][k]; nn += e[k] * e[k]; }
        nn = 1.0 / sqrt(nn);
        for (k = 0; k < 9; k++) models[count][k] = e[k] * nn;
        count++;
    }
    return count;
}

static int find_inliers(const double* m1, const double* m2, int n, const double* E, float thresh, uint8_t* mask) {
    int good = 0;
    for (int i = 0; i < n; i++) {
        double x1 = m1[2 * i], y1 = m1[2 * i + 1], x2 = m2[2 * i], y2 = m2[2 * i + 1];
        double a0 = E[0] * x1 + E[1] * y1 + E[2], a1 = E[3] * x1 + E[4] * y1 + E[5], a2 = E[6] * x1 + E[7] * y1 + E[8];
        double b0 = E[0] * x2 + E[3] * y2 + E[6], b1 = E[1] * x2 + E[4] * y2 + E[7];
        double s = x2 * a0 + y2 * a1 + a2;
        float err = (float)(s * s / (a0 * a0 + a1 * a1 + b0 * b0 + b1 * b1));
        mask[i] = err <= thresh;
        good += mask[i];
    }
    return good;
}

static int update_iters(double p, double ep, int model_points, int max_iters) {   /* RANSACUpdateNumIters */
    p = p < 0 ? 0 : (p > 1 ? 1 : p);
    ep = ep < 0 ? 0 : (ep > 1 ? 1 : ep);
    double num = 1 - p > DBL_MIN ? 1 - p : DBL_MIN, denom = 1 - pow(1 - ep, model_points);
    if (denom < DBL_MIN) return 0;
    num = log(num); denom = log(denom);
    return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : (int)nearbyint(num / denom);
}

/* m1, m2: n already-normalised points (float64 x, y); returns the inlier count of the winner (0: no model), E row-major */
API int essential_ransac(const double* m1, const double* m2, int n, double threshold, double prob, int max_iters,
                         double* E_out, uint8_t* mask_out, uint8_t* scratch_mask, int* iters_out) {
    init_tables();
    double models[10][9];
    float thresh = (float)(threshold * threshold);
    int niters = max_iters > 1 ? max_iters : 1, best = 0, it = 0, idx[5];
    uint64_t st = 0xFFFFFFFFFFFFFFFFull;
    if (iters_out) *iters_out = 0;
    if (n < 5) return 0;
    if (n == 5) {
        for (int i = 0; i < 5; i++) idx[i] = i;
        if (run_kernel(m1, m2, idx, models) <= 0) return 0;
        memcpy(E_out, models[0], sizeof(double) * 9);
        memset(mask_out, 1, 5);
        return 5;
    }
    for (; it < niters; it++) {
        for (int i = 0; i < 5;) {
            int v = (int)(rng_next(&st) % (uint32_t)n), j;
            for (j = 0; j < i; j++) if (idx[j] == v) break;
            if (j < i) continue;
            idx[i++] = v;
        }
        int nm = run_kernel(m1, m2, idx, models);
        for (int k = 0; k < nm; k++) {
            int good = find_inliers(m1, m2, n, models[k], thresh, scratch_mask);
            if (good > (best > 4 ? best : 4)) {
                best = good;
                memcpy(mask_out, scratch_mask, n);
                memcpy(E_out, models[k], sizeof(double) * 9);
                niters = update_iters(prob, (double)(n - good) / n, 5, niters);
            }
        }
    }
    if (iters_out) *iters_out = it;
    return best;
}

/* the solver alone, for unit tests: returns the number of models */
API int essential_five_point(const double* m1, const double* m2, double* models_out) {
    init_tables();
    int idx[5] = {0, 1, 2, 3, 4};
    return run_kernel(m1, m2, idx, (double(*)[9])models_out);
}
