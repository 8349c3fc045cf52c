/*
 * oracle/bf_oracle.c — TEST INFRASTRUCTURE ONLY (checker, never the product path).
 *
 * Plain-C CPU restatement of the arithmetic the reference delegates to OpenCV:
 *   cv2.BFMatcher(norm, crossCheck).knnMatch(k=2) / .match()
 * as called at /root/reference/FeatureMatchingExtraction/feature_matchers.py:197 (construct),
 * :210 (knnMatch k=2) and :227 (match).  The arithmetic itself lives in a third-party dependency
 * that is NOT under /root/reference: OpenCV `cv::BFMatcher` -> `cv::batchDistance`
 * (modules/core/src/batch_distance.cpp), pinned by the reference as `opencv-python>=4.5.0`
 * (FeatureMatchingExtraction/requirements.txt:2, setup.py:31; Dockerfile:20 pins 4.12.0.88).
 * The published algorithm restated here (and pinned against cv2 4.13.0 in tests/test_oracle.py):
 *   - Hamming: popcount(xor) over all descriptor bytes, exact int.
 *   - L2: sqrt_f32( float32( sum (a-b)^2 ) ); for integer-valued inputs the sum is exact in fp32
 *     (< 2^24), so the result is bit-equal no matter the summation order.  For general floats the
 *     oracle accumulates in double and rounds once (cv2's SIMD order is implementation defined;
 *     tests compare with rel. tol 1e-5).
 *   - k-NN: ascending distance; among equal distances the LOWEST train index comes first
 *     (batchDistance inserts a candidate only when it is strictly smaller).
 *   - k=2 with a single train row yields one neighbour (idx2 = -1 here).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may load
 * this file's shared object.  OpenMP is used so the CPU baseline can use every host core.
 */
#include <stdint.h>
#include <stddef.h>
#include <math.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_API __attribute__((visibility("default")))

ORACLE_API int oracle_version(void) { return 1; }

ORACLE_API int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

ORACLE_API void oracle_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

static inline int hamming_bytes(const uint8_t* a, const uint8_t* b, int D) {
    int d = 0, k = 0;
    for (; k + 8 <= D; k += 8) {
        uint64_t x, y;
        memcpy(&x, a + k, 8);
        memcpy(&y, b + k, 8);
        d += __builtin_popcountll(x ^ y);
    }
    for (; k < D; ++k) d += __builtin_popcount((unsigned)(a[k] ^ b[k]));
    return d;
}

/* knn k=2, NORM_HAMMING.  q: [N, D] uint8, t: [M, D] uint8 (row-major, contiguous).
 * idx1/idx2: best / second-best train index per query (idx2 = -1 when M < 2),
 * d1/d2: their integer distances (d2 = -1 when M < 2). */
ORACLE_API int oracle_knn2_hamming(const uint8_t* q, int N, const uint8_t* t, int M, int D,
                                   int32_t* idx1, int32_t* idx2, int32_t* d1, int32_t* d2) {
    if (N < 0 || M < 0 || D <= 0) return -1;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i) {
        const uint8_t* a = q + (size_t)i * D;
        int b1 = -1, b2 = -1, v1 = 0x7fffffff, v2 = 0x7fffffff;
        for (int j = 0; j < M; ++j) {
            int d = hamming_bytes(a, t + (size_t)j * D, D);
            if (d < v1) { v2 = v1; b2 = b1; v1 = d; b1 = j; }
            else if (d < v2) { v2 = d; b2 = j; }
        }
        idx1[i] = b1; idx2[i] = b2;
        d1[i] = b1 >= 0 ? v1 : -1;
        d2[i] = b2 >= 0 ? v2 : -1;
    }
    return 0;
}

/* knn k=2, NORM_L2 on float32 rows.  Squared distance accumulated in double then rounded to
 * float32 once; the reported distance is sqrtf of that float32 (cv::batchDistance's
 * `std::sqrt(normL2Sqr)` in float).  Exact for integer-valued inputs whose sum < 2^24. */
ORACLE_API int oracle_knn2_l2(const float* q, int N, const float* t, int M, int D,
                              int32_t* idx1, int32_t* idx2, float* d1, float* d2) {
    if (N < 0 || M < 0 || D <= 0) return -1;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i) {
        const float* a = q + (size_t)i * D;
        int b1 = -1, b2 = -1;
        float v1 = INFINITY, v2 = INFINITY;
        for (int j = 0; j < M; ++j) {
            const float* b = t + (size_t)j * D;
            double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
            int k = 0;
            for (; k + 4 <= D; k += 4) {
                double e0 = (double)a[k] - (double)b[k];
                double e1 = (double)a[k + 1] - (double)b[k + 1];
                double e2 = (double)a[k + 2] - (double)b[k + 2];
                double e3 = (double)a[k + 3] - (double)b[k + 3];
                s0 += e0 * e0; s1 += e1 * e1; s2 += e2 * e2; s3 += e3 * e3;
            }
            for (; k < D; ++k) { double e = (double)a[k] - (double)b[k]; s0 += e * e; }
            float d = (float)((s0 + s1) + (s2 + s3));
            if (d < v1 || b1 < 0) { if (b1 >= 0) { v2 = v1; b2 = b1; } v1 = d; b1 = j; }
            else if (d < v2 || b2 < 0) { v2 = d; b2 = j; }
        }
        idx1[i] = b1; idx2[i] = b2;
        d1[i] = b1 >= 0 ? sqrtf(v1) : -1.0f;
        d2[i] = b2 >= 0 ? sqrtf(v2) : -1.0f;
    }
    return 0;
}

/* Full N x M distance matrices for tiny cases (used by tests to cross-check tie handling). */
ORACLE_API int oracle_dist_hamming(const uint8_t* q, int N, const uint8_t* t, int M, int D,
                                   int32_t* out) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i)
        for (int j = 0; j < M; ++j)
            out[(size_t)i * M + j] = hamming_bytes(q + (size_t)i * D, t + (size_t)j * D, D);
    return 0;
}
