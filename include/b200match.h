/*
 * b200match.h — C-ABI of the B200-native exhaustive descriptor matcher.
 *
 * Drop-in boundary for ONE hot path of Tarboucha/3DReconstruction: `EnhancedBFMatcher`
 * (/root/reference/FeatureMatchingExtraction/feature_matchers.py:184-252), i.e. the arithmetic the
 * reference delegates to cv2.BFMatcher at :197 (construct), :210 (knnMatch k=2) and :227 (match),
 * plus the ratio test (:212-224), the cross-check semantics of `crossCheck=True` (:225-238) and
 * the final stable sort by distance (:241).
 *
 * Conventions
 *   - plain C: pointers + sizes, no C++/torch types.  `stream` is a cudaStream_t passed as void*.
 *   - every data pointer is a DEVICE pointer owned by the caller unless the name ends in `_host`.
 *   - no allocation inside, no exceptions: each entry point returns B200M_OK (0) or a negative
 *     B200M_E* code; b200m_error_string() names it; b200m_last_cuda_error() returns the CUDA
 *     error text of the last failing call on this thread.
 *   - work is enqueued on `stream` and is asynchronous; outputs are valid once the stream syncs.
 *   - there is NO CPU fallback: without an sm_100 device the compute entry points return
 *     B200M_ENODEVICE.
 */
#ifndef B200MATCH_H
#define B200MATCH_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200M_VERSION 200

#if defined(__GNUC__)
#define B200M_API __attribute__((visibility("default")))
#else
#define B200M_API
#endif

/* error codes */
#define B200M_OK          0
#define B200M_EINVAL     -1   /* bad argument (null pointer, negative size, width 0, ...)      */
#define B200M_EWORKSPACE -2   /* workspace too small: query b200m_workspace_bytes()             */
#define B200M_ECUDA      -3   /* a CUDA runtime call / launch failed                            */
#define B200M_ENODEVICE  -4   /* no sm_100 device: this library has no CPU path                 */
#define B200M_EUNSUPPORTED -5 /* shape outside what the kernels implement (stated per call)     */
#define B200M_ESCRATCH   -6   /* b200m_match_pairs_static only: the batch needs more than
                                 B200M_STATIC_SCRATCH_BYTES of workspace                        */

/* The three behaviours of EnhancedBFMatcher.match. */
#define B200M_MODE_RATIO  0   /* feature_matchers.py:208-224: knn k=2 + `d1 < ratio*d2` in fp64 */
#define B200M_MODE_CROSS  1   /* :225-238 with crossCheck=True: strict mutual first-argmin      */
#define B200M_MODE_NN     2   /* :225-238 with crossCheck=False, ratio None: plain 1-NN         */

/* Descriptor kinds (what cv2.BFMatcher's normType + dtype select, feature_matchers.py:197). */
#define B200M_KIND_HAMMING 0  /* uint8 rows, NORM_HAMMING: __popc INT-pipe kernel               */
#define B200M_KIND_L2_I8   1  /* int8 rows (= integer-valued 0..255 float descriptors shifted by
                                 -128) + int32 row norms: tcgen05 kind::i8, exact              */
#define B200M_KIND_L2_F32  2  /* float32 rows: tcgen05 kind::tf32 candidates + fp32 re-rank     */
#define B200M_KIND_HAMMING_TC 3 /* uint8 rows, NORM_HAMMING, on the tensor cores: every bit is a +-1 int8
                                 dimension (b200m_prep_hamming_tc), a.b = Dbits - 2 h exactly; tcgen05
                                 kind::i8 finds each row's two best 16-column chunks, a POPC pass over the
                                 packed bits resolves them to the exact neighbours.  Same results as
                                 B200M_KIND_HAMMING, D <= 64 bytes                              */
#define B200M_KIND_HAMMING_FP4 4 /* the same with every bit as one e2m1 NIBBLE (+1.0 / -1.0, b200m_prep_hamming_fp4)
                                 and tcgen05 kind::mxf4.block_scale with all scale factors 1.0: twice the MACs
                                 per instruction, half the operand bytes; fp32 sums of +-1 are exact         */

#define B200M_KIND_L2_F16C 5     /* float32 rows as L2_F32, candidates from tcgen05 kind::f16 over a scaled fp16 copy of the
                                 rows (b200m_prep_l2_f16): operands rounded to 11 significant bits instead of TF32's
                                 truncation to 11, twice the MMA rate, half the operand bytes; the same proven error
                                 window + exact fp32 re-rank, so the results are those of L2_F32                  */

/* One image pair of a batch: rows [q_off, q_off+n) of the query store against rows
 * [t_off, t_off+m) of the train store; its matches are written at out_off.. (capacity n). */
typedef struct b200m_pair {
    int32_t q_off, n;
    int32_t t_off, m;
    int64_t out_off;
} b200m_pair;

B200M_API int         b200m_version(void);
B200M_API const char* b200m_error_string(int code);
B200M_API const char* b200m_last_cuda_error(void);
/* sm count / compute capability of the current device; B200M_ENODEVICE when none. */
B200M_API int         b200m_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* ------------------------------------------------------------------------------------------
 * Descriptor preparation (replaces nothing in the reference: cv2 consumes float32 directly;
 * this is the exact re-encoding SURVEY §7.1 step 4 calls the pre-pass).
 *   src   float32 [rows, D] row-major
 *   dst   int8    [rows, Dpad]  (Dpad = D rounded up to 128, zero padded), value = src - 128
 *   aux   b200m_l2_i8_aux_bytes(rows) bytes, 256-byte aligned.  Starts with int32 norms[rows]
 *         = sum over D of (src-128)^2; behind them the "folded norm" digits (32 int8 per row) that
 *         let MODE_RATIO carry the train norm through the tensor cores, and 8 info words.  This
 *         block is what b200m_match_pairs* / b200m_knn2_pairs take as q_norm / t_norm for L2_I8,
 *         together with the same `rows` as q_rows / t_rows.
 *   not_integer  int32[1]       set to 1 if any element is not an integer in 0..255
 *                               (caller must then use the F32 kind); caller zeroes it first.
 */
B200M_API int b200m_l2_i8_aux_bytes(int64_t rows, size_t* bytes);
B200M_API int b200m_prep_l2_i8(const float* src, int64_t rows, int D, int8_t* dst, int Dpad,
                     void* aux, int32_t* not_integer, void* stream);
/* Incremental form, for a store that is uploaded in pieces while matching already runs on the rows that
 * are resident: b200m_l2_i8_aux_init() once for a store of total_rows rows of width D, then
 * b200m_prep_l2_i8_rows() per piece: src = float32 [row_count, D] (the piece only), written to rows
 * [row_begin, row_begin + row_count) of dst / aux.  b200m_prep_l2_i8 == init + one piece of all rows. */
B200M_API int b200m_l2_i8_aux_init(void* aux, int64_t total_rows, int D, void* stream);
/* Norms, fold digits and info words of rows [row_begin, row_begin + row_count) recomputed from int8 rows that are ALREADY
 * in dst (they were encoded on another GPU and arrived by NCCL all-gather): every rank derives identical aux data from
 * the rows themselves, so the descriptor rows are the only thing that crosses NVLink.  After b200m_l2_i8_aux_init. */
B200M_API int b200m_l2_i8_aux_rows(const int8_t* dst, int64_t row_begin, int64_t row_count, int64_t total_rows, int Dpad,
                     void* aux, void* stream);
B200M_API int b200m_prep_l2_i8_rows(const float* src, int64_t row_begin, int64_t row_count, int64_t total_rows,
                     int D, int8_t* dst, int Dpad, void* aux, int32_t* not_integer, void* stream);

/* +-1 expansion for B200M_KIND_HAMMING_TC.
 *   src   uint8 [rows, D] packed descriptor bits (what cv2 consumes: ORB 32, BRISK 64, AKAZE 61 bytes)
 *   dst   int8  [rows, Dpad], Dpad from b200m_hamming_tc_width(D) (8 D rounded up to 128; B200M_EUNSUPPORTED when
 *         D > 64): dimension 8k+b = +1 if bit b of byte k is set, else -1; dimensions past 8 D are 0.  16-byte aligned.
 * Stateless per row, so a store can be expanded in pieces (offset src/dst by whole rows).  The expanded store is
 * what b200m_match_pairs* / b200m_knn2_pairs take as q_norm / t_norm for this kind, next to the packed bits. */
B200M_API int b200m_hamming_tc_width(int D, int* Dpad);
B200M_API int b200m_prep_hamming_tc(const uint8_t* src, int64_t rows, int D, int8_t* dst, int Dpad, void* stream);
/* e2m1 expansion for B200M_KIND_HAMMING_FP4: dst uint8 [rows, Dpad4], Dpad4 from b200m_hamming_fp4_width(D) (4 D bytes
 * rounded up to 128; B200M_EUNSUPPORTED when D > 64): nibble 8k+b = 0x2 (+1.0) if bit b of byte k is set, else 0xA
 * (-1.0); nibbles past 8 D are 0.  Passed as q_norm / t_norm like the int8 expansion. */
B200M_API int b200m_hamming_fp4_width(int D, int* Dpad4);
B200M_API int b200m_prep_hamming_fp4(const uint8_t* src, int64_t rows, int D, uint8_t* dst, int Dpad4, void* stream);

/* Aux block for B200M_KIND_L2_F16C: b200m_l2_f16_aux_bytes(rows, D) bytes, 256-byte aligned, passed as q_norm / t_norm:
 * float32 norms[rows] (as b200m_prep_l2_f32) | fp16 copy [rows, D rounded up to 64] of src * s | info (s, 1/s, max |x|);
 * s is the power of two with max |x| * s in (1/2, 1].  Two launches, no host synchronisation. */
B200M_API int b200m_l2_f16_aux_bytes(int64_t rows, int D, size_t* bytes);
B200M_API int b200m_prep_l2_f16(const float* src, int64_t rows, int D, void* aux, void* stream);

/* Row squared norms for the F32 kind: norms[r] = sum_k src[r,k]^2 (fp32). */
B200M_API int b200m_prep_l2_f32(const float* src, int64_t rows, int D, float* norms, void* stream);

/* ------------------------------------------------------------------------------------------
 * Workspace query for b200m_match_pairs / b200m_knn2_pairs.
 *   total_n = sum of pair.n, total_m = sum of pair.m, max_n/max_m = largest single pair.
 */
B200M_API int b200m_workspace_bytes(int kind, int mode, int num_pairs, int64_t total_n, int64_t total_m,
                          int max_n, int max_m, int D, size_t* bytes);

/* ------------------------------------------------------------------------------------------
 * The hot path, batched over image pairs (a single pair is num_pairs == 1).
 * Replaces, per pair, cv2.BFMatcher.knnMatch/match + ratio/cross-check + sort
 * (feature_matchers.py:208-241).
 *
 *   kind           B200M_KIND_*
 *   q_desc/t_desc  descriptor stores (may be the same pointer for all-pairs over one store):
 *                    HAMMING: uint8 [rows, D]      (D = bytes per descriptor, any D >= 1)
 *                    HAMMING_TC, HAMMING_FP4: the same packed uint8 [rows, D], D <= 64
 *                    L2_I8  : int8  [rows, D]      (D multiple of 128, from b200m_prep_l2_i8)
 *                    L2_F32, L2_F16C : float [rows, D]      (D multiple of 32 floats, D <= 256)
 *   q_norm/t_norm  L2_I8: the aux block of b200m_prep_l2_i8 (int32 norms first); L2_F32: float [rows]; L2_F16C: the aux block of b200m_prep_l2_f16;
 *                  HAMMING: NULL; HAMMING_TC: int8 [rows, Dpad] from b200m_prep_hamming_tc;
 *                  HAMMING_FP4: uint8 [rows, Dpad4] from b200m_prep_hamming_fp4
 *   q_rows/t_rows  number of rows in each store (bounds for TMA maps)
 *   pairs_host     HOST array of num_pairs descriptors (copied to the device workspace)
 *   mode, ratio    B200M_MODE_*; ratio used by MODE_RATIO only (fp64, as Python does at :216)
 *   out_q/out_t/out_d  int32/int32/float32 arrays, capacity = max(pair.out_off + pair.n):
 *                  pair p's matches occupy [out_off, out_off + out_count[p]) sorted by
 *                  (distance asc, queryIdx asc) — the reference's stable sort (:241)
 *   out_count      int32 [num_pairs]
 */
B200M_API int b200m_match_pairs(int kind, const void* q_desc, const void* q_norm, int64_t q_rows,
                      const void* t_desc, const void* t_norm, int64_t t_rows, int D,
                      const b200m_pair* pairs_host, int num_pairs, int mode, double ratio,
                      int32_t* out_q, int32_t* out_t, float* out_d, int32_t* out_count,
                      void* workspace, size_t workspace_bytes, void* stream);

/* Same as b200m_match_pairs but with a TIGHT CSR result: pair.out_off is ignored; pair p's matches
 * occupy [out_ptr[p], out_ptr[p+1]) of out_q/out_t/out_d (capacity sum(pair.n)); out_ptr is
 * int64 [num_pairs + 1] on the device.  Used by the all-pairs driver so that only real matches
 * cross PCIe / NVLink. */
B200M_API int b200m_match_pairs_csr(int kind, const void* q_desc, const void* q_norm, int64_t q_rows,
                          const void* t_desc, const void* t_norm, int64_t t_rows, int D,
                          const b200m_pair* pairs_host, int num_pairs, int mode, double ratio,
                          int32_t* out_q, int32_t* out_t, float* out_d, int32_t* out_count,
                          int64_t* out_ptr, void* workspace, size_t workspace_bytes, void* stream);

/* b200m_match_pairs with NO caller workspace and no allocation of any kind (no cudaMalloc / cudaFree, no allocating
 * library routine): for callers that cannot allocate -- code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it.  This entry point has no counterpart in the reference.
 * Same arguments, results and order as b200m_match_pairs (the caller still passes every input and output); the workspace
 * is a file-scope `static __device__` array of B200M_STATIC_SCRATCH_BYTES (32 MiB) inside the library, which every GPU
 * that loads the library holds until the process exits.
 *   - a batch whose plan needs more than that returns B200M_ESCRATCH (b200m_last_cuda_error() says so); nothing is
 *     enqueued.  b200m_workspace_bytes() gives an upper bound of the need: one 8192 x 8192 SIFT pair 4.4 MB, one ORB
 *     10k x 10k cross-check pair 1.5 MB, one 4096 x 4096 x 256 float pair with mutual check 18.2 MB.
 *   - every call on a device shares the ONE array, which keeps the last call's contents (the plan re-initialises what it
 *     needs on `stream`).  Calls that overlap on the GPU -- different streams, or different host threads on one device --
 *     RACE on it: serialise them (one stream, or events between them). */
#define B200M_STATIC_SCRATCH_BYTES (32u << 20)
B200M_API int b200m_match_pairs_static(int kind, const void* q_desc, const void* q_norm, int64_t q_rows,
                             const void* t_desc, const void* t_norm, int64_t t_rows, int D,
                             const b200m_pair* pairs_host, int num_pairs, int mode, double ratio,
                             int32_t* out_q, int32_t* out_t, float* out_d, int32_t* out_count, void* stream);

/* The same with ONE record per match instead of three arrays, so that a block of matches is one contiguous byte range
 * (one device-to-host copy, one NCCL message) and the 1000-image binary workload moves 6 instead of 12 bytes per match
 * (SURVEY 7.2(6)).  Replaces the per-match EnhancedDMatch objects of feature_matchers.py:213-238 as a wire format.
 *   rec_format  B200M_REC_Q32T32D32: {int32 queryIdx, int32 trainIdx, float32 distance}           12 bytes
 *               B200M_REC_Q16T16D32: {uint16 queryIdx, uint16 trainIdx, float32 distance}          8 bytes, n, m <= 65536
 *               B200M_REC_Q16T16D16: {uint16 queryIdx, uint16 trainIdx, uint16 distance}           6 bytes, Hamming kinds only
 *                                    (the distance is a bit count), n, m <= 65536; else B200M_EUNSUPPORTED
 *   out_rec     records of pair p at [out_ptr[p], out_ptr[p+1]) (record index), capacity sum(pair.n), 8-byte aligned
 * Order inside a pair: (distance asc, queryIdx asc), as b200m_match_pairs. */
#define B200M_REC_Q32T32D32 1
#define B200M_REC_Q16T16D32 2
#define B200M_REC_Q16T16D16 3
B200M_API int b200m_rec_bytes(int rec_format);
B200M_API int b200m_match_pairs_rec(int kind, const void* q_desc, const void* q_norm, int64_t q_rows,
                          const void* t_desc, const void* t_norm, int64_t t_rows, int D,
                          const b200m_pair* pairs_host, int num_pairs, int mode, double ratio,
                          int rec_format, void* out_rec, int32_t* out_count, int64_t* out_ptr,
                          void* workspace, size_t workspace_bytes, void* stream);

/* Raw k=2 neighbours per query row (what cv2.BFMatcher.knnMatch(k=2) returns at :210), for
 * parity checks of the knn stage itself.  Arrays of length sum(pair.n), pair p at out_off:
 * idx2 = -1 / d2 = -1 when the pair has a single train row (cv2 returns a length-1 list). */
B200M_API int b200m_knn2_pairs(int kind, const void* q_desc, const void* q_norm, int64_t q_rows,
                     const void* t_desc, const void* t_norm, int64_t t_rows, int D,
                     const b200m_pair* pairs_host, int num_pairs,
                     int32_t* idx1, int32_t* idx2, float* d1, float* d2,
                     void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * Matches -> N x 4 correspondences (x1, y1, x2, y2) for the structure-from-motion stage: the rows the reference
 * builds one Python object at a time at result_converters.py:763-766 and CameraPoseEstimation reads as
 * matches_data[(img_i, img_j)]['correspondences'] (pair_selector.py:54-67, correspondence_manager.py:78-119).
 *   kp_xy       float32 [rows, 2]: cv2.KeyPoint.pt of every row of the descriptor store (same row order)
 *   pair_q_off / pair_t_off   int32 [num_pairs] (device): first store row of each pair's query / train image
 *   ptr, q_idx, t_idx         the CSR b200m_match_pairs_csr produced (device)
 *   out_xyxy    float32 [ptr[num_pairs], 4], 16-byte aligned: row k = (kp[q_off + q_idx[k]], kp[t_off + t_idx[k]])
 */
B200M_API int b200m_gather_correspondences(const float* kp_xy, int64_t rows, const int32_t* pair_q_off,
                     const int32_t* pair_t_off, const int64_t* ptr, int num_pairs, const int32_t* q_idx,
                     const int32_t* t_idx, float* out_xyxy, void* stream);

/* Geometric verification of matched pairs: what, per image pair,
 *     H, mask = cv2.findHomography(pts1, pts2, cv2.RANSAC, thr, confidence=confidence, maxIters=max_iters)
 * returns at /root/reference/FeatureMatchingExtraction/utils.py:141 (enhanced_filter_matches_with_homography, :118-153),
 * for every pair of a CSR in one launch (one CTA per pair; cv2's own random sequence, see csrc/ransac.cu).
 *   corr_xyxy   float32 [ptr[num_pairs], 4] from b200m_gather_correspondences (device, 16-byte aligned)
 *   ptr         int64 [num_pairs + 1] (device)
 *   H_out       float64 [num_pairs, 9] row-major 3x3, h33 = 1; NaN where cv2 returns None
 *   mask_out    uint8 [ptr[num_pairs]]: 1 = inlier of the refined H (cv2's mask.ravel()), 0 elsewhere
 *   inliers_out int32 [num_pairs]: number of inliers, or -1 where the reference yields ([], None): fewer than 4
 *               matches (utils.py:135) or no model
 *   iters_out   int32 [num_pairs] or NULL: RANSAC iterations run (diagnostic)
 * confidence must lie in (0, 1) (cv2 asserts the same).  No workspace. */
B200M_API int b200m_ransac_homography(const float* corr_xyxy, const int64_t* ptr, int num_pairs, double thr,
                     double confidence, int max_iters, double* H_out, uint8_t* mask_out, int32_t* inliers_out,
                     int32_t* iters_out, void* stream);

/* Essential matrix of matched pairs: what, per image pair,
 *     E, mask = cv2.findEssentialMat(pts1_norm, pts2_norm, focal=1.0, pp=(0, 0), method=cv2.RANSAC, prob=prob,
 *                                    threshold=threshold, maxIters=max_iters)
 * returns at /root/reference/CameraPoseEstimation/essential_estimation.py:118-126 on pts_norm = cv2.undistortPoints(pts, K, None)
 * (:98-108), for every pair of a CSR in one launch (one CTA per pair; cv2's own random sequence and its order of the up to 10
 * candidate models per sample, see csrc/essential.cu).
 *   corr_xyxy   float32 [ptr[num_pairs], 4] PIXEL coordinates from b200m_gather_correspondences (device, 16-byte aligned)
 *   ptr         int64 [num_pairs + 1] (device)
 *   cam         float64 [num_pairs, 8] (device): fx1, fy1, cx1, cy1, fx2, fy2, cx2, cy2 of the pair's two cameras; the kernel
 *               normalises x' = (x - cx) * (1 / fx) in float64, which is undistortPoints without distortion bit for bit.
 *               NULL: the coordinates are already normalised.
 *   threshold   in NORMALISED units (the reference passes focal = 1); prob in (0, 1)
 *   E_out       float64 [num_pairs, 9] row-major 3x3, unit Frobenius norm; NaN where cv2 returns None
 *   mask_out    uint8 [ptr[num_pairs]]: 1 = inlier of E (cv2's mask.ravel())
 *   inliers_out int32 [num_pairs]: number of inliers, or -1 when there is no model (fewer than 5 points included)
 *   iters_out   int32 [num_pairs] or NULL: RANSAC iterations run (diagnostic)
 * No workspace. */
B200M_API int b200m_ransac_essential(const float* corr_xyxy, const int64_t* ptr, int num_pairs, const double* cam,
                     double threshold, double prob, int max_iters, double* E_out, uint8_t* mask_out, int32_t* inliers_out,
                     int32_t* iters_out, void* stream);

/* Measurement hooks for bench.py's roofline: when enabled, CUDA events are recorded on the
 * launching stream around every knn-stage kernel; collect() synchronises them and returns the summed
 * device time (ms) and the number of knn launches since the last collect/enable. */
B200M_API int b200m_timing_enable(int on);
B200M_API int b200m_timing_collect(double* knn_ms, int* knn_launches);

/* L2_F32 kind: number of query rows whose TF32 candidate list was not provably sufficient and were
 * re-scanned exhaustively in fp32 (and the distance evaluations that cost) since the last reset.
 * Synchronises the device. */
B200M_API int b200m_f32_stats(int64_t* fallback_rows, int64_t* fallback_evals, int reset);

/* Counters for bench.py's `gpu_launches`: kernels launched by this library since the last reset. */
B200M_API int64_t b200m_launch_count(void);
B200M_API void    b200m_reset_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* B200MATCH_H */
