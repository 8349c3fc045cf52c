// common.cuh — shared device/host definitions for the b200match kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/b200match.h"

typedef unsigned long long u64;
typedef unsigned int u32;

// Per-row result of the knn stage.  key = (value << 32) | train_index, so that u64 `min`
// orders by (value asc, index asc) == cv2's "ties -> lowest trainIdx" (SURVEY F8-i).
// value: HAMMING -> bit count; L2_I8 -> exact integer squared distance;
//        L2_F32  -> IEEE bits of the (non-negative) fp32 squared distance.
struct Top2 { u64 k1, k2; };
#define B200M_SENTINEL 0xFFFFFFFFFFFFFFFFull
#define B200M_I8_GROUPS 4    // independent epilogue groups of the L2_I8 kernel = Top2 slices per column split
#define B200M_I8_CHUNK 64    // columns per key chunk of the L2_I8 kernel (runner-up refined inside the winner's chunk)
// L2_I8 "folded norm" block (l2_i8.cu, fold kernel): 32 extra int8 K-dims per descriptor row.  The query side
// of the block is the constant vector c = {-128 x15, -1} x2, the train side holds digits t with
// sum c.t = -(h - bias), h = ||b'||^2 >> 1, so the int32 accumulator is E = a'.b' - h + bias and
// d^2 = ||a'||^2 + (||b'||^2 & 1) + 2 (bias - E): the epilogue needs no per-column constant at all.
// 30 coarse digits (weight 128) and one fine digit (weight 1) cover h - bias in [XMIN, XMIN + SPREAD].
#define B200M_FOLD_K 32
#define B200M_FOLD_SLICES 1          // Top2 slices per column split the fold kernel writes: its epilogue warps merge in shared memory
                                     // (the keyed kernel writes B200M_I8_GROUPS)
#ifndef B200M_FOLD_CHUNK
#define B200M_FOLD_CHUNK 16
#endif
//          // columns per chunk of the fold kernel (finalize recomputes one chunk per undecided row)
#define B200M_FOLD_XMIN (-491520)     // -128 * 30 * 128
#define B200M_FOLD_SPREAD 979327      // 127 * 30 * 128 + 127 - XMIN
#define B200M_FOLD_VBIAS (1 << 24)    // Top2 value of a fold-kernel result = VBIAS - E  (|E| < 2^23)
// Auxiliary block b200m_prep_l2_i8 writes into the caller's `norms` buffer (b200m_l2_i8_aux_bytes(rows) bytes):
//   int32 norms[rows] | int8 fold[rows][32] at fold_off | int32 info[8] at info_off  (both 256-byte aligned)
//   info: [0] min h, [1] max h over the store, [2] bias, [3] 1 if every row's h - bias fits the digits
__host__ __device__ inline size_t b200m_aux_fold_off(long long rows) { return ((size_t)rows * 4 + 255) / 256 * 256; }
__host__ __device__ inline size_t b200m_aux_info_off(long long rows) {
    return b200m_aux_fold_off(rows) + ((size_t)rows * B200M_FOLD_K + 255) / 256 * 256;
}
__host__ __device__ inline size_t b200m_aux_bytes(long long rows) { return b200m_aux_info_off(rows) + 256; }
// Auxiliary block of b200m_prep_l2_f16 (kind L2_F16C): float norms[rows] | __half copy[rows][Dh] at half_off |
// float info[64] at info_off: [0] scale s (a power of two, max |x| * s <= 1), [1] 1 / s, [2] max |x|.  Dh = D rounded up to 64.
__host__ __device__ inline int b200m_f16_width(int D) { return (D + 63) / 64 * 64; }
__host__ __device__ inline size_t b200m_f16_half_off(long long rows) { return ((size_t)rows * 4 + 255) / 256 * 256; }
__host__ __device__ inline size_t b200m_f16_info_off(long long rows, int D) {
    return b200m_f16_half_off(rows) + ((size_t)rows * b200m_f16_width(D) * 2 + 255) / 256 * 256;
}
// ... | "folded norm" rows x[rows][16 halves] at x_off: {hi, mid, lo, 0 x13} with hi + mid + lo = -(s^2 / 2) ||b||^2 (three
// fp16 terms hold the fp32 value exactly); one extra K = 16 MMA step against the constant query-side row {r, r, r, 0 ...},
// r = s_q / s_t, adds -(s_q s_t / 2) ||b||^2 to the accumulator (l2_f32.cu, FOLD kernels)
__host__ __device__ inline size_t b200m_f16_x_off(long long rows, int D) { return b200m_f16_info_off(rows, D) + 256; }
__host__ __device__ inline size_t b200m_f16_aux_bytes(long long rows, int D) {
    return b200m_f16_x_off(rows, D) + ((size_t)rows * 32 + 255) / 256 * 256;
}
#define B200M_HTC_CHUNK 16   // columns per key chunk of the tensor-core Hamming kernel (hamming_tc.cu); resolve recomputes two chunks per row
#define B200M_F32_CAND 32    // TF32 candidate list capacity per (query row, column half, column split) (l2_f32.cu)

// Device-side pair record (host b200m_pair + offsets into the per-row workspaces).
struct DevPair {
    int q_off, n, t_off, m;
    long long out_off;   // first output slot (capacity n)
    long long fwd_off;   // first row of this pair in the forward Top2 array (prefix sum of n)
    long long bwd_off;   // first row in the backward Top2 array (prefix sum of m), MODE_CROSS only
    long long sort_off;  // first slot in the padded sort buffer (prefix sum of pow2ceil(n))
    int npad;            // pow2ceil(n)
    int _pad;
};

struct MatchPlan {       // host-side carve-up of the caller's workspace
    int32_t* flags;      // [64] per-call device words: [0] = snapshot of the train store's "fold digits usable" word, taken
                         // on the call's stream before the knn launches, so that the keyed kernel, the fold kernel and
                         // finalize all act on ONE decision even if another stream re-encodes rows of the store meanwhile
    DevPair* pairs;      // [P]
    Top2* fwd;           // [slices][total_n]
    Top2* bwd;           // [slices][total_m]   (MODE_CROSS)
    u32* colmin;         // HAMMING MODE_CROSS, fused single pass: [total_m] best (dist << 19 | query row) per train row; aliases bwd
    u64* sortbuf;        // [sum npad]
    int32_t* tbest;      // [total_n]
    void* cand;          // L2_F32: [2S][rows][B200M_F32_CAND] (estimate, column) + [2S][rows] counts, rows = max(total_n, total_m)
    float* pmax;         // L2_F32: [P] max train-row norm of the direction being processed
    int S;               // column splits per direction
    int slices;          // Top2 slices per row the knn kernel writes (S, or S * B200M_I8_GROUPS for L2_I8)
    long long total_n, total_m, total_pad;
    size_t bytes;
};

__device__ __forceinline__ void top2_insert(u64& k1, u64& k2, u64 k) {
    u64 hi = k1 > k ? k1 : k;
    k1 = k1 < k ? k1 : k;
    k2 = k2 < hi ? k2 : hi;
}
__device__ __forceinline__ void top2_insert32(u32& k1, u32& k2, u32 k) {
    k2 = min(k2, max(k1, k));
    k1 = min(k1, k);
}

// host helpers implemented in capi.cu
void b200m_count_launch(int n = 1);
// One-time setup that is a fact of a DEVICE, not of the process (cudaFuncSetAttribute applies to the current device's
// context only; cluster occupancy and SM counts differ between devices): every launcher keeps one of these per kernel
// family, indexed by the ordinal of the current device, so an Engine("cuda:0") and an Engine("cuda:1") in one process
// each get their own opt-in.  Two threads racing on the same device both run the (idempotent) setup.
#define B200M_MAX_DEVICES 64
struct B200mPerDevice {
    int done[B200M_MAX_DEVICES];     // 0 = setup not run on this device yet
    int val[B200M_MAX_DEVICES][12];  // whatever the setup measured (cluster sizes, resident cluster counts)
};
// ordinal and SM count of the current device; returns false when there is none (or the ordinal is out of range)
bool b200m_current_device(int* dev, int* sm_count);
int b200m_set_cuda_error(cudaError_t e, const char* where);
#define B200M_CUDA_CHECK(expr, where)                                   \
    do {                                                                \
        cudaError_t _e = (expr);                                        \
        if (_e != cudaSuccess) return b200m_set_cuda_error(_e, where);  \
    } while (0)

// kernel launchers (one per .cu)
int launch_hamming_top2(const uint8_t* q_desc, const uint8_t* t_desc, int D, const DevPair* pairs,
                        int num_pairs, int max_n, int max_m, int S, bool backward, Top2* out,
                        long long total_rows, u32* colmin, cudaStream_t st);
// Tensor-core Hamming (hamming_tc.cu): knn over the +-1 int8 stores, then the exact per-row resolve on the packed bits
int launch_hamming_tc_top2(const void* q_store, int64_t q_rows, const void* t_store, int64_t t_rows, int Dpad, int dbits,
                           bool fp4, const DevPair* pairs, int num_pairs, int max_n, int max_m, int S, bool backward,
                           Top2* out, long long total_rows, cudaStream_t st);
int launch_hamming_tc_resolve(const uint8_t* q_bits, const uint8_t* t_bits, int D, const DevPair* pairs, int num_pairs,
                              int max_n, int max_m, int S, bool backward, bool best_only, Top2* arr, long long total_rows,
                              cudaStream_t st);
#define B200M_HAMMING_IDX_BITS 19   // key32 = dist << 19 | index (hamming.cu); fused column minima need n <= 2^19
int launch_l2_i8_top2(const int8_t* q_desc, const int32_t* q_norm, int64_t q_rows,
                      const int8_t* t_desc, const int32_t* t_norm, int64_t t_rows, int D,
                      const DevPair* pairs, int num_pairs, int max_n, int max_m, int S, bool backward,
                      const int32_t* fold_flag, Top2* out, long long total_rows, cudaStream_t st);
int launch_l2_f32_top2(const float* q_desc, const float* q_norm, int64_t q_rows,
                       const float* t_desc, const float* t_norm, int64_t t_rows, int D,
                       const DevPair* pairs, int num_pairs, int max_n, int max_m, int S, bool backward,
                       bool need2, void* cand_ws, float* pmax, Top2* out, long long total_rows, cudaStream_t st, bool h16 = false);
int b200m_f32_stats_read(unsigned long long out[2], bool reset);
// L2_I8 descriptors for the fused exact runner-up refinement (null for the other kinds)
struct RefineI8 {
    const int8_t* q_desc; const int32_t* q_norm; const int8_t* t_desc; const int32_t* t_norm; int D;
    const int32_t* t_info;      // info words of the train store's aux block ([2] = fold bias, a constant of the width)
    const int32_t* fold_flag;   // MatchPlan::flags: fold kernel results iff fold_flag[0] (null: keyed kernel results)
};
int launch_finalize(int kind, int mode, double ratio, const MatchPlan& plan, int num_pairs, int max_n,
                    int32_t* counts, const RefineI8& rf, cudaStream_t st);
int launch_scan_offsets(const MatchPlan& plan, const int32_t* counts, int num_pairs, long long* out_ptr,
                        cudaStream_t st);
int launch_sort_emit(int mode, const MatchPlan& plan, int num_pairs, int max_npad, int32_t* out_q,
                     int32_t* out_t, float* out_d, int32_t* out_count, const int32_t* counts, cudaStream_t st,
                     int rec_format = 0, void* out_rec = nullptr);
int launch_knn2_emit(int kind, const MatchPlan& plan, int num_pairs, int max_n, int32_t* idx1, int32_t* idx2,
                     float* d1, float* d2, const RefineI8& rf, cudaStream_t st);
