// prep.cu — descriptor re-encoding ahead of the tensor-core L2 kernels.
//   prep_l2_i8 : OpenCV SIFT emits integer-valued float32 in 0..255
//                (/root/reference/FeatureMatchingExtraction/traditional_detectors.py:42).  They are
//                shifted by -128 into int8 (exactly representable) and the int32 row norm
//                sum (x-128)^2 is produced in the same pass; (a-128)-(b-128) == a-b, so
//                ||a'||^2 + ||b'||^2 - 2 a'.b' is the exact integer sum (a-b)^2 that cv2's fp32
//                accumulation also yields exactly (< 2^24).  A flag reports any non-integer or
//                out-of-range element so the caller can route to the fp32 path instead.
//   prep_l2_f32: fp32 row norms for the TF32 candidate kernel.
// HBM-bound streaming kernels: one warp per row, coalesced 128-bit loads.
#include <cuda_fp16.h>
#include "common.cuh"

namespace {

// One warp per row.  `row0` = first row of this call inside the store (incremental uploads prepare row ranges);
// src holds only the call's rows.  Besides the int8 row and its norm the warp writes the row's folded-norm digits
// (common.cuh): t with sum_k c_k t_k = -(h - bias), c = {-128 x15, -1} x2, h = norm >> 1.  bias is a constant of
// the descriptor width (info[2], set by fold_init_kernel): the digit range is anchored at the largest possible h,
// so it covers every row whose norm is above ~13 % of the maximum (SIFT-like rows always are); a row below that
// clears info[3] and the matcher then uses the keyed kernel for this store.
// FROM_I8: the int8 rows are already in `dst` (they arrived from another rank, allpairs.DistributedAllPairs); only the norms,
// digits and info words are (re)computed from them -- every rank derives identical aux data, no second collective.
template <bool FROM_I8>
__global__ void __launch_bounds__(256)
prep_l2_i8_kernel(const float* __restrict__ src, long long row0, long long rows, int D, int8_t* __restrict__ dst, int Dpad,
                  int32_t* __restrict__ norms, int8_t* __restrict__ fold, int32_t* __restrict__ not_integer,
                  int32_t* __restrict__ info) {
    __shared__ int s_min[8], s_max[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    long long r = (long long)blockIdx.x * (blockDim.x >> 5) + warp;
    const bool valid = r < rows;
    if (!valid) r = rows - 1;                         // recompute the last row: keeps the block barrier simple
    const float* s = src + r * D;
    const long long row = row0 + r;
    int8_t* d = dst + row * Dpad;
    int acc = 0, bad = 0;
    // Dpad is a multiple of 128: each lane converts 4 consecutive elements per step.
    for (int k = lane * 4; k < Dpad; k += 128) {
        if constexpr (FROM_I8) {
            const u32 w = *reinterpret_cast<const u32*>(d + k);
            acc = __dp4a((int)w, (int)w, acc);          // sum of squares of four int8
            continue;
        }
        int v[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            int kk = k + e;
            if (kk < D) {
                float f = __ldg(s + kk);
                float rr = rintf(f);
                bad |= !(f == rr && rr >= 0.0f && rr <= 255.0f);
                int iv = (int)fminf(fmaxf(rr, 0.0f), 255.0f) - 128;
                v[e] = iv;
                acc += iv * iv;
            } else {
                v[e] = 0;                       // zero padding adds 0 to dot products and norms
            }
        }
        u32 packed = (u32)(v[0] & 0xff) | ((u32)(v[1] & 0xff) << 8) | ((u32)(v[2] & 0xff) << 16) |
                     ((u32)(v[3] & 0xff) << 24);
        *reinterpret_cast<u32*>(d + k) = packed;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        acc += __shfl_xor_sync(0xffffffffu, acc, off);
        bad |= __shfl_xor_sync(0xffffffffu, bad, off);
    }
    // fold digits: lanes 0..7 each build one 32-bit word (4 digits) of the row's 32-byte block
    const int h = acc >> 1;
    int x = h - __ldg(info + 2);
    const bool fits = x >= B200M_FOLD_XMIN && x <= B200M_FOLD_XMIN + B200M_FOLD_SPREAD;
    x = min(max(x, B200M_FOLD_XMIN), B200M_FOLD_XMIN + B200M_FOLD_SPREAD);     // clamped digits are never used (flag)
    const int S = x >> 7, F = x & 127;                 // x = 128 S + F, S in [-3840, 3810]
    int a = S / 30, b = S - a * 30;
    if (b < 0) { b += 30; a -= 1; }                    // floor division: S = 30 a + b, 0 <= b < 30
    if (lane < 8) {
        u32 word = 0;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int k = 4 * lane + e;                // K position; coarse digit index skips positions 15 and 31
            int v;
            if (k == 15) v = F;
            else if (k == 31) v = 0;
            else { const int ci = k < 15 ? k : k - 1; v = a + (ci < b ? 1 : 0); }
            word |= (u32)(v & 0xff) << (8 * e);
        }
        reinterpret_cast<u32*>(fold + row * B200M_FOLD_K)[lane] = word;
    }
    if (lane == 0) {
        if (valid) norms[row] = acc;
        if (!FROM_I8 && bad) atomicOr(not_integer, 1);
        if (!fits) atomicAnd(info + 3, 0);
        s_min[warp] = s_max[warp] = h;
    }
    __syncthreads();
    if (threadIdx.x == 0) {                           // range of h over the store (diagnostics, tests)
        int lo = s_min[0], hi = s_max[0];
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) { lo = min(lo, s_min[w]); hi = max(hi, s_max[w]); }
        atomicMin(info + 0, lo);
        atomicMax(info + 1, hi);
    }
}

// info words of a fresh aux block: [0] min h, [1] max h, [2] bias, [3] digits usable.  D == 0: an empty store (unusable).
__global__ void fold_init_kernel(int32_t* __restrict__ info, int D) {
    info[0] = 0x7fffffff; info[1] = (int)0x80000000;
    info[2] = D * 8192 - (B200M_FOLD_XMIN + B200M_FOLD_SPREAD);     // largest possible h = D * 128^2 / 2 maps to the top digit value
    info[3] = D > 0 ? 1 : 0;
}

__global__ void __launch_bounds__(256)
prep_l2_f32_kernel(const float* __restrict__ src, long long rows, int D, float* __restrict__ norms) {
    const int lane = threadIdx.x & 31;
    long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const float* s = src + row * D;
    float acc = 0.0f;
    for (int k = lane; k < D; k += 32) { float f = __ldg(s + k); acc = fmaf(f, f, acc); }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
    if (lane == 0) norms[row] = acc;
}

// L2_F16C, pass 1: row norms (fp32) and the store's max |x| (atomicMax on the bits of a non-negative float)
__global__ void __launch_bounds__(256)
prep_l2_f16_norm_kernel(const float* __restrict__ src, long long rows, int D, float* __restrict__ norms,
                        unsigned int* __restrict__ maxbits) {
    const int lane = threadIdx.x & 31;
    long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const float* s = src + row * D;
    float acc = 0.0f, mx = 0.0f;
    for (int k = lane; k < D; k += 32) { float f = __ldg(s + k); acc = fmaf(f, f, acc); mx = fmaxf(mx, fabsf(f)); }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        acc += __shfl_xor_sync(0xffffffffu, acc, off);
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    }
    if (lane == 0) {
        norms[row] = acc;
        if (mx == mx && mx < 3.0e38f) atomicMax(maxbits, __float_as_uint(mx));
    }
}

// pass 2: dst = half(src * s), s = 2^-ceil(log2 max|x|) (exact scaling; max |x| * s <= 1, so nothing overflows and the
// fp32 accumulator of a 256-term dot product stays below 256); columns D..Dh-1 are zero
__global__ void __launch_bounds__(256)
prep_l2_f16_convert_kernel(const float* __restrict__ src, long long rows, int D, int Dh, __half* __restrict__ dst,
                           float* __restrict__ info, const float* __restrict__ norms, __half* __restrict__ xrows) {
    const float mx = __uint_as_float(*(const unsigned int*)(info + 2));
    int e = 0;
    if (mx > 0.0f) { frexpf(mx, &e); if (ldexpf(1.0f, e - 1) == mx) e -= 1; }     // mx <= 2^e, tight
    e = max(-100, min(100, e));
    const float s = ldexpf(1.0f, -e);
    if (blockIdx.x == 0 && threadIdx.x == 0) { info[0] = s; info[1] = ldexpf(1.0f, e); }
    const long long total = rows * (long long)(Dh / 2);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / (Dh / 2);
        const int k = (int)(i - r * (Dh / 2)) * 2;
        const float a = k < D ? __ldg(src + r * D + k) * s : 0.0f, b = k + 1 < D ? __ldg(src + r * D + k + 1) * s : 0.0f;
        *reinterpret_cast<__half2*>(dst + r * Dh + k) = __floats2half2_rn(a, b);
        if (k == 0) {                               // the row's folded norm: three fp16 terms of w = -(s^2 / 2) ||b||^2
            const float w = -0.5f * s * s * norms[r];
            const __half hi = __float2half_rn(w);
            const float r1 = w - __half2float(hi);
            const __half mid = __float2half_rn(r1);
            const __half lo = __float2half_rn(r1 - __half2float(mid));
            __half2* x = reinterpret_cast<__half2*>(xrows + r * 16);
            x[0] = __halves2half2(hi, mid);
            x[1] = __halves2half2(lo, __float2half_rn(0.0f));
#pragma unroll
            for (int j = 2; j < 8; ++j) x[j] = __floats2half2_rn(0.0f, 0.0f);
        }
    }
}

}  // namespace

extern "C" {

int b200m_l2_f16_aux_bytes(int64_t rows, int D, size_t* bytes) {
    if (rows < 0 || D <= 0 || !bytes) return B200M_EINVAL;
    *bytes = b200m_f16_aux_bytes(rows, D);
    return B200M_OK;
}

int b200m_prep_l2_f16(const float* src, int64_t rows, int D, void* aux, void* stream) {
    if (rows < 0 || D <= 0) return B200M_EINVAL;
    if (!aux || ((uintptr_t)aux & 255)) return B200M_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    float* info = (float*)((char*)aux + b200m_f16_info_off(rows, D));
    B200M_CUDA_CHECK(cudaMemsetAsync(info, 0, 256, st), "l2_f16 info memset");
    if (rows == 0) return B200M_OK;
    if (!src) return B200M_EINVAL;
    long long blocks = (rows + 7) / 8;
    if (blocks > 0x7fffffffLL) return B200M_EUNSUPPORTED;
    prep_l2_f16_norm_kernel<<<(unsigned)blocks, 256, 0, st>>>(src, rows, D, (float*)aux, (unsigned int*)(info + 2));
    B200M_CUDA_CHECK(cudaGetLastError(), "prep_l2_f16_norm_kernel launch");
    const int Dh = b200m_f16_width(D);
    long long want = (rows * (long long)(Dh / 2) + 255) / 256;
    prep_l2_f16_convert_kernel<<<(unsigned)(want < 148 * 16 ? want : 148 * 16), 256, 0, st>>>(
        src, rows, D, Dh, (__half*)((char*)aux + b200m_f16_half_off(rows)), info, (const float*)aux,
        (__half*)((char*)aux + b200m_f16_x_off(rows, D)));
    B200M_CUDA_CHECK(cudaGetLastError(), "prep_l2_f16_convert_kernel launch");
    b200m_count_launch(2);
    return B200M_OK;
}

int b200m_l2_i8_aux_bytes(int64_t rows, size_t* bytes) {
    if (rows < 0 || !bytes) return B200M_EINVAL;
    *bytes = b200m_aux_bytes(rows);
    return B200M_OK;
}

int b200m_l2_i8_aux_init(void* aux, int64_t total_rows, int D, void* stream) {
    if (!aux || total_rows < 0 || D <= 0 || ((uintptr_t)aux & 255)) return B200M_EINVAL;
    fold_init_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((int32_t*)((char*)aux + b200m_aux_info_off(total_rows)),
                                                        total_rows > 0 ? D : 0);
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "fold_init_kernel launch");
    return B200M_OK;
}

int b200m_prep_l2_i8_rows(const float* src, int64_t row_begin, int64_t row_count, int64_t total_rows, int D,
                          int8_t* dst, int Dpad, void* aux, int32_t* not_integer, void* stream) {
    if (row_begin < 0 || row_count < 0 || row_begin + row_count > total_rows || D <= 0 || Dpad < D || Dpad % 128 != 0 ||
        ((uintptr_t)aux & 255))
        return B200M_EINVAL;
    if (row_count == 0) return B200M_OK;
    if (!src || !dst || !aux || !not_integer) return B200M_EINVAL;
    long long blocks = (row_count + 7) / 8;
    if (blocks > 0x7fffffffLL) return B200M_EUNSUPPORTED;
    prep_l2_i8_kernel<false><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        src, row_begin, row_count, D, dst, Dpad, (int32_t*)aux, (int8_t*)aux + b200m_aux_fold_off(total_rows), not_integer,
        (int32_t*)((char*)aux + b200m_aux_info_off(total_rows)));
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "prep_l2_i8_kernel launch");
    return B200M_OK;
}

int b200m_l2_i8_aux_rows(const int8_t* dst, int64_t row_begin, int64_t row_count, int64_t total_rows, int Dpad, void* aux,
                         void* stream) {
    if (row_begin < 0 || row_count < 0 || row_begin + row_count > total_rows || Dpad <= 0 || Dpad % 128 != 0 ||
        ((uintptr_t)aux & 255) || ((uintptr_t)dst & 3))
        return B200M_EINVAL;
    if (row_count == 0) return B200M_OK;
    if (!dst || !aux) return B200M_EINVAL;
    long long blocks = (row_count + 7) / 8;
    if (blocks > 0x7fffffffLL) return B200M_EUNSUPPORTED;
    prep_l2_i8_kernel<true><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        nullptr, row_begin, row_count, Dpad, const_cast<int8_t*>(dst), Dpad, (int32_t*)aux,
        (int8_t*)aux + b200m_aux_fold_off(total_rows), nullptr, (int32_t*)((char*)aux + b200m_aux_info_off(total_rows)));
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "prep_l2_i8_kernel<from int8> launch");
    return B200M_OK;
}

int b200m_prep_l2_i8(const float* src, int64_t rows, int D, int8_t* dst, int Dpad, void* aux,
                     int32_t* not_integer, void* stream) {
    if (rows < 0 || D <= 0 || Dpad < D || Dpad % 128 != 0 || ((uintptr_t)aux & 255)) return B200M_EINVAL;
    if (rows > 0 && (!src || !dst || !aux || !not_integer)) return B200M_EINVAL;
    if (aux) {
        int rc = b200m_l2_i8_aux_init(aux, rows, D, stream);
        if (rc != B200M_OK) return rc;
    }
    return b200m_prep_l2_i8_rows(src, 0, rows, rows, D, dst, Dpad, aux, not_integer, stream);
}

int b200m_prep_l2_f32(const float* src, int64_t rows, int D, float* norms, void* stream) {
    if (rows < 0 || D <= 0) return B200M_EINVAL;
    if (rows == 0) return B200M_OK;
    if (!src || !norms) return B200M_EINVAL;
    long long blocks = (rows + 7) / 8;
    if (blocks > 0x7fffffffLL) return B200M_EUNSUPPORTED;
    prep_l2_f32_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, rows, D, norms);
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "prep_l2_f32_kernel launch");
    return B200M_OK;
}

}  // extern "C"
