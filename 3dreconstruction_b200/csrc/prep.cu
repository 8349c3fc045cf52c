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
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256)
prep_l2_i8_kernel(const float* __restrict__ src, long long rows, int D, int8_t* __restrict__ dst, int Dpad,
                  int32_t* __restrict__ norms, int32_t* __restrict__ not_integer) {
    const int lane = threadIdx.x & 31;
    long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const float* s = src + row * D;
    int8_t* d = dst + row * Dpad;
    int acc = 0, bad = 0;
    // Dpad is a multiple of 128: each lane converts 4 consecutive elements per step.
    for (int k = lane * 4; k < Dpad; k += 128) {
        int v[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            int kk = k + e;
            if (kk < D) {
                float f = __ldg(s + kk);
                float r = rintf(f);
                bad |= !(f == r && r >= 0.0f && r <= 255.0f);
                int iv = (int)fminf(fmaxf(r, 0.0f), 255.0f) - 128;
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
    if (lane == 0) {
        norms[row] = acc;
        if (bad) atomicOr(not_integer, 1);
    }
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

}  // namespace

extern "C" {

int b200m_prep_l2_i8(const float* src, int64_t rows, int D, int8_t* dst, int Dpad, int32_t* norms,
                     int32_t* not_integer, void* stream) {
    if (rows < 0 || D <= 0 || Dpad < D || Dpad % 128 != 0) return B200M_EINVAL;
    if (rows == 0) return B200M_OK;
    if (!src || !dst || !norms || !not_integer) return B200M_EINVAL;
    long long blocks = (rows + 7) / 8;
    if (blocks > 0x7fffffffLL) return B200M_EUNSUPPORTED;
    prep_l2_i8_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, rows, D, dst, Dpad, norms, not_integer);
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "prep_l2_i8_kernel launch");
    return B200M_OK;
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
