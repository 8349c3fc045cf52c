// adapter.cu — matches -> correspondences for the structure-from-motion stage.
// The reference builds, per image pair and per match, pts1 = keypoints1[m.queryIdx].pt, pts2 = keypoints2[m.trainIdx].pt
// and stacks them into an N x 4 array (/root/reference/FeatureMatchingExtraction/result_converters.py:763-766), which
// is what CameraPoseEstimation consumes (pair_selector.py:54-67, correspondence_manager.py:78-119).  Here the same
// rows come from ONE gather over the CSR the matcher already produced: keypoint coordinates live next to the
// descriptor store in HBM (float32 [rows, 2], cv2.KeyPoint.pt is float32), one CTA per pair, one thread per match,
// a 16-byte store per match.  HBM-bound: 8 B of indices + 16 B of coordinates read, 16 B written per match.
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(128)
gather_correspondences_kernel(const float2* __restrict__ kp_xy, const int32_t* __restrict__ pair_q_off,
                              const int32_t* __restrict__ pair_t_off, const long long* __restrict__ ptr,
                              const int32_t* __restrict__ q_idx, const int32_t* __restrict__ t_idx,
                              float4* __restrict__ out) {
    const int p = blockIdx.x;
    const long long s = ptr[p], e = ptr[p + 1];
    const float2* q = kp_xy + pair_q_off[p];
    const float2* t = kp_xy + pair_t_off[p];
    for (long long k = s + threadIdx.x; k < e; k += blockDim.x) {
        const float2 a = __ldg(q + q_idx[k]), b = __ldg(t + t_idx[k]);
        out[k] = make_float4(a.x, a.y, b.x, b.y);
    }
}

}  // namespace

extern "C" {

int b200m_gather_correspondences(const float* kp_xy, int64_t rows, const int32_t* pair_q_off, const int32_t* pair_t_off,
                                 const int64_t* ptr, int num_pairs, const int32_t* q_idx, const int32_t* t_idx,
                                 float* out_xyxy, void* stream) {
    if (num_pairs < 0 || rows < 0) return B200M_EINVAL;
    if (num_pairs == 0) return B200M_OK;
    if (!kp_xy || !pair_q_off || !pair_t_off || !ptr || !q_idx || !t_idx || !out_xyxy) return B200M_EINVAL;
    if (((uintptr_t)kp_xy & 7) || ((uintptr_t)out_xyxy & 15)) return B200M_EINVAL;
    gather_correspondences_kernel<<<(unsigned)num_pairs, 128, 0, (cudaStream_t)stream>>>(
        (const float2*)kp_xy, pair_q_off, pair_t_off, (const long long*)ptr, q_idx, t_idx, (float4*)out_xyxy);
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "gather_correspondences_kernel launch");
    return B200M_OK;
}

}  // extern "C"
