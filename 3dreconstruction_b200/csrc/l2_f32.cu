// l2_f32.cu — placeholder until the TF32 candidate + fp32 re-rank kernel lands.
#include "common.cuh"
int launch_l2_f32_top2(const float*, const float*, int64_t, const float*, const float*, int64_t, int,
                       const DevPair*, int, int, int, int, bool, Top2*, long long, cudaStream_t) {
    return B200M_EUNSUPPORTED;
}
