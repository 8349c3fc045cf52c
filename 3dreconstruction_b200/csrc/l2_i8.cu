// l2_i8.cu — placeholder until the tcgen05 kind::i8 kernel lands.
#include "common.cuh"
int launch_l2_i8_top2(const int8_t*, const int32_t*, int64_t, const int8_t*, const int32_t*, int64_t, int,
                      const DevPair*, int, int, int, int, bool, Top2*, long long, cudaStream_t) {
    return B200M_EUNSUPPORTED;
}
