// tools/tmem_layout.cu — prints which (TMEM lane, column) each thread/register receives from
// tcgen05.ld.16x256b (the layout the l2_i8 epilogue relies on).  Values are written with tcgen05.st.32x32b
// (lane = row, register = column, known layout) as row*1000 + col.
#include <cstdio>
#include <cuda_runtime.h>
#include "../3dreconstruction_b200/csrc/sm100.cuh"

__global__ void __launch_bounds__(128, 1) probe(int* out) {
    __shared__ uint32_t tptr;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) sm100::tmem_alloc<64>(&tptr);
    sm100::tc_fence_before(); __syncthreads(); sm100::tc_fence_after();
    const uint32_t tbase = tptr;
    // every warp writes its own lane quarter: 32 columns
    {
        uint32_t v[32];
        for (int c = 0; c < 32; ++c) v[c] = (warp * 32 + lane) * 1000 + c;
        const uint32_t taddr = tbase + ((uint32_t)(warp * 32) << 16);
        asm volatile(
            "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
            "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
            "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
            "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
            "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]),
            "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]),
            "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
            : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    sm100::tc_fence_before(); __syncthreads(); sm100::tc_fence_after();
    // warp 1 reads its quarter (lanes 32..63) with 16x256b.x2 at lane offsets 0 and 16
    if (warp == 1) {
        for (int half = 0; half < 2; ++half) {
            uint32_t r[8];
            const uint32_t taddr = tbase + ((uint32_t)(32 + half * 16) << 16);
            asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                         : "r"(taddr) : "memory");
            sm100::tmem_ld_wait();
            for (int k = 0; k < 8; ++k) out[(half * 32 + lane) * 8 + k] = (int)r[k];
        }
    }
    sm100::tc_fence_before(); __syncthreads();
    if (warp == 0) sm100::tmem_dealloc<64>(tbase);
}

int main() {
    int* d; cudaMalloc(&d, 64 * 8 * 4);
    probe<<<1, 128>>>(d);
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    int h[64 * 8]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    for (int half = 0; half < 2; ++half)
        for (int lane = 0; lane < 32; ++lane) {
            printf("half %d lane %2d:", half, lane);
            for (int k = 0; k < 8; ++k) printf(" r%d=(row %d,col %d)", k, h[(half * 32 + lane) * 8 + k] / 1000, h[(half * 32 + lane) * 8 + k] % 1000);
            printf("\n");
        }
    return 0;
}
