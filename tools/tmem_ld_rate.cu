// tools/tmem_ld_rate.cu — microbenchmark: tcgen05.ld (32x32b.x32) throughput per SM sub-partition, alone and while the
// tensor pipe runs back-to-back kind::i8 M128 N256 MMAs into the same tensor memory.  Answers how long the epilogue
// of a 128 x 256 int32 accumulator tile must hold a TMEM stage (the floor of l2_i8_fold_kernel's stage turnaround).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tmem_ld_rate tools/tmem_ld_rate.cu && tools/tmem_ld_rate
#include <cstdio>
#include <cuda_runtime.h>
#include "../3dreconstruction_b200/csrc/sm100.cuh"

// warps 0..15 read (warp w: lane quarter w % 4), warp 16 allocates and (optionally) issues MMAs
__global__ void __launch_bounds__(17 * 32, 1) ld_rate_kernel(int iters, int nwarps, int with_mma, long long* cycles, uint32_t* sink) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t tptr;
    __shared__ int stop;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < (48 * 1024) / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0x01010101u;
    if (threadIdx.x == 0) { sm100::mbar_init(&bar, 1); sm100::fence_barrier_init(); stop = 0; }
    if (warp == 16) sm100::tmem_alloc<512>(&tptr);
    sm100::fence_proxy_async();
    sm100::tc_fence_before(); __syncthreads(); sm100::tc_fence_after();
    const uint32_t tbase = tptr;
    if (warp == 16) {
        if (with_mma && lane == 0) {
            const uint32_t idesc = sm100::make_idesc(2, 1, 1, 128, 256);
            uint64_t ad = sm100::make_smem_desc_sw128(sm100::smem_u32(smem));
            uint64_t bd = sm100::make_smem_desc_sw128(sm100::smem_u32(smem + 16384));
            int n = 0;
            while (*(volatile int*)&stop < nwarps && n < (1 << 22)) {        // until every reading warp is done
#pragma unroll
                for (int k = 0; k < 4; ++k) sm100::mma_i8(tbase + ((n & 1) ? 256 : 0), ad + 2 * k, bd + 2 * k, idesc, 1);
                ++n;
                if ((n & 7) == 0) { sm100::mma_commit(&bar); sm100::mbar_wait(&bar, ((n >> 3) - 1) & 1); }   // bounded queue depth
            }
        }
    } else if (warp < nwarps) {
        const uint32_t taddr = tbase + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 64;
        uint32_t a[32], b[32], x = 0;
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            sm100::tmem_ld32(taddr + ((it & 1) ? 256 : 0), a);
            sm100::tmem_ld32(taddr + ((it & 1) ? 256 : 0) + 32, b);
            sm100::tmem_ld_wait();
            x ^= a[0] ^ a[31] ^ b[5] ^ b[17];
        }
        long long t1 = clock64();
        if (blockIdx.x == 0 && lane == 0) cycles[warp] = t1 - t0;
        if (x == 0x12345678u) sink[0] = x;
        if (lane == 0) atomicAdd(&stop, 1);
    }
    sm100::tc_fence_before(); __syncthreads();
    if (warp == 16) sm100::tmem_dealloc<512>(tbase);
}

int main() {
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    long long* dcyc; cudaMalloc(&dcyc, 16 * 8);
    uint32_t* sink; cudaMalloc(&sink, 4);
    cudaFuncSetAttribute(ld_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    const int iters = 4000;
    for (int with_mma = 0; with_mma < 2; ++with_mma)
        for (int nw = 4; nw <= 16; nw *= 2) {
            cudaMemset(dcyc, 0, 16 * 8);
            ld_rate_kernel<<<sms, 17 * 32, 64 * 1024>>>(iters, nw, with_mma, dcyc, sink);
            cudaError_t err = cudaDeviceSynchronize();
            long long c[16]; cudaMemcpy(c, dcyc, sizeof(c), cudaMemcpyDeviceToHost);
            long long mx = 0; for (int i = 0; i < nw; ++i) mx = c[i] > mx ? c[i] : mx;
            const double per_ld = (double)mx / (iters * 2.0) / (nw / 4);       // port time per x32 load on one sub-partition
            printf("%s: %2d reading warps (%d per SMSP)%s: %.1f cycles per 32x32b.x32 load per SMSP = %.1f B/clk/SMSP; a 128x256 tile = %.0f cycles\n",
                   cudaGetErrorString(err), nw, nw / 4, with_mma ? ", MMAs running" : "", per_ld, 4096.0 / per_ld, per_ld * 8);
        }
    return 0;
}
