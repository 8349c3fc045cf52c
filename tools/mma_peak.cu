// tools/mma_peak.cu — microbenchmark: raw tcgen05.mma issue rate per .kind on this GPU (no epilogue, no TMA).
// One CTA per SM; one thread issues back-to-back M=128 N=256 K=32-byte MMAs on (uninitialised) smem operands
// into two TMEM accumulators.  Prints TOP/s per kind: the honest tensor-pipe roofline for each data type.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/mma_peak tools/mma_peak.cu && tools/mma_peak
#include <cstring>
#include <cstdlib>
#include <cstdio>
#include <cuda_runtime.h>
#include "../3dreconstruction_b200/csrc/sm100.cuh"

#define MMA_KIND(NAME, KINDSTR)                                                                          \
    __device__ __forceinline__ void NAME(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) { \
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"                                   \
                     "tcgen05.mma.cta_group::1.kind::" KINDSTR " [%0], %1, %2, %3, p;\n\t}" ::"r"(d),    \
                     "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");                                   \
    }
MMA_KIND(mma_k_i8, "i8")
MMA_KIND(mma_k_f16, "f16")
MMA_KIND(mma_k_tf32, "tf32")
MMA_KIND(mma_k_f8, "f8f6f4")

template <int KIND>
__global__ void __launch_bounds__(128, 1) peak_kernel(int iters, uint32_t idesc, long long* cycles, int random_fill = 0) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t tptr;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (48 * 1024) / 4; i += blockDim.x) {
        uint32_t v = 0x3c003c00u;                                  // benign values
        if (random_fill) { v = (uint32_t)i * 2654435761u + blockIdx.x * 40503u; v ^= v >> 15; v *= 2246822519u; v ^= v >> 13; }
        ((uint32_t*)smem)[i] = v;                                  // random bytes: realistic operand toggling for power
    }
    if (threadIdx.x == 0) { sm100::mbar_init(&bar, 1); sm100::fence_barrier_init(); }
    if (warp == 0) sm100::tmem_alloc<512>(&tptr);
    sm100::fence_proxy_async();
    sm100::tc_fence_before(); __syncthreads(); sm100::tc_fence_after();
    uint32_t tbase = tptr;
    if (threadIdx.x == 32) {
        uint64_t ad = sm100::make_smem_desc_sw128(sm100::smem_u32(smem));
        uint64_t bd = sm100::make_smem_desc_sw128(sm100::smem_u32(smem + 16384));
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                uint32_t d = tbase + ((it & 1) ? 256 : 0);
                if (KIND == 0) mma_k_i8(d, ad + 2 * k, bd + 2 * k, idesc, 1);
                if (KIND == 1) mma_k_f16(d, ad + 2 * k, bd + 2 * k, idesc, 1);
                if (KIND == 2) mma_k_tf32(d, ad + 2 * k, bd + 2 * k, idesc, 1);
                if (KIND == 3) mma_k_f8(d, ad + 2 * k, bd + 2 * k, idesc, 1);
            }
        }
        sm100::mma_commit(&bar);
        sm100::mbar_wait(&bar, 0);
        long long t1 = clock64();
        if (blockIdx.x == 0) *cycles = t1 - t0;
    }
    sm100::tc_fence_before(); __syncthreads();
    if (warp == 0) sm100::tmem_dealloc<512>(tbase);
}

// kind::mxf4 (block-scaled E2M1, K = 64 elements = 32 bytes per MMA): accumulators of N columns at TMEM columns 0 and 256,
// scale factors (UE8M0 1.0 = 0x7F everywhere, so their layout does not matter) at columns 240..255.
__device__ __forceinline__ void mma_k_mxf4(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t sfa, uint32_t sfb, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %6, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%4], [%5], p;\n\t}" ::"r"(d),
                 "l"(a), "l"(b), "r"(idesc), "r"(sfa), "r"(sfb), "r"(acc) : "memory");
}
__global__ void __launch_bounds__(128, 1) peak_mxf4_kernel(int iters, uint32_t idesc, long long* cycles, int random_fill) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t tptr;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (48 * 1024) / 4; i += blockDim.x) {
        uint32_t v = 0x2A2A2A2Au;                                  // nibbles +1 / -1
        if (random_fill) { v = (uint32_t)i * 2654435761u + blockIdx.x * 40503u; v ^= v >> 15; v *= 2246822519u; v ^= v >> 13;
                           v = (v & 0x88888888u) | 0x22222222u; }  // random signs, magnitude 1.0
        ((uint32_t*)smem)[i] = v;
    }
    if (threadIdx.x == 0) { sm100::mbar_init(&bar, 1); sm100::fence_barrier_init(); }
    if (warp == 0) sm100::tmem_alloc<512>(&tptr);
    sm100::fence_proxy_async();
    sm100::tc_fence_before(); __syncthreads(); sm100::tc_fence_after();
    uint32_t tbase = tptr;
    {   // scale factors: every warp fills its 32 TMEM lanes x 16 columns with 0x7F7F7F7F
        uint32_t taddr = tbase + ((uint32_t)(warp * 32) << 16) + 240;
        uint32_t v = 0x7F7F7F7Fu;
        asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};"
                     ::"r"(taddr), "r"(v) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    sm100::tc_fence_before(); __syncthreads(); sm100::tc_fence_after();
    if (threadIdx.x == 32) {
        uint64_t ad = sm100::make_smem_desc_sw128(sm100::smem_u32(smem));
        uint64_t bd = sm100::make_smem_desc_sw128(sm100::smem_u32(smem + 16384));
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                mma_k_mxf4(tbase + ((it & 1) ? 256 : 0), ad + 2 * k, bd + 2 * k, idesc, tbase + 240, tbase + 240, 1);
        }
        sm100::mma_commit(&bar);
        sm100::mbar_wait(&bar, 0);
        long long t1 = clock64();
        if (blockIdx.x == 0) *cycles = t1 - t0;
    }
    sm100::tc_fence_before(); __syncthreads();
    if (warp == 0) sm100::tmem_dealloc<512>(tbase);
}
// idesc of kind::mxf4: a/b format E2M1 = 1 at bits 7 and 10, N >> 3 at 17, scale format UE8M0 at bit 23, M >> 4 at 24
static uint32_t idesc_mxf4(int N) { return (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | (1u << 23) | (8u << 24); }
static void run_mxf4(int N, int sms, double sustained_s) {
    long long* dcyc; cudaMalloc(&dcyc, 8);
    cudaFuncSetAttribute(peak_mxf4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = sustained_s > 0 ? 100000 : 20000;
    peak_mxf4_kernel<<<sms, 128, 64 * 1024>>>(200, idesc_mxf4(N), dcyc, 0);
    double elapsed = 0; int n = 0;
    do {
        cudaEventRecord(e0);
        peak_mxf4_kernel<<<sms, 128, 64 * 1024>>>(iters, idesc_mxf4(N), dcyc, sustained_s > 0 ? 1 : 0);
        cudaEventRecord(e1);
        cudaError_t err = cudaDeviceSynchronize();
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        long long cyc = 0; cudaMemcpy(&cyc, dcyc, 8, cudaMemcpyDeviceToHost);
        double macs = (double)sms * iters * 4.0 * 128 * N * 64;
        elapsed += ms; ++n;
        if (sustained_s <= 0 || n % 8 == 0)
            printf("mxf4 N%d %s t=%.2fs: %.3f ms, %.1f TOP/s, %.1f cycles per MMA (M128 K=64), %.0f MAC/clk/SM, SM clock %.0f MHz\n", N,
                   cudaGetErrorString(err), elapsed / 1e3, ms, 2 * macs / (ms * 1e-3) / 1e12, (double)cyc / (iters * 4.0),
                   128.0 * N * 64 / ((double)cyc / (iters * 4.0)), (double)cyc / ms / 1e3);
        if (err != cudaSuccess) break;
    } while (elapsed < sustained_s * 1e3);
    cudaFree(dcyc);
}

template <int KIND>
void run(const char* name, uint32_t idesc, int k_elems_per_mma, int sms) {
    long long* dcyc; cudaMalloc(&dcyc, 8);
    cudaFuncSetAttribute(peak_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    const int iters = 20000;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    peak_kernel<KIND><<<sms, 128, 64 * 1024>>>(200, idesc, dcyc);
    cudaEventRecord(e0);
    peak_kernel<KIND><<<sms, 128, 64 * 1024>>>(iters, idesc, dcyc);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    long long cyc = 0; cudaMemcpy(&cyc, dcyc, 8, cudaMemcpyDeviceToHost);
    double macs = (double)sms * iters * 4.0 * 128 * 256 * k_elems_per_mma;
    printf("%-8s %s: %.3f ms, %.1f TOP/s, %.1f cycles per MMA (M128 N256 K=%d), %.0f MAC/clk/SM\n", name,
           cudaGetErrorString(err), ms, 2 * macs / (ms * 1e-3) / 1e12, (double)cyc / (iters * 4.0), k_elems_per_mma,
           128.0 * 256 * k_elems_per_mma / ((double)cyc / (iters * 4.0)));
    cudaFree(dcyc);
}


// Variant: tile = 4 MMAs (one 128-byte K block) into stage (it % nstage), then `commits` tcgen05.commit on scratch barriers.
__global__ void __launch_bounds__(128, 1) tile_kernel(int iters, uint32_t idesc, int ncols, int nstage, int commits, long long* cycles) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar, scratch[2];
    __shared__ uint32_t tptr;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (48 * 1024) / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0x01010101u;
    if (threadIdx.x == 0) { sm100::mbar_init(&bar, 1); sm100::mbar_init(&scratch[0], 1); sm100::mbar_init(&scratch[1], 1); sm100::fence_barrier_init(); }
    if (warp == 0) sm100::tmem_alloc<512>(&tptr);
    sm100::fence_proxy_async();
    sm100::tc_fence_before(); __syncthreads(); sm100::tc_fence_after();
    uint32_t tbase = tptr;
    if (threadIdx.x == 32) {
        uint64_t ad = sm100::make_smem_desc_sw128(sm100::smem_u32(smem));
        uint64_t bd = sm100::make_smem_desc_sw128(sm100::smem_u32(smem + 16384));
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            uint32_t d = tbase + (it % nstage) * ncols;
#pragma unroll
            for (int k = 0; k < 4; ++k) mma_k_i8(d, ad + 2 * k, bd + 2 * k, idesc, k ? 1 : 0);
            for (int c = 0; c < commits; ++c) sm100::mma_commit(&scratch[c]);
        }
        sm100::mma_commit(&bar);
        sm100::mbar_wait(&bar, 0);
        long long t1 = clock64();
        if (blockIdx.x == 0) *cycles = t1 - t0;
    }
    sm100::tc_fence_before(); __syncthreads();
    if (warp == 0) sm100::tmem_dealloc<512>(tbase);
}

void run_tile(int N, int nstage, int commits, int sms) {
    long long* dcyc; cudaMalloc(&dcyc, 8);
    cudaFuncSetAttribute(tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    const int iters = 20000;
    uint32_t idesc = sm100::make_idesc(2, 1, 1, 128, N);
    tile_kernel<<<sms, 128, 64 * 1024>>>(200, idesc, N, nstage, commits, dcyc);
    tile_kernel<<<sms, 128, 64 * 1024>>>(iters, idesc, N, nstage, commits, dcyc);
    cudaError_t err = cudaDeviceSynchronize();
    long long cyc = 0; cudaMemcpy(&cyc, dcyc, 8, cudaMemcpyDeviceToHost);
    printf("i8 tile M128 N%-3d stages %d commits/tile %d: %s, %.1f cycles per tile (floor %d)\n", N, nstage, commits,
           cudaGetErrorString(err), (double)cyc / iters, N * 2);
    cudaFree(dcyc);
}

// Contention probe: the MMA thread issues N=256 tiles (4 MMAs + 2 commits) while 16 other warps hammer
// TMEM (tcgen05.ld of the OTHER stage) and/or shared memory (broadcast LDS.128), like the epilogue does.
__global__ void __launch_bounds__(576, 1) contend_kernel(int iters, uint32_t idesc, int act, long long* cycles, int* sink) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar, scratch[2];
    __shared__ uint32_t tptr;
    __shared__ volatile int done;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < (56 * 1024) / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0x01010101u;
    if (threadIdx.x == 0) { done = 0; sm100::mbar_init(&bar, 1); sm100::mbar_init(&scratch[0], 1); sm100::mbar_init(&scratch[1], 1); sm100::fence_barrier_init(); }
    if (warp == 16) sm100::tmem_alloc<512>(&tptr);
    sm100::fence_proxy_async();
    sm100::tc_fence_before(); __syncthreads(); sm100::tc_fence_after();
    uint32_t tbase = tptr;
    if (warp == 17) {
        if (lane == 0) {
            uint64_t ad = sm100::make_smem_desc_sw128(sm100::smem_u32(smem));
            uint64_t bd = sm100::make_smem_desc_sw128(sm100::smem_u32(smem + 16384));
            long long t0 = clock64();
            for (int it = 0; it < iters; ++it) {
                uint32_t d = tbase + (it & 1) * 256;
#pragma unroll
                for (int k = 0; k < 4; ++k) mma_k_i8(d, ad + 2 * k, bd + 2 * k, idesc, k ? 1 : 0);
                sm100::mma_commit(&scratch[0]);
                sm100::mma_commit(&scratch[1]);
            }
            sm100::mma_commit(&bar);
            sm100::mbar_wait(&bar, 0);
            long long t1 = clock64();
            if (blockIdx.x == 0) *cycles = t1 - t0;
            done = 1;
        }
    } else if (warp < 16) {
        const uint32_t taddr = tbase + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 64;
        const uint32_t cb = sm100::smem_u32(smem + 49152);
        int acc_x = 0;
        long long n = 0;
        while (!done) {
            for (int rep = 0; rep < 16; ++rep) {
                if (act & 1) {
                    uint32_t v[32];
                    sm100::tmem_ld32(taddr + ((rep & 1) ? 256 : 0), v);
                    sm100::tmem_ld32(taddr + ((rep & 1) ? 256 : 0) + 32, v);
                    sm100::tmem_ld_wait();
                    acc_x ^= v[lane & 31];
                }
                if (act & 2) {
#pragma unroll
                    for (int q = 0; q < 16; ++q) { int4 c = sm100::lds128(cb + q * 16); acc_x += c.x ^ c.y ^ c.z ^ c.w; }
                }
                if (act & 4) {
#pragma unroll
                    for (int q = 0; q < 64; ++q) acc_x = acc_x * -128 + q;
                }
                ++n;
            }
        }
        if (acc_x == 0x7654321) sink[0] = acc_x;
        if (blockIdx.x == 0 && threadIdx.x == 0) cycles[1] = n;
    }
    sm100::tc_fence_before(); __syncthreads();
    if (warp == 16) sm100::tmem_dealloc<512>(tbase);
}

void run_contend(int act, int sms) {
    long long* dcyc; cudaMalloc(&dcyc, 16); int* sink; cudaMalloc(&sink, 4);
    cudaFuncSetAttribute(contend_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    const int iters = 20000;
    uint32_t idesc = sm100::make_idesc(2, 1, 1, 128, 256);
    contend_kernel<<<sms, 576, 64 * 1024>>>(iters, idesc, act, dcyc, sink);
    cudaError_t err = cudaDeviceSynchronize();
    long long cyc[2] = {0, 0}; cudaMemcpy(cyc, dcyc, 16, cudaMemcpyDeviceToHost);
    printf("contend act=%d (1=LDTM 2=LDS.128 bcast 4=IMAD): %s, %.1f cycles per N256 tile (floor 512, alone 590); side-warp iterations per tile %.2f\n",
           act, cudaGetErrorString(err), (double)cyc[0] / iters, (double)cyc[1] / iters);
    cudaFree(dcyc); cudaFree(sink);
}

// cycles per MMA instruction as a function of N (no commits inside the loop), i8 and tf32
template <int KIND>
__global__ void __launch_bounds__(128, 1) nsweep_kernel(int iters, uint32_t idesc, long long* cycles) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t tptr;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (48 * 1024) / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0x01010101u;
    if (threadIdx.x == 0) { sm100::mbar_init(&bar, 1); sm100::fence_barrier_init(); }
    if (warp == 0) sm100::tmem_alloc<512>(&tptr);
    sm100::fence_proxy_async();
    sm100::tc_fence_before(); __syncthreads(); sm100::tc_fence_after();
    uint32_t tbase = tptr;
    if (threadIdx.x == 32) {
        uint64_t ad = sm100::make_smem_desc_sw128(sm100::smem_u32(smem));
        uint64_t bd = sm100::make_smem_desc_sw128(sm100::smem_u32(smem + 16384));
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            uint32_t d = tbase + ((it & 1) ? 256 : 0);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (KIND == 0) mma_k_i8(d, ad + 2 * k, bd + 2 * k, idesc, 1);
                else mma_k_tf32(d, ad + 2 * k, bd + 2 * k, idesc, 1);
            }
        }
        sm100::mma_commit(&bar);
        sm100::mbar_wait(&bar, 0);
        long long t1 = clock64();
        if (blockIdx.x == 0) *cycles = t1 - t0;
    }
    sm100::tc_fence_before(); __syncthreads();
    if (warp == 0) sm100::tmem_dealloc<512>(tbase);
}
void run_nsweep(int sms) {
    long long* dcyc; cudaMalloc(&dcyc, 8);
    cudaFuncSetAttribute(nsweep_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    cudaFuncSetAttribute(nsweep_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    const int iters = 10000;
    for (int kind = 0; kind < 2; ++kind)
        for (int N = 32; N <= 256; N += 32) {
            uint32_t idesc = kind == 0 ? sm100::make_idesc(2, 1, 1, 128, N) : sm100::make_idesc(1, 2, 2, 128, N);
            if (kind == 0) nsweep_kernel<0><<<sms, 128, 64 * 1024>>>(iters, idesc, dcyc);
            else nsweep_kernel<1><<<sms, 128, 64 * 1024>>>(iters, idesc, dcyc);
            cudaError_t err = cudaDeviceSynchronize();
            long long cyc = 0; cudaMemcpy(&cyc, dcyc, 8, cudaMemcpyDeviceToHost);
            printf("%s M128 N%-3d K=32B: %s, %.1f cycles per MMA (ideal N/2 = %d)\n", kind ? "tf32" : "i8  ", N,
                   cudaGetErrorString(err), (double)cyc / (iters * 4.0), N / 2);
        }
    cudaFree(dcyc);
}

// Sustained int8 issue rate under the power cap: the i8 peak kernel (operands resident in shared memory, random bytes,
// no TMA, no epilogue) back to back for `seconds`; prints each launch's TOP/s and SM cycles/ms so the settled clock shows.
static void run_sustained(int sms, double seconds) {
    long long* dcyc; cudaMalloc(&dcyc, 8);
    cudaFuncSetAttribute(peak_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    const uint32_t idesc = sm100::make_idesc(2, 1, 1, 128, 256);
    const int iters = 100000;                                       // ~26 ms per launch at 1.965 GHz
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    double elapsed = 0; int n = 0;
    while (elapsed < seconds * 1e3) {
        cudaEventRecord(e0);
        peak_kernel<0><<<sms, 128, 64 * 1024>>>(iters, idesc, dcyc, 1);
        cudaEventRecord(e1);
        cudaDeviceSynchronize();
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        long long cyc = 0; cudaMemcpy(&cyc, dcyc, 8, cudaMemcpyDeviceToHost);
        double macs = (double)sms * iters * 4.0 * 128 * 256 * 32;
        elapsed += ms; ++n;
        if (n % 8 == 0)
            printf("sustained i8 t=%.2fs: %.1f TOP/s, SM clock %.0f MHz (cycles/ms), %.1f cycles per MMA\n", elapsed / 1e3,
                   2 * macs / (ms * 1e-3) / 1e12, (double)cyc / ms / 1e3, (double)cyc / (iters * 4.0));
    }
    cudaFree(dcyc);
}

// ---------------------------------------------------------------- cta_group::2: one MMA spans a CTA pair
// Cluster of 2 CTAs; the leader's thread 32 issues M=256 (128 rows of A per CTA) x N (N/2 rows of B per CTA) MMAs.
// Answers: does one M256 N256 K32 kind::i8 instruction take the 128 cycles of the M128 N256 one (2x per-SM-pair rate)
// or 256 (no gain over two cta_group::1 streams)?
#define MMA_KIND2(NAME, KINDSTR)                                                                         \
    __device__ __forceinline__ void NAME(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) { \
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"                                   \
                     "tcgen05.mma.cta_group::2.kind::" KINDSTR " [%0], %1, %2, %3, p;\n\t}" ::"r"(d),    \
                     "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");                                   \
    }
MMA_KIND2(mma2_k_i8, "i8")
MMA_KIND2(mma2_k_f16, "f16")

template <int KIND>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
pair_kernel(int iters, uint32_t idesc, int ncols, long long* cycles) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t tptr;
    const int warp = threadIdx.x >> 5;
    const uint32_t rank = sm100::cluster_ctarank();
    for (int i = threadIdx.x; i < (48 * 1024) / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0x01010101u;
    if (threadIdx.x == 0) { sm100::mbar_init(&bar, 1); sm100::fence_barrier_init(); }
    sm100::cluster_sync_all();
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sm100::smem_u32(&tptr)),
                     "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    sm100::fence_proxy_async();
    sm100::tc_fence_before(); sm100::cluster_sync_all(); sm100::tc_fence_after();
    uint32_t tbase = tptr;
    if (rank == 0 && threadIdx.x == 32) {
        uint64_t ad = sm100::make_smem_desc_sw128(sm100::smem_u32(smem));
        uint64_t bd = sm100::make_smem_desc_sw128(sm100::smem_u32(smem + 16384));
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            uint32_t d = tbase + ((it & 1) ? ncols : 0);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (KIND == 0) mma2_k_i8(d, ad + 2 * k, bd + 2 * k, idesc, 1);
                else mma2_k_f16(d, ad + 2 * k, bd + 2 * k, idesc, 1);
            }
        }
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                     ::"r"(sm100::smem_u32(&bar)), "h"((uint16_t)3) : "memory");
        sm100::mbar_wait(&bar, 0);
        long long t1 = clock64();
        if (blockIdx.x == 0) *cycles = t1 - t0;
    } else if (threadIdx.x == 32) {
        sm100::mbar_wait(&bar, 0);
    }
    sm100::tc_fence_before(); sm100::cluster_sync_all();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tbase), "n"(512) : "memory");
}

template <int KIND>
static void run_pair(const char* name, int N, int k_elems_per_mma, int sms, uint32_t cfmt, uint32_t abfmt) {
    const int iters = 20000;
    long long* dcyc; cudaMalloc(&dcyc, 8); cudaMemset(dcyc, 0, 8);
    cudaFuncSetAttribute(pair_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    const uint32_t idesc = sm100::make_idesc(cfmt, abfmt, abfmt, 256, N);
    const int grid = sms & ~1;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    pair_kernel<KIND><<<grid, 128, 64 * 1024>>>(200, idesc, N > 256 ? 256 : N, dcyc);
    cudaEventRecord(e0);
    pair_kernel<KIND><<<grid, 128, 64 * 1024>>>(iters, idesc, N > 256 ? 256 : N, dcyc);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    long long cyc = 0; cudaMemcpy(&cyc, dcyc, 8, cudaMemcpyDeviceToHost);
    double macs = (double)(grid / 2) * iters * 4.0 * 256 * N * k_elems_per_mma;
    printf("pair %-5s %s: %.3f ms, %.1f TOP/s, %.1f cycles per MMA (cta_group::2 M256 N%d K=%d), %.0f MAC/clk/SM\n", name,
           cudaGetErrorString(err), ms, 2 * macs / (ms * 1e-3) / 1e12, (double)cyc / (iters * 4.0), N, k_elems_per_mma,
           256.0 * N * k_elems_per_mma / ((double)cyc / (iters * 4.0)) / 2);
    cudaFree(dcyc);
}

int main(int argc, char** argv) {
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    printf("SMs: %d\n", sms);
    if (argc > 1 && !strcmp(argv[1], "mxf4")) {
        run_mxf4(256, sms, 0); run_mxf4(240, sms, 0); run_mxf4(128, sms, 0);
        run_mxf4(240, sms, argc > 2 ? atof(argv[2]) : 3.0);
        return 0;
    }
    if (argc > 1 && !strcmp(argv[1], "pair")) {
        run<0>("i8", sm100::make_idesc(2, 1, 1, 128, 256), 32, sms);
        run_pair<0>("i8", 256, 32, sms, 2, 1); run_pair<0>("i8", 128, 32, sms, 2, 1); run_pair<0>("i8", 64, 32, sms, 2, 1);
        run<1>("f16", sm100::make_idesc(1, 0, 0, 128, 256), 16, sms);
        run_pair<1>("f16", 256, 16, sms, 1, 0); run_pair<1>("f16", 128, 16, sms, 1, 0);
        return 0;
    }
    if (argc > 1 && !strcmp(argv[1], "sustained")) { run_sustained(sms, argc > 2 ? atof(argv[2]) : 4.0); return 0; }
    run<0>("i8", sm100::make_idesc(2, 1, 1, 128, 256), 32, sms);
    run<1>("f16", sm100::make_idesc(1, 0, 0, 128, 256), 16, sms);
    run<1>("bf16", sm100::make_idesc(1, 1, 1, 128, 256), 16, sms);
    run<2>("tf32", sm100::make_idesc(1, 2, 2, 128, 256), 8, sms);
    run<3>("e4m3", sm100::make_idesc(1, 0, 0, 128, 256), 32, sms);
    run_tile(256, 2, 0, sms); run_tile(256, 2, 2, sms); run_tile(128, 4, 0, sms); run_tile(128, 4, 1, sms); run_tile(128, 4, 2, sms);
    run_tile(64, 8, 2, sms); run_tile(128, 1, 2, sms);
    for (int act = 0; act < 4; ++act) run_contend(act, sms);
    run_nsweep(sms);
    return 0;
}
