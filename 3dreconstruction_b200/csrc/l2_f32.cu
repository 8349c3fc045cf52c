// l2_f32.cu — L2 knn (k=2) for non-integer float32 descriptors (SuperPoint 256-d, DISK/ALIKED 128-d;
// /root/reference/FeatureMatchingExtraction/deep_learning_detectors.py:124,202,282) replacing
// cv2.BFMatcher(NORM_L2).knnMatch(k=2)/match as called at feature_matchers.py:210,227.
//
// Two stages:
//   1. l2_f32_cand_kernel — TF32 CANDIDATES on the tensor cores.  Same persistent, warp-specialised
//      pipeline as l2_i8.cu (TMA SWIZZLE_128B -> tcgen05.mma kind::tf32 M128 x N128 x K8, fp32
//      accumulators double-buffered in TMEM).  The epilogue forms the estimates
//      e_j = ||b_j||^2 - 2 a.b_j, keeps a branch-free running top-2 (v1, v2) of 16-column minima per
//      query row, and APPENDS every column with e_j <= v2 + window to a small per-row list.
//   2. refine_f32_kernel — exact fp32 RE-RANK.  The TF32 estimate of d^2 is off by at most
//      eps = C_EPS * ||a|| * max_j ||b_j||  (operands truncated to 10 mantissa bits: relative 2^-10
//      each, |sum a_k b_k| <= ||a|| ||b||, d^2 carries 2x the dot error; C_EPS = 0.0042 > 2^-8 covers
//      the fp32 accumulation and norm rounding as well), window = 2 eps.  The true top-2 lie among
//      the columns with e <= e_s + window, e_s the second smallest estimate of the row.  At any time
//      v2 (second smallest of minima of disjoint column groups seen so far) is >= e_s, so every such
//      column passes the append test when it is visited: the list is complete by construction
//      unless it overflows (counted -> the row is re-scanned exhaustively in fp32).  The few
//      survivors of the final bound get the direct fp32 sum (a-b)^2 — what cv2 computes.
// Result: Top2 keys (IEEE bits of the exact fp32 d^2 << 32 | column), ties -> lowest column.
#include <cuda_fp16.h>
#include <math_constants.h>
#include <stdlib.h>
#include "common.cuh"
#include "sm100.cuh"

namespace {

constexpr int TM = 128;            // query rows per tile  (UMMA M)
constexpr int TN = 128;            // train columns per tile (UMMA N)
constexpr int KB_BYTES = 128;      // one swizzle-128B K block = 32 floats
constexpr int A_KB_BYTES = TM * KB_BYTES;   // 16 KiB
constexpr int B_KB_BYTES = TN * KB_BYTES;   // 16 KiB
constexpr int MAX_NKB = 8;         // D <= 256 floats: the whole query tile stays resident (128 KiB)
constexpr int NB = 6;              // train K-block ring (as deep as 227 KB of shared memory allow)
constexpr int EPI_THREADS = 256;
constexpr int THREADS = EPI_THREADS + 64;
constexpr int KC = B200M_F32_CAND; // list capacity per (row, column half, column split)
constexpr float C_EPS = 0.0042f;
// fp16 candidates (kind L2_F16C): operands ROUNDED to 11 significant bits (relative 2^-11 each): |error of a.b| <=
// (2^-10 + 2^-22) ||a|| ||b||, d^2 carries twice that = 0.001953; C_H16 adds the same 7.5 % margin for the fp32
// accumulation and the norms.  Scaled elements below 2^-14 are fp16 subnormals with ABSOLUTE error <= 2^-25: the dot
// product moves by at most 2^-25 sqrt(Dh) (||b|| / s_a + ||a|| / s_b) in unscaled units (Cauchy-Schwarz on the 1-norm),
// d^2 by twice that -- the `sub` term of the window.
constexpr float C_H16 = 0.0021f;

struct Window {                 // eps(a) = c * ||a|| * bmax + sub_b * bmax + sub_a * ||a|| + quad * bmax^2;  window = 2 eps
    float c, sub_a, sub_b, quad;
    __device__ __forceinline__ float operator()(float a_norm2, float bmax) const {
        const float an = sqrtf(a_norm2);
        return 2.0f * (c * an * bmax + sub_b * bmax + sub_a * an + quad * bmax * bmax);
    }
};
// q_info / t_info: info words of the two stores' aux blocks (scale, 1 / scale, max |x|), null for the TF32 kind
__device__ __forceinline__ Window make_window(const float* q_info, const float* t_info, int Dh) {
    Window w;
    if (!q_info) { w.c = C_EPS; w.sub_a = 0.0f; w.sub_b = 0.0f; w.quad = 0.0f; return w; }
    const float g = 5.9604645e-8f * sqrtf((float)Dh);        // 2 * 2^-25 sqrt(Dh)
    w.c = C_H16;
    w.quad = 4.8e-7f;       // 2^-21: fp32 rounding of -(||b||^2 / 2) riding in the accumulator (FOLD kernels), with margin
    w.sub_b = g * __ldg(q_info + 1);                          // ||b|| / s_a
    w.sub_a = g * __ldg(t_info + 1);                          // ||a|| / s_b
    return w;
}
constexpr float BIG_F = 3.0e38f;   // ||b||^2 of a masked column: never appended, never a minimum that matters

struct Item {
    int q_off, n, t_off, m, r0, c_begin, c_end, split, pair;
    long long row_base;
    bool valid;
};

__device__ __forceinline__ Item decode_item(long long item, const DevPair* __restrict__ pairs, int tiles_max, int S,
                                            int backward, int rows_per_item = TM, int tn = TN) {
    Item it;
    int tile = (int)(item % tiles_max); item /= tiles_max;
    it.split = (int)(item % S);
    int p = (int)(item / S);
    it.pair = p;
    DevPair pr = pairs[p];
    if (!backward) { it.q_off = pr.q_off; it.n = pr.n; it.t_off = pr.t_off; it.m = pr.m; it.row_base = pr.fwd_off; }
    else           { it.q_off = pr.t_off; it.n = pr.m; it.t_off = pr.q_off; it.m = pr.n; it.row_base = pr.bwd_off; }
    it.r0 = tile * rows_per_item;
    int cols_per = ((it.m + S - 1) / S + tn - 1) / tn * tn;
    it.c_begin = it.split * cols_per;
    it.c_end = min(it.m, it.c_begin + cols_per);
    it.valid = it.r0 < it.n;
    return it;
}

constexpr int NB_FOLD = 5;         // ring depth of the FOLD kernels (20 KiB go to the folded-norm tiles)
constexpr int X_BYTES = TN * 32;   // folded-norm rows of one train tile (SWIZZLE_32B, 16 halves per row)
constexpr int AX_BYTES = TM * 32;  // the constant query-side block

struct SmemLayout {
    uint8_t* a;        // [nabuf][nkb][16 KiB]
    uint8_t* b;        // [nb][16 KiB]
    uint8_t* x;        // FOLD: [2][X_BYTES]
    uint8_t* ax;       // FOLD: [AX_BYTES]
    float* cb;         // [2][TN]   ||b||^2 of the tile's columns (+inf when masked)
    uint64_t* a_full; uint64_t* a_empty;   // [2]
    uint64_t* b_full; uint64_t* b_empty;   // [NB]
    uint64_t* acc_full; uint64_t* acc_empty;   // [2]
    uint32_t* tmem_ptr;
};

__host__ __device__ constexpr int nabuf_of(int nkb) { return nkb <= MAX_NKB / 2 ? 2 : 1; }

__device__ __forceinline__ SmemLayout carve(uint8_t* base, int nkb, bool fold = false) {
    SmemLayout s;
    s.a = base;
    s.b = s.a + nabuf_of(nkb) * nkb * A_KB_BYTES;
    uint8_t* p = s.b + (fold ? NB_FOLD : NB) * B_KB_BYTES;
    s.x = p; s.ax = p;
    if (fold) { s.ax = p + 2 * X_BYTES; p += 2 * X_BYTES + AX_BYTES; }
    s.cb = (float*)p; p += 2 * TN * sizeof(float);
    s.a_full = (uint64_t*)p; p += 2 * 8;
    s.a_empty = (uint64_t*)p; p += 2 * 8;
    s.b_full = (uint64_t*)p; p += NB * 8;
    s.b_empty = (uint64_t*)p; p += NB * 8;
    s.acc_full = (uint64_t*)p; p += 2 * 8;
    s.acc_empty = (uint64_t*)p; p += 2 * 8;
    s.tmem_ptr = (uint32_t*)p;
    return s;
}

__host__ __device__ constexpr size_t smem_bytes(int nkb, bool fold = false) {
    return 1024 + (size_t)nabuf_of(nkb) * nkb * A_KB_BYTES + (size_t)(fold ? NB_FOLD : NB) * B_KB_BYTES +
           (fold ? 2 * X_BYTES + AX_BYTES : 0) + 2 * TN * sizeof(float) + (8 + 2 * NB) * 8 + 16;
}

// H16: the tensor maps address the scaled fp16 copies (K block = 64 halves), the MMA is kind::f16 and the accumulator
// is s_q s_t a.b; everything else (pipeline, epilogue, lists) is shared with the TF32 kind.
// MT = query row tiles per work item (1 or 2).  With MT = 2 the two buffers that otherwise double-buffer one query tile
// hold two row tiles, and every train K block is issued against both, into two TMEM accumulators per stage pair: half
// the train bytes from L2 per MMA.  Used when the query tile is small enough (nkb <= 4: fp16 up to D = 256) and the
// batch still fills the SMs.  Measured: 2.46 -> 2.37 ms per 276-pair launch only -- the epilogue's instruction count
// (8 warps, ~0.5 IPC per scheduler; adding instructions slows the kernel proportionally), not the L2 -> SM path,
// bounds the fp16 kernel.
// FOLD (fp16 only): -(||b||^2 / 2) rides in the accumulator through one extra K = 16 step against a constant query-side
// row (the aux block's folded-norm rows, see common.cuh), so the epilogue compares RAW accumulators -- larger = closer --
// with no per-column constant: no shared-memory norm staging, no barrier per tile, one FMNMX instead of FMA + LDS + FMNMX
// per element.  fold_mode: 0 = run; 1 = run only if r = s_q / s_t is a small power of two (the FOLD kernel's own gate);
// 2 = return if it is (the plain kernel launched right behind a FOLD kernel: device-side dispatch, no host sync).
// PAIR (fp16 FOLD kernel, launched as clusters of two): the cluster is a tcgen05 CTA pair.  An item is still two query row
// tiles, but each CTA holds ONE of them and its own 128 rows of a 256-column train tile; one cta_group::2 MMA of M = 256,
// N = 256 covers both.  Why: an M128 x N128 instruction reads 8 KiB of operands in 64 cycles -- the whole 128 B/clk of shared
// memory -- so with the TMA fill on top the MT = 2 kernel sat at 0.76 of the f16 issue rate (156 B/clk demanded); per CTA of
// a pair it is 8 KiB per 128 cycles + 16 KiB of fill per 512 = 96 B/clk.  Protocol as in l2_i8_fold_kernel: leader-issued
// MMAs, both CTAs' TMA bytes on the leader's "full" barriers, multicast commits, remote TMEM-release arrives.
template <bool H16, int MT, bool FOLD, bool PAIR = false>
__global__ void __launch_bounds__(THREADS, 1)
l2_f32_cand_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_t,
                   const __grid_constant__ CUtensorMap tmap_x, int fold_mode,
                   const float* __restrict__ t_norm, const DevPair* __restrict__ pairs, long long total_items,
                   int tiles_max, int S, int backward, int nkb, const float* __restrict__ q_norm,
                   const float* __restrict__ pmax, float2* __restrict__ cand, int* __restrict__ cand_cnt,
                   long long total_rows, const float* __restrict__ q_info, const float* __restrict__ t_info) {
    static_assert(!PAIR || (H16 && FOLD && MT == 1), "the CTA-pair variant exists for the fp16 FOLD kernel");
    constexpr int KB_ELEMS = H16 ? KB_BYTES / 2 : KB_BYTES / 4;
    constexpr int NBK = FOLD ? NB_FOLD : NB;            // train K-block ring depth
    constexpr int TNK = PAIR ? 2 * TN : TN;             // train columns per tile (a CTA of a pair supplies TN of them)
    constexpr uint16_t MC_MASK = 3;
    const uint32_t crank = PAIR ? sm100::cluster_ctarank() : 0;
    const long long first_item = PAIR ? (long long)sm100::cluster_id_x() : (long long)blockIdx.x;
    const long long item_stride = PAIR ? (long long)sm100::cluster_count_x() : (long long)gridDim.x;
    if (fold_mode) {
        const float r = __ldg(q_info) * __ldg(t_info + 1);          // s_q / s_t, a power of two
        const bool foldable = r >= 0.00390625f && r <= 256.0f;
        if ((fold_mode == 1) != foldable) return;
    }
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const SmemLayout sm = carve(smem, nkb, FOLD);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nabuf = MT == 2 ? 1 : nabuf_of(nkb);      // MT == 2: both buffers hold the item's two row tiles
    constexpr int ROWS_ITEM = PAIR ? 2 * TM : TM * MT;  // PAIR: one row tile per CTA
    constexpr uint32_t ACC_COLS = PAIR ? TNK : TN * MT; // TMEM columns of one stage (MT accumulators side by side)

    if (warp == 8 && lane == 0) {
        sm100::tma_prefetch_desc(&tmap_q);
        sm100::tma_prefetch_desc(&tmap_t);
        for (int i = 0; i < 2; ++i) { sm100::mbar_init(&sm.a_full[i], 1); sm100::mbar_init(&sm.a_empty[i], 1); }
        for (int i = 0; i < NBK; ++i) { sm100::mbar_init(&sm.b_full[i], 1); sm100::mbar_init(&sm.b_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { sm100::mbar_init(&sm.acc_full[i], 1); sm100::mbar_init(&sm.acc_empty[i], (PAIR ? 2 : 1) * (EPI_THREADS / 32)); }
        sm100::fence_barrier_init();
        if constexpr (FOLD) sm100::tma_prefetch_desc(&tmap_x);
    }
    if constexpr (FOLD) {
        if (warp == 0) {
            // query side of the folded-norm step: every row {r, r, r, 0 x5} twice -- both 16-byte halves equal, so the tile is
            // invariant under the 16-byte chunk permutation of SWIZZLE_32B
            const uint32_t rb = (uint32_t)__half_as_ushort(__float2half_rn(__ldg(q_info) * __ldg(t_info + 1)));
            const uint32_t ax = sm100::smem_u32(sm.ax);
            for (int w = lane; w < AX_BYTES / 4; w += 32)
                sm100::sts32(ax + 4 * w, (w & 3) == 0 ? (int)(rb | (rb << 16)) : (w & 3) == 1 ? (int)rb : 0);
            sm100::fence_proxy_async();            // generic-proxy stores -> visible to tcgen05.mma (async proxy)
        }
    }
    if constexpr (PAIR) {
        sm100::cluster_sync_all();                          // both CTAs of the pair are running before the paired allocation
        if (warp == 9) sm100::tmem_alloc_2sm<2 * ACC_COLS>(sm.tmem_ptr);
    } else {
        if (warp == 9) sm100::tmem_alloc<2 * ACC_COLS>(sm.tmem_ptr);
    }
    sm100::tc_fence_before();
    __syncthreads();
    if constexpr (PAIR) sm100::cluster_sync_all();         // the peer's barriers exist before anything signals them
    sm100::tc_fence_after();
    const uint32_t tmem_base = *sm.tmem_ptr;

    if (warp == 8) {
        // ===================== TMA producer: lockstep warp, one elected lane issues =====================
        uint32_t ab = 0, a_par = 1, bs = 0, b_par = 1;     // parities to wait for on a_empty / b_empty
        uint32_t xs = 0;                                   // folded-norm slot = tile parity (the MMA warp's `as`)
        for (long long item = first_item; item < total_items; item += item_stride) {
            Item it = decode_item(item, pairs, tiles_max, S, backward, ROWS_ITEM, TNK);
            if (!it.valid) continue;
            sm100::mbar_wait(&sm.a_empty[ab], a_par);
            if (sm100::elect_one()) {
                if constexpr (PAIR) {                   // this CTA's row tile; both tiles are counted on the leader's barrier
                    if (crank == 0) sm100::mbar_arrive_expect_tx(&sm.a_full[ab], (uint32_t)(2 * nkb * A_KB_BYTES));
                    const uint32_t lbar = sm100::mapa_u32(sm100::smem_u32(&sm.a_full[ab]), 0);
                    for (int kb = 0; kb < nkb; ++kb)
                        sm100::tma_load_2d_2sm(sm.a + (size_t)(ab * nkb + kb) * A_KB_BYTES, &tmap_q, lbar, kb * KB_ELEMS,
                                               it.q_off + it.r0 + (int)crank * TM);
                } else {
                    sm100::mbar_arrive_expect_tx(&sm.a_full[ab], (uint32_t)(MT * nkb * A_KB_BYTES));
                    for (int t = 0; t < MT; ++t)
                        for (int kb = 0; kb < nkb; ++kb)
                            sm100::tma_load_2d(sm.a + (size_t)((ab * MT + t) * nkb + kb) * A_KB_BYTES, &tmap_q, &sm.a_full[ab],
                                               kb * KB_ELEMS, it.q_off + it.r0 + t * TM);
                }
            }
            __syncwarp();
            if (++ab == (uint32_t)nabuf) { ab = 0; a_par ^= 1; }
            for (int c0 = it.c_begin; c0 < it.c_end; c0 += TNK) {
                for (int kb = 0; kb < nkb; ++kb) {
                    sm100::mbar_wait(&sm.b_empty[bs], b_par);
                    if (sm100::elect_one()) {
                        const bool withx = FOLD && kb == nkb - 1;      // the tile's folded-norm rows travel with its last K block
                        if constexpr (PAIR) {          // this CTA's TN train rows of the 2 TN-column tile stay here
                            if (crank == 0) sm100::mbar_arrive_expect_tx(&sm.b_full[bs], (uint32_t)(2 * (B_KB_BYTES + (withx ? X_BYTES : 0))));
                            const uint32_t lbar = sm100::mapa_u32(sm100::smem_u32(&sm.b_full[bs]), 0);
                            const int trow = it.t_off + c0 + (int)crank * TN;
                            sm100::tma_load_2d_2sm(sm.b + (size_t)bs * B_KB_BYTES, &tmap_t, lbar, kb * KB_ELEMS, trow);
                            if (withx) sm100::tma_load_2d_2sm(sm.x + (size_t)xs * X_BYTES, &tmap_x, lbar, 0, trow);
                        } else {
                        sm100::mbar_arrive_expect_tx(&sm.b_full[bs], (uint32_t)(B_KB_BYTES + (withx ? X_BYTES : 0)));
                        sm100::tma_load_2d(sm.b + (size_t)bs * B_KB_BYTES, &tmap_t, &sm.b_full[bs], kb * KB_ELEMS,
                                           it.t_off + c0);
                        if (withx) sm100::tma_load_2d(sm.x + (size_t)xs * X_BYTES, &tmap_x, &sm.b_full[bs], 0, it.t_off + c0);
                        }
                    }
                    __syncwarp();
                    if (++bs == NBK) { bs = 0; b_par ^= 1; }
                }
                xs ^= 1;
            }
        }
    } else if (warp == 9 && (!PAIR || crank == 0)) {
        // ===================== MMA issuer: lockstep warp, one elected lane issues (PAIR: the even CTA only) =====================
        // (a single thread with per-tile modulo arithmetic and 64-bit descriptor maths was the bottleneck of
        // the int8 kernel; descriptors here are a constant high word and a running 32-bit low word)
        constexpr uint32_t idesc = H16 ? sm100::make_idesc(/*F32*/ 1, /*f16*/ 0, /*f16*/ 0, PAIR ? 2 * TM : TM, TNK)
                                       : sm100::make_idesc(/*F32*/ 1, /*tf32*/ 2, /*tf32*/ 2, TM, TN);
        auto mma16 = [](uint32_t d, uint64_t a, uint64_t b, uint32_t id, uint32_t acc) {
            if constexpr (PAIR) {
                asm volatile(
                    "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                    "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                    ::"r"(d), "l"(a), "l"(b), "r"(id), "r"(acc) : "memory");
            } else {
                sm100::mma_f16(d, a, b, id, acc);
            }
        };
        const uint32_t a_lo0 = sm100::smem_desc_lo_sw128(sm100::smem_u32(sm.a));
        const uint32_t b_lo0 = sm100::smem_desc_lo_sw128(sm100::smem_u32(sm.b));
        const uint32_t x_lo0 = sm100::smem_desc_lo_sw128(sm100::smem_u32(sm.x));      // same low-word format
        const uint64_t ax_desc = sm100::smem_desc_join_hi(sm100::SMEM_DESC_HI_SW32, sm100::smem_desc_lo_sw128(sm100::smem_u32(sm.ax)));
        uint32_t ab = 0, a_par = 0, bs = 0, b_par = 0, as = 0, acc_par = 3;
        for (long long item = first_item; item < total_items; item += item_stride) {
            Item it = decode_item(item, pairs, tiles_max, S, backward, ROWS_ITEM, TNK);
            if (!it.valid) continue;
            sm100::mbar_wait(&sm.a_full[ab], a_par);
            const uint32_t a_lo = a_lo0 + ab * (uint32_t)(MT * nkb * (A_KB_BYTES >> 4));
            for (int c0 = it.c_begin; c0 < it.c_end; c0 += TNK) {
                sm100::mbar_wait(&sm.acc_empty[as], (acc_par >> as) & 1);
                acc_par ^= 1u << as;
                sm100::tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * ACC_COLS;
                for (int kb = 0; kb < nkb; ++kb) {
                    sm100::mbar_wait(&sm.b_full[bs], b_par);
                    sm100::tc_fence_after();
                    if (sm100::elect_one()) {
                        const uint32_t bl = b_lo0 + bs * (B_KB_BYTES >> 4);
#pragma unroll
                        for (int t = 0; t < MT; ++t) {               // every train K block against each resident row tile
                            const uint32_t al = a_lo + (t * nkb + kb) * (A_KB_BYTES >> 4);
#pragma unroll
                            for (int k = 0; k < KB_BYTES / 32; ++k) {    // UMMA K = 8 tf32 or 16 halves = 32 bytes; +2 = 32 B >> 4
                                if constexpr (H16)
                                    mma16(d_tmem + t * TN, sm100::smem_desc_join(al + 2 * k), sm100::smem_desc_join(bl + 2 * k),
                                          idesc, (kb | k) ? 1u : 0u);
                                else
                                    sm100::mma_tf32(d_tmem + t * TN, sm100::smem_desc_join(al + 2 * k), sm100::smem_desc_join(bl + 2 * k),
                                                    idesc, (kb | k) ? 1u : 0u);
                            }
                        }
                        if constexpr (FOLD) {
                            if (kb == nkb - 1) {                     // + r (hi + mid + lo) = -(s_q s_t / 2) ||b||^2, into every accumulator
                                const uint64_t xd = sm100::smem_desc_join_hi(sm100::SMEM_DESC_HI_SW32, x_lo0 + as * (X_BYTES >> 4));
#pragma unroll
                                for (int t = 0; t < MT; ++t) mma16(d_tmem + t * TN, ax_desc, xd, idesc, 1u);
                            }
                        }
                        if constexpr (PAIR) {
                            sm100::mma_commit_2sm_mc(&sm.b_empty[bs], MC_MASK);
                            if (kb == nkb - 1) sm100::mma_commit_2sm_mc(&sm.acc_full[as], MC_MASK);
                        } else {
                            sm100::mma_commit(&sm.b_empty[bs]);
                            if (kb == nkb - 1) sm100::mma_commit(&sm.acc_full[as]);
                        }
                    }
                    __syncwarp();
                    if (++bs == NBK) { bs = 0; b_par ^= 1; }
                }
                as ^= 1;
            }
            if (sm100::elect_one()) {
                if constexpr (PAIR) sm100::mma_commit_2sm_mc(&sm.a_empty[ab], MC_MASK);
                else sm100::mma_commit(&sm.a_empty[ab]);
            }
            __syncwarp();
            if (++ab == (uint32_t)nabuf) { ab = 0; a_par ^= 1; }
        }
    } else if (warp < 8) {
        // ===================== epilogue: 8 warps, TMEM lane quarter = warp % 4, column half = warp / 4 =====================
        const int e = threadIdx.x;                 // 0..255
        const int quarter = warp & 3, half = warp >> 2;
        const int row_in_tile = quarter * 32 + lane;
        constexpr int HC = TNK / 2;                // columns per half
        constexpr int NCH = HC / 32;               // 32-column tcgen05.ld per warp and tile (PAIR: 4, all issued before any arithmetic)
        uint32_t acc_cnt = 0;
        const Window window = make_window(H16 ? q_info : nullptr, t_info, nkb * KB_ELEMS);
        const float neg2 = H16 ? -2.0f * __ldg(q_info + 1) * __ldg(t_info + 1) : -2.0f;     // -2 / (s_q s_t), exact
        for (long long item = first_item; item < total_items; item += item_stride) {
            Item it = decode_item(item, pairs, tiles_max, S, backward, ROWS_ITEM, TNK);
            if (!it.valid) continue;
            // per-row state (one set per resident row tile): running top-2 of 16-column minima, append window, list cursor
            bool row_ok[MT];
            float v1[MT], v2[MT], win[MT];
            long long lrow[MT];
            float2* list[MT];
            int cnt[MT];
#pragma unroll
            for (int t = 0; t < MT; ++t) {
                const int row = it.r0 + (PAIR ? (int)crank : t) * TM + row_in_tile;
                row_ok[t] = row < it.n;
                v1[t] = CUDART_INF_F; v2[t] = CUDART_INF_F;
                win[t] = row_ok[t] ? window(__ldg(q_norm + it.q_off + row), __ldg(pmax + it.pair)) : -CUDART_INF_F;
                lrow[t] = (long long)(it.split * 2 + half) * total_rows + it.row_base + (row_ok[t] ? row : 0);
                list[t] = cand + lrow[t] * KC;
                cnt[t] = 0;
            }
            if constexpr (FOLD) {
                // raw accumulators y = s_q s_t (a.b - ||b||^2 / 2): larger = closer, estimate e = neg2 * y
                constexpr float NEG_BIG = -3.0e38f;
                const float half_ss = -1.0f / neg2;                  // s_q s_t / 2: the window in accumulator units
#pragma unroll
                for (int t = 0; t < MT; ++t) { v1[t] = NEG_BIG; v2[t] = NEG_BIG; win[t] *= half_ss; }
                for (int c0 = it.c_begin; c0 < it.c_end; c0 += TNK) {
                    const uint32_t as = acc_cnt & 1;
                    const bool tail = c0 + TNK > it.c_end;           // columns past the pair's last train row: other images' rows
                    sm100::mbar_wait(&sm.acc_full[as], (acc_cnt >> 1) & 1);
                    sm100::tc_fence_after();
#pragma unroll
                    for (int t = 0; t < MT; ++t) {
                        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + as * ACC_COLS + t * TN + half * HC;
                        uint32_t acc[NCH][32];
#pragma unroll
                        for (int ch = 0; ch < NCH; ++ch) sm100::tmem_ld32(taddr + 32 * ch, acc[ch]);
                        sm100::tmem_ld_wait();
                        if (t == MT - 1) {
                            sm100::tc_fence_before();
                            __syncwarp();
                            if (lane == 0) {
                                if constexpr (PAIR) sm100::mbar_arrive_cluster(sm100::mapa_u32(sm100::smem_u32(&sm.acc_empty[as]), 0));
                                else sm100::mbar_arrive(&sm.acc_empty[as]);
                            }
                        }
#pragma unroll
                        // (1) all 16-column group maxima of this warp's share of the tile FIRST, and the running top-2 over all
                        //     of them, so that the tile's groups are tested against the tightest threshold available (any subset
                        //     of groups gives a valid bound; the first tiles of a scan used to append most of their columns);
                        // (2) a group that passes is walked in four 4-column steps, each behind its own maximum: one lane's
                        //     row passing makes the whole warp walk, ~30 instructions now instead of ~50.
                        if (tail) {
#pragma unroll
                            for (int ch = 0; ch < NCH; ++ch)
#pragma unroll
                                for (int i = 0; i < 32; ++i)
                                    acc[ch][i] = c0 + half * HC + ch * 32 + i < it.c_end ? acc[ch][i] : __float_as_uint(NEG_BIG);
                        }
                        float hm[2 * NCH];
#pragma unroll
                        for (int ch = 0; ch < NCH; ++ch) {
                            float h0 = __uint_as_float(acc[ch][0]), h1 = __uint_as_float(acc[ch][16]);
#pragma unroll
                            for (int i = 1; i < 16; ++i) {
                                h0 = fmaxf(h0, __uint_as_float(acc[ch][i]));
                                h1 = fmaxf(h1, __uint_as_float(acc[ch][16 + i]));
                            }
                            hm[2 * ch] = h0; hm[2 * ch + 1] = h1;
                            v2[t] = fmaxf(v2[t], fminf(v1[t], h0)); v1[t] = fmaxf(v1[t], h0);
                            v2[t] = fmaxf(v2[t], fminf(v1[t], h1)); v1[t] = fmaxf(v1[t], h1);
                        }
                        // rows past the pair (window -inf) never append; masked columns (NEG_BIG) never pass
                        const float tau = row_ok[t] ? fmaxf(v2[t] - win[t], 0.5f * NEG_BIG) : CUDART_INF_F;
#pragma unroll
                        for (int ch = 0; ch < NCH; ++ch) {
#pragma unroll
                            for (int g = 0; g < 2; ++g) {
                                if (hm[2 * ch + g] >= tau) {
                                    const int colg = c0 + half * HC + ch * 32 + g * 16;
#pragma unroll
                                    for (int q = 0; q < 4; ++q) {
                                        const float x0 = __uint_as_float(acc[ch][g * 16 + 4 * q]), x1 = __uint_as_float(acc[ch][g * 16 + 4 * q + 1]);
                                        const float x2 = __uint_as_float(acc[ch][g * 16 + 4 * q + 2]), x3 = __uint_as_float(acc[ch][g * 16 + 4 * q + 3]);
                                        if (fmaxf(fmaxf(x0, x1), fmaxf(x2, x3)) >= tau) {
#pragma unroll
                                            for (int i = 0; i < 4; ++i) {
                                                const float x = i == 0 ? x0 : i == 1 ? x1 : i == 2 ? x2 : x3;
                                                const bool pass = x >= tau;
                                                if (pass & (cnt[t] < KC)) list[t][cnt[t]] = make_float2(neg2 * x, __int_as_float(colg + 4 * q + i));
                                                cnt[t] += pass;
                                            }
                                        }
                                    }
                                }
                            }
                        }
                    }
                    ++acc_cnt;
                }
            } else {
            // ||b||^2 of the NEXT tile's columns is fetched while this tile is reduced (the load's latency used to sit in
            // front of every tile: 12 % of the stall samples at the barrier below)
            float nb = e < TN ? (it.c_begin + e < it.c_end ? __ldg(t_norm + it.t_off + it.c_begin + e) : BIG_F) : 0.0f;
            for (int c0 = it.c_begin; c0 < it.c_end; c0 += TN) {
                const uint32_t as = acc_cnt & 1;
                if (e < TN) {
                    sm.cb[as * TN + e] = nb;
                    const int col = c0 + TN + e;
                    nb = col < it.c_end ? __ldg(t_norm + it.t_off + col) : BIG_F;
                }
                asm volatile("bar.sync 1, 256;" ::: "memory");
                sm100::mbar_wait(&sm.acc_full[as], (acc_cnt >> 1) & 1);
                sm100::tc_fence_after();
                const float* cbv = sm.cb + as * TN + half * HC;
#pragma unroll
                for (int t = 0; t < MT; ++t) {
                    const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + as * ACC_COLS + t * TN + half * HC;
                    uint32_t acc[2][32];
                    sm100::tmem_ld32(taddr, acc[0]);
                    sm100::tmem_ld32(taddr + 32, acc[1]);
                    sm100::tmem_ld_wait();
                    if (t == MT - 1) {
                        sm100::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) sm100::mbar_arrive(&sm.acc_empty[as]);    // the stage's accumulators are in registers
                    }
#pragma unroll
                    for (int ch = 0; ch < 2; ++ch) {
                        float ev[32];
                        float h0 = CUDART_INF_F, h1 = CUDART_INF_F;
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            ev[i] = fmaf(neg2, __uint_as_float(acc[ch][i]), cbv[ch * 32 + i]);
                            if (i < 16) h0 = fminf(h0, ev[i]); else h1 = fminf(h1, ev[i]);
                        }
                        // top-2 of group minima first (the tighter, still valid bound), then the append test
                        v2[t] = fminf(v2[t], fmaxf(v1[t], h0)); v1[t] = fminf(v1[t], h0);
                        v2[t] = fminf(v2[t], fmaxf(v1[t], h1)); v1[t] = fminf(v1[t], h1);
                        const float tau = fminf(v2[t] + win[t], BIG_F * 0.5f);      // masked columns (||b||^2 = BIG_F) never pass
                        const int colb = c0 + half * HC + ch * 32;
                        // per 16-column group, and only where the group's minimum passes; the body is predicated, not
                        // branched (one lane's row passing makes the whole warp walk the group)
#pragma unroll
                        for (int g = 0; g < 2; ++g) {
                            if ((g ? h1 : h0) <= tau) {
#pragma unroll
                                for (int i = 0; i < 16; ++i) {
                                    const float x = ev[g * 16 + i];
                                    const bool pass = x <= tau;
                                    if (pass & (cnt[t] < KC)) list[t][cnt[t]] = make_float2(x, __int_as_float(colb + g * 16 + i));
                                    cnt[t] += pass;
                                }
                            }
                        }
                    }
                }
                ++acc_cnt;
            }
            }
#pragma unroll
            for (int t = 0; t < MT; ++t)
                if (row_ok[t]) cand_cnt[lrow[t]] = cnt[t];
        }
    }

    sm100::tc_fence_before();
    __syncthreads();
    if constexpr (PAIR) sm100::cluster_sync_all();         // the leader may still commit to this CTA's barriers until it is done too
    if (warp == 9) {
        if constexpr (PAIR) sm100::tmem_dealloc_2sm<2 * ACC_COLS>(tmem_base);
        else sm100::tmem_dealloc<2 * ACC_COLS>(tmem_base);
    }
}

// max ||b|| over the train rows of each pair (for the TF32 error window)
__global__ void __launch_bounds__(256)
pair_max_norm_kernel(const DevPair* __restrict__ pairs, const float* __restrict__ t_norm, int backward,
                     float* __restrict__ pmax) {
    __shared__ float red[8];
    DevPair pr = pairs[blockIdx.x];
    const int t_off = backward ? pr.q_off : pr.t_off, m = backward ? pr.n : pr.m;
    float mx = 0.0f;
    for (int j = threadIdx.x; j < m; j += blockDim.x) mx = fmaxf(mx, __ldg(t_norm + t_off + j));
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) mx = fmaxf(mx, red[w]);
        pmax[blockIdx.x] = sqrtf(mx);
    }
}

__device__ unsigned long long g_f32_stats[2];   // [0] rows re-scanned exhaustively (list not provably sufficient), [1] their evaluations

// One warp per query row: exact fp32 re-rank of the TF32 candidates (see the file header).
__global__ void __launch_bounds__(256)
refine_f32_kernel(int backward, int need2, const DevPair* __restrict__ pairs, int blocks_per_pair,
                  const float2* __restrict__ cand, const int* __restrict__ cand_cnt, long long total_rows, int S,
                  const float* __restrict__ q_desc,
                  const float* __restrict__ q_norm, const float* __restrict__ t_desc, int D,
                  const float* __restrict__ pmax, Top2* __restrict__ out, u64* __restrict__ fbq, int* __restrict__ fb_count,
                  const float* __restrict__ q_info, const float* __restrict__ t_info, int Dh) {
    const int lane = threadIdx.x & 31;
    const int p = blockIdx.x / blocks_per_pair;
    const int i = (blockIdx.x - p * blocks_per_pair) * 8 + (threadIdx.x >> 5);
    DevPair pr = pairs[p];
    int q_off = pr.q_off, n = pr.n, t_off = pr.t_off, m = pr.m;
    long long row_base = pr.fwd_off;
    if (backward) { q_off = pr.t_off; n = pr.m; t_off = pr.q_off; m = pr.n; row_base = pr.bwd_off; }
    if (i >= n) return;
    const long long row = row_base + i;
    const int nf4 = D >> 2;                                  // D is a multiple of 32 floats, <= 256
    const float4* a4 = reinterpret_cast<const float4*>(q_desc + (long long)(q_off + i) * D);
    float4 ar[2];
    ar[0] = lane < nf4 ? __ldg(a4 + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
    ar[1] = lane + 32 < nf4 ? __ldg(a4 + lane + 32) : make_float4(0.f, 0.f, 0.f, 0.f);
    auto exact = [&](int col) -> float {                     // direct sum (a-b)^2 in fp32, as cv2 does
        const float4* b4 = reinterpret_cast<const float4*>(t_desc + (long long)(t_off + col) * D);
        float s = 0.0f;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (lane + 32 * h < nf4) {
                float4 b = __ldg(b4 + lane + 32 * h);
                float dx = ar[h].x - b.x, dy = ar[h].y - b.y, dz = ar[h].z - b.z, dw = ar[h].w - b.w;
                s = fmaf(dx, dx, s); s = fmaf(dy, dy, s); s = fmaf(dz, dz, s); s = fmaf(dw, dw, s);
            }
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
        return s;
    };
    // Walk the 2 * S lists of this row (<= KC entries each): second smallest estimate, then the survivors.
    const int nlist = 2 * S;
    float m1 = CUDART_INF_F, m2 = CUDART_INF_F;
    bool overflow = false;
    // One column split (the usual case): both lists are fetched whole and unconditionally, together with the counts and the
    // row's norm, so the warp's dependent global loads are two deep (lists, then the survivors' rows) instead of four
    // (count -> entries -> entries again -> rows); this kernel is bound by that latency chain, not by bytes.
    const bool two = nlist == 2;
    float2 v0 = make_float2(CUDART_INF_F, 0.0f), v1 = v0;
    int c0 = 0, c1 = 0;
    if (two) {
        c0 = __ldg(cand_cnt + row); c1 = __ldg(cand_cnt + total_rows + row);
        const float2 w0 = cand[row * KC + lane], w1 = cand[(total_rows + row) * KC + lane];
        overflow = c0 > KC || c1 > KC;
        if (lane < min(c0, KC)) v0 = w0;
        if (lane < min(c1, KC)) v1 = w1;
        m1 = fminf(v0.x, v1.x); m2 = fmaxf(v0.x, v1.x);
    } else {
        for (int l = 0; l < nlist; ++l) {
            const long long lrow = (long long)l * total_rows + row;
            const int c = __ldg(cand_cnt + lrow);
            overflow |= c > KC;
            float e = lane < min(c, KC) ? cand[lrow * KC + lane].x : CUDART_INF_F;
            float hi = fmaxf(m1, e);
            m1 = fminf(m1, e);
            m2 = fminf(m2, hi);
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        float o1 = __shfl_xor_sync(0xffffffffu, m1, off), o2 = __shfl_xor_sync(0xffffffffu, m2, off);
        float hi = fmaxf(m1, o1);
        m1 = fminf(m1, o1);
        m2 = fminf(fminf(m2, o2), hi);
    }
    u64 k1 = B200M_SENTINEL, k2 = B200M_SENTINEL;
    const bool fallback = overflow;
    if (m > 0 && !fallback) {
        const float win = make_window(q_info, t_info, Dh)(__ldg(q_norm + q_off + i), __ldg(pmax + p));
        const float bound = (need2 ? m2 : m1) + win;       // m2 == +inf with a single column: everything passes
        for (int l = 0; l < nlist; ++l) {
            float2 v;
            int c;
            if (two) {
                v = l ? v1 : v0; c = min(l ? c1 : c0, KC);   // (bound is +inf when the pair has a single train row)
            } else {
                const long long lrow = (long long)l * total_rows + row;
                c = min(__ldg(cand_cnt + lrow), KC);
                v = lane < c ? cand[lrow * KC + lane] : make_float2(CUDART_INF_F, 0.0f);
            }
            unsigned mask = __ballot_sync(0xffffffffu, lane < c && v.x <= bound);
            while (mask) {
                int src = __ffs(mask) - 1;
                mask &= mask - 1;
                int col = __shfl_sync(0xffffffffu, __float_as_int(v.y), src);
                float d2 = exact(col);
                top2_insert(k1, k2, ((u64)__float_as_uint(d2) << 32) | (u64)(u32)col);
            }
        }
    } else if (m > 0 && lane == 0) {
        // list overflow: the row goes to the exhaustive fp32 re-scan (fallback_f32_kernel, one CTA per row)
        fbq[atomicAdd(fb_count, 1)] = ((u64)(u32)p << 32) | (u64)(u32)i;
    }
    if (lane == 0) {
        Top2 t; t.k1 = k1; t.k2 = k2;
        out[row] = t;
        Top2 none; none.k1 = B200M_SENTINEL; none.k2 = B200M_SENTINEL;
        for (int s = 1; s < S; ++s) out[(long long)s * total_rows + row] = none;
    }
}

// Exhaustive fp32 re-scan of the queued rows (candidate list overflowed): FB_SLICES CTAs per row (a handful of rows per
// launch is the normal case -- one CTA per row left the launch as long as one CTA needs for m kilobyte-sized gathers,
// 0.22 ms for 8 rows of 4096 columns, 7 % of the float knn stage), each warp takes a strided share of the columns,
// 8 columns per step so that 16 independent 512-byte loads are in flight per warp.  The slices merge into the row's
// Top2 (which refine_f32_kernel left at the sentinel) with two 64-bit atomicMin: k1 keeps the minimum, every key that
// loses there -- each key except the final minimum loses exactly once -- competes for k2.
constexpr int FB_SLICES = 16;
__global__ void __launch_bounds__(256)
fallback_f32_kernel(int backward, const DevPair* __restrict__ pairs, const u64* __restrict__ fbq,
                    const int* __restrict__ fb_count, const float* __restrict__ q_desc,
                    const float* __restrict__ t_desc, int D, Top2* __restrict__ out) {
    __shared__ u64 red[8][2];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int count = *fb_count;
    const int nf4 = D >> 2;
    for (int item = blockIdx.x; item < count * FB_SLICES; item += gridDim.x) {
        const int q = item / FB_SLICES, g = item - q * FB_SLICES;
        const u64 ent = fbq[q];
        const int p = (int)(ent >> 32), i = (int)(u32)ent;
        DevPair pr = pairs[p];
        int q_off = pr.q_off, t_off = pr.t_off, m = pr.m;
        long long row_base = pr.fwd_off;
        if (backward) { q_off = pr.t_off; t_off = pr.q_off; m = pr.n; row_base = pr.bwd_off; }
        const float4* a4 = reinterpret_cast<const float4*>(q_desc + (long long)(q_off + i) * D);
        float4 ar[2];
        ar[0] = lane < nf4 ? __ldg(a4 + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
        ar[1] = lane + 32 < nf4 ? __ldg(a4 + lane + 32) : make_float4(0.f, 0.f, 0.f, 0.f);
        u64 k1 = B200M_SENTINEL, k2 = B200M_SENTINEL;
        for (int c0 = (g * 8 + w) * 8; c0 < m; c0 += 64 * FB_SLICES) {
            float s[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int col = min(c0 + u, m - 1);
                const float4* b4 = reinterpret_cast<const float4*>(t_desc + (long long)(t_off + col) * D);
                float acc = 0.0f;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    if (lane + 32 * h < nf4) {
                        float4 b = __ldg(b4 + lane + 32 * h);
                        float dx = ar[h].x - b.x, dy = ar[h].y - b.y, dz = ar[h].z - b.z, dw = ar[h].w - b.w;
                        acc = fmaf(dx, dx, acc); acc = fmaf(dy, dy, acc); acc = fmaf(dz, dz, acc); acc = fmaf(dw, dw, acc);
                    }
                }
                s[u] = acc;
            }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1)
#pragma unroll
                for (int u = 0; u < 8; ++u) s[u] += __shfl_xor_sync(0xffffffffu, s[u], off);
#pragma unroll
            for (int u = 0; u < 8; ++u)
                if (c0 + u < m) top2_insert(k1, k2, ((u64)__float_as_uint(s[u]) << 32) | (u64)(u32)(c0 + u));
        }
        if (lane == 0) { red[w][0] = k1; red[w][1] = k2; }
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int ww = 1; ww < 8; ++ww) { top2_insert(k1, k2, red[ww][0]); top2_insert(k1, k2, red[ww][1]); }
            unsigned long long* o = reinterpret_cast<unsigned long long*>(out + row_base + i);      // {k1, k2}
            static_assert(sizeof(Top2) == 16, "Top2 is two 64-bit keys");
            for (int r = 0; r < 2; ++r) {
                const u64 key = r ? k2 : k1;
                if (key == B200M_SENTINEL) break;
                const u64 old = atomicMin(o, (unsigned long long)key);
                atomicMin(o + 1, (unsigned long long)(old > key ? old : key));
            }
        }
        __syncthreads();
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && count > 0) {
        atomicAdd(&g_f32_stats[0], (unsigned long long)count);
        unsigned long long ev = 0;
        for (int q = 0; q < count; ++q) {
            DevPair pr = pairs[(int)(fbq[q] >> 32)];
            ev += (unsigned long long)(backward ? pr.n : pr.m);
        }
        atomicAdd(&g_f32_stats[1], ev);
    }
}

}  // namespace


// ---------------------------------------------------------------- host side
int b200m_f32_stats_read(unsigned long long out[2], bool reset) {
    B200M_CUDA_CHECK(cudaMemcpyFromSymbol(out, g_f32_stats, sizeof(unsigned long long) * 2), "f32 stats read");
    if (reset) {
        unsigned long long z[2] = {0, 0};
        B200M_CUDA_CHECK(cudaMemcpyToSymbol(g_f32_stats, z, sizeof(z)), "f32 stats reset");
    }
    return B200M_OK;
}

int launch_l2_f32_top2(const float* q_desc, const float* q_norm, int64_t q_rows, const float* t_desc,
                       const float* t_norm, int64_t t_rows, int D, const DevPair* pairs, int num_pairs, int max_n,
                       int max_m, int S, bool backward, bool need2, void* cand_ws, float* pmax, Top2* out,
                       long long total_rows, cudaStream_t st, bool h16) {
    if (D % (KB_BYTES / 4) != 0) return B200M_EINVAL;
    const int Dh = h16 ? b200m_f16_width(D) : D;        // width of the operand rows the tensor cores read
    const int nkb = h16 ? Dh / (KB_BYTES / 2) : D / (KB_BYTES / 4);
    if (D > 256 || nkb > MAX_NKB) return B200M_EUNSUPPORTED;       // query tile must stay resident in shared memory
    if (!cand_ws || !pmax) return B200M_EINVAL;
    if (backward) {
        const float* d = q_desc; q_desc = t_desc; t_desc = d;
        const float* nn = q_norm; q_norm = t_norm; t_norm = nn;
        int64_t r = q_rows; q_rows = t_rows; t_rows = r;
        int mx = max_n; max_n = max_m; max_m = mx;
    }
    // L2_F16C: q_norm / t_norm are aux blocks (norms first, then the fp16 copy and the info words)
    const float* q_info = h16 ? (const float*)((const char*)q_norm + b200m_f16_info_off(q_rows, D)) : nullptr;
    const float* t_info = h16 ? (const float*)((const char*)t_norm + b200m_f16_info_off(t_rows, D)) : nullptr;
    CUtensorMap tq, tt;
    int rc;
    if (h16) {
        if (((uintptr_t)q_norm & 255) || ((uintptr_t)t_norm & 255)) return B200M_EINVAL;
        rc = b200m_make_tmap_2d(&tq, (const char*)q_norm + b200m_f16_half_off(q_rows), (uint64_t)q_rows, (uint64_t)Dh * 2, TM,
                                CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2);
        if (rc != B200M_OK) return rc;
        rc = b200m_make_tmap_2d(&tt, (const char*)t_norm + b200m_f16_half_off(t_rows), (uint64_t)t_rows, (uint64_t)Dh * 2, TN,
                                CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2);
    } else {
        rc = b200m_make_tmap_2d(&tq, q_desc, (uint64_t)q_rows, (uint64_t)D * 4, TM, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4);
        if (rc != B200M_OK) return rc;
        rc = b200m_make_tmap_2d(&tt, t_desc, (uint64_t)t_rows, (uint64_t)D * 4, TN, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4);
    }
    if (rc != B200M_OK) return rc;
    static B200mPerDevice once;
    int dev = 0, sms = 0;
    if (!b200m_current_device(&dev, &sms)) return B200M_ENODEVICE;
    if (!once.done[dev]) {
#define B200M_CAND_ATTR(H, M, F)                                                                                             \
        B200M_CUDA_CHECK(cudaFuncSetAttribute(l2_f32_cand_kernel<H, M, F>, cudaFuncAttributeMaxDynamicSharedMemorySize,     \
                                              (int)smem_bytes(MAX_NKB)), "l2_f32 smem attr")
        B200M_CAND_ATTR(false, 1, false); B200M_CAND_ATTR(false, 2, false); B200M_CAND_ATTR(true, 1, false);
        B200M_CAND_ATTR(true, 2, false); B200M_CAND_ATTR(true, 1, true); B200M_CAND_ATTR(true, 2, true);
#undef B200M_CAND_ATTR
        B200M_CUDA_CHECK(cudaFuncSetAttribute(l2_f32_cand_kernel<true, 1, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)smem_bytes(MAX_NKB)), "l2_f32 smem attr");
        // CTA pairs for the fp16 FOLD kernel: only if 2-CTA clusters of it cover every SM
        cudaLaunchConfig_t qc = {};
        cudaLaunchAttribute qa[1];
        qa[0].id = cudaLaunchAttributeClusterDimension;
        qa[0].val.clusterDim.x = 2; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
        qc.gridDim = dim3((unsigned)(sms & ~1)); qc.blockDim = dim3(THREADS); qc.dynamicSmemBytes = smem_bytes(MAX_NKB / 2, true);
        qc.attrs = qa; qc.numAttrs = 1;
        int nc = 0;
        if (cudaOccupancyMaxActiveClusters(&nc, l2_f32_cand_kernel<true, 1, true, true>, &qc) == cudaSuccess && nc * 2 >= sms &&
            !getenv("B200M_NO_CLUSTER") && !getenv("B200M_NO_PAIR"))
            once.val[dev][0] = nc;
        else
            cudaGetLastError();
        once.done[dev] = 1;
    }
    const int pair_clusters = once.val[dev][0];
    // two resident row tiles per work item when both fit (nkb <= 4) and the pairs are tall enough to fill the SMs that way
    static const int mt_env = getenv("B200M_F32_MT") ? atoi(getenv("B200M_F32_MT")) : 0;
    const int MTr = mt_env ? mt_env : 2;
    int MT = nabuf_of(nkb) == 2 && MTr == 2 && max_n > TM ? 2 : 1;
    if (MT == 2 && !mt_env && (long long)num_pairs * S * ((max_n + 2 * TM - 1) / (2 * TM)) < sms) MT = 1;   // too few items to fill the SMs
    int tiles_max = (max_n + TM * MT - 1) / (TM * MT);
    long long total_items = (long long)num_pairs * S * tiles_max;
    if (total_items <= 0) return B200M_OK;
    unsigned grid = (unsigned)(total_items < sms ? total_items : sms);
    float2* cand = (float2*)cand_ws;
    int* cand_cnt = (int*)((char*)cand_ws + 8ull * KC * 2 * (size_t)S * (size_t)total_rows);
    u64* fbq = (u64*)(cand_cnt + 2 * (size_t)S * (size_t)total_rows + (((2 * (size_t)S * (size_t)total_rows) & 1) ? 1 : 0));
    int* fb_count = (int*)(fbq + total_rows);
    B200M_CUDA_CHECK(cudaMemsetAsync(fb_count, 0, sizeof(int), st), "fallback counter");
    pair_max_norm_kernel<<<num_pairs, 256, 0, st>>>(pairs, t_norm, backward ? 1 : 0, pmax);
    B200M_CUDA_CHECK(cudaGetLastError(), "pair_max_norm_kernel launch");
    // fp16 kind with at least 3 K blocks (the folded-norm slots are reused every second tile: 2 nkb >= ring depth): the FOLD
    // kernel, gated on the device by the two stores' scale ratio, with the plain kernel behind it for the other case
    static const int fold_env = getenv("B200M_F32_FOLD") ? atoi(getenv("B200M_F32_FOLD")) : 1;
    const bool fold = h16 && fold_env && 2 * nkb >= NB_FOLD;
    CUtensorMap tx = tt;
    if (fold) {
        rc = b200m_make_tmap_2d_ex(&tx, (const char*)t_norm + b200m_f16_x_off(t_rows, D), (uint64_t)t_rows, 32, TN,
                                   CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, 32, CU_TENSOR_MAP_SWIZZLE_32B);
        if (rc != B200M_OK) return rc;
    }
#define B200M_LAUNCH_CAND(H, M, F, MODE, QI, TI)                                                                                 \
    l2_f32_cand_kernel<H, M, F><<<grid, THREADS, smem_bytes(nkb, F), st>>>(tq, tt, tx, MODE, t_norm, pairs, total_items, tiles_max, S, \
                                                                          backward ? 1 : 0, nkb, q_norm, pmax, cand, cand_cnt, total_rows, QI, TI)
    if (h16) {
        if (fold && MT == 2 && pair_clusters > 0) {
            // the two row tiles of an item go to the two CTAs of a pair (same item count as MT == 2)
            cudaLaunchConfig_t cfg = {};
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.gridDim = dim3((unsigned)((total_items < pair_clusters ? total_items : pair_clusters) * 2));
            cfg.blockDim = dim3(THREADS);
            cfg.dynamicSmemBytes = smem_bytes(nkb, true);
            cfg.stream = st;
            cfg.attrs = at; cfg.numAttrs = 1;
            const int fmode = 1, bw = backward ? 1 : 0;
            cudaError_t e = cudaLaunchKernelEx(&cfg, l2_f32_cand_kernel<true, 1, true, true>, tq, tt, tx, fmode, t_norm, pairs, total_items,
                                               tiles_max, S, bw, nkb, q_norm, pmax, cand, cand_cnt, total_rows, q_info, t_info);
            B200M_CUDA_CHECK(e, "l2_f32_cand_kernel (CTA pairs) launch");
        } else
        if (fold) { if (MT == 2) B200M_LAUNCH_CAND(true, 2, true, 1, q_info, t_info); else B200M_LAUNCH_CAND(true, 1, true, 1, q_info, t_info); }
        const int mode = fold ? 2 : 0;
        if (MT == 2) B200M_LAUNCH_CAND(true, 2, false, mode, q_info, t_info); else B200M_LAUNCH_CAND(true, 1, false, mode, q_info, t_info);
    } else {
        if (MT == 2) B200M_LAUNCH_CAND(false, 2, false, 0, nullptr, nullptr); else B200M_LAUNCH_CAND(false, 1, false, 0, nullptr, nullptr);
    }
#undef B200M_LAUNCH_CAND
    B200M_CUDA_CHECK(cudaGetLastError(), "l2_f32_cand_kernel launch");
    int bpp = (max_n + 7) / 8;
    long long rgrid = (long long)num_pairs * bpp;
    if (rgrid > 0x7fffffffLL) return B200M_EUNSUPPORTED;
    refine_f32_kernel<<<(unsigned)rgrid, 256, 0, st>>>(backward ? 1 : 0, need2 ? 1 : 0, pairs, bpp, cand, cand_cnt, total_rows, S,
                                                       q_desc, q_norm, t_desc, D, pmax, out, fbq, fb_count, q_info, t_info, Dh);
    B200M_CUDA_CHECK(cudaGetLastError(), "refine_f32_kernel launch");
    fallback_f32_kernel<<<4 * sms, 256, 0, st>>>(backward ? 1 : 0, pairs, fbq, fb_count, q_desc, t_desc, D, out);
    B200M_CUDA_CHECK(cudaGetLastError(), "fallback_f32_kernel launch");
    b200m_count_launch(4);
    return B200M_OK;
}
