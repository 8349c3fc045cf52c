// hamming_tc.cu — NORM_HAMMING knn (k=2) on the 5th-gen tensor cores.
// Same contract as hamming.cu (cv2.BFMatcher(NORM_HAMMING).knnMatch(k=2)/match as called at
// /root/reference/FeatureMatchingExtraction/feature_matchers.py:210,227 for ORB/AKAZE/BRISK rows), bit-exact,
// but the N x M distance evaluations run as an int8 GEMM instead of XOR+POPC:
//   every descriptor bit becomes one int8 dimension, +1 for a set bit and -1 for a clear one (prep_hamming_tc_kernel);
//   then a.b = (#agreeing bits) - (#differing bits) = Dbits - 2 h, i.e. h = (Dbits - a.b) / 2 exactly.
// POPC issues 16 lanes/clk/SM: a 256-bit distance costs 8 POPC, so the INT pipe tops out at 582 Gdist/s (hamming.cu).
// tcgen05.mma kind::i8 retires 8192 MAC/clk/SM: the same distance is 256 MACs = 1/32 clk/SM, 9.3 Tdist/s at the
// same clock -- 16x the POPC bound, at the price of an 8x larger operand stream (256 B instead of 32 B per row).
//
// Kernel = the pipeline of l2_i8_fold_kernel (l2_i8.cu) without the norm: persistent warp-specialised CTAs in
// clusters of two -- CTA pairs issuing cta_group::2 MMAs of M = 256 where the device co-schedules them everywhere (each
// CTA keeps its half of the train tile), else TMA-multicast clusters or single CTAs with M = 128 -- N = 256 (240 for
// e2m1) tiles, NKB = Dbits / 128 K-blocks of 128 int8, int32 accumulators double-buffered in TMEM.  All rows have the same norm, so the epilogue is a bare
// 3-input MAX tree over the raw accumulators (larger a.b == closer).  Per 16-column chunk only the maximum
// survives, packed with the chunk index into ONE int32 key ((a.b + 512) << 20 | ~chunk): the running top-2 of chunk
// keys is two integer max/min instructions and orders chunks by (distance asc, first column asc) -- cv2's tie rule.
// hamming_tc_resolve_kernel then recomputes the 2 x 16 distances of a row's two best chunks with POPC on the packed
// bits and writes the exact (distance, column) top-2 in the format the POPC kernel produces; tail.cu is unchanged.
//   Exactness: the best column lies in the best chunk (ties: the lowest chunk holds the lowest column).  The
//   runner-up is either another column of that chunk or the best column of the best OTHER chunk, which is what the
//   second key names (again lowest chunk first); both chunks are recomputed, so index and distance are exact.
#include <cstdio>
#include <cstdlib>
#include "common.cuh"
#include "sm100.cuh"

namespace {

constexpr int TM = 128, KB_BYTES = 128;
constexpr int A_KB_BYTES = TM * KB_BYTES;      // 16 KiB
// Two operand formats (template flag FP4):
//   int8  : one byte per bit (+1 / -1), tcgen05.mma kind::i8, K = 32 per instruction, int32 accumulators, N = 256 per tile;
//   e2m1  : one NIBBLE per bit (+1.0 = 0x2, -1.0 = 0xA), tcgen05.mma kind::mxf4.block_scale with every UE8M0 scale factor
//           = 1.0, K = 64 per instruction at the same instruction rate = twice the MACs (tools/mma_peak mxf4: 16384
//           MAC/clk/SM, 7.8 PFLOP/s sustained), fp32 accumulators (sums of +-1: exact).  The scale factors live in TMEM, and
//           two 256-column accumulators already fill its 512 columns, so tiles are N = 240 wide: accumulator stage s at
//           column 256 s, 16 columns of 0x7F bytes at column 240 serve as both scale-factor operands (all bytes equal, so
//           their layout does not matter).
__host__ __device__ constexpr int tn_of(bool fp4) { return fp4 ? 240 : 256; }          // train columns per tile (UMMA N)
__host__ __device__ constexpr int b_kb_bytes(bool fp4) { return tn_of(fp4) * KB_BYTES; }  // one K block of a train tile
constexpr int SF_COL = 240;
constexpr uint32_t IDESC_MXF4 = (1u << 7) | (1u << 10) | ((uint32_t)(240 >> 3) << 17) | (1u << 23) | ((uint32_t)(TM >> 4) << 24);
constexpr int NBF = 4;                         // train K-block ring depth
constexpr int CTRL_WARPS = 4, EPI_WARPS = 16;
constexpr int THREADS = (CTRL_WARPS + EPI_WARPS) * 32;
constexpr int CHUNK = B200M_HTC_CHUNK;         // columns per key chunk
constexpr int KEY_SHIFT = 20;                  // key = (a.b + 512) << 20 | (0xFFFFF - chunk index)
constexpr int KEY_LOW = (1 << KEY_SHIFT) - 1;
constexpr int MASKED = -1024;                  // accumulator value of a column that does not exist: key < 0
static_assert(CHUNK == 16, "the epilogue reduces 16-column chunks");

__host__ __device__ constexpr int nabuf_of(int nkb) { return nkb <= 2 ? 2 : 1; }      // query tile buffers
// "accumulator ready" can ride on the ring-stage release of the tile's last K block (saving one tcgen05.commit per tile)
// only while an epilogue group meets every phase of the barriers it waits on in order: a parity wait that trails by two
// phases passes early.  With NKB <= 2 a group always waits on the same stage, once per use; with NKB = 3 it alternates
// between two stages and skips uses, with NKB = 4 the stage is reused by the very next tile -- both get their own commit
// (one per 1536 / 2048 cycles of MMA).
__host__ __device__ constexpr bool sep_acc_of(int nkb) { return nkb > 2; }

struct Item {
    int q_off, n, t_off, m, r0, c_begin, c_end, split, ntiles;
    long long row_base;
    bool valid;
};

__device__ __forceinline__ Item decode_tile(int p, int split, int tile, const DevPair* __restrict__ pairs, int S, int backward,
                                            int TN) {
    Item it;
    it.split = split;
    DevPair pr = pairs[p];
    if (!backward) { it.q_off = pr.q_off; it.n = pr.n; it.t_off = pr.t_off; it.m = pr.m; it.row_base = pr.fwd_off; }
    else           { it.q_off = pr.t_off; it.n = pr.m; it.t_off = pr.q_off; it.m = pr.n; it.row_base = pr.bwd_off; }
    it.r0 = tile * TM;
    int cols_per = ((it.m + S - 1) / S + TN - 1) / TN * TN;
    it.c_begin = it.split * cols_per;
    it.c_end = min(it.m, it.c_begin + cols_per);
    it.valid = it.r0 < it.n;
    it.ntiles = it.c_end > it.c_begin ? (it.c_end - it.c_begin + TN - 1) / TN : 0;
    return it;
}

struct Smem {
    uint8_t *a, *b;
    int2* xch;            // [2][EPI_WARPS - 4][32] partial top-2 keys of the non-merging epilogue warps
    uint64_t *a_full, *a_empty, *b_full, *b_empty, *acc_full, *acc_empty;
    uint32_t* tmem_ptr;
};
// pair == true (cta_group::2): a CTA keeps only its own half of every train tile, so a ring stage is half as large and
// the ring twice as deep (see l2_i8_fold_kernel).
__host__ __device__ constexpr int nbf_of(bool pair) { return pair ? 2 * NBF : NBF; }
__device__ __forceinline__ Smem carve(uint8_t* base, int nkb, bool fp4, bool pair = false) {
    const int nbf = nbf_of(pair);
    Smem s;
    s.a = base;
    s.b = s.a + nabuf_of(nkb) * nkb * A_KB_BYTES;
    s.xch = (int2*)(s.b + nbf * (b_kb_bytes(fp4) / (pair ? 2 : 1)));
    uint8_t* p = (uint8_t*)(s.xch + 2 * (EPI_WARPS - 4) * 32);
    s.a_full = (uint64_t*)p; p += 2 * 8;
    s.a_empty = (uint64_t*)p; p += 2 * 8;
    s.b_full = (uint64_t*)p; p += nbf * 8;
    s.b_empty = (uint64_t*)p; p += nbf * 8;
    s.acc_full = (uint64_t*)p; p += 2 * 8;
    s.acc_empty = (uint64_t*)p; p += 2 * 8;
    s.tmem_ptr = (uint32_t*)p;
    return s;
}
__host__ __device__ constexpr size_t smem_bytes(int nkb, bool fp4 = false, bool pair = false) {
    return 1024 + (size_t)nabuf_of(nkb) * nkb * A_KB_BYTES + (size_t)nbf_of(pair) * (b_kb_bytes(fp4) / (pair ? 2 : 1)) +
           2 * (EPI_WARPS - 4) * 32 * sizeof(int2) + (4 + 2 * nbf_of(pair) + 4) * 8 + 16;
}
static_assert(smem_bytes(1) <= 232448 && smem_bytes(2) <= 232448 && smem_bytes(3) <= 232448 && smem_bytes(4) <= 232448,
              "shared memory (int8)");
static_assert(smem_bytes(1, false, true) <= 232448 && smem_bytes(2, false, true) <= 232448 &&
              smem_bytes(3, false, true) <= 232448 && smem_bytes(4, false, true) <= 232448 &&
              smem_bytes(1, true, true) <= 232448 && smem_bytes(2, true, true) <= 232448, "shared memory (CTA pairs)");

__device__ __forceinline__ int max3(int a, int b, int c) { return max(max(a, b), c); }
template <int O>
__device__ __forceinline__ int max16(const uint32_t (&v)[32]) {
    int a0 = max3((int)v[O + 0], (int)v[O + 1], (int)v[O + 2]), a1 = max3((int)v[O + 3], (int)v[O + 4], (int)v[O + 5]);
    int a2 = max3((int)v[O + 6], (int)v[O + 7], (int)v[O + 8]), a3 = max3((int)v[O + 9], (int)v[O + 10], (int)v[O + 11]);
    int a4 = max3((int)v[O + 12], (int)v[O + 13], (int)v[O + 14]);
    return max(max3(a0, a1, a2), max3(a3, a4, (int)v[O + 15]));
}
__device__ __forceinline__ float fmax3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }    // one FMNMX3
template <int O>
__device__ __forceinline__ int max16f(const uint32_t (&v)[32]) {          // fp32 accumulators holding exact integers
#define F(i) __uint_as_float(v[O + (i)])
    float a0 = fmax3(F(0), F(1), F(2)), a1 = fmax3(F(3), F(4), F(5)), a2 = fmax3(F(6), F(7), F(8));
    float a3 = fmax3(F(9), F(10), F(11)), a4 = fmax3(F(12), F(13), F(14));
    return __float2int_rn(fmaxf(fmax3(a0, a1, a2), fmax3(a3, a4, F(15))));
#undef F
}
template <int O, bool FP4>
__device__ __forceinline__ int chunk_max(const uint32_t (&v)[32]) {
    if constexpr (FP4) return max16f<O>(v);
    else return max16<O>(v);
}
template <bool FP4>
__device__ __forceinline__ void mask32(uint32_t (&v)[32], int lim) {
    constexpr uint32_t M = FP4 ? 0xC4800000u /* -1024.0f */ : (uint32_t)MASKED;
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = i < lim ? v[i] : M;
}
__device__ __forceinline__ void mma_mxf4(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t sf,
                                         uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%4], [%4], p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(sf), "r"(accumulate)
        : "memory");
}

__device__ __forceinline__ void mma_mxf4_2sm(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t sf,
                                             uint32_t accumulate) {      // M = 256 over a CTA pair; issued by the even CTA
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%4], [%4], p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(sf), "r"(accumulate)
        : "memory");
}
// m * kmul + low as ONE IMAD on the FMA pipe (kmul in a register so that ptxas cannot turn it into ALU-pipe shift + add)
__device__ __forceinline__ int pack_key(int m, int kmul, int low) {
    int r;
    asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(r) : "r"(m), "r"(kmul), "r"(low));
    return r;
}

// CS = CTAs per cluster (1 or 2): the two CTAs of a cluster take adjacent query row tiles of the same pair, each fetches
// half of every train tile and TMA-multicasts it to both (see l2_i8_fold_kernel for the protocol).
// PAIR (CS == 2): the cluster is a tcgen05 CTA pair instead -- one cta_group::2 MMA of M = 256 spans both CTAs' query
// tiles and each CTA keeps only its own half of the train tile.  Per CTA and e2m1 tile that is 62 KiB of operand reads +
// 30 KiB of TMA fill per 960 cycles (97 B/clk) instead of 92 + 60 KiB (160 B/clk against the 128 B/clk of shared memory,
// which is what held the multicast version at 0.77 of the mxf4 issue rate: sm__mem_tensor_cycles_active 92 %).
template <int NKB, int CS, bool FP4, bool PAIR = false>
__global__ void __launch_bounds__(THREADS, 1)
hamming_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_t,
                  const DevPair* __restrict__ pairs, long long total_items, int tiles_max, int S, int backward,
                  int dbits, Top2* __restrict__ out, long long total_rows, int kmul) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    static_assert(!PAIR || CS == 2, "a CTA pair is a cluster of two");
    const Smem sm = carve(smem, NKB, FP4, PAIR);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int W_MMA = 0, W_PROD = 1;
    constexpr int TN = tn_of(FP4), B_KB_BYTES = b_kb_bytes(FP4);
    constexpr int NBFK = nbf_of(PAIR), B_STAGE = PAIR ? B_KB_BYTES / 2 : B_KB_BYTES;
    static_assert(!FP4 || NKB <= 2, "e2m1: 256 bits per K block, |a.b| <= 512");
    constexpr int NABUF = nabuf_of(NKB);
    constexpr bool SEP_ACC = sep_acc_of(NKB);
    constexpr uint16_t MC_MASK = (1u << CS) - 1;
    const uint32_t crank = CS > 1 ? sm100::cluster_ctarank() : 0;
    const long long first_item = CS > 1 ? (long long)sm100::cluster_id_x() : (long long)blockIdx.x;
    const long long item_stride = CS > 1 ? (long long)sm100::cluster_count_x() : (long long)gridDim.x;
    auto decode = [&](long long item) -> Item {
        int tg = (int)(item % tiles_max); item /= tiles_max;     // tiles_max = tile groups per (pair, split)
        Item it = decode_tile((int)(item / S), (int)(item % S), tg * CS + (int)crank, pairs, S, backward, TN);
        it.valid = tg * CS * TM < it.n;        // a CTA whose own row tile lies past the pair still streams for its peer
        return it;
    };

    if (warp == W_PROD && lane == 0) {
        sm100::tma_prefetch_desc(&tmap_q);
        sm100::tma_prefetch_desc(&tmap_t);
        for (int i = 0; i < NABUF; ++i) { sm100::mbar_init(&sm.a_full[i], 1); sm100::mbar_init(&sm.a_empty[i], 1); }
        // PAIR: one multicast commit of the leader frees a stage in both CTAs; a TMEM stage is free once the 16 epilogue
        // warps of BOTH CTAs have drained it (they arrive on the leader's barrier)
        for (int i = 0; i < NBFK; ++i) { sm100::mbar_init(&sm.b_full[i], 1); sm100::mbar_init(&sm.b_empty[i], PAIR ? 1 : CS); }
        for (int i = 0; i < 2; ++i) { sm100::mbar_init(&sm.acc_full[i], 1); sm100::mbar_init(&sm.acc_empty[i], (PAIR ? 2 : 1) * EPI_WARPS); }
        sm100::fence_barrier_init();
    }
    if constexpr (PAIR) {
        sm100::cluster_sync_all();                         // both CTAs of the pair are running before the paired allocation
        if (warp == W_MMA) sm100::tmem_alloc_2sm<512>(sm.tmem_ptr);
    } else {
        if (warp == W_MMA) sm100::tmem_alloc<512>(sm.tmem_ptr);
    }
    sm100::tc_fence_before();
    __syncthreads();
    if constexpr (CS > 1) sm100::cluster_sync_all();
    sm100::tc_fence_after();
    const uint32_t tmem_base = *sm.tmem_ptr;
    constexpr int HALF_ROWS = TN / CS;
    if constexpr (FP4) {      // scale factors 1.0 everywhere: warps 4..7 own the four TMEM lane quarters
        if (warp >= CTRL_WARPS && warp < CTRL_WARPS + 4) {
            const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + SF_COL;
            const uint32_t v = 0x7F7F7F7Fu;
            asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};"
                         ::"r"(taddr), "r"(v) : "memory");
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        }
        sm100::tc_fence_before();
        __syncthreads();
        sm100::tc_fence_after();
    }

    if (warp == W_PROD) {
        uint32_t ab = 0, a_par = 1, bs = 0, b_par = 1;
        for (long long item = first_item; item < total_items; item += item_stride) {
            Item it = decode(item);
            if (!it.valid) continue;
            sm100::mbar_wait(&sm.a_empty[ab], a_par);
            if (sm100::elect_one()) {
                if constexpr (PAIR) {                  // both CTAs' query tiles are counted on the leader's barrier
                    if (crank == 0) sm100::mbar_arrive_expect_tx(&sm.a_full[ab], (uint32_t)(2 * NKB * A_KB_BYTES));
                    const uint32_t lbar = sm100::mapa_u32(sm100::smem_u32(&sm.a_full[ab]), 0);
#pragma unroll
                    for (int kb = 0; kb < NKB; ++kb)
                        sm100::tma_load_2d_2sm(sm.a + (size_t)(ab * NKB + kb) * A_KB_BYTES, &tmap_q, lbar,
                                               kb * KB_BYTES, it.q_off + it.r0);
                } else {
                    sm100::mbar_arrive_expect_tx(&sm.a_full[ab], (uint32_t)(NKB * A_KB_BYTES));
#pragma unroll
                    for (int kb = 0; kb < NKB; ++kb)
                        sm100::tma_load_2d(sm.a + (size_t)(ab * NKB + kb) * A_KB_BYTES, &tmap_q, &sm.a_full[ab],
                                           kb * KB_BYTES, it.q_off + it.r0);
                }
            }
            __syncwarp();
            if (++ab == NABUF) { ab = 0; a_par ^= 1; }
            const int row0 = it.t_off + it.c_begin + (int)crank * HALF_ROWS;
            for (int j = 0; j < it.ntiles; ++j) {
#pragma unroll
                for (int kb = 0; kb < NKB; ++kb) {
                    sm100::mbar_wait(&sm.b_empty[bs], b_par);
                    if (sm100::elect_one()) {
                        if constexpr (PAIR) {          // this CTA's half of the tile stays here; both halves count on the leader's barrier
                            if (crank == 0) sm100::mbar_arrive_expect_tx(&sm.b_full[bs], (uint32_t)(2 * B_STAGE));
                            sm100::tma_load_2d_2sm(sm.b + (size_t)bs * B_STAGE, &tmap_t, sm100::mapa_u32(sm100::smem_u32(&sm.b_full[bs]), 0),
                                                   kb * KB_BYTES, row0 + j * TN);
                        } else {
                            sm100::mbar_arrive_expect_tx(&sm.b_full[bs], (uint32_t)B_KB_BYTES);
                            uint8_t* bdst = sm.b + (size_t)bs * B_KB_BYTES + (size_t)crank * (B_KB_BYTES / CS);
                            if constexpr (CS > 1) sm100::tma_load_2d_mc(bdst, &tmap_t, &sm.b_full[bs], kb * KB_BYTES, row0 + j * TN, MC_MASK);
                            else sm100::tma_load_2d(bdst, &tmap_t, &sm.b_full[bs], kb * KB_BYTES, row0 + j * TN);
                        }
                    }
                    __syncwarp();
                    if (++bs == NBFK) { bs = 0; b_par ^= 1; }
                }
            }
        }
    } else if (warp == W_MMA && (!PAIR || crank == 0)) {
        constexpr uint32_t idesc = sm100::make_idesc(/*S32*/ 2, /*s8*/ 1, /*s8*/ 1, PAIR ? 2 * TM : TM, 256);
        constexpr uint32_t idesc4 = PAIR ? (IDESC_MXF4 & ~(0x1Fu << 24)) | ((uint32_t)(2 * TM >> 4) << 24) : IDESC_MXF4;
        const uint32_t a_lo0 = sm100::smem_desc_lo_sw128(sm100::smem_u32(sm.a));
        const uint32_t b_lo0 = sm100::smem_desc_lo_sw128(sm100::smem_u32(sm.b));
        uint32_t ab = 0, a_par = 0, bs = 0, b_par = 0, as = 0, acc_par = 3;
        for (long long item = first_item; item < total_items; item += item_stride) {
            Item it = decode(item);
            if (!it.valid) continue;
            sm100::mbar_wait(&sm.a_full[ab], a_par);
            const uint32_t a_lo = a_lo0 + ab * (NKB * A_KB_BYTES >> 4);
            for (int j = 0; j < it.ntiles; ++j) {
                sm100::mbar_wait(&sm.acc_empty[as], (acc_par >> as) & 1);
                acc_par ^= 1u << as;
                sm100::tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * 256;
#pragma unroll
                for (int kb = 0; kb < NKB; ++kb) {
                    sm100::mbar_wait(&sm.b_full[bs], b_par);
                    sm100::tc_fence_after();
                    if (sm100::elect_one()) {
                        const uint32_t al = a_lo + kb * (A_KB_BYTES >> 4), bl = b_lo0 + bs * (B_STAGE >> 4);
#pragma unroll
                        for (int k = 0; k < KB_BYTES / 32; ++k) {
                            const uint64_t ad = sm100::smem_desc_join(al + 2 * k), bd = sm100::smem_desc_join(bl + 2 * k);
                            const uint32_t accf = (kb | k) ? 1u : 0u;
                            if constexpr (FP4 && PAIR) mma_mxf4_2sm(d_tmem, ad, bd, idesc4, tmem_base + SF_COL, accf);
                            else if constexpr (FP4) mma_mxf4(d_tmem, ad, bd, idesc4, tmem_base + SF_COL, accf);
                            else if constexpr (PAIR) sm100::mma_i8_2sm(d_tmem, ad, bd, idesc, accf);
                            else sm100::mma_i8(d_tmem, ad, bd, idesc, accf);
                        }
                        if constexpr (PAIR) sm100::mma_commit_2sm_mc(&sm.b_empty[bs], MC_MASK);
                        else if constexpr (CS > 1) sm100::mma_commit_mc(&sm.b_empty[bs], MC_MASK);
                        else sm100::mma_commit(&sm.b_empty[bs]);
                        if (SEP_ACC && kb == NKB - 1) {
                            if constexpr (PAIR) sm100::mma_commit_2sm_mc(&sm.acc_full[as], MC_MASK);
                            else sm100::mma_commit(&sm.acc_full[as]);
                        }
                    }
                    __syncwarp();
                    if (++bs == NBFK) { bs = 0; b_par ^= 1; }
                }
                as ^= 1;
            }
            if (sm100::elect_one()) {
                if constexpr (PAIR) sm100::mma_commit_2sm_mc(&sm.a_empty[ab], MC_MASK);   // both CTAs' query tiles are free
                else sm100::mma_commit(&sm.a_empty[ab]);
            }
            __syncwarp();
            if (++ab == NABUF) { ab = 0; a_par ^= 1; }
        }
    } else if (warp >= CTRL_WARPS) {
        // epilogue: ALL 16 warps drain EVERY tile -- warp (quarter, cb) reads its 32 TMEM lanes x 64 columns with two
        // 32-column tcgen05.ld issued back to back and hands the stage back before any arithmetic (l2_i8_fold_kernel has
        // the measurements behind this shape).  Column block 3 of an e2m1 tile ends at column 240: its last chunk would be
        // the scale factors and is skipped.
        const int e = warp - CTRL_WARPS;
        const int cb = e >> 2, quarter = warp & 3;
        constexpr int CBW = 64, CB_CHUNKS = CBW / CHUNK;
        const int width = FP4 && cb == 3 ? TN - 3 * CBW : CBW;
        constexpr int KOFF = (512 << KEY_SHIFT) + KEY_LOW;
        uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + cb * CBW;     // this tile's stage: toggles by 256 columns
        uint32_t acc_bar = PAIR ? sm100::mapa_u32(sm100::smem_u32(&sm.acc_empty[0]), 0) : sm100::smem_u32(&sm.acc_empty[0]);
        uint32_t ready_bar = sm100::smem_u32(SEP_ACC ? &sm.acc_full[0] : &sm.b_empty[0]);
        // "accumulator ready": SEP_ACC -> acc_full[tile & 1], parity (tile >> 1) & 1; else the ring stage of the tile's last K block
        uint32_t as8 = 0, rs = SEP_ACC ? 0 : (NKB - 1) * 8, rp = 0, nitem = 0;
        int mtn = -(TN / CHUNK);
        asm volatile("" : "+r"(taddr), "+r"(acc_bar), "+r"(ready_bar), "+r"(mtn));
        for (long long item = first_item; item < total_items; item += item_stride) {
            Item it = decode(item);
            if (!it.valid) continue;
            int v1 = (int)0x80000000, v2 = (int)0x80000000;
            auto insert = [&](int key) {                   // running top-2 of packed chunk keys
                v2 = max(v2, min(v1, key));
                v1 = max(v1, key);
            };
            auto push2 = [&](int ka, int kb) {             // top-2 of {v1, v2, ka, kb}: 5 min/max instead of 6
                const int mx = max(ka, kb), mn = min(ka, kb);
                const int t = min(v1, mx);
                v1 = max(v1, mx);
                v2 = max3(t, v2, mn);
            };
            // key of chunk c of this warp's 4 in tile j = (max << 20) + lowc_c - j * (TN / 16)
            int lowc0 = KOFF - ((it.c_begin + cb * CBW) >> 4), lowc1 = lowc0 - 1, lowc2 = lowc0 - 2, lowc3 = lowc0 - 3;
            asm volatile("" : "+r"(lowc0), "+r"(lowc1), "+r"(lowc2), "+r"(lowc3));
            int rem = it.c_end - it.c_begin - cb * CBW;      // existing columns from this warp's first one on
            for (int j = 0; j < it.ntiles; ++j) {
                sm100::mbar_wait_addr(ready_bar + rs, rp);
                if constexpr (SEP_ACC) {
                    rs ^= 8; rp ^= (rs == 0);
                } else {
                    rs += NKB * 8;
                    if (rs >= (uint32_t)(NBFK * 8)) { rs -= NBFK * 8; rp ^= 1; }
                }
                sm100::tc_fence_after();
                uint32_t acc[2][32];
                sm100::tmem_ld32(taddr, acc[0]);
                sm100::tmem_ld32(taddr + 32, acc[1]);
                sm100::tmem_ld_wait();                        // everything is in registers: hand the TMEM stage back first
                sm100::tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if constexpr (PAIR) sm100::mbar_arrive_cluster(acc_bar + as8);
                    else sm100::mbar_arrive_addr(acc_bar + as8);
                }
                taddr ^= 256; as8 ^= 8;
                if (rem < width) { mask32<FP4>(acc[0], rem); mask32<FP4>(acc[1], rem - 32); }     // tail tile
                rem -= TN;
                const int m0 = chunk_max<0, FP4>(acc[0]), m1 = chunk_max<16, FP4>(acc[0]);
                const int m2 = chunk_max<0, FP4>(acc[1]);
                push2(pack_key(m0, kmul, pack_key(j, mtn, lowc0)), pack_key(m1, kmul, pack_key(j, mtn, lowc1)));
                if (FP4 && cb == 3) {          // columns 240..255 of an e2m1 tile are the scale factors, not distances
                    insert(pack_key(m2, kmul, pack_key(j, mtn, lowc2)));
                } else {
                    const int m3 = chunk_max<16, FP4>(acc[1]);
                    push2(pack_key(m2, kmul, pack_key(j, mtn, lowc2)), pack_key(m3, kmul, pack_key(j, mtn, lowc3)));
                }
            }
            // the four warps of a TMEM lane quarter hold partial top-2s of the same 32 query rows: merge in shared memory
            int2* xbuf = sm.xch + (size_t)(nitem & 1) * (EPI_WARPS - 4) * 32;
            ++nitem;
            const bool merger = e < 4;
            if (!merger) xbuf[(e - 4) * 32 + lane] = make_int2(v1, v2);
            asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32) : "memory");
            if (merger) {
#pragma unroll
                for (int w = 1; w < 4; ++w) {
                    const int2 o = xbuf[((w - 1) * 4 + quarter) * 32 + lane];
                    insert(o.x); insert(o.y);
                }
                const int row = it.r0 + quarter * 32 + lane;
                if (row < it.n) {
                    auto mk = [&](int key) -> u64 {        // (Hamming distance of the chunk's best column, chunk's first column)
                        if (key < 0) return B200M_SENTINEL;
                        const int dot = (key >> KEY_SHIFT) - 512;
                        const int c0 = (KEY_LOW - (key & KEY_LOW)) << 4;
                        return ((u64)(u32)((dbits - dot) >> 1) << 32) | (u64)(u32)c0;
                    };
                    Top2 t; t.k1 = mk(v1); t.k2 = mk(v2);
                    out[(long long)it.split * total_rows + it.row_base + row] = t;
                }
            }
        }
    }

    sm100::tc_fence_before();
    __syncthreads();
    if constexpr (CS > 1) sm100::cluster_sync_all();
    if (warp == W_MMA) {
        if constexpr (PAIR) sm100::tmem_dealloc_2sm<512>(tmem_base);
        else sm100::tmem_dealloc<512>(tmem_base);
    }
}

// ---------------------------------------------------------------------------------------------------------
// Hamming distance between two D-byte rows (any D, any alignment class named by `vec`: 16, 4 or 1 bytes per load).
__device__ __forceinline__ int popc_row(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, int D, int vec) {
    int d = 0;
    if (vec == 16) {
        for (int k = 0; k < D; k += 16) {
            uint4 x = __ldg(reinterpret_cast<const uint4*>(a + k)), y = __ldg(reinterpret_cast<const uint4*>(b + k));
            d += __popc(x.x ^ y.x) + __popc(x.y ^ y.y) + __popc(x.z ^ y.z) + __popc(x.w ^ y.w);
        }
    } else if (vec == 4) {
        for (int k = 0; k < D; k += 4)
            d += __popc(__ldg(reinterpret_cast<const u32*>(a + k)) ^ __ldg(reinterpret_cast<const u32*>(b + k)));
    } else {
        for (int k = 0; k < D; ++k) d += __popc((u32)(__ldg(a + k) ^ __ldg(b + k)));
    }
    return d;
}

// One warp per query row: merge the column splits' chunk-level top-2, recompute both chunks (lane = chunk x column)
// with POPC on the packed bits, write the exact (distance << 32 | column) top-2 into slice 0.
// BEST_ONLY (cross-check and plain 1-NN consume only the nearest neighbour): half a warp per row recomputes the best
// chunk alone -- half the gathered train rows (the kernel is bound by those 64-byte gathers from L2) and a single-key
// min reduction instead of the top-2 butterfly.
template <bool BEST_ONLY>
__global__ void __launch_bounds__(256)
hamming_tc_resolve_kernel(const uint8_t* __restrict__ q_bits, const uint8_t* __restrict__ t_bits, int D, int vec,
                          const DevPair* __restrict__ pairs, int blocks_per_pair, Top2* __restrict__ arr,
                          long long total_rows, int S, int backward) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int p = blockIdx.x / blocks_per_pair;
    const int i = BEST_ONLY ? (blockIdx.x - p * blocks_per_pair) * 16 + warp * 2 + (lane >> 4)
                            : (blockIdx.x - p * blocks_per_pair) * 8 + warp;
    const DevPair pr = pairs[p];
    int q_off = pr.q_off, n = pr.n, t_off = pr.t_off, m = pr.m;
    long long row_base = pr.fwd_off;
    const uint8_t *qb = q_bits, *tb = t_bits;
    if (backward) { q_off = pr.t_off; n = pr.m; t_off = pr.q_off; m = pr.n; row_base = pr.bwd_off; qb = t_bits; tb = q_bits; }
    if constexpr (BEST_ONLY) {
        static_assert(CHUNK == 16, "half a warp per chunk");
        const bool live = i < n;                     // the other half-warp of this warp may still have a row
        const long long row = row_base + (live ? i : 0);
        u64 ck = B200M_SENTINEL;
        if (live) {
            ck = arr[row].k1;
            for (int s = 1; s < S; ++s) { const u64 o = arr[(long long)s * total_rows + row].k1; ck = o < ck ? o : ck; }
        }
        const int col = (int)(u32)ck + (lane & (CHUNK - 1));
        u64 k1 = B200M_SENTINEL;
        if (ck != B200M_SENTINEL && col < m) {
            const int d = popc_row(qb + (long long)(q_off + i) * D, tb + (long long)(t_off + col) * D, D, vec);
            k1 = ((u64)(u32)d << 32) | (u64)(u32)col;
        }
#pragma unroll
        for (int off = 8; off > 0; off >>= 1) { const u64 o = __shfl_xor_sync(0xffffffffu, k1, off); k1 = o < k1 ? o : k1; }
        if (live && (lane & 15) == 0) { Top2 r; r.k1 = k1; r.k2 = B200M_SENTINEL; arr[row] = r; }
        return;
    }
    if (i >= n) return;
    const long long row = row_base + i;
    Top2 t = arr[row];
    for (int s = 1; s < S; ++s) {
        Top2 o = arr[(long long)s * total_rows + row];
        top2_insert(t.k1, t.k2, o.k1);
        top2_insert(t.k1, t.k2, o.k2);
    }
    const u64 ck = lane < CHUNK ? t.k1 : t.k2;
    const int col = (int)(u32)ck + (lane & (CHUNK - 1));
    u64 k1 = B200M_SENTINEL, k2 = B200M_SENTINEL;
    if (ck != B200M_SENTINEL && col < m) {
        const int d = popc_row(qb + (long long)(q_off + i) * D, tb + (long long)(t_off + col) * D, D, vec);
        k1 = ((u64)(u32)d << 32) | (u64)(u32)col;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        u64 o1 = __shfl_xor_sync(0xffffffffu, k1, off), o2 = __shfl_xor_sync(0xffffffffu, k2, off);
        top2_insert(k1, k2, o1);
        top2_insert(k1, k2, o2);
    }
    if (lane == 0) { Top2 r; r.k1 = k1; r.k2 = k2; arr[row] = r; }
}

// bits -> +-1 int8 dimensions.  One thread per source byte: 8 output bytes (dimension 8k + b = bit b of byte k);
// dimensions past 8 D are 0 (add nothing to any dot product).
__global__ void __launch_bounds__(256)
prep_hamming_tc_kernel(const uint8_t* __restrict__ src, long long rows, int D, int8_t* __restrict__ dst, int Dpad) {
    const int per_row = Dpad >> 3;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= rows * per_row) return;
    const long long r = idx / per_row;
    const int k = (int)(idx - r * per_row);
    uint2 o = make_uint2(0u, 0u);
    if (k < D) {
        const u32 x = __ldg(src + r * D + k);
        auto spread = [](u32 nib) -> u32 {             // 4 bits -> 4 bytes of 0x01 (set) / 0xFF (clear)
            u32 w = (nib & 1u) | ((nib & 2u) << 7) | ((nib & 4u) << 14) | ((nib & 8u) << 21);
            return w | ((w ^ 0x01010101u) * 0xFFu);
        };
        o.x = spread(x & 15u);
        o.y = spread(x >> 4);
    }
    *reinterpret_cast<uint2*>(dst + r * Dpad + 8 * k) = o;
}

// bits -> e2m1 nibbles: +1.0 = 0x2 for a set bit, -1.0 = 0xA for a clear one.  One thread per source byte: one 32-bit word
// (nibble b of the word = bit b of the byte); nibbles past 8 D are 0.0.
__global__ void __launch_bounds__(256)
prep_hamming_fp4_kernel(const uint8_t* __restrict__ src, long long rows, int D, uint8_t* __restrict__ dst, int Dpad4) {
    const int per_row = Dpad4 >> 2;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= rows * per_row) return;
    const long long r = idx / per_row;
    const int k = (int)(idx - r * per_row);
    u32 o = 0u;
    if (k < D) {
        const u32 x = __ldg(src + r * D + k);
        u32 w = 0u;
#pragma unroll
        for (int b = 0; b < 8; ++b) w |= ((x >> b) & 1u) << (4 * b + 3);
        o = 0xAAAAAAAAu ^ w;                           // set bit: 0xA ^ 0x8 = 0x2
    }
    *reinterpret_cast<u32*>(dst + r * Dpad4 + 4 * k) = o;
}

}  // namespace

static_assert(smem_bytes(1, true) <= 232448 && smem_bytes(2, true) <= 232448, "shared memory (e2m1)");

// fp4 == false: stores are int8 [rows, Dpad], Dpad = 8 D rounded up to 128 (b200m_prep_hamming_tc);
// fp4 == true : stores are packed e2m1 [rows, Dpad], Dpad = 4 D rounded up to 128 bytes (b200m_prep_hamming_fp4).
int launch_hamming_tc_top2(const void* q_store, int64_t q_rows, const void* t_store, int64_t t_rows, int Dpad, int dbits,
                           bool fp4, const DevPair* pairs, int num_pairs, int max_n, int max_m, int S, bool backward,
                           Top2* out, long long total_rows, cudaStream_t st) {
    if (Dpad % KB_BYTES != 0 || Dpad * (fp4 ? 2 : 1) < dbits) return B200M_EINVAL;
    const int nkb = Dpad / KB_BYTES;
    if (nkb < 1 || nkb > (fp4 ? 2 : 4)) return B200M_EUNSUPPORTED;           // |a.b| <= 512 is what the packed key holds
    if (backward) {
        const void* d = q_store; q_store = t_store; t_store = d;
        int64_t r = q_rows; q_rows = t_rows; t_rows = r;
        int mx = max_n; max_n = max_m; max_m = mx;
    }
    const int TN = tn_of(fp4);
    if (((long long)max_m + TN) >> 4 > KEY_LOW) return B200M_EUNSUPPORTED;
    static B200mPerDevice once;          // val[f4 * 5 + nkb] = max active 2-CTA clusters per format and K-block count (0: single CTAs)
    int dev = 0, sms = 0;
    if (!b200m_current_device(&dev, &sms)) return B200M_ENODEVICE;
    int (&cs_ok)[12] = once.val[dev];
    // cs_ == 3 names the CTA-pair (cta_group::2) instantiation of the 2-CTA cluster
    auto kern = [](int nkb_, int cs_, bool f4) -> const void* {
        if (cs_ == 3) {
            if (f4) return nkb_ == 1 ? (const void*)hamming_tc_kernel<1, 2, true, true> : (const void*)hamming_tc_kernel<2, 2, true, true>;
            switch (nkb_) {
                case 1: return (const void*)hamming_tc_kernel<1, 2, false, true>; case 2: return (const void*)hamming_tc_kernel<2, 2, false, true>;
                case 3: return (const void*)hamming_tc_kernel<3, 2, false, true>; default: return (const void*)hamming_tc_kernel<4, 2, false, true>;
            }
        }
        if (f4) {
            switch (nkb_ * 2 + cs_ - 1) {
                case 2: return (const void*)hamming_tc_kernel<1, 1, true>; case 3: return (const void*)hamming_tc_kernel<1, 2, true>;
                case 4: return (const void*)hamming_tc_kernel<2, 1, true>; default: return (const void*)hamming_tc_kernel<2, 2, true>;
            }
        }
        switch (nkb_ * 2 + cs_ - 1) {
            case 2: return (const void*)hamming_tc_kernel<1, 1, false>; case 3: return (const void*)hamming_tc_kernel<1, 2, false>;
            case 4: return (const void*)hamming_tc_kernel<2, 1, false>; case 5: return (const void*)hamming_tc_kernel<2, 2, false>;
            case 6: return (const void*)hamming_tc_kernel<3, 1, false>; case 7: return (const void*)hamming_tc_kernel<3, 2, false>;
            case 8: return (const void*)hamming_tc_kernel<4, 1, false>; default: return (const void*)hamming_tc_kernel<4, 2, false>;
        }
    };
    if (!once.done[dev]) {
        for (int f4 = 0; f4 < 2; ++f4)
            for (int nk = 1; nk <= (f4 ? 2 : 4); ++nk) {
                for (int cs = 1; cs <= 3; ++cs)
                    B200M_CUDA_CHECK(cudaFuncSetAttribute(kern(nk, cs, f4), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                          (int)smem_bytes(nk, f4, cs == 3)), "hamming_tc smem attr");
                cudaLaunchConfig_t qc = {};
                cudaLaunchAttribute qa[1];
                qa[0].id = cudaLaunchAttributeClusterDimension;
                qa[0].val.clusterDim.x = 2; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
                qc.gridDim = dim3((unsigned)(sms & ~1)); qc.blockDim = dim3(THREADS); qc.dynamicSmemBytes = smem_bytes(nk, f4, true);
                qc.attrs = qa; qc.numAttrs = 1;
                int nc = 0;
                if (cudaOccupancyMaxActiveClusters(&nc, kern(nk, 3, f4), &qc) == cudaSuccess && nc * 2 >= sms && !getenv("B200M_NO_CLUSTER"))
                    cs_ok[f4 * 5 + nk] = nc;
                else
                    cudaGetLastError();
            }
        cs_ok[11] = getenv("B200M_NO_PAIR") ? 0 : 1;       // A/B switch: multicast clusters of single-CTA MMAs instead of CTA pairs
        once.done[dev] = 1;
    }
    CUtensorMap tq, tt;
    const int cs = cs_ok[(fp4 ? 5 : 0) + nkb] > 0 ? 2 : 1;
    const bool pair = cs == 2 && cs_ok[11];
    int rc = b200m_make_tmap_2d(&tq, q_store, (uint64_t)q_rows, (uint64_t)Dpad, TM, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1);
    if (rc != B200M_OK) return rc;
    rc = b200m_make_tmap_2d(&tt, t_store, (uint64_t)t_rows, (uint64_t)Dpad, TN / cs, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1);
    if (rc != B200M_OK) return rc;
    const int tiles_max = (max_n + TM - 1) / TM;
    int groups = (tiles_max + cs - 1) / cs;
    long long items = (long long)num_pairs * S * groups;
    if (items <= 0) return B200M_OK;
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    const long long max_units = cs == 2 ? cs_ok[(fp4 ? 5 : 0) + nkb] : sms;
    cfg.gridDim = dim3((unsigned)((items < max_units ? items : max_units) * cs));
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = smem_bytes(nkb, fp4, pair);
    cfg.stream = st;
    cfg.attrs = at; cfg.numAttrs = 1;
    int bw = backward ? 1 : 0, kmul = 1 << KEY_SHIFT;
    void* args[] = {(void*)&tq, (void*)&tt, (void*)&pairs, (void*)&items, (void*)&groups, (void*)&S, (void*)&bw,
                    (void*)&dbits, (void*)&out, (void*)&total_rows, (void*)&kmul};
    cudaError_t e = cudaLaunchKernelExC(&cfg, kern(nkb, pair ? 3 : cs, fp4), args);
    b200m_count_launch();
    B200M_CUDA_CHECK(e, "hamming_tc_kernel launch");
    return B200M_OK;
}

int launch_hamming_tc_resolve(const uint8_t* q_bits, const uint8_t* t_bits, int D, const DevPair* pairs, int num_pairs,
                              int max_n, int max_m, int S, bool backward, bool best_only, Top2* arr, long long total_rows,
                              cudaStream_t st) {
    const int max_rows = backward ? max_m : max_n;
    const int rpb = best_only ? 16 : 8;            // rows per 256-thread block
    const int bpp = (max_rows + rpb - 1) / rpb;
    const long long grid = (long long)num_pairs * bpp;
    if (grid <= 0) return B200M_OK;
    if (grid > 0x7fffffffLL) return B200M_EUNSUPPORTED;
    const uintptr_t al = (uintptr_t)q_bits | (uintptr_t)t_bits | (uintptr_t)D;
    const int vec = (al & 15) == 0 ? 16 : ((al & 3) == 0 ? 4 : 1);
    if (best_only)
        hamming_tc_resolve_kernel<true><<<(unsigned)grid, 256, 0, st>>>(q_bits, t_bits, D, vec, pairs, bpp, arr, total_rows, S,
                                                                        backward ? 1 : 0);
    else
        hamming_tc_resolve_kernel<false><<<(unsigned)grid, 256, 0, st>>>(q_bits, t_bits, D, vec, pairs, bpp, arr, total_rows, S,
                                                                         backward ? 1 : 0);
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "hamming_tc_resolve_kernel launch");
    return B200M_OK;
}

extern "C" {

int b200m_hamming_tc_width(int D, int* Dpad) {
    if (D <= 0 || !Dpad) return B200M_EINVAL;
    const int w = (8 * D + KB_BYTES - 1) / KB_BYTES * KB_BYTES;
    if (w > 4 * KB_BYTES) return B200M_EUNSUPPORTED;
    *Dpad = w;
    return B200M_OK;
}

int b200m_hamming_fp4_width(int D, int* Dpad4) {
    if (D <= 0 || !Dpad4) return B200M_EINVAL;
    const int w = (4 * D + KB_BYTES - 1) / KB_BYTES * KB_BYTES;
    if (w > 2 * KB_BYTES) return B200M_EUNSUPPORTED;
    *Dpad4 = w;
    return B200M_OK;
}

int b200m_prep_hamming_tc(const uint8_t* src, int64_t rows, int D, int8_t* dst, int Dpad, void* stream) {
    int want = 0;
    int rc = b200m_hamming_tc_width(D, &want);
    if (rc != B200M_OK) return rc;
    if (rows < 0 || Dpad != want) return B200M_EINVAL;
    if (rows == 0) return B200M_OK;
    if (!src || !dst || ((uintptr_t)dst & 15)) return B200M_EINVAL;
    const long long work = rows * (Dpad >> 3);
    const long long blocks = (work + 255) / 256;
    if (blocks > 0x7fffffffLL) return B200M_EUNSUPPORTED;
    prep_hamming_tc_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, rows, D, dst, Dpad);
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "prep_hamming_tc_kernel launch");
    return B200M_OK;
}

int b200m_prep_hamming_fp4(const uint8_t* src, int64_t rows, int D, uint8_t* dst, int Dpad4, void* stream) {
    int want = 0;
    int rc = b200m_hamming_fp4_width(D, &want);
    if (rc != B200M_OK) return rc;
    if (rows < 0 || Dpad4 != want) return B200M_EINVAL;
    if (rows == 0) return B200M_OK;
    if (!src || !dst || ((uintptr_t)dst & 15)) return B200M_EINVAL;
    const long long work = rows * (Dpad4 >> 2);
    const long long blocks = (work + 255) / 256;
    if (blocks > 0x7fffffffLL) return B200M_EUNSUPPORTED;
    prep_hamming_fp4_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, rows, D, dst, Dpad4);
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "prep_hamming_fp4_kernel launch");
    return B200M_OK;
}

}  // extern "C"
