// l2_i8.cu — exact integer L2 knn (k=2) for SIFT-style descriptors on the 5th-gen tensor cores.
// Replaces cv2.BFMatcher(NORM_L2).knnMatch(k=2)/match as called at
// /root/reference/FeatureMatchingExtraction/feature_matchers.py:210,227 for integer-valued 0..255
// float32 descriptors (traditional_detectors.py:42), bit-exactly:
//   d^2(i,j) = ||a'_i||^2 + ||b'_j||^2 - 2 a'_i . b'_j   with a' = a-128, b' = b-128 (int8),
// the dot product from tcgen05.mma kind::i8 (int32 accumulators in TMEM), all terms exact integers.
//
// Persistent, warp-specialised CTA (one per SM), 608 threads:
//   warps 0-15  epilogue : TMEM lane quarter = warp % 4, 64-column slice of the 256-column tile = warp / 4
//                          (every SM sub-partition hosts four epilogue warps).  Per tile and warp: two
//                          tcgen05.ld of 32 columns, key = (||b'||^2*64 + c) - 128*acc (one IMAD per element),
//                          3-input-min tree over the 64-column chunk, running top-2 of chunk minima.  The only
//                          synchronisation is mbarrier waits (accumulator full, column constants staged).
//   warp  16    producer : TMA (cp.async.bulk.tensor, SWIZZLE_128B) query tile (kept for the whole work
//                          item, double buffered across items) and a ring of train tiles
//   warp  17    MMA      : one thread issues tcgen05.mma M=128 N=256 K=32 x 4 per 128-byte K block into a
//                          double-buffered TMEM accumulator (2 x 256 columns = all 512).  N=256 because
//                          N=128 issues at 2/3 rate and every tcgen05.commit drains the pipe (tools/mma_peak).
//   warp  18    stager   : column constants ||b'||^2*64 + (column & 63) (or a mask for out-of-range columns)
//                          of upcoming tiles into an 8-deep shared-memory ring, so the epilogue warps carry
//                          no global loads, no stores and no CTA barrier.
// No N x M matrix reaches HBM: per query row and column slice only {best, runner-up candidate} leave the
// SM (B200M_I8_GROUPS slices of Top2 per column split; finalize merges them like column splits).
//
// Exactness of the runner-up: the running top-2 is over CHUNK MINIMA (64 consecutive train columns,
// aligned to 64), so k1 is the exact best (ties -> lowest column: the column sits in the low key bits,
// a group visits its chunks in ascending order with a strict compare, and slices merge by u64 min)
// and k2 is the best among all OTHER chunks.  The true runner-up is min(k2, runner-up inside the
// winner's chunk); tail.cu recomputes those <= 63 distances exactly where the result depends on it.
#include <cstdio>
#include <cstdlib>
#include "common.cuh"
#include "sm100.cuh"

// (The fold kernel's epilogue warps spin on mbarrier.try_wait; a suspend-time hint and a __nanosleep back-off were measured
// equal, profiles/r02_notes.md.)

namespace {

constexpr int TM = 128;            // query rows per tile  (UMMA M)
constexpr int TN = 256;            // train columns per tile (UMMA N) = one TMEM stage
constexpr int KB_BYTES = 128;      // one swizzle-128B K block
constexpr int A_KB_BYTES = TM * KB_BYTES;   // 16 KiB
constexpr int B_KB_BYTES = TN * KB_BYTES;   // 32 KiB
constexpr int NABUF = 2;
constexpr int NSLICE = B200M_I8_GROUPS;     // 64-column slices of a tile = Top2 slices per split
constexpr int EPI_WARPS = 4 * NSLICE;
constexpr int EPI_THREADS = EPI_WARPS * 32;
constexpr int CTRL_WARPS = 4;                // MMA, producer, stager (+1 idle) take the lowest warp ids
constexpr int THREADS = EPI_THREADS + CTRL_WARPS * 32;
constexpr int CHUNK = B200M_I8_CHUNK;       // columns per key chunk (6 index bits) = slice width
constexpr int CBR = 4;                      // column-constant ring depth (tiles)
constexpr int MASK_CB = 0x60000000;         // column-constant of a masked column: key stays in [2^30, 2^31)
constexpr int INVALID_VAL = 1 << 24;        // (masked key >> 6) >= 2^24 > any valid value
static_assert(TN == NSLICE * CHUNK, "one chunk per slice per tile");

struct Item {
    int q_off, n, t_off, m, r0, c_begin, c_end, split, ntiles;
    long long row_base;
    bool valid;
};

__device__ __forceinline__ Item decode_tile(int p, int split, int tile, const DevPair* __restrict__ pairs, int S, int backward);

__device__ __forceinline__ Item decode_item(long long item, const DevPair* __restrict__ pairs, int tiles_max, int S,
                                            int backward) {
    int tile = (int)(item % tiles_max); item /= tiles_max;
    return decode_tile((int)(item / S), (int)(item % S), tile, pairs, S, backward);
}

__device__ __forceinline__ Item decode_tile(int p, int split, int tile, const DevPair* __restrict__ pairs, int S, int backward) {
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

__device__ __forceinline__ int min3(int a, int b, int c) { return min(min(a, b), c); }

struct SmemLayout {
    uint8_t* a;        // [NABUF][nkb][16 KiB]
    uint8_t* b;        // [NB][32 KiB]
    int* cb;           // [CBR][TN]
    uint64_t* a_full;  // [NABUF]
    uint64_t* a_empty; // [NABUF]
    uint64_t* b_full;  // [NB]  (NB <= 8)
    uint64_t* b_empty; // [NB]
    uint64_t* acc_full;   // [2]
    uint64_t* acc_empty;  // [2]
    uint64_t* cb_full;    // [CBR]
    uint64_t* cb_empty;   // [CBR]
    uint32_t* tmem_ptr;
};

__host__ __device__ constexpr int nb_of(int nkb) { return nkb == 1 ? 5 : 4; }   // train K-block ring depth

__device__ __forceinline__ SmemLayout carve(uint8_t* base, int nkb) {
    SmemLayout s;
    s.a = base;
    s.b = s.a + NABUF * nkb * A_KB_BYTES;
    uint8_t* p = s.b + nb_of(nkb) * B_KB_BYTES;
    s.cb = (int*)p; p += CBR * TN * sizeof(int);
    s.a_full = (uint64_t*)p; p += NABUF * 8;
    s.a_empty = (uint64_t*)p; p += NABUF * 8;
    s.b_full = (uint64_t*)p; p += 8 * 8;
    s.b_empty = (uint64_t*)p; p += 8 * 8;
    s.acc_full = (uint64_t*)p; p += 2 * 8;
    s.acc_empty = (uint64_t*)p; p += 2 * 8;
    s.cb_full = (uint64_t*)p; p += CBR * 8;
    s.cb_empty = (uint64_t*)p; p += CBR * 8;
    s.tmem_ptr = (uint32_t*)p;
    return s;
}

__host__ __device__ constexpr size_t smem_bytes(int nkb) {
    return 1024 /*alignment slack*/ + (size_t)NABUF * nkb * A_KB_BYTES + (size_t)nb_of(nkb) * B_KB_BYTES +
           CBR * TN * sizeof(int) + (2 * NABUF + 16 + 4 + 2 * CBR) * 8 + 16;
}

template <int NKB>
__global__ void __launch_bounds__(THREADS, 1)
l2_i8_top2_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_t,
                  const int32_t* __restrict__ q_norm, const int32_t* __restrict__ t_norm,
                  const DevPair* __restrict__ pairs, long long total_items, int tiles_max, int S, int backward,
                  Top2* __restrict__ out, long long total_rows, const int32_t* __restrict__ fold_flag) {
    if (fold_flag && __ldg(fold_flag)) return;          // MODE_RATIO with usable fold digits (the call's snapshot): l2_i8_fold_kernel runs instead
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);   // SWIZZLE_128B: 1 KiB alignment
    const SmemLayout sm = carve(smem, NKB);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int NB = nb_of(NKB);
    // Control roles take the lowest warp ids (the scheduler favours the oldest ready warp).
    constexpr int W_MMA = 0, W_PROD = 1, W_STAGE = 2;

    if (warp == W_PROD && lane == 0) {
        sm100::tma_prefetch_desc(&tmap_q);
        sm100::tma_prefetch_desc(&tmap_t);
        for (int i = 0; i < NABUF; ++i) { sm100::mbar_init(&sm.a_full[i], 1); sm100::mbar_init(&sm.a_empty[i], 1); }
        for (int i = 0; i < NB; ++i) { sm100::mbar_init(&sm.b_full[i], 1); sm100::mbar_init(&sm.b_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { sm100::mbar_init(&sm.acc_full[i], 1); sm100::mbar_init(&sm.acc_empty[i], EPI_WARPS / 2); }
        for (int i = 0; i < CBR; ++i) { sm100::mbar_init(&sm.cb_full[i], 1); sm100::mbar_init(&sm.cb_empty[i], EPI_WARPS / 2); }
        sm100::fence_barrier_init();
    }
    if (warp == W_MMA) sm100::tmem_alloc<512>(sm.tmem_ptr);
    sm100::tc_fence_before();
    __syncthreads();
    sm100::tc_fence_after();
    const uint32_t tmem_base = *sm.tmem_ptr;

    if (warp == W_PROD) {
        // ===================== TMA producer: the warp runs in lockstep, one elected lane issues =====================
        uint32_t ab = 0, a_par = 1;            // a_par: parity to wait for on a_empty[ab] (fresh barrier passes parity 1)
        uint32_t bs = 0, b_par = 1;
        for (long long item = blockIdx.x; item < total_items; item += gridDim.x) {
            Item it = decode_item(item, pairs, tiles_max, S, backward);
            if (!it.valid) continue;
            sm100::mbar_wait(&sm.a_empty[ab], a_par);
            if (sm100::elect_one()) {
                sm100::mbar_arrive_expect_tx(&sm.a_full[ab], (uint32_t)(NKB * A_KB_BYTES));
#pragma unroll
                for (int kb = 0; kb < NKB; ++kb)
                    sm100::tma_load_2d(sm.a + (size_t)(ab * NKB + kb) * A_KB_BYTES, &tmap_q, &sm.a_full[ab],
                                       kb * KB_BYTES, it.q_off + it.r0);
            }
            __syncwarp();
            if (++ab == NABUF) { ab = 0; a_par ^= 1; }
            const int row0 = it.t_off + it.c_begin;
            for (int j = 0; j < it.ntiles; ++j) {
#pragma unroll
                for (int kb = 0; kb < NKB; ++kb) {
                    sm100::mbar_wait(&sm.b_empty[bs], b_par);
                    if (sm100::elect_one()) {
                        sm100::mbar_arrive_expect_tx(&sm.b_full[bs], (uint32_t)B_KB_BYTES);
                        sm100::tma_load_2d(sm.b + (size_t)bs * B_KB_BYTES, &tmap_t, &sm.b_full[bs], kb * KB_BYTES,
                                           row0 + j * TN);
                    }
                    __syncwarp();
                    if (++bs == NB) { bs = 0; b_par ^= 1; }
                }
            }
        }
    } else if (warp == W_MMA) {
        // ===================== MMA issuer: lockstep warp, one elected lane issues =====================
        // Everything per tile is a handful of 32-bit adds: descriptors are kept as (constant hi, running lo).
        constexpr uint32_t idesc = sm100::make_idesc(/*S32*/ 2, /*s8*/ 1, /*s8*/ 1, TM, TN);
        const uint32_t a_lo0 = sm100::smem_desc_lo_sw128(sm100::smem_u32(sm.a));
        const uint32_t b_lo0 = sm100::smem_desc_lo_sw128(sm100::smem_u32(sm.b));
        uint32_t ab = 0, a_par = 0;            // parity to wait for on a_full[ab]
        uint32_t bs = 0, b_par = 0;
        uint32_t as = 0, acc_par = 3;          // bit s: parity to wait for on acc_empty[s]
        for (long long item = blockIdx.x; item < total_items; item += gridDim.x) {
            Item it = decode_item(item, pairs, tiles_max, S, backward);
            if (!it.valid) continue;
            sm100::mbar_wait(&sm.a_full[ab], a_par);
            const uint32_t a_lo = a_lo0 + ab * (NKB * A_KB_BYTES >> 4);
            for (int j = 0; j < it.ntiles; ++j) {
                sm100::mbar_wait(&sm.acc_empty[as], (acc_par >> as) & 1);
                acc_par ^= 1u << as;
                sm100::tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * TN;
#pragma unroll
                for (int kb = 0; kb < NKB; ++kb) {
                    sm100::mbar_wait(&sm.b_full[bs], b_par);
                    sm100::tc_fence_after();
                    if (sm100::elect_one()) {
                        const uint32_t al = a_lo + kb * (A_KB_BYTES >> 4), bl = b_lo0 + bs * (B_KB_BYTES >> 4);
#pragma unroll
                        for (int k = 0; k < KB_BYTES / 32; ++k)      // UMMA K = 32 bytes of int8; +2 = 32 B >> 4
                            sm100::mma_i8(d_tmem, sm100::smem_desc_join(al + 2 * k), sm100::smem_desc_join(bl + 2 * k),
                                          idesc, (kb | k) ? 1u : 0u);
                        sm100::mma_commit(&sm.b_empty[bs]);          // smem stage free once these MMAs retire
                        if (kb == NKB - 1) sm100::mma_commit(&sm.acc_full[as]);
                    }
                    __syncwarp();
                    if (++bs == NB) { bs = 0; b_par ^= 1; }
                }
                as ^= 1;
            }
            if (sm100::elect_one()) sm100::mma_commit(&sm.a_empty[ab]);
            __syncwarp();
            if (++ab == NABUF) { ab = 0; a_par ^= 1; }
        }
    } else if (warp == W_STAGE) {
        // ===================== column-constant stager (whole warp) =====================
        // Software pipelined: the norms of the next two tiles are in flight while one tile is written to the ring.
        struct TileIter { long long item; Item it; int j; bool ok; };
        auto advance = [&](TileIter& ti, bool first) {
            if (!first && ++ti.j < ti.it.ntiles) return;
            for (;;) {
                ti.item = first ? (long long)blockIdx.x : ti.item + gridDim.x;
                first = false;
                if (ti.item >= total_items) { ti.ok = false; return; }
                ti.it = decode_item(ti.item, pairs, tiles_max, S, backward);
                if (ti.it.valid && ti.it.ntiles > 0) { ti.j = 0; ti.ok = true; return; }
            }
        };
        auto load = [&](const TileIter& ti, int (&nt)[TN / 32]) {
            const int c0 = ti.it.c_begin + ti.j * TN;
#pragma unroll
            for (int k = 0; k < TN / 32; ++k) {
                int col = c0 + k * 32 + lane;
                nt[k] = (ti.ok && col < ti.it.c_end) ? __ldg(t_norm + ti.it.t_off + col) : -1;
            }
        };
        TileIter L; L.item = 0; L.j = 0; L.ok = false;
        advance(L, true);
        int q0[TN / 32], q1[TN / 32];
        bool v0 = L.ok;
        load(L, q0);
        if (L.ok) advance(L, false);
        bool v1 = L.ok;
        load(L, q1);
        if (L.ok) advance(L, false);
        uint32_t r = 0, par = 1;               // parity to wait for on cb_empty[r]
        while (v0) {
            sm100::mbar_wait(&sm.cb_empty[r], par);
            int* dst = sm.cb + r * TN;
#pragma unroll
            for (int k = 0; k < TN / 32; ++k)
                dst[k * 32 + lane] = q0[k] >= 0 ? q0[k] * CHUNK + ((k * 32 + lane) & (CHUNK - 1)) : MASK_CB;
            __syncwarp();
            if (lane == 0) sm100::mbar_arrive(&sm.cb_full[r]);   // release: the stores above are visible to waiters
            if (++r == CBR) { r = 0; par ^= 1; }
#pragma unroll
            for (int k = 0; k < TN / 32; ++k) q0[k] = q1[k];
            v0 = v1;
            v1 = L.ok;
            if (L.ok) { load(L, q1); advance(L, false); }
        }
    } else if (warp >= CTRL_WARPS) {
        // ===================== epilogue: two GROUPS of 8 warps on alternate tiles =====================
        // group g = (warp - 4) / 8 owns TMEM stage g and drains the tiles with (global tile counter & 1) == g;
        // inside a group: TMEM lane quarter = warp % 4, 128-column half of the tile = ((warp - 4) / 4) & 1.
        // While one group waits for its accumulator and its tcgen05.ld, the other group's math keeps the issue
        // slots busy (with all 16 warps on the same tile those latencies were exposed once per tile for everybody).
        const int e = warp - CTRL_WARPS;
        const int g = e >> 3, half = (e >> 2) & 1, quarter = warp & 3;
        const int slice = g * 2 + half;                                    // Top2 slice of this warp's partial result
        const uint32_t cb_s = sm100::smem_u32(sm.cb) + half * 2 * CHUNK * 4;
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + g * TN + half * 2 * CHUNK;
        uint32_t k = 0;                        // tiles this group has drained: acc_full[g] parity = k & 1,
                                               // constant ring slot = g + 2 * (k & 1), its parity = (k >> 1) & 1
        uint32_t tc_base = 0;                  // global tile counter at the start of the item
        for (long long item = blockIdx.x; item < total_items; item += gridDim.x) {
            Item it = decode_item(item, pairs, tiles_max, S, backward);
            if (!it.valid) continue;
            int v1 = 0x7fffffff, v2 = 0x7fffffff, i1 = -1, i2 = -1;
            auto insert = [&](int pm, int base) {
                // insert a 64-column chunk minimum.  Compare VALUES only (key >> 6): the low bits are the column inside
                // the chunk and must not let a later chunk win a tie; later chunks have larger columns.
                bool lt1 = pm < (v1 & ~(CHUNK - 1)), lt2 = pm < (v2 & ~(CHUNK - 1));
                v2 = lt1 ? v1 : (lt2 ? pm : v2);
                i2 = lt1 ? i1 : (lt2 ? base : i2);
                v1 = lt1 ? pm : v1;
                i1 = lt1 ? base : i1;
            };
            auto sub_min = [&](const uint32_t(&a32)[32], uint32_t cbv) -> int {      // min key of 32 columns
                int key[32];
#pragma unroll
                for (int q4 = 0; q4 < 8; ++q4) {
                    int4 c4 = sm100::lds128(cbv + q4 * 16);                  // warp-uniform address: broadcast
                    key[4 * q4 + 0] = c4.x - 2 * CHUNK * (int)a32[4 * q4 + 0];
                    key[4 * q4 + 1] = c4.y - 2 * CHUNK * (int)a32[4 * q4 + 1];
                    key[4 * q4 + 2] = c4.z - 2 * CHUNK * (int)a32[4 * q4 + 2];
                    key[4 * q4 + 3] = c4.w - 2 * CHUNK * (int)a32[4 * q4 + 3];
                }
                int a[11];
#pragma unroll
                for (int t = 0; t < 10; ++t) a[t] = min3(key[3 * t], key[3 * t + 1], key[3 * t + 2]);
                a[10] = min(key[30], key[31]);
                return min(min3(min3(a[0], a[1], a[2]), min3(a[3], a[4], a[5]), min3(a[6], a[7], a[8])), min(a[9], a[10]));
            };
            for (int j = (int)((g - tc_base) & 1); j < it.ntiles; j += 2, ++k) {
                const uint32_t slot = g + 2 * (k & 1);
                const uint32_t cbv = cb_s + slot * TN * 4;
                const int base = it.c_begin + j * TN + half * 2 * CHUNK;
                sm100::mbar_wait(&sm.acc_full[g], k & 1);
                sm100::tc_fence_after();
                uint32_t acc[2][32];
                sm100::tmem_ld32(taddr, acc[0]);
                sm100::tmem_ld32(taddr + 32, acc[1]);
                sm100::mbar_wait(&sm.cb_full[slot], (k >> 1) & 1);
                sm100::tmem_ld_wait();
                int m0 = sub_min(acc[0], cbv);
                sm100::tmem_ld32(taddr + 64, acc[0]);                        // overlaps the math below
                int m1 = sub_min(acc[1], cbv + 32 * 4);
                sm100::tmem_ld32(taddr + 96, acc[1]);
                insert(min(m0, m1), base);
                sm100::tmem_ld_wait();                                       // whole half-tile read:
                sm100::tc_fence_before();                                    // free the TMEM stage before the rest of the math
                __syncwarp();
                if (lane == 0) sm100::mbar_arrive(&sm.acc_empty[g]);
                m0 = sub_min(acc[0], cbv + 64 * 4);
                m1 = sub_min(acc[1], cbv + 96 * 4);
                __syncwarp();
                if (lane == 0) sm100::mbar_arrive(&sm.cb_empty[slot]);       // constants consumed
                insert(min(m0, m1), base + CHUNK);
            }
            tc_base += (uint32_t)it.ntiles;
            // publish this warp's partial top-2 (Top2 slice = split * NSLICE + slice); finalize merges the slices
            const int row = it.r0 + quarter * 32 + lane;
            if (row < it.n) {
                const int nq = __ldg(q_norm + it.q_off + row);
                auto mk = [&](int v, int i) -> u64 {
                    int val = v >> 6;                                    // CHUNK == 64
                    return (i < 0 || val >= INVALID_VAL) ? B200M_SENTINEL
                                                         : (((u64)(u32)(val + nq)) << 32) | (u64)(u32)(i + (v & (CHUNK - 1)));
                };
                Top2 t; t.k1 = mk(v1, i1); t.k2 = mk(v2, i2);
                out[(long long)(it.split * NSLICE + slice) * total_rows + it.row_base + row] = t;
            }
        }
    }

    sm100::tc_fence_before();
    __syncthreads();
    if (warp == W_MMA) sm100::tmem_dealloc<512>(tmem_base);
}

// =====================================================================================================
// Folded-norm variant (MODE_RATIO): a fifth K step of 32 int8 dims carries -(||b'||^2 >> 1) + bias into the
// accumulator (common.cuh, prep.cu), so the epilogue is a bare 3-input MAX tree over the TMEM values: no
// column constants, no IMAD, no shared-memory loads.  Per 16-column chunk only the maximum E = a'.b' - h + bias
// survives (fold chunks are B200M_FOLD_CHUNK = 16 columns); the running top-2 of chunk maxima (value, chunk base) leaves the SM.  tail.cu recomputes the
// winner's chunk exactly (dp4a) for the rows whose ratio test is not already decided by the bounds
//   d1 >= na + 2 (bias - E1),   d2 <= na + 1 + 2 (bias - E2).
constexpr int X_BYTES = TN * B200M_FOLD_K;       // 8 KiB: fold digits of one train tile (SWIZZLE_32B rows)
constexpr int AX_BYTES = TM * B200M_FOLD_K;      // 4 KiB: the constant query-side block
static_assert(B200M_FOLD_CHUNK == 16, "the epilogue reduces 16-column chunks");
constexpr int NBF = 4;
constexpr int FOLD_EPI_WARPS = EPI_WARPS;
constexpr int FOLD_THREADS = (CTRL_WARPS + FOLD_EPI_WARPS) * 32;                           // train K-block ring depth of the fold kernel

struct SmemFold {
    uint8_t *a, *b, *x, *ax;
    int4* xch;            // [2][EPI_WARPS - 4][32] partial top-2s of the non-merging epilogue warps, double buffered by item parity
    uint64_t *a_full, *a_empty, *b_full, *b_empty, *acc_full, *acc_empty;
    uint32_t* tmem_ptr;
};
// pair == true (cta_group::2): a CTA holds only ITS half of every train tile (128 rows), so a ring stage is half as
// large and the ring twice as deep.
__host__ __device__ constexpr int nbf_of(bool pair) { return pair ? 2 * NBF : NBF; }
__device__ __forceinline__ SmemFold carve_fold(uint8_t* base, int nkb, bool pair = false) {
    const int nbf = nbf_of(pair), bstage = pair ? B_KB_BYTES / 2 : B_KB_BYTES, xstage = pair ? X_BYTES / 2 : X_BYTES;
    SmemFold s;
    s.a = base;
    s.b = s.a + NABUF * nkb * A_KB_BYTES;
    s.x = s.b + nbf * bstage;
    s.ax = s.x + (nbf / nkb) * xstage;
    s.xch = (int4*)(s.ax + AX_BYTES);
    uint8_t* p = (uint8_t*)(s.xch + 2 * (EPI_WARPS - 4) * 32);
    s.a_full = (uint64_t*)p; p += NABUF * 8;
    s.a_empty = (uint64_t*)p; p += NABUF * 8;
    s.b_full = (uint64_t*)p; p += nbf * 8;
    s.b_empty = (uint64_t*)p; p += nbf * 8;
    s.acc_full = (uint64_t*)p; p += 2 * 8;
    s.acc_empty = (uint64_t*)p; p += 2 * 8;
    s.tmem_ptr = (uint32_t*)p;
    return s;
}
__host__ __device__ constexpr size_t smem_bytes_fold(int nkb, bool pair = false) {
    return 1024 + (size_t)NABUF * nkb * A_KB_BYTES + (size_t)nbf_of(pair) * (pair ? B_KB_BYTES / 2 : B_KB_BYTES) +
           (size_t)(nbf_of(pair) / nkb) * (pair ? X_BYTES / 2 : X_BYTES) + AX_BYTES +
           2 * (EPI_WARPS - 4) * 32 * sizeof(int4) + (2 * NABUF + 2 * nbf_of(pair) + 4) * 8 + 16;
}

static_assert(smem_bytes_fold(1) <= 232448 && smem_bytes_fold(2) <= 232448 && smem_bytes_fold(1, true) <= 232448 &&
              smem_bytes_fold(2, true) <= 232448, "fold kernel shared memory");

__device__ __forceinline__ int max3(int a, int b, int c) { return max(max(a, b), c); }
// maximum of 16 consecutive accumulator columns v[O .. O+15] = one fold chunk: 8 three-input max instructions
template <int O>
__device__ __forceinline__ int max16(const uint32_t (&v)[32]) {
    int a0 = max3((int)v[O + 0], (int)v[O + 1], (int)v[O + 2]), a1 = max3((int)v[O + 3], (int)v[O + 4], (int)v[O + 5]);
    int a2 = max3((int)v[O + 6], (int)v[O + 7], (int)v[O + 8]), a3 = max3((int)v[O + 9], (int)v[O + 10], (int)v[O + 11]);
    int a4 = max3((int)v[O + 12], (int)v[O + 13], (int)v[O + 14]);
    return max(max3(a0, a1, a2), max3(a3, a4, (int)v[O + 15]));
}
template <int FILL>
__device__ __forceinline__ void mask32(uint32_t (&v)[32], int lim) {     // columns >= lim do not exist
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = i < lim ? v[i] : (uint32_t)FILL;
}
// E * kmul + low on the FMA pipe (IMAD); kmul arrives as a kernel argument so that ptxas cannot turn it into an ALU-pipe LEA
__device__ __forceinline__ int pack_key(int e, int kmul, int low) {
    int r;
    asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(r) : "r"(e), "r"(kmul), "r"(low));
    return r;
}

// CS = CTAs per cluster (1 or 2).  With CS == 2 the two CTAs of a cluster work on two adjacent query row tiles of
// the same pair, so they stream the SAME train tiles: each CTA fetches half of every train tile (128 of the 256 rows,
// and half of the fold digits) and TMA multicasts it into both CTAs' shared memory -- L2 -> SM read traffic per CTA
// halves.  A ring stage may be refilled only when BOTH CTAs' MMAs have consumed it: the stage's empty barrier counts
// CS arrivals and every MMA issuer commits to the barrier of both CTAs (tcgen05.commit ... multicast::cluster).
// Work items are (pair, split, tile pair) over clusters; the MMAs themselves stay single-CTA (cta_group::1).
//
// PAIR (needs CS == 2) turns the cluster into a tcgen05 CTA pair: ONE cta_group::2 MMA of M = 256 covers both CTAs' query
// row tiles, and each CTA supplies only its own 128 rows of the 256-row train tile -- nothing is multicast, a CTA's shared
// memory sees half the TMA writes (20 instead of 40 KiB per tile) and two thirds of the operand reads (40 instead of
// 60 KiB).  That traffic, 100 KiB per 640-cycle tile against 128 B/clk, was what held the multicast version at ~790
// cycles per tile (sm__mem_tensor_cycles_active 82 %, profiles/r01_l2i8_v9_fold_cluster_ncu_full.md); the issue rate of
// the pair instruction is the same 128 cycles (profiles/r02_mma_pair_issue_rate.txt).  The even CTA (leader) issues
// every MMA; both CTAs' TMA loads report to the leader's "full" barriers, the leader's commits are multicast to both
// CTAs' "empty" barriers, and both CTAs' epilogue warps hand TMEM stages back on the leader's acc_empty barriers.
template <int NKB, int CS, bool PAIR = false>
__global__ void __launch_bounds__(FOLD_THREADS, 1)
l2_i8_fold_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_t,
                  const __grid_constant__ CUtensorMap tmap_x, const int32_t* __restrict__ fold_flag,
                  const DevPair* __restrict__ pairs, long long total_items, int tiles_max, int S,
                  Top2* __restrict__ out, long long total_rows, int kmul) {
    // packed running top-2 of the epilogue: index field wide enough for one window of train tiles, value field for |E|
    // (|a'.b'| <= 2^21 per 128-byte K block, |h - bias| < 2^19: |E| < 2^22 for D = 128, < 2^23 for D = 256)
    constexpr int IDX_BITS = NKB == 1 ? 9 : 8, IDX_MASK = (1 << IDX_BITS) - 1, EMIN = -(1 << (31 - IDX_BITS));
    constexpr int WIN_SHIFT = IDX_BITS - 4, WIN_TILES = 1 << WIN_SHIFT;     // (2^IDX_BITS chunks) * 16 columns / TN
    static_assert(TN / B200M_FOLD_CHUNK == 16, "window arithmetic assumes 16 chunks per tile");
    if (__ldg(fold_flag) == 0) return;             // digits do not cover this store (the call's snapshot): the keyed kernel runs instead
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    static_assert(!PAIR || CS == 2, "a CTA pair is a cluster of two");
    const SmemFold sm = carve_fold(smem, NKB, PAIR);
    constexpr int NBFK = nbf_of(PAIR);                                  // train K-block ring depth
    constexpr int B_STAGE = PAIR ? B_KB_BYTES / 2 : B_KB_BYTES, X_STAGE = PAIR ? X_BYTES / 2 : X_BYTES;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int W_MMA = 0, W_PROD = 1, W_CONST = 2;
    constexpr int backward = 0;
    constexpr uint16_t MC_MASK = (1u << CS) - 1;
    const uint32_t crank = CS > 1 ? sm100::cluster_ctarank() : 0;
    // work items: total_items counts (pair, split, tile group of CS row tiles); this CTA takes row tile CS * group + crank
    const long long first_item = CS > 1 ? (long long)sm100::cluster_id_x() : (long long)blockIdx.x;
    const long long item_stride = CS > 1 ? (long long)sm100::cluster_count_x() : (long long)gridDim.x;
    auto decode = [&](long long item) -> Item {
        int tg = (int)(item % tiles_max); item /= tiles_max;     // tiles_max = tile groups per (pair, split)
        Item it = decode_tile((int)(item / S), (int)(item % S), tg * CS + (int)crank, pairs, S, backward);
        // a CTA whose own row tile lies past the pair still streams its half of the train tiles for its peer
        it.valid = tg * CS * TM < it.n;
        return it;
    };

    if (warp == W_PROD && lane == 0) {
        sm100::tma_prefetch_desc(&tmap_q);
        sm100::tma_prefetch_desc(&tmap_t);
        sm100::tma_prefetch_desc(&tmap_x);
        for (int i = 0; i < NABUF; ++i) { sm100::mbar_init(&sm.a_full[i], 1); sm100::mbar_init(&sm.a_empty[i], 1); }
        // PAIR: one multicast commit of the leader frees a stage in both CTAs; a TMEM stage is free again once the 16
        // epilogue warps of BOTH CTAs have drained it
        for (int i = 0; i < NBFK; ++i) { sm100::mbar_init(&sm.b_full[i], 1); sm100::mbar_init(&sm.b_empty[i], PAIR ? 1 : CS); }
        for (int i = 0; i < 2; ++i) { sm100::mbar_init(&sm.acc_full[i], 1); sm100::mbar_init(&sm.acc_empty[i], (PAIR ? 2 : 1) * EPI_WARPS); }
        sm100::fence_barrier_init();
    }
    if (warp == W_CONST) {
        // query side of the fold block: every row is c = {-128 x15, -1} x2.  Both 16-byte halves of a row are equal,
        // so the tile is invariant under the 16-byte chunk permutation of SWIZZLE_32B (and so is sum c.t).
        const uint32_t ax = sm100::smem_u32(sm.ax);
        for (int w = lane; w < AX_BYTES / 4; w += 32) sm100::sts32(ax + 4 * w, (w & 3) == 3 ? (int)0xFF808080 : (int)0x80808080);
        sm100::fence_proxy_async();                // generic-proxy stores -> visible to tcgen05.mma (async proxy)
    }
    if constexpr (PAIR) {
        sm100::cluster_sync_all();                         // both CTAs of the pair are running before the paired allocation
        if (warp == W_MMA) sm100::tmem_alloc_2sm<512>(sm.tmem_ptr);
    } else {
        if (warp == W_MMA) sm100::tmem_alloc<512>(sm.tmem_ptr);
    }
    sm100::tc_fence_before();
    __syncthreads();
    if constexpr (CS > 1) sm100::cluster_sync_all();      // the peer's barriers exist before anything is multicast to them
    sm100::tc_fence_after();
    const uint32_t tmem_base = *sm.tmem_ptr;
    constexpr int HALF_ROWS = TN / CS;                     // train rows this CTA fetches per tile

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
                    sm100::mbar_wait(&sm.b_empty[bs], b_par);          // CS arrivals: every CTA of the cluster is done with it
                    if (sm100::elect_one()) {
                        const bool last = kb == NKB - 1;       // the fold digits travel with the tile's last K block
                        uint8_t* bdst = sm.b + (size_t)bs * B_STAGE + (PAIR ? 0 : (size_t)crank * (B_KB_BYTES / CS));
                        uint8_t* xdst = sm.x + (size_t)(bs / NKB) * X_STAGE + (PAIR ? 0 : (size_t)crank * (X_BYTES / CS));
                        if constexpr (PAIR) {          // this CTA's 128 train rows stay here; both halves count on the leader's barrier
                            if (crank == 0)
                                sm100::mbar_arrive_expect_tx(&sm.b_full[bs], (uint32_t)(2 * (B_STAGE + (last ? X_STAGE : 0))));
                            const uint32_t lbar = sm100::mapa_u32(sm100::smem_u32(&sm.b_full[bs]), 0);
                            sm100::tma_load_2d_2sm(bdst, &tmap_t, lbar, kb * KB_BYTES, row0 + j * TN);
                            if (last) sm100::tma_load_2d_2sm(xdst, &tmap_x, lbar, 0, row0 + j * TN);
                        } else if constexpr (CS > 1) {
                            sm100::mbar_arrive_expect_tx(&sm.b_full[bs], (uint32_t)(B_KB_BYTES + (last ? X_BYTES : 0)));
                            sm100::tma_load_2d_mc(bdst, &tmap_t, &sm.b_full[bs], kb * KB_BYTES, row0 + j * TN, MC_MASK);
                            if (last) sm100::tma_load_2d_mc(xdst, &tmap_x, &sm.b_full[bs], 0, row0 + j * TN, MC_MASK);
                        } else {
                            sm100::mbar_arrive_expect_tx(&sm.b_full[bs], (uint32_t)(B_KB_BYTES + (last ? X_BYTES : 0)));
                            sm100::tma_load_2d(bdst, &tmap_t, &sm.b_full[bs], kb * KB_BYTES, row0 + j * TN);
                            if (last) sm100::tma_load_2d(xdst, &tmap_x, &sm.b_full[bs], 0, row0 + j * TN);
                        }
                    }
                    __syncwarp();
                    if (++bs == NBFK) { bs = 0; b_par ^= 1; }
                }
            }
        }
    } else if (warp == W_MMA && (!PAIR || crank == 0)) {
        constexpr uint32_t idesc = sm100::make_idesc(/*S32*/ 2, /*s8*/ 1, /*s8*/ 1, PAIR ? 2 * TM : TM, TN);
        const uint32_t a_lo0 = sm100::smem_desc_lo_sw128(sm100::smem_u32(sm.a));
        const uint32_t b_lo0 = sm100::smem_desc_lo_sw128(sm100::smem_u32(sm.b));
        const uint32_t x_lo0 = sm100::smem_desc_lo_sw128(sm100::smem_u32(sm.x));      // same low-word format
        const uint64_t ax_desc = sm100::smem_desc_join_hi(sm100::SMEM_DESC_HI_SW32,
                                                          sm100::smem_desc_lo_sw128(sm100::smem_u32(sm.ax)));
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
                const uint32_t d_tmem = tmem_base + as * TN;
#pragma unroll
                for (int kb = 0; kb < NKB; ++kb) {
                    sm100::mbar_wait(&sm.b_full[bs], b_par);
                    sm100::tc_fence_after();
                    if (sm100::elect_one()) {
                        const uint32_t al = a_lo + kb * (A_KB_BYTES >> 4), bl = b_lo0 + bs * (B_STAGE >> 4);
                        const uint64_t x_desc = sm100::smem_desc_join_hi(sm100::SMEM_DESC_HI_SW32, x_lo0 + (bs / NKB) * (X_STAGE >> 4));
#pragma unroll
                        for (int k = 0; k < KB_BYTES / 32; ++k) {
                            if constexpr (PAIR)
                                sm100::mma_i8_2sm(d_tmem, sm100::smem_desc_join(al + 2 * k), sm100::smem_desc_join(bl + 2 * k),
                                                  idesc, (kb | k) ? 1u : 0u);
                            else
                                sm100::mma_i8(d_tmem, sm100::smem_desc_join(al + 2 * k), sm100::smem_desc_join(bl + 2 * k),
                                              idesc, (kb | k) ? 1u : 0u);
                        }
                        if (kb == NKB - 1) {                    // + c . t = -(h - bias)
                            if constexpr (PAIR) sm100::mma_i8_2sm(d_tmem, ax_desc, x_desc, idesc, 1u);
                            else sm100::mma_i8(d_tmem, ax_desc, x_desc, idesc, 1u);
                        }
                        // ONE commit per K block: it frees the ring stage (in both CTAs of a cluster) and, for the tile's last
                        // K block, is also what the epilogue waits on -- every tcgen05.commit drains the MMA pipe for ~40 cycles
                        // (tools/mma_peak), so a separate "accumulator full" commit per tile would cost ~5 % of the tensor time.
                        if constexpr (PAIR) sm100::mma_commit_2sm_mc(&sm.b_empty[bs], MC_MASK);
                        else if constexpr (CS > 1) sm100::mma_commit_mc(&sm.b_empty[bs], MC_MASK);
                        else sm100::mma_commit(&sm.b_empty[bs]);
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
        // 32-column tcgen05.ld issued back to back, so no load waits for arithmetic and a TMEM stage is held only for the
        // time its 128 KiB take to cross the 64 B/clk/SMSP read port (~512 cycles) plus the wake-up.  (Two groups of 8
        // warps on alternate tiles, each warp re-using its registers for four loads per tile, held a stage ~960 cycles:
        // with two stages the MMA of tile t+2 then starts 640 + 960 after the MMA of tile t -- 800 cycles per tile.)
        const int e = warp - CTRL_WARPS;
        const int cb = e >> 2, quarter = warp & 3;                 // cb: 64-column block of the tile
        constexpr int CBW = TN / 4, CB_CHUNKS = CBW / B200M_FOLD_CHUNK;
        uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + cb * CBW;     // this tile's stage: toggles by TN columns
        uint32_t acc_bar = PAIR ? sm100::mapa_u32(sm100::smem_u32(&sm.acc_empty[0]), 0) : sm100::smem_u32(&sm.acc_empty[0]);
        uint32_t ready_bar = sm100::smem_u32(&sm.b_empty[0]);
        uint32_t as8 = 0, rs = (NKB - 1) * 8, rp = 0, nitem = 0;   // acc_empty[as8 / 8]; "ready" = b_empty[rs / 8], parity rp
        int lowc0 = IDX_MASK - cb * CB_CHUNKS, lowc1 = lowc0 - 1, lowc2 = lowc0 - 2, lowc3 = lowc0 - 3, m16 = -(TN / B200M_FOLD_CHUNK);
        asm volatile("" : "+r"(taddr), "+r"(acc_bar), "+r"(ready_bar));
        asm volatile("" : "+r"(lowc0), "+r"(lowc1), "+r"(lowc2), "+r"(lowc3), "+r"(m16));
        for (long long item = first_item; item < total_items; item += item_stride) {
            Item it = decode(item);
            if (!it.valid) continue;
            int v1 = (int)0x80000000, v2 = (int)0x80000000, i1 = -1, i2 = -1;
            auto insert = [&](int pm, int base) {           // top-2 of chunk maxima; ties keep the earlier chunk
                bool gt1 = pm > v1, gt2 = pm > v2;
                v2 = gt1 ? v1 : (gt2 ? pm : v2);
                i2 = gt1 ? i1 : (gt2 ? base : i2);
                v1 = gt1 ? pm : v1;
                i1 = gt1 ? base : i1;
            };
            // Inside a window of WIN_TILES train tiles the running top-2 lives in ONE word per entry:
            //   key = E * 2^IDX_BITS + (2^IDX_BITS - 1 - chunk index in the window),   |E| < 2^(31 - IDX_BITS),
            // so an update is min/max only (2.5 ALU-pipe instructions per chunk instead of 7; the packing is one IMAD on
            // the otherwise idle FMA pipe).  At a window's end the two keys are unpacked into (value, column) and merged
            // as before, so any m works.
            int k1 = (int)0x80000000, k2 = (int)0x80000000;
            int win = 0;
            auto flush = [&]() {
                auto put = [&](int key) {
                    const int ev = key >> IDX_BITS;
                    if (ev != EMIN) insert(ev, it.c_begin + win * (WIN_TILES * TN) + (IDX_MASK - (key & IDX_MASK)) * B200M_FOLD_CHUNK);
                };
                put(k1); put(k2);                              // k1 first: on equal values the earlier chunk stays ahead
                k1 = k2 = (int)0x80000000;
            };
            auto push2 = [&](int ka, int kb) {               // top-2 of {k1, k2, ka, kb}: 5 min/max
                const int mx = max(ka, kb), mn = min(ka, kb);
                const int t = min(k1, mx);
                k1 = max(k1, mx);
                k2 = max3(t, k2, mn);
            };
            // Per tile the loop carries only: the TMEM address (toggling between the two stages), the "ready" barrier's
            // address and parity, the columns left for this warp, and the tile's position in the window.  Everything
            // thread-constant was made opaque above so that ptxas keeps it in a register instead of re-deriving it from
            // %tid every tile (the ALU pipe is the epilogue's budget: 2 cycles per instruction per sub-partition).
            int rem = it.c_end - it.c_begin - cb * CBW;          // existing columns from this warp's first one on
            for (int j0 = 0; j0 < it.ntiles; j0 += WIN_TILES, ++win) {
                const int jn = min(it.ntiles - j0, WIN_TILES);
                for (int jw = 0; jw < jn; ++jw) {
                    // accumulator ready == the ring stage of this tile's last K block has been released (see the MMA warp)
                    sm100::mbar_wait_addr(ready_bar + rs, rp);
                    rs += NKB * 8;
                    if (rs >= (uint32_t)(NBFK * 8)) { rs -= NBFK * 8; rp ^= 1; }
                    sm100::tc_fence_after();
                    uint32_t acc[2][32];
                    sm100::tmem_ld32(taddr, acc[0]);
                    sm100::tmem_ld32(taddr + 32, acc[1]);
                    sm100::tmem_ld_wait();                    // everything is in registers: hand the TMEM stage back first
                    sm100::tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        if constexpr (PAIR) sm100::mbar_arrive_cluster(acc_bar + as8);
                        else sm100::mbar_arrive_addr(acc_bar + as8);
                    }
                    taddr ^= TN; as8 ^= 8;
                    if (rem < CBW) { mask32<EMIN>(acc[0], rem); mask32<EMIN>(acc[1], rem - 32); }     // tail tile
                    rem -= TN;
                    const int m0 = max16<0>(acc[0]), m1 = max16<16>(acc[0]);
                    const int m2 = max16<0>(acc[1]), m3 = max16<16>(acc[1]);
                    // chunk index inside the window, counted down from the top of the index field: (low_c) - 16 jw, on the FMA pipe
                    push2(pack_key(m0, kmul, pack_key(jw, m16, lowc0)), pack_key(m1, kmul, pack_key(jw, m16, lowc1)));
                    push2(pack_key(m2, kmul, pack_key(jw, m16, lowc2)), pack_key(m3, kmul, pack_key(jw, m16, lowc3)));
                }
                flush();
            }
            // The four warps of a TMEM lane quarter hold partial top-2s of the same 32 query rows (4 column blocks):
            // merge them through shared memory so that ONE Top2 per row and split leaves the SM.
            int4* xbuf = sm.xch + (size_t)(nitem & 1) * (EPI_WARPS - 4) * 32;
            ++nitem;
            const bool merger = e < 4;                      // cb == 0
            if (!merger) xbuf[(e - 4) * 32 + lane] = make_int4(v1, i1, v2, i2);
            asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32) : "memory");     // epilogue warps only
            if (merger) {
#pragma unroll
                for (int w = 1; w < 4; ++w) {
                    const int4 o = xbuf[((w - 1) * 4 + quarter) * 32 + lane];
                    if (o.y >= 0) insert(o.x, o.y);
                    if (o.w >= 0) insert(o.z, o.w);
                }
                const int row = it.r0 + quarter * 32 + lane;
                if (row < it.n) {
                    auto mk = [&](int v, int i) -> u64 {
                        return (i < 0 || v == (int)0x80000000) ? B200M_SENTINEL
                                                               : (((u64)(u32)(B200M_FOLD_VBIAS - v)) << 32) | (u64)(u32)i;
                    };
                    Top2 t; t.k1 = mk(v1, i1); t.k2 = mk(v2, i2);
                    out[(long long)it.split * total_rows + it.row_base + row] = t;
                }
            }
        }
    }

    sm100::tc_fence_before();
    __syncthreads();
    if constexpr (CS > 1) sm100::cluster_sync_all();      // the peer may still commit to this CTA's barriers until it is done too
    if (warp == W_MMA) {
        if constexpr (PAIR) sm100::tmem_dealloc_2sm<512>(tmem_base);
        else sm100::tmem_dealloc<512>(tmem_base);
    }
}

}  // namespace

// ---------------------------------------------------------------- host side
static b200m_encode_tiled_fn get_encode_fn() {
    static b200m_encode_tiled_fn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = (b200m_encode_tiled_fn)p;
    }
    return fn;
}

int b200m_make_tmap_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t row_bytes, uint32_t box_rows,
                       CUtensorMapDataType dtype, uint32_t elem_bytes) {
    return b200m_make_tmap_2d_ex(out, base, rows, row_bytes, box_rows, dtype, elem_bytes, 128, CU_TENSOR_MAP_SWIZZLE_128B);
}

int b200m_make_tmap_2d_ex(CUtensorMap* out, const void* base, uint64_t rows, uint64_t row_bytes, uint32_t box_rows,
                          CUtensorMapDataType dtype, uint32_t elem_bytes, uint32_t inner_bytes, CUtensorMapSwizzle swz) {
    b200m_encode_tiled_fn enc = get_encode_fn();
    if (!enc) return b200m_set_cuda_error(cudaErrorNotSupported, "cuTensorMapEncodeTiled entry point");
    if (((uintptr_t)base & 15) || (row_bytes & 15) || rows == 0) return B200M_EINVAL;
    cuuint64_t gdim[2] = {row_bytes / elem_bytes, rows};
    cuuint64_t gstride[1] = {row_bytes};
    cuuint32_t box[2] = {inner_bytes / elem_bytes, box_rows};  // inner box = one swizzle span
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(out, dtype, 2, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return b200m_set_cuda_error(cudaErrorInvalidValue, "cuTensorMapEncodeTiled");
    return B200M_OK;
}

int launch_l2_i8_top2(const int8_t* q_desc, const int32_t* q_norm, int64_t q_rows, const int8_t* t_desc,
                      const int32_t* t_norm, int64_t t_rows, int D, const DevPair* pairs, int num_pairs, int max_n,
                      int max_m, int S, bool backward, const int32_t* fold_flag, Top2* out, long long total_rows, cudaStream_t st) {
    bool allow_fold = fold_flag != nullptr;        // MatchPlan::flags: [0] = this call's snapshot of "fold digits usable"
    if (D % KB_BYTES != 0) return B200M_EINVAL;
    const int nkb = D / KB_BYTES;
    if (nkb > 2) return B200M_EUNSUPPORTED;      // d^2 must stay < 2^24 (exact in fp32, as in cv2) and keys in int32
    if (backward) {                              // roles swapped: train rows look for their best query row
        const int8_t* d = q_desc; q_desc = t_desc; t_desc = d;
        const int32_t* nn = q_norm; q_norm = t_norm; t_norm = nn;
        int64_t r = q_rows; q_rows = t_rows; t_rows = r;
        int mx = max_n; max_n = max_m; max_m = mx;
        allow_fold = false;
    }
    CUtensorMap tq, tt;
    int rc = b200m_make_tmap_2d(&tq, q_desc, (uint64_t)q_rows, (uint64_t)D, TM, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1);
    if (rc != B200M_OK) return rc;
    rc = b200m_make_tmap_2d(&tt, t_desc, (uint64_t)t_rows, (uint64_t)D, TN, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1);
    if (rc != B200M_OK) return rc;
    const size_t smem = smem_bytes(nkb);
    static B200mPerDevice once;                      // val[0] = CTAs per cluster of the fold kernel (2 when the device co-schedules pairs),
    int dev = 0, sms = 0;                            // val[1] = resident 2-CTA clusters
    if (!b200m_current_device(&dev, &sms)) return B200M_ENODEVICE;
    if (!once.done[dev]) {
        int fold_cs = 1, fold_clusters = 0;
        B200M_CUDA_CHECK(cudaFuncSetAttribute(l2_i8_top2_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)smem_bytes(1)), "l2_i8 smem attr");
        B200M_CUDA_CHECK(cudaFuncSetAttribute(l2_i8_top2_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)smem_bytes(2)), "l2_i8 smem attr");
        B200M_CUDA_CHECK(cudaFuncSetAttribute(l2_i8_fold_kernel<1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)smem_bytes_fold(1)), "l2_i8 fold smem attr");
        B200M_CUDA_CHECK(cudaFuncSetAttribute(l2_i8_fold_kernel<2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)smem_bytes_fold(2)), "l2_i8 fold smem attr");
        B200M_CUDA_CHECK(cudaFuncSetAttribute(l2_i8_fold_kernel<1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)smem_bytes_fold(1)), "l2_i8 fold smem attr");
        B200M_CUDA_CHECK(cudaFuncSetAttribute(l2_i8_fold_kernel<2, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)smem_bytes_fold(2)), "l2_i8 fold smem attr");
        B200M_CUDA_CHECK(cudaFuncSetAttribute(l2_i8_fold_kernel<1, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)smem_bytes_fold(1, true)), "l2_i8 fold smem attr");
        B200M_CUDA_CHECK(cudaFuncSetAttribute(l2_i8_fold_kernel<2, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)smem_bytes_fold(2, true)), "l2_i8 fold smem attr");
        // How many 2-CTA clusters of the fold kernel can be resident at once?  Use them only if they cover every SM.
        cudaLaunchConfig_t qc = {};
        cudaLaunchAttribute qa[1];
        qa[0].id = cudaLaunchAttributeClusterDimension;
        qa[0].val.clusterDim.x = 2; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
        qc.gridDim = dim3((unsigned)(sms & ~1)); qc.blockDim = dim3(FOLD_THREADS); qc.dynamicSmemBytes = smem_bytes_fold(1, true);
        qc.attrs = qa; qc.numAttrs = 1;
        int nc = 0;
        if (cudaOccupancyMaxActiveClusters(&nc, l2_i8_fold_kernel<1, 2, true>, &qc) == cudaSuccess && nc * 2 >= sms &&
            !getenv("B200M_NO_CLUSTER")) {
            fold_cs = 2;
            fold_clusters = nc;
        } else {
            cudaGetLastError();
        }
        if (getenv("B200M_VERBOSE")) fprintf(stderr, "b200match: device %d: fold kernel clusters of %d (max active 2-CTA clusters %d, SMs %d)\n", dev, fold_cs, nc, sms);
        once.val[dev][0] = fold_cs; once.val[dev][1] = fold_clusters;
        once.val[dev][2] = getenv("B200M_NO_PAIR") ? 0 : 1;          // A/B switch: multicast clusters of single-CTA MMAs instead of CTA pairs
        once.done[dev] = 1;
    }
    const int fold_cs = once.val[dev][0], fold_clusters = once.val[dev][1];
    int tiles_max = (max_n + TM - 1) / TM;
    long long total_items = (long long)num_pairs * S * tiles_max;
    if (total_items <= 0) return B200M_OK;
    unsigned grid = (unsigned)(total_items < sms ? total_items : sms);
    // Device-side dispatch (no host sync): with allow_fold both kernels are enqueued and exactly one of them
    // finds its precondition in t_info[3]; the other returns at once.
    const int32_t* fold_info = allow_fold ? fold_flag : nullptr;
    if (nkb == 1)
        l2_i8_top2_kernel<1><<<grid, THREADS, smem, st>>>(tq, tt, q_norm, t_norm, pairs, total_items, tiles_max, S,
                                                          backward ? 1 : 0, out, total_rows, fold_info);
    else
        l2_i8_top2_kernel<2><<<grid, THREADS, smem, st>>>(tq, tt, q_norm, t_norm, pairs, total_items, tiles_max, S,
                                                          backward ? 1 : 0, out, total_rows, fold_info);
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "l2_i8_top2_kernel launch");
    if (allow_fold) {
        const int cs = fold_cs;
        CUtensorMap tth, tx;                          // per-CTA share of a train tile: TN / cs rows
        rc = b200m_make_tmap_2d(&tth, t_desc, (uint64_t)t_rows, (uint64_t)D, TN / cs, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1);
        if (rc != B200M_OK) return rc;
        rc = b200m_make_tmap_2d_ex(&tx, (const char*)t_norm + b200m_aux_fold_off(t_rows), (uint64_t)t_rows, B200M_FOLD_K, TN / cs,
                                   CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, B200M_FOLD_K, CU_TENSOR_MAP_SWIZZLE_32B);
        if (rc != B200M_OK) return rc;
        const int groups = (tiles_max + cs - 1) / cs;               // tile groups of cs row tiles per (pair, split)
        const long long fitems = (long long)num_pairs * S * groups;
        cudaLaunchConfig_t cfg = {};
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = (unsigned)cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        const long long max_units = cs == 2 ? fold_clusters : sms;
        cfg.gridDim = dim3((unsigned)((fitems < max_units ? fitems : max_units) * cs));
        cfg.blockDim = dim3(FOLD_THREADS);
        const bool pair = cs == 2 && once.val[dev][2];
        cfg.dynamicSmemBytes = smem_bytes_fold(nkb, pair);
        cfg.stream = st;
        cfg.attrs = at; cfg.numAttrs = 1;
        const int kmul = nkb == 1 ? 512 : 256;        // 2^IDX_BITS of the kernel's packed top-2 keys
        cudaError_t e;
        if (pair)
            e = nkb == 1 ? cudaLaunchKernelEx(&cfg, l2_i8_fold_kernel<1, 2, true>, tq, tth, tx, fold_flag, pairs, fitems, groups, S, out, total_rows, kmul)
                         : cudaLaunchKernelEx(&cfg, l2_i8_fold_kernel<2, 2, true>, tq, tth, tx, fold_flag, pairs, fitems, groups, S, out, total_rows, kmul);
        else if (cs == 2)
            e = nkb == 1 ? cudaLaunchKernelEx(&cfg, l2_i8_fold_kernel<1, 2>, tq, tth, tx, fold_flag, pairs, fitems, groups, S, out, total_rows, kmul)
                         : cudaLaunchKernelEx(&cfg, l2_i8_fold_kernel<2, 2>, tq, tth, tx, fold_flag, pairs, fitems, groups, S, out, total_rows, kmul);
        else
            e = nkb == 1 ? cudaLaunchKernelEx(&cfg, l2_i8_fold_kernel<1, 1>, tq, tth, tx, fold_flag, pairs, fitems, groups, S, out, total_rows, kmul)
                         : cudaLaunchKernelEx(&cfg, l2_i8_fold_kernel<2, 1>, tq, tth, tx, fold_flag, pairs, fitems, groups, S, out, total_rows, kmul);
        b200m_count_launch();
        B200M_CUDA_CHECK(e, "l2_i8_fold_kernel launch");
    }
    return B200M_OK;
}
