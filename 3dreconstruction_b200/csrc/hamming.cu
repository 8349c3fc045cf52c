// hamming.cu — NORM_HAMMING knn (k=2) on the INT pipe: 128-bit loads, shared-memory tiles, __popc,
// fused per-row top-2.  Replaces cv2.BFMatcher(NORM_HAMMING).knnMatch(k=2)/match as called at
// /root/reference/FeatureMatchingExtraction/feature_matchers.py:210,227 (ORB/AKAZE/BRISK rows).
//
// Roofline: POPC issues at 16 lanes/clk/SM, a D-byte descriptor pair costs D/4 POPC, so the bound
// is 148 * 16 * f_clk / (D/4) distances/s (582 Gdist/s for 256-bit ORB at 1.965 GHz).  Everything
// else (XOR, IADD3, top-2) runs on the 64-lane ALU pipe underneath that.
//
// CTA = 256 threads = 16 (tx: train columns) x 16 (ty: query rows); thread tile 8 x 8, CTA tile
// 128 query rows x 128 train columns; train tiles stream through shared memory.  Shared layout is
// [16-byte chunk][row] so that a quarter-warp's LDS.128 touches consecutive 16-byte words
// (conflict free) and the query operand is a warp-half broadcast.
#include "common.cuh"

namespace {

constexpr int TM = 128, TN = 128, KB = 8;   // KB: 16-byte chunks per K-block (128 bytes)
constexpr int IDX_BITS = 19;                // key32 = dist << 19 | column relative to split start
constexpr u32 IDX_MASK = (1u << IDX_BITS) - 1u;

// Copy rows [row0, row0+rows_valid) x 16-byte chunks [c0, c0+nc) of a [*, D]-byte matrix into
// smem[chunk][row]; bytes past D and rows past rows_valid are zero (adds 0 to every distance).
__device__ __forceinline__ void load_tile(uint4* __restrict__ smem, const uint8_t* __restrict__ base,
                                          long long row0, int rows_valid, int D, int c0, int nc,
                                          bool fast16) {
    for (int idx = threadIdx.x; idx < TM * nc; idx += blockDim.x) {
        int row = idx / nc, ch = idx - row * nc;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (row < rows_valid) {
            const uint8_t* p = base + (row0 + row) * (long long)D + (long long)(c0 + ch) * 16;
            int rem = D - (c0 + ch) * 16;
            if (fast16 && rem >= 16) {
                v = __ldg(reinterpret_cast<const uint4*>(p));
            } else if (rem > 0) {
                u32 w[4] = {0, 0, 0, 0};
                int nb = rem < 16 ? rem : 16;
                for (int b = 0; b < nb; ++b) w[b >> 2] |= (u32)__ldg(p + b) << ((b & 3) * 8);
                v = make_uint4(w[0], w[1], w[2], w[3]);
            }
        }
        smem[ch * TM + row] = v;
    }
}

// COLMIN: also reduce every train COLUMN's best query row (key = dist << 19 | query row, ties -> lowest row)
// into colmin[] with atomicMin, so that the mutual cross-check (feature_matchers.py:225-238 with
// crossCheck=True) needs ONE pass over the N x M distances instead of a second, role-swapped pass.
template <bool COLMIN>
__global__ void __launch_bounds__(256)
hamming_top2_kernel(const uint8_t* __restrict__ q_desc, const uint8_t* __restrict__ t_desc, int D,
                    const DevPair* __restrict__ pairs, int tiles_max, int S, int backward,
                    Top2* __restrict__ out, long long total_rows, u32* __restrict__ colmin) {
    __shared__ uint4 sq[KB * TM];
    __shared__ uint4 st[KB * TN];
    __shared__ u32 scol[TN];
    if (COLMIN && threadIdx.x < TN) scol[threadIdx.x] = 0xFFFFFFFFu;

    // work item = (pair, split, row tile)
    int item = blockIdx.x;
    int tile = item % tiles_max; item /= tiles_max;
    int split = item % S;
    int p = item / S;
    DevPair pr = pairs[p];
    const uint8_t* qb = q_desc; const uint8_t* tb = t_desc;
    int q_off = pr.q_off, n = pr.n, t_off = pr.t_off, m = pr.m;
    long long row_base = pr.fwd_off;
    if (backward) {   // roles swapped: rows of the train image look for their best query row
        qb = t_desc; tb = q_desc; q_off = pr.t_off; n = pr.m; t_off = pr.q_off; m = pr.n;
        row_base = pr.bwd_off;
    }
    int r0 = tile * TM;
    if (r0 >= n) return;
    // this CTA's column range
    int cols_per = ((m + S - 1) / S + TN - 1) / TN * TN;
    int c_begin = split * cols_per;
    int c_end = min(m, c_begin + cols_per);

    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int nchunks = (D + 15) >> 4;
    const bool fast16 = (D % 16 == 0) && ((((uintptr_t)qb | (uintptr_t)tb) & 15) == 0);
    const bool single_kb = nchunks <= KB;

    u32 k1[8], k2[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) { k1[r] = 0xFFFFFFFFu; k2[r] = 0xFFFFFFFFu; }

    if (single_kb) load_tile(sq, qb, (long long)q_off + r0, min(TM, n - r0), D, 0, nchunks, fast16);

    for (int c0 = c_begin; c0 < c_end; c0 += TN) {
        int acc[8][8];
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int c = 0; c < 8; ++c) acc[r][c] = 0;

        for (int kb0 = 0; kb0 < nchunks; kb0 += KB) {
            int nc = min(KB, nchunks - kb0);
            __syncthreads();   // previous tile fully consumed
            if (!single_kb) load_tile(sq, qb, (long long)q_off + r0, min(TM, n - r0), D, kb0, nc, fast16);
            load_tile(st, tb, (long long)t_off + c0, min(TN, c_end - c0), D, kb0, nc, fast16);
            __syncthreads();
            for (int ch = 0; ch < nc; ++ch) {
                uint4 qa[8];
#pragma unroll
                for (int r = 0; r < 8; ++r) qa[r] = sq[ch * TM + ty * 8 + r];
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    uint4 b = st[ch * TN + tx + 16 * c];
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        acc[r][c] += __popc(qa[r].x ^ b.x) + __popc(qa[r].y ^ b.y);
                        acc[r][c] += __popc(qa[r].z ^ b.z) + __popc(qa[r].w ^ b.w);
                    }
                }
            }
        }
        if (COLMIN) {
            // column minima of this 128 x 128 tile: 8 rows in registers, the partner half-warp by shuffle, the 8
            // warps through shared-memory atomics, then one global atomicMin per column
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                u32 best = 0xFFFFFFFFu;
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    int row = r0 + ty * 8 + r;
                    u32 key = row < n ? (((u32)acc[r][c] << IDX_BITS) | (u32)row) : 0xFFFFFFFFu;
                    best = min(best, key);
                }
                best = min(best, __shfl_xor_sync(0xffffffffu, best, 16));
                if ((threadIdx.x & 16) == 0) atomicMin(&scol[tx + 16 * c], best);
            }
            __syncthreads();
            if (threadIdx.x < TN) {
                int col = c0 + threadIdx.x;
                if (col < c_end) atomicMin(&colmin[pr.bwd_off + col], scol[threadIdx.x]);
                scol[threadIdx.x] = 0xFFFFFFFFu;
            }
        }
        // fused top-2 (ties -> lowest column, because the column index sits in the low key bits)
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            int col = c0 + tx + 16 * c;
            u32 rel = (u32)(col - c_begin);
            bool ok = col < c_end;
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                u32 key = ok ? (((u32)acc[r][c] << IDX_BITS) | rel) : 0xFFFFFFFFu;
                top2_insert32(k1[r], k2[r], key);
            }
        }
    }

    // merge the 16 tx lanes that share a query row (they sit in one half-warp)
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
        for (int off = 1; off < 16; off <<= 1) {
            u32 o1 = __shfl_xor_sync(0xffffffffu, k1[r], off);
            u32 o2 = __shfl_xor_sync(0xffffffffu, k2[r], off);
            u32 hi = max(k1[r], o1);
            k1[r] = min(k1[r], o1);
            k2[r] = min(min(k2[r], o2), hi);
        }
        int row = r0 + ty * 8 + r;
        if (tx == 0 && row < n) {
            Top2 t;
            t.k1 = k1[r] == 0xFFFFFFFFu ? B200M_SENTINEL
                                         : ((u64)(k1[r] >> IDX_BITS) << 32) | (u64)(c_begin + (k1[r] & IDX_MASK));
            t.k2 = k2[r] == 0xFFFFFFFFu ? B200M_SENTINEL
                                         : ((u64)(k2[r] >> IDX_BITS) << 32) | (u64)(c_begin + (k2[r] & IDX_MASK));
            out[(long long)split * total_rows + row_base + row] = t;
        }
    }
}

}  // namespace

int launch_hamming_top2(const uint8_t* q_desc, const uint8_t* t_desc, int D, const DevPair* pairs,
                        int num_pairs, int max_n, int max_m, int S, bool backward, Top2* out,
                        long long total_rows, u32* colmin, cudaStream_t stream) {
    if (D * 8 >= (1 << (32 - IDX_BITS))) return B200M_EUNSUPPORTED;       // distance must fit 13 bits
    int max_rows = backward ? max_m : max_n;
    int max_cols = backward ? max_n : max_m;
    if (((max_cols + S - 1) / S + TN) > (1 << IDX_BITS)) return B200M_EUNSUPPORTED;
    int tiles_max = (max_rows + TM - 1) / TM;
    long long grid = (long long)num_pairs * S * tiles_max;
    if (grid <= 0) return B200M_OK;
    if (grid > 0x7fffffffLL) return B200M_EUNSUPPORTED;
    if (colmin) {
        if (backward || max_n > (1 << IDX_BITS)) return B200M_EINVAL;    // query row must fit the key's index bits
        hamming_top2_kernel<true><<<(unsigned)grid, 256, 0, stream>>>(q_desc, t_desc, D, pairs, tiles_max, S, 0, out,
                                                                      total_rows, colmin);
    } else {
        hamming_top2_kernel<false><<<(unsigned)grid, 256, 0, stream>>>(q_desc, t_desc, D, pairs, tiles_max, S,
                                                                       backward ? 1 : 0, out, total_rows, nullptr);
    }
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "hamming_top2_kernel launch");
    return B200M_OK;
}
