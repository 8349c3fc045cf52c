// tail.cu — everything after the knn stage, on the device:
//   finalize : merge column-split partial top-2s, convert to cv2's float32 distance, apply the
//              mode test of EnhancedBFMatcher.match
//              (/root/reference/FeatureMatchingExtraction/feature_matchers.py:212-224 ratio test in
//              fp64 as Python evaluates `m.distance < ratio * n.distance`; :225-238 cross-check ==
//              strict mutual first-argmin; or plain 1-NN)
//   sort_emit: the reference's stable `sort(key=score)` (:241) == order by (distance, queryIdx);
//              one CTA per pair runs a bitonic network over 64-bit keys
//              (float32 distance bits << 32 | queryIdx); rejected rows carry a sentinel and sink.
//   knn2_emit: raw k=2 neighbours for knnMatch-level parity checks.
#include "common.cuh"

namespace {

__device__ __forceinline__ float value_to_distance(int kind, u32 v) {
    if (kind == B200M_KIND_HAMMING) return (float)v;                     // exact integer as float
    if (kind == B200M_KIND_L2_I8) return __fsqrt_rn((float)v);          // v < 2^24: exact -> IEEE sqrt
    return __fsqrt_rn(__uint_as_float(v));                              // fp32 squared distance bits
}

__device__ __forceinline__ Top2 merge_splits(const Top2* __restrict__ arr, long long total, int S, long long row) {
    Top2 r = arr[row];
    for (int s = 1; s < S; ++s) {
        Top2 o = arr[(long long)s * total + row];
        top2_insert(r.k1, r.k2, o.k1);
        top2_insert(r.k1, r.k2, o.k2);
    }
    return r;
}


// Exact runner-up for the tensor-core L2_I8 kernel (see l2_i8.cu header): k2 so far is the best among
// all chunks other than the winner's 64-column chunk; recompute the winner's chunk exactly (int8 dp4a,
// the whole warp works on one query row at a time, one lane per column) and fold it in.  `need` is this
// lane's row predicate: in MODE_RATIO a row is refined only if it passes the ratio test with the
// candidate (the true runner-up is never larger, so a row that fails with the candidate fails anyway).
// Must be called by all 32 lanes of a warp whose rows belong to the same pair.
__device__ __forceinline__ void refine_i8_warp(bool need, int i, Top2& t, const DevPair& pr, const RefineI8& rf) {
    const int lane = threadIdx.x & 31;
    unsigned mask = __ballot_sync(0xffffffffu, need);
    while (mask) {
        const int src = __ffs(mask) - 1;
        mask &= mask - 1;
        const int ri = __shfl_sync(0xffffffffu, i, src);
        const int j1 = __shfl_sync(0xffffffffu, (int)(u32)t.k1, src);
        const int4* a = reinterpret_cast<const int4*>(rf.q_desc + (long long)(pr.q_off + ri) * rf.D);
        const int nq = __ldg(rf.q_norm + pr.q_off + ri);
        u64 key = B200M_SENTINEL;
#pragma unroll
        for (int h = 0; h < B200M_I8_CHUNK / 32; ++h) {
            const int col = (j1 & ~(B200M_I8_CHUNK - 1)) + h * 32 + lane;
            if (col < pr.m && col != j1) {
                const int4* b = reinterpret_cast<const int4*>(rf.t_desc + (long long)(pr.t_off + col) * rf.D);
                int dot = 0;
                for (int k = 0; k < rf.D / 16; ++k) {
                    int4 x = __ldg(a + k), y = __ldg(b + k);
                    dot = __dp4a(x.x, y.x, dot); dot = __dp4a(x.y, y.y, dot);
                    dot = __dp4a(x.z, y.z, dot); dot = __dp4a(x.w, y.w, dot);
                }
                int d2i = nq + __ldg(rf.t_norm + pr.t_off + col) - 2 * dot;
                u64 k = ((u64)(u32)d2i << 32) | (u64)(u32)col;
                key = k < key ? k : key;
            }
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            u64 o = __shfl_xor_sync(0xffffffffu, key, off);
            key = o < key ? o : key;
        }
        if (lane == src && key < t.k2) t.k2 = key;
    }
}


// ---- exact int8 distances for the folded-norm kernel's results (l2_i8.cu): whole-warp helpers -------------
__device__ __forceinline__ void warp_merge_top2(u64& k1, u64& k2) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        u64 o1 = __shfl_xor_sync(0xffffffffu, k1, off), o2 = __shfl_xor_sync(0xffffffffu, k2, off);
        top2_insert(k1, k2, o1);
        top2_insert(k1, k2, o2);
    }
}
// exact top-2 of query row ri over train columns [c0, c1) of its pair (every lane returns the same pair of keys).
// Eight lanes share a column (16 bytes of the 128-byte row each: a load instruction reads four whole rows,
// fully coalesced), partial dot products meet by three butterfly shuffles.  D == 128 * nkb, nkb <= 2.
__device__ __forceinline__ void warp_exact_top2_i8(int ri, int c0, int c1, const DevPair& pr, const RefineI8& rf,
                                                   u64& k1, u64& k2) {
    const int lane = threadIdx.x & 31, part = lane & 7, sub = lane >> 3;
    const int8_t* qrow = rf.q_desc + (long long)(pr.q_off + ri) * rf.D;
    const int nq = __ldg(rf.q_norm + pr.q_off + ri);
    const int nkb = rf.D >> 7;
    int4 a0 = __ldg(reinterpret_cast<const int4*>(qrow) + part), a1 = make_int4(0, 0, 0, 0);
    if (nkb > 1) a1 = __ldg(reinterpret_cast<const int4*>(qrow + 128) + part);
    k1 = B200M_SENTINEL; k2 = B200M_SENTINEL;
    for (int cb = c0; cb < c1; cb += 4) {
        const int col = cb + sub;
        const bool ok = col < c1;
        const int8_t* trow = rf.t_desc + (long long)(pr.t_off + (ok ? col : c0)) * rf.D;
        int4 y = __ldg(reinterpret_cast<const int4*>(trow) + part);
        int dot = __dp4a(a0.x, y.x, 0);
        dot = __dp4a(a0.y, y.y, dot); dot = __dp4a(a0.z, y.z, dot); dot = __dp4a(a0.w, y.w, dot);
        if (nkb > 1) {
            y = __ldg(reinterpret_cast<const int4*>(trow + 128) + part);
            dot = __dp4a(a1.x, y.x, dot); dot = __dp4a(a1.y, y.y, dot);
            dot = __dp4a(a1.z, y.z, dot); dot = __dp4a(a1.w, y.w, dot);
        }
        dot += __shfl_xor_sync(0xffffffffu, dot, 4);
        dot += __shfl_xor_sync(0xffffffffu, dot, 2);
        dot += __shfl_xor_sync(0xffffffffu, dot, 1);
        if (ok && part == 0) {
            const int d2i = nq + __ldg(rf.t_norm + pr.t_off + col) - 2 * dot;
            top2_insert(k1, k2, ((u64)(u32)d2i << 32) | (u64)(u32)col);
        }
    }
    warp_merge_top2(k1, k2);
}

// The same for one fold chunk [c0, c0 + B200M_FOLD_CHUNK) clipped to the pair: the chunk's B200M_FOLD_CHUNK / 4 row
// loads are issued back to back (rows past the end re-read row c0 and are dropped), so a row costs one memory round trip.
__device__ __forceinline__ void warp_exact_top2_chunk_i8(int ri, int c0, const DevPair& pr, const RefineI8& rf,
                                                         u64& k1, u64& k2) {
    constexpr int IT = B200M_FOLD_CHUNK / 4;
    const int lane = threadIdx.x & 31, part = lane & 7, sub = lane >> 3;
    const int8_t* qrow = rf.q_desc + (long long)(pr.q_off + ri) * rf.D;
    const int nkb = rf.D >> 7;
    int4 a0 = __ldg(reinterpret_cast<const int4*>(qrow) + part), a1 = make_int4(0, 0, 0, 0);
    if (nkb > 1) a1 = __ldg(reinterpret_cast<const int4*>(qrow + 128) + part);
    int4 y0[IT], y1[IT];
    int nt[IT];
#pragma unroll
    for (int t = 0; t < IT; ++t) {
        const int col = c0 + 4 * t + sub;
        const int8_t* trow = rf.t_desc + (long long)(pr.t_off + (col < pr.m ? col : c0)) * rf.D;
        y0[t] = __ldg(reinterpret_cast<const int4*>(trow) + part);
        y1[t] = nkb > 1 ? __ldg(reinterpret_cast<const int4*>(trow + 128) + part) : make_int4(0, 0, 0, 0);
        nt[t] = __ldg(rf.t_norm + pr.t_off + (col < pr.m ? col : c0));
    }
    const int nq = __ldg(rf.q_norm + pr.q_off + ri);
    k1 = B200M_SENTINEL; k2 = B200M_SENTINEL;
#pragma unroll
    for (int t = 0; t < IT; ++t) {
        int dot = __dp4a(a0.x, y0[t].x, 0);
        dot = __dp4a(a0.y, y0[t].y, dot); dot = __dp4a(a0.z, y0[t].z, dot); dot = __dp4a(a0.w, y0[t].w, dot);
        dot = __dp4a(a1.x, y1[t].x, dot); dot = __dp4a(a1.y, y1[t].y, dot);
        dot = __dp4a(a1.z, y1[t].z, dot); dot = __dp4a(a1.w, y1[t].w, dot);
        dot += __shfl_xor_sync(0xffffffffu, dot, 4);
        dot += __shfl_xor_sync(0xffffffffu, dot, 2);
        dot += __shfl_xor_sync(0xffffffffu, dot, 1);
        const int col = c0 + 4 * t + sub;
        if (col < pr.m && part == 0) top2_insert(k1, k2, ((u64)(u32)(nq + nt[t] - 2 * dot) << 32) | (u64)(u32)col);
    }
    warp_merge_top2(k1, k2);
}

__device__ __forceinline__ bool ratio_pass_i8(double ratio, int v1, int v2) {      // the test of finalize, on exact d^2
    float c1 = __fsqrt_rn((float)v1), c2 = __fsqrt_rn((float)v2);
    return (double)c1 < ratio * (double)c2;
}

// MODE_RATIO on fold-kernel results.  t holds the top-2 CHUNKS by E = a'.b' - h + bias (value = VBIAS - E, index =
// first column of the 16-column chunk).  With p = ||b'||^2 & 1 of the column, d^2 = nq + p + 2 (bias - E), so a chunk
// value brackets its best column's distance within {lb, lb + 1}.  Rows are decided, in this order, by
//   (1) bounds only: d1 >= lb(E1), d2 <= lb(E2) + 1 and the test is monotone -> most rows fail here, untouched;
//   (2) exact recomputation of the winner's chunk (E1 > E2: the best column is in it): d1 exact, and d2 is either
//       the chunk's exact runner-up (if <= lb(E2)) or one of {lb(E2), lb(E2) + 1} -- enough unless the test
//       flips between those two;
//   (3) an exhaustive exact scan of the row (chunk-level tie E1 == E2, or the flip in (2)): rare.
// Returns keep; t.k1 becomes the exact (d^2 << 32 | column) key of the best neighbour when keep is true.
// Must be called by all 32 lanes of a warp whose rows belong to the same pair.
__device__ __forceinline__ bool fold_ratio_warp(bool has, int i, Top2& t, double ratio, const DevPair& pr,
                                                const RefineI8& rf) {
    const int lane = threadIdx.x & 31;
    const int bias = __ldg(rf.t_info + 2);
    int nq = 0, lb1 = 0, lb2 = 0;
    bool need = false;
    const bool other = t.k2 != B200M_SENTINEL;
    if (has) {
        nq = __ldg(rf.q_norm + pr.q_off + i);
        lb1 = nq + 2 * (bias - (B200M_FOLD_VBIAS - (int)(u32)(t.k1 >> 32)));
        if (other) {
            lb2 = nq + 2 * (bias - (B200M_FOLD_VBIAS - (int)(u32)(t.k2 >> 32)));
            need = ratio_pass_i8(ratio, max(lb1, 0), lb2 + 1);
        } else {
            need = true;
        }
    }
    u64 a1 = B200M_SENTINEL, a2 = B200M_SENTINEL;
    unsigned mask = __ballot_sync(0xffffffffu, need);
    while (mask) {
        const int src = __ffs(mask) - 1;
        mask &= mask - 1;
        const int ri = __shfl_sync(0xffffffffu, i, src);
        const int c0 = __shfl_sync(0xffffffffu, (int)(u32)t.k1, src);
        u64 x1, x2;
        warp_exact_top2_chunk_i8(ri, c0, pr, rf, x1, x2);
        if (lane == src) { a1 = x1; a2 = x2; }
    }
    bool keep = false, full = false;
    if (need) {
        const int d1 = (int)(u32)(a1 >> 32);
        const bool has2 = a2 != B200M_SENTINEL;
        const int r2 = (int)(u32)(a2 >> 32);
        if (!other) {
            keep = has2 && ratio_pass_i8(ratio, d1, r2);       // a single chunk: everything is exact already
        } else if (lb1 == lb2) {
            full = true;                                        // two chunks tie on E: the winner may be in either (or a third)
        } else if (has2 && r2 <= lb2) {
            keep = ratio_pass_i8(ratio, d1, r2);
        } else if (ratio_pass_i8(ratio, d1, max(lb2, 0))) {
            keep = true;                                        // passes with the smallest possible runner-up
        } else if (ratio_pass_i8(ratio, d1, lb2 + 1)) {
            full = true;                                        // outcome depends on the parity of the other chunk's best column
        }
    }
    mask = __ballot_sync(0xffffffffu, full);
    while (mask) {
        const int src = __ffs(mask) - 1;
        mask &= mask - 1;
        const int ri = __shfl_sync(0xffffffffu, i, src);
        u64 x1, x2;
        warp_exact_top2_i8(ri, 0, pr.m, pr, rf, x1, x2);
        if (lane == src) {
            a1 = x1;
            keep = x2 != B200M_SENTINEL && ratio_pass_i8(ratio, (int)(u32)(x1 >> 32), (int)(u32)(x2 >> 32));
        }
    }
    if (keep) t.k1 = a1;
    return keep;
}

__global__ void __launch_bounds__(256)
finalize_kernel(int kind, int mode, double ratio, const DevPair* __restrict__ pairs, int blocks_per_pair,
                const Top2* __restrict__ fwd, long long total_n, const Top2* __restrict__ bwd,
                long long total_m, int S, u64* __restrict__ sortbuf, int32_t* __restrict__ tbest,
                int32_t* __restrict__ counts, RefineI8 rf, const u32* __restrict__ colmin, int rpw) {
    // rpw = rows per warp: 32 (one per lane), or fewer for calls with a handful of pairs -- the exact recomputations of the
    // fold path are whole-warp work done one row after the other (a memory round trip each), so a warp owning 32 rows of
    // which half need one is a 16-deep latency chain while most SMs have nothing to run: lanes >= rpw own no row and only
    // help (one SIFT 8192 x 8192 call: 34 -> ~10 us).
    int p = blockIdx.x / blocks_per_pair;
    const int lane_ = threadIdx.x & 31;
    int i = ((blockIdx.x - p * blocks_per_pair) * (blockDim.x >> 5) + (threadIdx.x >> 5)) * rpw + lane_;
    DevPair pr = pairs[p];
    const bool active = lane_ < rpw && i < pr.n;
    Top2 t; t.k1 = B200M_SENTINEL; t.k2 = B200M_SENTINEL;
    const bool fold = kind == B200M_KIND_L2_I8 && mode == B200M_MODE_RATIO && rf.fold_flag && __ldg(rf.fold_flag);
    // S counts Top2 slices: splits x B200M_I8_GROUPS for the keyed L2_I8 kernel, splits x B200M_FOLD_SLICES when the fold kernel ran
    if (active) t = merge_splits(fwd, total_n, fold ? S / B200M_I8_GROUPS * B200M_FOLD_SLICES : S, pr.fwd_off + i);
    bool keep_fold = false;
    if (fold) {                                                      // grid-uniform branch
        keep_fold = fold_ratio_warp(t.k1 != B200M_SENTINEL, i, t, ratio, pr, rf);
    } else if (kind == B200M_KIND_L2_I8 && mode == B200M_MODE_RATIO) {      // block-uniform branch
        bool need = false;
        if (t.k1 != B200M_SENTINEL) {
            need = ratio <= 0.0 || t.k2 == B200M_SENTINEL;
            if (!need) {
                float c1 = value_to_distance(B200M_KIND_L2_I8, (u32)(t.k1 >> 32));
                float c2 = value_to_distance(B200M_KIND_L2_I8, (u32)(t.k2 >> 32));
                need = (double)c1 < ratio * (double)c2;
            }
        }
        refine_i8_warp(need, i, t, pr, rf);
    }
    u64 key = B200M_SENTINEL;
    int32_t tb = -1;
    if (t.k1 != B200M_SENTINEL) {
        u32 j1 = (u32)t.k1;
        float d1 = value_to_distance(kind, (u32)(t.k1 >> 32));
        bool keep;
        if (fold) {
            keep = keep_fold;
        } else if (mode == B200M_MODE_RATIO) {
            keep = false;
            if (t.k2 != B200M_SENTINEL) {            // cv2 returns a single neighbour when m == 1
                float d2 = value_to_distance(kind, (u32)(t.k2 >> 32));
                keep = (double)d1 < ratio * (double)d2;
            }
        } else if (mode == B200M_MODE_CROSS) {
            if (colmin) {          // fused Hamming pass: packed (dist << 19 | best query row) per train row
                keep = (colmin[pr.bwd_off + j1] & ((1u << B200M_HAMMING_IDX_BITS) - 1u)) == (u32)i;
            } else {
                Top2 b = merge_splits(bwd, total_m, S, pr.bwd_off + j1);
                keep = (u32)b.k1 == (u32)i;
            }
        } else {
            keep = true;
        }
        tb = (int32_t)j1;
        if (keep) key = ((u64)__float_as_uint(d1) << 32) | (u64)(u32)i;
    }
    if (active) {
        sortbuf[pr.sort_off + i] = key;
        tbest[pr.fwd_off + i] = tb;
    }
    if (counts) {                                   // CSR mode: survivors per pair (a block never spans pairs);
        int c = __popc(__ballot_sync(0xffffffffu, key != B200M_SENTINEL));   // per warp: no block barrier, warps retire alone
        if ((threadIdx.x & 31) == 0 && c) atomicAdd(&counts[p], c);
    }
}

// CSR mode: exclusive scan of the per-pair survivor counts -> tight output offsets.
__global__ void __launch_bounds__(1024)
scan_offsets_kernel(DevPair* __restrict__ pairs, const int32_t* __restrict__ counts, int num_pairs,
                    long long* __restrict__ out_ptr) {
    __shared__ long long warp_sums[32];
    __shared__ long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int base = 0; base < num_pairs; base += blockDim.x) {
        int p = base + threadIdx.x;
        long long v = p < num_pairs ? counts[p] : 0, x = v;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            long long o = __shfl_up_sync(0xffffffffu, x, off);
            if (lane >= off) x += o;
        }
        if (lane == 31) warp_sums[w] = x;
        __syncthreads();
        if (w == 0) {
            long long y = lane < nw ? warp_sums[lane] : 0;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                long long o = __shfl_up_sync(0xffffffffu, y, off);
                if (lane >= off) y += o;
            }
            warp_sums[lane] = y;
        }
        __syncthreads();
        long long excl = carry + (w ? warp_sums[w - 1] : 0) + x - v;
        if (p < num_pairs) { pairs[p].out_off = excl; out_ptr[p] = excl; }
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) out_ptr[num_pairs] = carry;
}

__device__ __forceinline__ void bitonic(u64* a, int npad) {
    for (int k = 2; k <= npad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < (npad >> 1); t += blockDim.x) {
                int lo = ((t & ~(j - 1)) << 1) | (t & (j - 1));   // index with bit j clear
                int hi = lo | j;
                u64 x = a[lo], y = a[hi];
                bool up = (lo & k) == 0;
                if ((x > y) == up) { a[lo] = y; a[hi] = x; }
            }
            __syncthreads();
        }
    }
}

// Slot of this lane's element in a block-wide compaction: one atomicAdd per warp on the shared counter (thousands of
// single-lane atomics on one shared-memory word serialise).  Every lane of the warp must call.
__device__ __forceinline__ int warp_compact_slot(bool live, int* counter) {
    const unsigned m = __ballot_sync(0xffffffffu, live);
    const int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == 0 && m) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    return base + __popc(m & ((1u << lane) - 1u));
}

// one match (key = distance bits << 32 | query row) to output position o, in the caller's record format
__device__ __forceinline__ void emit_match(u64 k, long long o, const int32_t* __restrict__ tbest_pair, int32_t* __restrict__ out_q,
                                           int32_t* __restrict__ out_t, float* __restrict__ out_d, int rec_format,
                                           void* __restrict__ out_rec) {
    const int q = (int)(u32)k;
    const int tr = tbest_pair[q];
    const u32 dbits = (u32)(k >> 32);
    if (rec_format == 0) {                               // three arrays (b200m_match_pairs / _csr)
        out_q[o] = q;
        out_t[o] = tr;
        out_d[o] = __uint_as_float(dbits);
    } else if (rec_format == B200M_REC_Q32T32D32) {      // one record per match (b200m_match_pairs_rec)
        int32_t* r = (int32_t*)out_rec + 3 * o;
        r[0] = q; r[1] = tr; r[2] = (int32_t)dbits;
    } else if (rec_format == B200M_REC_Q16T16D32) {
        ((uint2*)out_rec)[o] = make_uint2((u32)q | ((u32)tr << 16), dbits);
    } else {                                             // B200M_REC_Q16T16D16: Hamming bit counts are small integers
        unsigned short* r = (unsigned short*)out_rec + 3 * o;
        r[0] = (unsigned short)q; r[1] = (unsigned short)tr; r[2] = (unsigned short)(int)__uint_as_float(dbits);
    }
}

__global__ void __launch_bounds__(1024)
sort_emit_kernel(const DevPair* __restrict__ pairs, u64* __restrict__ sortbuf,
                 const int32_t* __restrict__ tbest, int smem_keys, int32_t* __restrict__ out_q,
                 int32_t* __restrict__ out_t, float* __restrict__ out_d, int32_t* __restrict__ out_count,
                 const int32_t* __restrict__ counts, int lo_count, int hi_count, int rec_format, void* __restrict__ out_rec) {
    // This launch handles the pairs with lo_count < survivors <= hi_count (two launches with different block shapes
    // split the pairs between them); `counts` (CSR mode: finalize's survivor counts) lets the others leave at once.
    extern __shared__ u64 skeys[];
    __shared__ int s_count;
    if (counts) {
        const int c = counts[blockIdx.x];
        if (c <= lo_count || c > hi_count) return;
    }
    DevPair pr = pairs[blockIdx.x];
    u64* g = sortbuf + pr.sort_off;
    const int n = pr.n, npad = pr.npad;
    if (threadIdx.x == 0) s_count = 0;
    __syncthreads();
    u64* a;
    int sort_n;
    // compact the survivors into shared memory (order is irrelevant before the sort): the bitonic network then runs
    // over pow2ceil(#survivors) keys instead of pow2ceil(n), and the block needs only a small shared-memory slice
    for (int t0 = 0; t0 < n; t0 += blockDim.x) {
        const int t = t0 + threadIdx.x;
        const u64 k = t < n ? g[t] : B200M_SENTINEL;
        const int slot = warp_compact_slot(k != B200M_SENTINEL, &s_count);
        if (k != B200M_SENTINEL && slot < smem_keys) skeys[slot] = k;
    }
    __syncthreads();
    if (s_count <= lo_count || s_count > hi_count) return;
    if (s_count <= smem_keys) {
        a = skeys;
        sort_n = 1;
        while (sort_n < s_count) sort_n <<= 1;
        for (int t = s_count + threadIdx.x; t < sort_n; t += blockDim.x) a[t] = B200M_SENTINEL;
        __syncthreads();
    } else {                                            // more survivors than the slice holds: sort in place in global memory
        a = g;
        sort_n = npad;
        for (int t = n + threadIdx.x; t < npad; t += blockDim.x) a[t] = B200M_SENTINEL;
        __syncthreads();
    }
    if (sort_n > 1) bitonic(a, sort_n);
    __syncthreads();
    const int cnt = s_count;
    for (int t = threadIdx.x; t < cnt; t += blockDim.x)
        emit_match(a[t], pr.out_off + t, tbest + pr.fwd_off, out_q, out_t, out_d, rec_format, out_rec);
    if (threadIdx.x == 0) out_count[blockIdx.x] = cnt;
}

// ---- a handful of pairs (the plugin's one-pair call): one CTA per pair leaves the sort to a single SM -- 84 us of a
// 150 us ORB 10k x 10k call (tools/orb_pair_probe.py), more than both tensor passes.  SPLIT_G CTAs per pair instead:
// each compacts and sorts one segment of the pair's keys (sort_segments_kernel), then every CTA stages all SPLIT_G sorted
// runs in shared memory and places ITS run's keys by rank: position = index in its own run + sum over the other runs of
// the number of smaller keys (keys are unique: distance bits << 32 | query row), found by binary search.
constexpr int SPLIT_G = 8;

__global__ void __launch_bounds__(1024)
sort_segments_kernel(const DevPair* __restrict__ pairs, u64* __restrict__ sortbuf, int32_t* __restrict__ segcnt) {
    extern __shared__ u64 skeys[];
    __shared__ int s_count;
    const int p = blockIdx.x / SPLIT_G, sg = blockIdx.x - p * SPLIT_G;
    const DevPair pr = pairs[p];
    const int L = max(pr.npad / SPLIT_G, 1);                 // segment length (npad is a power of two)
    u64* g = sortbuf + pr.sort_off + (size_t)sg * L;
    const int lim = min(L, pr.n - sg * L);                   // rows of the pair inside this segment (may be <= 0)
    if (threadIdx.x == 0) s_count = 0;
    __syncthreads();
    for (int t0 = 0; t0 < lim; t0 += blockDim.x) {            // whole warps iterate together: one shared-memory atomic per warp
        const int t = t0 + threadIdx.x;
        const u64 k = t < lim ? g[t] : B200M_SENTINEL;
        const int slot = warp_compact_slot(k != B200M_SENTINEL, &s_count);
        if (k != B200M_SENTINEL) skeys[slot] = k;
    }
    __syncthreads();
    const int c = s_count;
    int sort_n = 1;
    while (sort_n < c) sort_n <<= 1;
    for (int t = c + threadIdx.x; t < sort_n; t += blockDim.x) skeys[t] = B200M_SENTINEL;
    __syncthreads();
    if (sort_n > 1) bitonic(skeys, sort_n);
    __syncthreads();
    for (int t = threadIdx.x; t < c; t += blockDim.x) g[t] = skeys[t];      // the sorted run, in place
    if (threadIdx.x == 0) segcnt[blockIdx.x] = c;
}

__global__ void __launch_bounds__(1024)
merge_emit_kernel(const DevPair* __restrict__ pairs, const u64* __restrict__ sortbuf, const int32_t* __restrict__ segcnt,
                  const int32_t* __restrict__ tbest, int32_t* __restrict__ out_q, int32_t* __restrict__ out_t,
                  float* __restrict__ out_d, int32_t* __restrict__ out_count, int rec_format, void* __restrict__ out_rec) {
    extern __shared__ u64 skeys[];                           // the pair's SPLIT_G runs, run r at r * L
    __shared__ int cnts[SPLIT_G];
    const int p = blockIdx.x / SPLIT_G, sg = blockIdx.x - p * SPLIT_G;
    const DevPair pr = pairs[p];
    const int L = max(pr.npad / SPLIT_G, 1);
    const u64* g = sortbuf + pr.sort_off;
    if (threadIdx.x < SPLIT_G) cnts[threadIdx.x] = segcnt[p * SPLIT_G + threadIdx.x];
    __syncthreads();
    for (int r = 0; r < SPLIT_G; ++r)
        for (int t = threadIdx.x; t < cnts[r]; t += blockDim.x) skeys[r * L + t] = g[(size_t)r * L + t];
    __syncthreads();
    const int mine = cnts[sg];
    for (int x = threadIdx.x; x < mine; x += blockDim.x) {
        const u64 k = skeys[sg * L + x];
        int rank = x;
#pragma unroll
        for (int r = 0; r < SPLIT_G; ++r) {
            if (r == sg) continue;
            const u64* run = skeys + r * L;
            int lo = 0, hi = cnts[r];                        // number of keys < k in run r
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (run[mid] < k) lo = mid + 1; else hi = mid;
            }
            rank += lo;
        }
        emit_match(k, pr.out_off + rank, tbest + pr.fwd_off, out_q, out_t, out_d, rec_format, out_rec);
    }
    if (sg == 0 && threadIdx.x == 0) {
        int total = 0;
        for (int r = 0; r < SPLIT_G; ++r) total += cnts[r];
        out_count[p] = total;
    }
}

__global__ void __launch_bounds__(256)
knn2_emit_kernel(int kind, const DevPair* __restrict__ pairs, int blocks_per_pair,
                 const Top2* __restrict__ fwd, long long total_n, int S, int32_t* __restrict__ idx1,
                 int32_t* __restrict__ idx2, float* __restrict__ d1, float* __restrict__ d2, RefineI8 rf) {
    int p = blockIdx.x / blocks_per_pair;
    int i = (blockIdx.x - p * blocks_per_pair) * blockDim.x + threadIdx.x;
    DevPair pr = pairs[p];
    const bool active = i < pr.n;
    Top2 t; t.k1 = B200M_SENTINEL; t.k2 = B200M_SENTINEL;
    if (active) t = merge_splits(fwd, total_n, S, pr.fwd_off + i);
    if (kind == B200M_KIND_L2_I8) refine_i8_warp(t.k1 != B200M_SENTINEL, i, t, pr, rf);
    if (!active) return;
    long long o = pr.out_off + i;
    bool h1 = t.k1 != B200M_SENTINEL, h2 = t.k2 != B200M_SENTINEL;
    idx1[o] = h1 ? (int32_t)(u32)t.k1 : -1;
    idx2[o] = h2 ? (int32_t)(u32)t.k2 : -1;
    d1[o] = h1 ? value_to_distance(kind, (u32)(t.k1 >> 32)) : -1.0f;
    d2[o] = h2 ? value_to_distance(kind, (u32)(t.k2 >> 32)) : -1.0f;
}

}  // namespace


int launch_scan_offsets(const MatchPlan& plan, const int32_t* counts, int num_pairs, long long* out_ptr,
                        cudaStream_t st) {
    scan_offsets_kernel<<<1, 256, 0, st>>>(plan.pairs, counts, num_pairs, out_ptr);
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "scan_offsets_kernel launch");
    return B200M_OK;
}

int launch_finalize(int kind, int mode, double ratio, const MatchPlan& plan, int num_pairs, int max_n,
                    int32_t* counts, const RefineI8& rf, cudaStream_t st) {
    // 128-thread blocks: small enough to share an SM with the persistent knn kernel of the next chunk
    const int threads = 128;
    int dev = 0, sms = 0;
    if (!b200m_current_device(&dev, &sms)) return B200M_ENODEVICE;
    // few rows in the whole call: spread them over more warps (see the kernel)
    const long long warps32 = (long long)num_pairs * ((max_n + 31) / 32);
    const int rpw = warps32 >= 16LL * sms ? 32 : (warps32 >= 4LL * sms ? 16 : 8);
    const int rows_per_block = (threads / 32) * rpw;
    int bpp = (max_n + rows_per_block - 1) / rows_per_block;
    long long grid = (long long)num_pairs * bpp;
    if (grid <= 0) return B200M_OK;
    if (grid > 0x7fffffffLL) return B200M_EUNSUPPORTED;
    finalize_kernel<<<(unsigned)grid, threads, 0, st>>>(kind, mode, ratio, plan.pairs, bpp, plan.fwd, plan.total_n,
                                                    plan.bwd, plan.total_m, plan.slices, plan.sortbuf, plan.tbest, counts, rf, plan.colmin, rpw);
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "finalize_kernel launch");
    return B200M_OK;
}

int launch_sort_emit(int mode, const MatchPlan& plan, int num_pairs, int max_npad, int32_t* out_q,
                     int32_t* out_t, float* out_d, int32_t* out_count, const int32_t* counts, cudaStream_t st,
                     int rec_format, void* out_rec) {
    if (num_pairs <= 0) return B200M_OK;
    static B200mPerDevice once;                           // the shared-memory opt-in is per device
    const int kMaxSmemKeys = 16384;                       // 128 KiB of 64-bit keys
    int dev = 0;
    if (!b200m_current_device(&dev, nullptr)) return B200M_ENODEVICE;
    if (!once.done[dev]) {
        B200M_CUDA_CHECK(cudaFuncSetAttribute(sort_emit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              kMaxSmemKeys * (int)sizeof(u64)), "sort_emit smem attr");
        once.done[dev] = 1;
    }
    // MODE_RATIO usually keeps a few percent of the rows: a 16 KiB slice (2048 keys) and 256 threads are enough, and such
    // a block fits beside a resident knn CTA of the next chunk.  Pairs that keep more (and the other modes, which keep
    // most rows) take the big shape: 128 KiB slice, 1024 threads.  A handful of pairs has nothing to overlap with: big only.
    if (num_pairs == 1 && max_npad >= 2048 && max_npad <= kMaxSmemKeys && !getenv("B200M_NO_SPLIT_SORT")) {
        // the one-pair call (the only shape this path has been validated on): SPLIT_G CTAs for the pair (segment sorts, then
        // placement by rank); segment counts live in the call's flag words [8, 8 + SPLIT_G)
        static_assert(8 + 4 * SPLIT_G <= 64, "segment counts fit the plan's flag words");
        int32_t* segcnt = plan.flags + 8;
        const int seg_keys = max_npad / SPLIT_G;
        const int seg_threads = seg_keys >= 1024 ? 1024 : (seg_keys >= 256 ? 256 : 64);
        if (!once.val[dev][0]) {
            B200M_CUDA_CHECK(cudaFuncSetAttribute(merge_emit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                  kMaxSmemKeys * (int)sizeof(u64)), "merge_emit smem attr");
            once.val[dev][0] = 1;
        }
        sort_segments_kernel<<<num_pairs * SPLIT_G, seg_threads, (size_t)seg_keys * sizeof(u64), st>>>(plan.pairs, plan.sortbuf, segcnt);
        B200M_CUDA_CHECK(cudaGetLastError(), "sort_segments_kernel launch");
        merge_emit_kernel<<<num_pairs * SPLIT_G, seg_threads, (size_t)max_npad * sizeof(u64), st>>>(
            plan.pairs, plan.sortbuf, segcnt, plan.tbest, out_q, out_t, out_d, out_count, rec_format, out_rec);
        B200M_CUDA_CHECK(cudaGetLastError(), "merge_emit_kernel launch");
        b200m_count_launch(2);
        return B200M_OK;
    }
    const bool few = mode == B200M_MODE_RATIO && num_pairs >= 64 && max_npad > 2048;
    const int big_keys = max_npad <= kMaxSmemKeys ? max_npad : kMaxSmemKeys;
    const int big_threads = max_npad >= 2048 ? 1024 : (max_npad >= 512 ? 256 : 64);
    if (few) {
        sort_emit_kernel<<<num_pairs, 256, 2048 * sizeof(u64), st>>>(plan.pairs, plan.sortbuf, plan.tbest, 2048, out_q, out_t,
                                                                     out_d, out_count, counts, -1, 2048, rec_format, out_rec);
        b200m_count_launch();
        B200M_CUDA_CHECK(cudaGetLastError(), "sort_emit_kernel launch");
    }
    sort_emit_kernel<<<num_pairs, big_threads, (size_t)big_keys * sizeof(u64), st>>>(
        plan.pairs, plan.sortbuf, plan.tbest, big_keys, out_q, out_t, out_d, out_count, counts, few ? 2048 : -1, 0x7fffffff,
        rec_format, out_rec);
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "sort_emit_kernel launch");
    return B200M_OK;
}

int launch_knn2_emit(int kind, const MatchPlan& plan, int num_pairs, int max_n, int32_t* idx1, int32_t* idx2,
                     float* d1, float* d2, const RefineI8& rf, cudaStream_t st) {
    int bpp = (max_n + 255) / 256;
    long long grid = (long long)num_pairs * bpp;
    if (grid <= 0) return B200M_OK;
    if (grid > 0x7fffffffLL) return B200M_EUNSUPPORTED;
    knn2_emit_kernel<<<(unsigned)grid, 256, 0, st>>>(kind, plan.pairs, bpp, plan.fwd, plan.total_n, plan.slices,
                                                     idx1, idx2, d1, d2, rf);
    b200m_count_launch();
    B200M_CUDA_CHECK(cudaGetLastError(), "knn2_emit_kernel launch");
    return B200M_OK;
}
