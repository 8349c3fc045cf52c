// capi.cu — the extern "C" boundary declared in include/b200match.h.  Host-side planning only:
// carve the caller's workspace, copy the pair table, enqueue the kernels.  No allocation, no
// exceptions, no CPU compute path.
#include <atomic>
#include <mutex>
#include <cstdio>
#include <cstring>
#include <vector>
#include "common.cuh"

static std::atomic<long long> g_launches{0};
static thread_local char g_cuda_err[256] = "";

void b200m_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int b200m_set_cuda_error(cudaError_t e, const char* where) {
    snprintf(g_cuda_err, sizeof(g_cuda_err), "%s: %s", where, cudaGetErrorString(e));
    return B200M_ECUDA;
}

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
static inline int pow2ceil(int n) { int p = 1; while (p < n) p <<= 1; return p; }

// Column splits per pair.  `bound` (workspace sizing, no device at hand): an upper limit of what any device would pick.
static int choose_splits(int kind, int num_pairs, int max_n, int max_m, int mode, int sm_count, bool bound = false) {
    // Enough CTAs for ~2 per SM; a single image pair has only ceil(n/128) row tiles.
    long long tiles = (long long)num_pairs * ((max_n + 127) / 128);
    if (mode == B200M_MODE_CROSS) tiles += (long long)num_pairs * ((max_m + 127) / 128);
    // keep >= 256 columns per split (>= 512 for L2_I8: four 128-column tiles, one per epilogue group)
    const int min_cols = kind == B200M_KIND_L2_I8 || kind == B200M_KIND_HAMMING_TC || kind == B200M_KIND_HAMMING_FP4 ? 512 : 256;
    int max_split = (((max_n < max_m ? max_n : max_m) + min_cols - 1) / min_cols);
    if (max_split > 8) max_split = 8;
    if (max_split < 1) max_split = 1;
    if (bound) return tiles >= 2LL * 256 ? 1 : max_split;              // no device has more than 256 SMs
    if (tiles >= 2LL * sm_count) return 1;
    const bool paired = (kind == B200M_KIND_L2_I8 && mode == B200M_MODE_RATIO) || kind == B200M_KIND_HAMMING_TC ||
                        kind == B200M_KIND_HAMMING_FP4;
    if (paired) {
        // The tensor kernels of these kinds are persistent: a cluster of two CTAs takes (pair, split, two row tiles) items in
        // turn, so a small call takes rounds x (train tiles per item + ~2 tiles of per-item work).  A few more or fewer splits
        // change the number of rounds: ORB 10k x 10k, 40 row-tile pairs on 74 clusters -- 2 splits are 2 rounds of 21 tiles,
        // 3 splits 2 rounds of 14.  Each direction of a cross-check is its own launch.
        const int units = sm_count / 2 > 0 ? sm_count / 2 : 1;
        const int tn = kind == B200M_KIND_HAMMING_FP4 ? 240 : 256;
        long long best_cost = -1;
        int best = 1;
        for (int S = 1; S <= max_split; ++S) {
            long long cost = 0;
            for (int dir = 0; dir < (mode == B200M_MODE_CROSS ? 2 : 1); ++dir) {
                const int rows = dir ? max_m : max_n, cols = dir ? max_n : max_m;
                const long long items = (long long)num_pairs * S * ((rows + 255) / 256);
                const long long rounds = (items + units - 1) / units;
                cost += rounds * (((cols + S - 1) / S + tn - 1) / tn + 2);
            }
            if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = S; }
        }
        return best;
    }
    long long want = 2LL * sm_count;
    int S = (int)((want + tiles - 1) / (tiles > 0 ? tiles : 1));
    if (S > max_split) S = max_split;
    if (S < 1) S = 1;
    return S;
}

// Optional device timing of the dominant (knn) kernels for bench.py's roofline: CUDA events recorded on
// the launching stream around each knn launch, summed by b200m_timing_collect().
// A measurement facility, not part of the data path: one process-wide list guarded by a mutex (the events themselves
// belong to whichever device was current when they were created; bench.py drives one device per process).
static std::mutex g_timing_mu;
static bool g_timing = false;
static std::vector<std::pair<cudaEvent_t, cudaEvent_t>> g_events;
static size_t g_events_used = 0;

static std::atomic<int> g_sm_count[B200M_MAX_DEVICES];      // per device ordinal; 0 = not queried yet
bool b200m_current_device(int* dev_out, int* sm_out) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= B200M_MAX_DEVICES) return false;
    int sm = g_sm_count[dev].load(std::memory_order_relaxed);
    if (sm == 0) {
        if (cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sm <= 0) return false;
        g_sm_count[dev].store(sm, std::memory_order_relaxed);
    }
    if (dev_out) *dev_out = dev;
    if (sm_out) *sm_out = sm;
    return true;
}
static int device_sm_count() {
    int sm = 0;
    return b200m_current_device(nullptr, &sm) ? sm : 0;
}

// Lay the workspace out.  base may be null (size query).
static void make_plan(MatchPlan& pl, char* base, int kind, int mode, int num_pairs, long long total_n,
                      long long total_m, long long total_pad, int S, bool need_bwd) {
    size_t off = 0;
    pl.S = S; pl.total_n = total_n; pl.total_m = total_m; pl.total_pad = total_pad;
    pl.slices = kind == B200M_KIND_L2_I8 ? S * B200M_I8_GROUPS : S;
    const size_t SL = (size_t)pl.slices;
    pl.flags = (int32_t*)(base + off); off = align_up(off + 64 * sizeof(int32_t), 256);
    pl.pairs = (DevPair*)(base + off); off = align_up(off + sizeof(DevPair) * (size_t)num_pairs, 256);
    pl.fwd = (Top2*)(base + off);      off = align_up(off + sizeof(Top2) * SL * total_n, 256);
    pl.bwd = (Top2*)(base + off);      off = align_up(off + (need_bwd ? sizeof(Top2) * SL * total_m : 0), 256);
    pl.sortbuf = (u64*)(base + off);   off = align_up(off + sizeof(u64) * (size_t)total_pad, 256);
    pl.tbest = (int32_t*)(base + off); off = align_up(off + sizeof(int32_t) * (size_t)total_n, 256);
    pl.cand = nullptr; pl.pmax = nullptr;
    if (kind == B200M_KIND_L2_F32 || kind == B200M_KIND_L2_F16C) {
        long long rows = need_bwd && total_m > total_n ? total_m : total_n;
        pl.cand = (void*)(base + off);  off = align_up(off + (8ull * B200M_F32_CAND + 4ull) * 2 * (size_t)S * rows + 8ull * rows + 64, 256);
        pl.pmax = (float*)(base + off); off = align_up(off + sizeof(float) * (size_t)num_pairs, 256);
    }
    pl.bytes = off;
    (void)mode;
}

extern "C" {

int b200m_version(void) { return B200M_VERSION; }

const char* b200m_error_string(int code) {
    switch (code) {
        case B200M_OK: return "ok";
        case B200M_EINVAL: return "invalid argument";
        case B200M_EWORKSPACE: return "workspace too small";
        case B200M_ECUDA: return "CUDA error";
        case B200M_ENODEVICE: return "no sm_100 device (this library has no CPU path)";
        case B200M_EUNSUPPORTED: return "unsupported shape";
        case B200M_ESCRATCH: return "problem exceeds the library's static scratch (B200M_STATIC_SCRATCH_BYTES): see b200m_last_cuda_error()";
        default: return "unknown error";
    }
}

const char* b200m_last_cuda_error(void) { return g_cuda_err; }

int b200m_device_info(int* sm_count, int* cc_major, int* cc_minor) {
    int dev = 0, n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) return B200M_ENODEVICE;
    if (cudaGetDevice(&dev) != cudaSuccess) return B200M_ENODEVICE;
    int sm = 0, ma = 0, mi = 0;
    cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&ma, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&mi, cudaDevAttrComputeCapabilityMinor, dev);
    if (sm_count) *sm_count = sm;
    if (cc_major) *cc_major = ma;
    if (cc_minor) *cc_minor = mi;
    return ma == 10 ? B200M_OK : B200M_ENODEVICE;
}

int b200m_timing_enable(int on) {
    std::lock_guard<std::mutex> lk(g_timing_mu);
    g_timing = on != 0;
    g_events_used = 0;
    return B200M_OK;
}

int b200m_timing_collect(double* knn_ms, int* knn_launches) {
    std::lock_guard<std::mutex> lk(g_timing_mu);
    double total = 0.0;
    for (size_t i = 0; i < g_events_used; ++i) {
        float ms = 0.f;
        B200M_CUDA_CHECK(cudaEventSynchronize(g_events[i].second), "event sync");
        B200M_CUDA_CHECK(cudaEventElapsedTime(&ms, g_events[i].first, g_events[i].second), "event elapsed");
        total += ms;
    }
    if (knn_ms) *knn_ms = total;
    if (knn_launches) *knn_launches = (int)g_events_used;
    g_events_used = 0;
    return B200M_OK;
}

int b200m_f32_stats(int64_t* fallback_rows, int64_t* fallback_evals, int reset) {
    unsigned long long v[2] = {0, 0};
    int rc = b200m_f32_stats_read(v, reset != 0);
    if (rc != B200M_OK) return rc;
    if (fallback_rows) *fallback_rows = (int64_t)v[0];
    if (fallback_evals) *fallback_evals = (int64_t)v[1];
    return B200M_OK;
}

int64_t b200m_launch_count(void) { return g_launches.load(); }
void b200m_reset_launch_count(void) { g_launches.store(0); }

int b200m_workspace_bytes(int kind, int mode, int num_pairs, int64_t total_n, int64_t total_m,
                          int max_n, int max_m, int D, size_t* bytes) {
    (void)D;
    if (!bytes || num_pairs < 0 || total_n < 0 || total_m < 0) return B200M_EINVAL;
    // Upper bounds that do not depend on the device: choose_splits(..., bound = true) is an upper limit of the split count
    // on any device (<= 256 SMs assumed), sum pow2ceil(n) <= 2 * total_n + P.
    MatchPlan pl;
    make_plan(pl, nullptr, kind, mode, num_pairs, total_n, total_m, 2 * total_n + num_pairs,
              choose_splits(kind, num_pairs, max_n, max_m, mode, 256, true), mode == B200M_MODE_CROSS);
    *bytes = pl.bytes;
    return B200M_OK;
}

}  // extern "C"

// Shared front half of match_pairs / knn2_pairs: validate, plan, upload pair table, run knn stage.
static int run_knn_stage(int kind, const void* q_desc, const void* q_norm, int64_t q_rows, const void* t_desc,
                         const void* t_norm, int64_t t_rows, int D, const b200m_pair* pairs_host,
                         int num_pairs, int mode, bool need2, void* workspace, size_t workspace_bytes,
                         cudaStream_t st, MatchPlan& pl, int& max_n, int& max_npad) {
    if (num_pairs < 0 || D <= 0 || (num_pairs > 0 && (!pairs_host || !q_desc || !t_desc || !workspace)))
        return B200M_EINVAL;
    if (kind != B200M_KIND_HAMMING && kind != B200M_KIND_L2_I8 && kind != B200M_KIND_L2_F32 && kind != B200M_KIND_HAMMING_TC &&
        kind != B200M_KIND_HAMMING_FP4 && kind != B200M_KIND_L2_F16C)
        return B200M_EINVAL;
    if (kind != B200M_KIND_HAMMING && (!q_norm || !t_norm) && num_pairs > 0) return B200M_EINVAL;
    int htc_pad = 0;                                  // HAMMING_TC: width of the +-1 stores
    const bool htc = kind == B200M_KIND_HAMMING_TC || kind == B200M_KIND_HAMMING_FP4;
    if (htc) {
        int rc = kind == B200M_KIND_HAMMING_FP4 ? b200m_hamming_fp4_width(D, &htc_pad) : b200m_hamming_tc_width(D, &htc_pad);
        if (rc != B200M_OK) return rc;
    }
    int sm = device_sm_count();
    if (sm == 0) return B200M_ENODEVICE;

    long long total_n = 0, total_m = 0, total_pad = 0;
    int max_m = 0;
    max_n = 0; max_npad = 1;
    std::vector<DevPair> dp((size_t)num_pairs);
    for (int p = 0; p < num_pairs; ++p) {
        const b200m_pair& s = pairs_host[p];
        if (s.n < 0 || s.m < 0 || s.q_off < 0 || s.t_off < 0 || s.out_off < 0) return B200M_EINVAL;
        if ((int64_t)s.q_off + s.n > q_rows || (int64_t)s.t_off + s.m > t_rows) return B200M_EINVAL;
        DevPair& d = dp[p];
        d.q_off = s.q_off; d.n = s.n; d.t_off = s.t_off; d.m = s.m; d.out_off = s.out_off;
        d.fwd_off = total_n; d.bwd_off = total_m; d.sort_off = total_pad;
        d.npad = pow2ceil(s.n > 0 ? s.n : 1); d._pad = 0;
        total_n += s.n; total_m += s.m; total_pad += d.npad;
        if (s.n > max_n) max_n = s.n;
        if (s.m > max_m) max_m = s.m;
        if (d.npad > max_npad) max_npad = d.npad;
    }
    const bool need_bwd = mode == B200M_MODE_CROSS;
    // Hamming cross-check: one pass yields both the row top-2 and the column minima (atomicMin on packed keys)
    const bool fused_cross = need_bwd && kind == B200M_KIND_HAMMING && max_n <= (1 << B200M_HAMMING_IDX_BITS);
    int S = choose_splits(kind, num_pairs, max_n, max_m, mode, sm);
    make_plan(pl, (char*)workspace, kind, mode, num_pairs, total_n, total_m, total_pad, S, need_bwd);
    if (pl.bytes > workspace_bytes) return B200M_EWORKSPACE;
    if (num_pairs == 0 || total_n == 0) return B200M_OK;
    B200M_CUDA_CHECK(cudaMemcpyAsync(pl.pairs, dp.data(), sizeof(DevPair) * (size_t)num_pairs,
                                     cudaMemcpyHostToDevice, st), "pair table upload");
    // pageable source: the copy is staged before the call returns, so `dp` may go out of scope.
    // MODE_RATIO on int8 rows: decide ONCE, on this stream, whether the fold kernel's digits cover the train store
    const bool fold_ok = kind == B200M_KIND_L2_I8 && mode == B200M_MODE_RATIO;
    if (fold_ok)
        B200M_CUDA_CHECK(cudaMemcpyAsync(pl.flags, (const char*)t_norm + b200m_aux_info_off(t_rows) + 3 * sizeof(int32_t),
                                         sizeof(int32_t), cudaMemcpyDeviceToDevice, st), "fold flag snapshot");
    pl.colmin = fused_cross ? (u32*)pl.bwd : nullptr;
    if (fused_cross)
        B200M_CUDA_CHECK(cudaMemsetAsync(pl.colmin, 0xFF, sizeof(u32) * (size_t)total_m, st), "colmin memset");
    for (int dir = 0; dir < (need_bwd && !fused_cross ? 2 : 1); ++dir) {
        bool back = dir == 1;
        Top2* out = back ? pl.bwd : pl.fwd;
        long long rows = back ? total_m : total_n;
        int rc;
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (g_timing) {
            std::lock_guard<std::mutex> lk(g_timing_mu);
            if (g_events_used == g_events.size()) {
                cudaEvent_t a, b;
                B200M_CUDA_CHECK(cudaEventCreate(&a), "event create");
                B200M_CUDA_CHECK(cudaEventCreate(&b), "event create");
                g_events.emplace_back(a, b);
            }
            e0 = g_events[g_events_used].first; e1 = g_events[g_events_used].second;
            ++g_events_used;
            cudaEventRecord(e0, st);
        }
        if (kind == B200M_KIND_HAMMING)
            rc = launch_hamming_top2((const uint8_t*)q_desc, (const uint8_t*)t_desc, D, pl.pairs, num_pairs,
                                     max_n, max_m, S, back, out, rows, pl.colmin, st);
        else if (htc) {
            rc = launch_hamming_tc_top2(q_norm, q_rows, t_norm, t_rows, htc_pad, 8 * D, kind == B200M_KIND_HAMMING_FP4,
                                        pl.pairs, num_pairs, max_n, max_m, S, back, out, rows, st);
            if (rc == B200M_OK)      // chunk-level top-2 -> exact (distance, column) top-2 in slice 0, the POPC kernel's format
                rc = launch_hamming_tc_resolve((const uint8_t*)q_desc, (const uint8_t*)t_desc, D, pl.pairs, num_pairs,
                                               max_n, max_m, S, back, !need2, out, rows, st);
        } else if (kind == B200M_KIND_L2_I8)
            rc = launch_l2_i8_top2((const int8_t*)q_desc, (const int32_t*)q_norm, q_rows, (const int8_t*)t_desc,
                                   (const int32_t*)t_norm, t_rows, D, pl.pairs, num_pairs, max_n, max_m, S,
                                   back, fold_ok && !back ? pl.flags : nullptr, out, rows, st);
        else
            rc = launch_l2_f32_top2((const float*)q_desc, (const float*)q_norm, q_rows, (const float*)t_desc,
                                    (const float*)t_norm, t_rows, D, pl.pairs, num_pairs, max_n, max_m, S,
                                    back, need2, pl.cand, pl.pmax, out, rows, st, kind == B200M_KIND_L2_F16C);
        if (e1) cudaEventRecord(e1, st);
        if (rc != B200M_OK) return rc;
    }
    if (htc) pl.slices = 1;     // resolved into slice 0
    return B200M_OK;
}

extern "C" {

static int match_pairs_impl(int kind, const void* q_desc, const void* q_norm, int64_t q_rows,
                            const void* t_desc, const void* t_norm, int64_t t_rows, int D,
                            const b200m_pair* pairs_host, int num_pairs, int mode, double ratio,
                            int32_t* out_q, int32_t* out_t, float* out_d, int32_t* out_count, int64_t* out_ptr,
                            void* workspace, size_t workspace_bytes, void* stream, int rec_format = 0, void* out_rec = nullptr) {
    if (mode != B200M_MODE_RATIO && mode != B200M_MODE_CROSS && mode != B200M_MODE_NN) return B200M_EINVAL;
    if (num_pairs > 0 && !out_count) return B200M_EINVAL;
    if (num_pairs > 0 && (rec_format ? !out_rec : (!out_q || !out_t || !out_d))) return B200M_EINVAL;
    if (rec_format == B200M_REC_Q16T16D32 || rec_format == B200M_REC_Q16T16D16) {
        for (int p = 0; p < num_pairs; ++p)                  // 16-bit row indices
            if (pairs_host && (pairs_host[p].n > 65536 || pairs_host[p].m > 65536)) return B200M_EUNSUPPORTED;
        const bool ham = kind == B200M_KIND_HAMMING || kind == B200M_KIND_HAMMING_TC || kind == B200M_KIND_HAMMING_FP4;
        if (rec_format == B200M_REC_Q16T16D16 && (!ham || D > 8191)) return B200M_EUNSUPPORTED;   // the distance must be a bit count < 2^16
    }
    cudaStream_t st = (cudaStream_t)stream;
    MatchPlan pl; int max_n = 0, max_npad = 1;
    int rc = run_knn_stage(kind, q_desc, q_norm, q_rows, t_desc, t_norm, t_rows, D, pairs_host, num_pairs,
                           mode, mode == B200M_MODE_RATIO, workspace, workspace_bytes, st, pl, max_n, max_npad);
    if (rc != B200M_OK) return rc;
    if (num_pairs == 0) {
        if (out_ptr) B200M_CUDA_CHECK(cudaMemsetAsync(out_ptr, 0, sizeof(int64_t), st), "ptr memset");
        return B200M_OK;
    }
    B200M_CUDA_CHECK(cudaMemsetAsync(out_count, 0, sizeof(int32_t) * (size_t)num_pairs, st), "count memset");
    if (pl.total_n == 0) {
        if (out_ptr) B200M_CUDA_CHECK(cudaMemsetAsync(out_ptr, 0, sizeof(int64_t) * ((size_t)num_pairs + 1), st), "ptr memset");
        return B200M_OK;
    }
    // L2_I8: the exact runner-up refinement (inside the winner's 64-column chunk) is fused into finalize.
    RefineI8 rf = {nullptr, nullptr, nullptr, nullptr, 0, nullptr, nullptr};
    if (kind == B200M_KIND_L2_I8)
        rf = RefineI8{(const int8_t*)q_desc, (const int32_t*)q_norm, (const int8_t*)t_desc, (const int32_t*)t_norm, D,
                      (const int32_t*)((const char*)t_norm + b200m_aux_info_off(t_rows)),
                      mode == B200M_MODE_RATIO ? pl.flags : nullptr};
    // CSR mode counts survivors in finalize (into out_count), scans them into tight offsets, then emits.
    const int tail_kind = kind == B200M_KIND_HAMMING_TC || kind == B200M_KIND_HAMMING_FP4 ? B200M_KIND_HAMMING    // integer bit counts either way
                          : kind == B200M_KIND_L2_F16C ? B200M_KIND_L2_F32 : kind;                               // exact fp32 d^2 bits either way
    rc = launch_finalize(tail_kind, mode, ratio, pl, num_pairs, max_n, out_ptr ? out_count : nullptr, rf, st);
    if (rc != B200M_OK) return rc;
    if (out_ptr) {
        rc = launch_scan_offsets(pl, out_count, num_pairs, (long long*)out_ptr, st);
        if (rc != B200M_OK) return rc;
    }
    return launch_sort_emit(mode, pl, num_pairs, max_npad, out_q, out_t, out_d, out_count, out_ptr ? out_count : nullptr, st,
                            rec_format, out_rec);
}

int b200m_match_pairs(int kind, const void* q_desc, const void* q_norm, int64_t q_rows,
                      const void* t_desc, const void* t_norm, int64_t t_rows, int D,
                      const b200m_pair* pairs_host, int num_pairs, int mode, double ratio,
                      int32_t* out_q, int32_t* out_t, float* out_d, int32_t* out_count,
                      void* workspace, size_t workspace_bytes, void* stream) {
    return match_pairs_impl(kind, q_desc, q_norm, q_rows, t_desc, t_norm, t_rows, D, pairs_host, num_pairs, mode,
                            ratio, out_q, out_t, out_d, out_count, nullptr, workspace, workspace_bytes, stream);
}

int b200m_match_pairs_csr(int kind, const void* q_desc, const void* q_norm, int64_t q_rows,
                          const void* t_desc, const void* t_norm, int64_t t_rows, int D,
                          const b200m_pair* pairs_host, int num_pairs, int mode, double ratio,
                          int32_t* out_q, int32_t* out_t, float* out_d, int32_t* out_count, int64_t* out_ptr,
                          void* workspace, size_t workspace_bytes, void* stream) {
    if (num_pairs >= 0 && !out_ptr) return B200M_EINVAL;
    return match_pairs_impl(kind, q_desc, q_norm, q_rows, t_desc, t_norm, t_rows, D, pairs_host, num_pairs, mode,
                            ratio, out_q, out_t, out_d, out_count, out_ptr, workspace, workspace_bytes, stream);
}

// ---- b200m_match_pairs_static: the workspace lives in this library, not in the caller --------------------------------
// File-scope scratch: every GPU that loads the library holds these bytes until the process exits.  In HOST code the name
// g_static_scratch is a host-side stand-in, not a device address; the address a kernel may use comes from
// cudaGetSymbolAddress, once per device, with that device current (static_scratch_ptr below).
static __device__ __align__(256) unsigned char g_static_scratch[B200M_STATIC_SCRATCH_BYTES];
static void* g_static_scratch_ptr[B200M_MAX_DEVICES];      // per device ordinal; null = not looked up yet

static int static_scratch_ptr(void** out) {
    int dev = 0;
    if (!b200m_current_device(&dev, nullptr)) return B200M_ENODEVICE;
    void* p = g_static_scratch_ptr[dev];
    if (!p) {
        B200M_CUDA_CHECK(cudaGetSymbolAddress(&p, g_static_scratch), "cudaGetSymbolAddress(static scratch)");
        if (!p || ((uintptr_t)p & 255)) return b200m_set_cuda_error(cudaErrorInvalidSymbol, "static scratch address");
        g_static_scratch_ptr[dev] = p;         // two threads racing here store the same address
    }
    *out = p;
    return B200M_OK;
}

int b200m_match_pairs_static(int kind, const void* q_desc, const void* q_norm, int64_t q_rows,
                             const void* t_desc, const void* t_norm, int64_t t_rows, int D,
                             const b200m_pair* pairs_host, int num_pairs, int mode, double ratio,
                             int32_t* out_q, int32_t* out_t, float* out_d, int32_t* out_count, void* stream) {
    void* ws = nullptr;
    int rc = static_scratch_ptr(&ws);
    if (rc != B200M_OK) return rc;
    // Nothing in the scratch has to start at zero across calls: the plan re-initialises what it reads before writing
    // (flag words, column minima, the caller's out_count) on `stream`, exactly as it does for a caller-owned workspace
    // that is reused dirty.
    rc = match_pairs_impl(kind, q_desc, q_norm, q_rows, t_desc, t_norm, t_rows, D, pairs_host, num_pairs, mode, ratio,
                          out_q, out_t, out_d, out_count, nullptr, ws, (size_t)B200M_STATIC_SCRATCH_BYTES, stream);
    if (rc == B200M_EWORKSPACE) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "b200m_match_pairs_static: this batch needs more than the %u-byte static scratch "
                 "(B200M_STATIC_SCRATCH_BYTES); use b200m_match_pairs with a caller-owned workspace", (unsigned)B200M_STATIC_SCRATCH_BYTES);
        return B200M_ESCRATCH;
    }
    return rc;
}

int b200m_rec_bytes(int rec_format) {
    return rec_format == B200M_REC_Q32T32D32 ? 12 : rec_format == B200M_REC_Q16T16D32 ? 8 : rec_format == B200M_REC_Q16T16D16 ? 6 : 0;
}

int b200m_match_pairs_rec(int kind, const void* q_desc, const void* q_norm, int64_t q_rows,
                          const void* t_desc, const void* t_norm, int64_t t_rows, int D,
                          const b200m_pair* pairs_host, int num_pairs, int mode, double ratio,
                          int rec_format, void* out_rec, int32_t* out_count, int64_t* out_ptr,
                          void* workspace, size_t workspace_bytes, void* stream) {
    if (b200m_rec_bytes(rec_format) == 0 || (num_pairs >= 0 && !out_ptr)) return B200M_EINVAL;
    if (((uintptr_t)out_rec & 7)) return B200M_EINVAL;
    return match_pairs_impl(kind, q_desc, q_norm, q_rows, t_desc, t_norm, t_rows, D, pairs_host, num_pairs, mode, ratio,
                            nullptr, nullptr, nullptr, out_count, out_ptr, workspace, workspace_bytes, stream, rec_format, out_rec);
}

int b200m_knn2_pairs(int kind, const void* q_desc, const void* q_norm, int64_t q_rows,
                     const void* t_desc, const void* t_norm, int64_t t_rows, int D,
                     const b200m_pair* pairs_host, int num_pairs,
                     int32_t* idx1, int32_t* idx2, float* d1, float* d2,
                     void* workspace, size_t workspace_bytes, void* stream) {
    if (num_pairs > 0 && (!idx1 || !idx2 || !d1 || !d2)) return B200M_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    MatchPlan pl; int max_n = 0, max_npad = 1;
    int rc = run_knn_stage(kind, q_desc, q_norm, q_rows, t_desc, t_norm, t_rows, D, pairs_host, num_pairs,
                           B200M_MODE_NN, true, workspace, workspace_bytes, st, pl, max_n, max_npad);
    if (rc != B200M_OK) return rc;
    if (num_pairs == 0 || pl.total_n == 0) return B200M_OK;
    RefineI8 rf = {nullptr, nullptr, nullptr, nullptr, 0, nullptr, nullptr};
    if (kind == B200M_KIND_L2_I8)
        rf = RefineI8{(const int8_t*)q_desc, (const int32_t*)q_norm, (const int8_t*)t_desc, (const int32_t*)t_norm, D, nullptr, nullptr};
    return launch_knn2_emit(kind == B200M_KIND_HAMMING_TC || kind == B200M_KIND_HAMMING_FP4 ? B200M_KIND_HAMMING
                            : kind == B200M_KIND_L2_F16C ? B200M_KIND_L2_F32 : kind, pl, num_pairs, max_n, idx1, idx2, d1, d2, rf, st);
}

}  // extern "C"
