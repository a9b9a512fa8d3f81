// dgnn_ctx.h -- internal: the context behind the opaque dgnn_ctx of include/dgnn_cabi.h and the small helpers shared by
// the translation units of the C ABI (dgnn_api.cu: context + build orchestration; dgnn_consumers.cu: host copies, RBF
// expansion, consumers of the CSR, DLPack export; dgnn_stream.cu uses the public ABI only).
#pragma once

#include <cuda_runtime.h>

#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <utility>
#include <vector>

#include "../../include/dgnn_cabi.h"
#include "dgnn_internal.h"

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};

struct BuildEvents {
    cudaEvent_t e[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
};

struct dgnn_ctx {
    int device = 0;
    uint64_t build_serial = 0;
    std::atomic<int> live_exports{0};  // DLPack tensors that borrow the arena (dgnn_collated_dlpack)
    std::atomic<bool> zombie{false};   // dgnn_destroy was called while exports were alive: the last deleter frees
    std::vector<std::pair<std::string, double>> opt_log;  // options applied so far (replayed on the streaming twin)
    dgnn_ctx* twin = nullptr;                              // second arena of dgnn_build_stream (created on first use)
    cudaStream_t stream = nullptr;
    int sm_count = 148;
    int smem_optin = 0;
    int smem_per_sm = 0;
    std::string err;
    // inputs staged from host
    DevBuf in_off, in_x, in_y, in_z, in_lat;
    // planning / binning
    DevBuf small, fa, fb, fc;
    DevBuf geoms, astruct, acell, ncells32, nitems32, cell_base, item_base, scratch;
    DevBuf cellcount, cellstart, cx, cy, cz, cj, cn;
    // results
    DevBuf deg, deg_raw, row_perm, cmask, cmask_tn, hits, hits_d2, hits_n, hit_thr, row_ptr, seo, dst, image, dist, rbf;
    // stand-alone rbf staging
    DevBuf tmp_in, tmp_out;
    DevBuf stats;
    DgStats* h_stats = nullptr;  // pinned
    BuildEvents ev;
    dgnn_build_info info{};
    // options
    double cell_frac = 0.5;
    int tile_target = 2048;
    int src_target = 512;
    int tile_limit = 0;  // 0: candidates straight from the L2-resident cloud (faster than CTA-staged tiles: no barriers); > 0: stage tiles up to this many points
    int traverse_grid_mult = 16;
    int direct_max_atoms = 2048;
    int direct_min_blocks = 2;
    int sort_buckets = 0;
    int fuse_rbf = 1;
    int hit_cap_opt = -1;                 // -1: auto (density estimate), 0: off, > 0: fixed slots per row
    double hits_max_bytes = 16e9;         // no hit lists beyond this much scratch
    cudaStream_t copy_stream = nullptr;   // host inputs: coordinates are uploaded in structure-aligned chunks on this
    cudaEvent_t copy_ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // stream ...
    long long upload_min_atoms = 1 << 18; // smaller batches go up in one piece
    cudaEvent_t copy_meta_ev = nullptr;
    cudaStream_t aux_stream = nullptr;    // split fill pass: the launch of the few large rows runs beside the main one
    cudaEvent_t aux_fork = nullptr, aux_join = nullptr;
    int64_t* sink_row_ptr = nullptr;      // dgnn_set_offset_sinks
    int64_t* sink_seo = nullptr;
    int upload_chunks = 4;                // ... while the per-atom prep + count pass of earlier chunks already run
    int fill_hits_bulk = 1;               // k_fill_hits feature stores: bulk-copy engine (1) or st.global.cs (0)
    int cell32 = 1;                       // cell-list structures: fp32 tiles staged in shared memory + fp64 decision in the
                                          // guard band (dgnn_cell32.cu); 0: the fp64 traversal with hit lists (A/B reference)
    int cell32_fill_warps = 16;
    int dyn_rows = 1;                     // small batches: dynamic per-warp row schedule in the direct kernels
    int hit_threshold = 1;                // truncated rows: hit lists hold only pairs below a per-structure d2 threshold
    int fill_hits_warps = 0;              // (tuning) warps per CTA of k_fill_hits, 0: most resident warps per SM
    int tie_probe = 1;                    // 1: sampled rows choose the tie ranking of the uncapped fill pass, 0 / 2: fixed
    int cell32_mask = 1;                  // fp32-tile count pass hands its candidate masks to the fill pass
    int split_rows = 1;                   // fill pass: rows beyond the full-occupancy row buffer get a launch of their own when few
    int sym_count = 1;                    // count pass sweeps j >= i only and decides both directions of a pair
    int rbf_recur = 1;                    // fused epilogue: Gaussian recurrence across a float4 (2 ex2 instead of 4)
    bool onepass_skip = false;            // (set while a build whose look-back timed out is being redone in two passes)
    int lb_fallbacks = 0;                 // builds redone that way since the context was created
    int lb_spin_limit = 0;                // (testing) look-back spins before giving up, 0: default
    int onepass_opt = -1;                 // single-pass build of direct-path batches: -1 auto, 0 off, 1 wherever legal
    DevBuf lb_status;                     // per-row look-back status words of the single-pass build
    // the last successful build, for the CSR consumers (collation, viewer arrays, neighbourhood clouds)
    struct {
        bool valid = false;
        long long S = 0, N = 0, E = 0;
        int max_deg = 0;
        const long long* off = nullptr;
        const double *x = nullptr, *y = nullptr, *z = nullptr;
    } last;
    DevBuf col_x, col_ei, col_batch, col_keys, col_lut, col_table, col_err;  // dgnn_collate outputs / staging
    DevBuf view_a, view_b;
    DevBuf tab_cnt, tab_ptr, tab_val;                                         // dgnn_distance_tables outputs / staging                                                    // edge views / clouds staging
    dgnn_consumer_info cinfo{};
    cudaEvent_t cev[2] = {nullptr, nullptr};
    int onepass_row_cap = 0;              // (testing) row buffer entries of the single-pass build, 0: density estimate
    double onepass_arena_scale = 1.06;    // (testing) edge arena = scale * density estimate of the edge total
};



inline int fail(dgnn_ctx* c, int code, const char* fmt, ...) {
    if (c) {
        char buf[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof(buf), fmt, ap);
        va_end(ap);
        c->err = buf;
    }
    return code;
}

#define CU(call)                                                                                       \
    do {                                                                                               \
        cudaError_t _e = (call);                                                                       \
        if (_e != cudaSuccess)                                                                         \
            return fail(ctx, _e == cudaErrorMemoryAllocation ? DGNN_ERR_OOM : DGNN_ERR_CUDA,           \
                        "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__);   \
    } while (0)

inline int ensure(dgnn_ctx* ctx, DevBuf& b, size_t bytes) {
    if (bytes <= b.cap) return DGNN_OK;
    if (b.p) {
        CU(cudaStreamSynchronize(ctx->stream));
        CU(cudaFree(b.p));
        b.p = nullptr;
        b.cap = 0;
    }
    size_t want = bytes < (size_t(64) << 20) ? bytes + bytes / 2 + 256 : bytes + bytes / 16;
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) {
        cudaGetLastError();
        want = bytes;
        e = cudaMalloc(&p, want);
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(ctx, DGNN_ERR_OOM, "device allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
    }
    b.p = p;
    b.cap = want;
    return DGNN_OK;
}

#define ENSURE(buf, bytes)                          \
    do {                                            \
        int _r = ensure(ctx, (buf), (bytes));       \
        if (_r != DGNN_OK) return _r;               \
    } while (0)

inline void release(DevBuf& b) {
    if (b.p) cudaFree(b.p);
    b.p = nullptr;
    b.cap = 0;
}

inline int status_to_code(int st) {
    switch (st) {
        case DG_ST_SINGULAR: return DGNN_ERR_SINGULAR;
        case DG_ST_IMAGE_RANGE: return DGNN_ERR_IMAGE_RANGE;
        case DG_ST_GEOMETRY: return DGNN_ERR_GEOMETRY;
        case DG_ST_UNWRAPPED: return DGNN_ERR_UNWRAPPED;
        default: return DGNN_ERR_UNSUPPORTED;
    }
}

inline bool is_device_ptr(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

inline int fetch_stats(dgnn_ctx* ctx) {
    CU(cudaMemcpyAsync(ctx->h_stats, ctx->stats.p, sizeof(DgStats), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return DGNN_OK;
}

inline int round_up(int v, int m) { return (v + m - 1) / m * m; }
inline long long max_nb_for_est(uint64_t mn) { return mn > 0x7fffffffULL ? 0x7fffffffLL : (long long)mn; }

// The near-cutoff band  fabs(sqrt(d2) - r_c) <= 1e-6  (evaluated in fp64 exactly like the oracle) as an interval of
// d2: sqrt and the subtraction are monotone, so the band is [lo, hi] in the ordering of non-negative doubles; both ends
// are found by bisection on the bit pattern.  Lets the count pass test the band without a square root.
inline void near_band_d2(double rc, double* lo_out, double* hi_out) {
    auto f = [rc](double d2) { return std::sqrt(d2) - rc; };
    auto from_bits = [](unsigned long long u) { double d; memcpy(&d, &u, 8); return d; };
    auto to_bits = [](double d) { unsigned long long u; memcpy(&u, &d, 8); return u; };
    const unsigned long long top = to_bits((rc + 1.0) * (rc + 1.0));
    // lo = smallest d2 >= 0 with f(d2) >= -1e-6
    unsigned long long a = 0, b = top;
    while (a < b) {
        const unsigned long long mid = a + (b - a) / 2;
        if (f(from_bits(mid)) >= -1e-6) b = mid; else a = mid + 1;
    }
    *lo_out = from_bits(a);
    // hi = largest d2 with f(d2) <= 1e-6
    a = 0; b = top;
    while (a < b) {
        const unsigned long long mid = a + (b - a + 1) / 2;
        if (f(from_bits(mid)) <= 1e-6) a = mid; else b = mid - 1;
    }
    *hi_out = from_bits(a);
}


