// dgnn_api.cu -- C ABI of libdgnn_b200.so (include/dgnn_cabi.h): context, device arena, build orchestration (options:
// dgnn_options.cu).
// No CPU fallback: every entry point needs a CUDA device; failures are reported as status codes + message.
#include "dgnn_ctx.h"

extern "C" {

int dgnn_abi_version(void) { return DGNN_ABI_VERSION; }

const char* dgnn_strerror(int code) {
    switch (code) {
        case DGNN_OK: return "ok";
        case DGNN_ERR_INVALID_ARG: return "invalid argument";
        case DGNN_ERR_CUDA: return "CUDA runtime error";
        case DGNN_ERR_OOM: return "device out of memory";
        case DGNN_ERR_SINGULAR: return "singular or non-finite lattice";
        case DGNN_ERR_IMAGE_RANGE: return "reference image range exceeds 127";
        case DGNN_ERR_GEOMETRY: return "cell too thin for the supported stencil reach";
        case DGNN_ERR_UNWRAPPED: return "atoms too far outside the unit cell";
        case DGNN_ERR_DEGREE: return "row degree exceeds the supported maximum";
        case DGNN_ERR_TOO_LARGE: return "batch too large: split it";
        case DGNN_ERR_UNSUPPORTED: return "unsupported";
        case DGNN_ERR_BUSY: return "DLPack exports of this context are still alive";
        default: return "unknown error";
    }
}

const char* dgnn_last_error(const dgnn_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int dgnn_create(int device, size_t arena_bytes, dgnn_ctx** out) {
    if (!out) return DGNN_ERR_INVALID_ARG;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        cudaGetLastError();
        return DGNN_ERR_CUDA;  // no CPU fallback by design
    }
    if (device < 0 || device >= ndev) return DGNN_ERR_INVALID_ARG;
    dgnn_ctx* ctx = new (std::nothrow) dgnn_ctx();
    if (!ctx) return DGNN_ERR_OOM;
    ctx->device = device;
    auto bail = [&](int code) {
        dgnn_destroy(ctx);
        return code;
    };
    if (cudaSetDevice(device) != cudaSuccess) return bail(DGNN_ERR_CUDA);
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return bail(DGNN_ERR_CUDA);
    if (prop.major < 10) return bail(DGNN_ERR_UNSUPPORTED);  // sm_100a cubin only
    ctx->sm_count = prop.multiProcessorCount;
    ctx->smem_optin = (int)prop.sharedMemPerBlockOptin;
    ctx->smem_per_sm = (int)prop.sharedMemPerMultiprocessor;
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) return bail(DGNN_ERR_CUDA);
    if (cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess) return bail(DGNN_ERR_CUDA);
    for (auto& e : ctx->copy_ev)
        if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return bail(DGNN_ERR_CUDA);
    if (cudaEventCreateWithFlags(&ctx->copy_meta_ev, cudaEventDisableTiming) != cudaSuccess) return bail(DGNN_ERR_CUDA);
    if (cudaStreamCreateWithFlags(&ctx->aux_stream, cudaStreamNonBlocking) != cudaSuccess) return bail(DGNN_ERR_CUDA);
    if (cudaEventCreateWithFlags(&ctx->aux_fork, cudaEventDisableTiming) != cudaSuccess) return bail(DGNN_ERR_CUDA);
    if (cudaEventCreateWithFlags(&ctx->aux_join, cudaEventDisableTiming) != cudaSuccess) return bail(DGNN_ERR_CUDA);
    for (auto& e : ctx->ev.e)
        if (cudaEventCreate(&e) != cudaSuccess) return bail(DGNN_ERR_CUDA);
    if (cudaMallocHost(&ctx->h_stats, sizeof(DgStats)) != cudaSuccess) return bail(DGNN_ERR_CUDA);
    if (ensure(ctx, ctx->stats, sizeof(DgStats)) != DGNN_OK) return bail(DGNN_ERR_OOM);
    if (arena_bytes > 0) {
        // split the hint the way a K = 40 build uses it: 160 B features + 16 B edge arrays per edge
        const size_t edges = arena_bytes / 176;
        if (ensure(ctx, ctx->rbf, edges * 160) != DGNN_OK || ensure(ctx, ctx->dist, edges * 8) != DGNN_OK ||
            ensure(ctx, ctx->dst, edges * 4) != DGNN_OK || ensure(ctx, ctx->image, edges * 4) != DGNN_OK)
            return bail(DGNN_ERR_OOM);
    }
    *out = ctx;
    return DGNN_OK;
}

void dgnn_destroy(dgnn_ctx* ctx) {
    if (!ctx) return;
    if (ctx->live_exports.load() > 0 && !ctx->zombie.exchange(true)) return;  // the last DLPack deleter comes back here
    if (ctx->twin) {
        dgnn_destroy(ctx->twin);
        ctx->twin = nullptr;
    }
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    DevBuf* bufs[] = {&ctx->in_off, &ctx->in_x, &ctx->in_y, &ctx->in_z, &ctx->in_lat, &ctx->small, &ctx->fa, &ctx->fb, &ctx->fc, &ctx->geoms, &ctx->astruct,
                      &ctx->acell, &ctx->ncells32, &ctx->nitems32, &ctx->cell_base, &ctx->item_base, &ctx->scratch,
                      &ctx->cellcount, &ctx->cellstart, &ctx->cx, &ctx->cy, &ctx->cz, &ctx->cj, &ctx->cn,
                      &ctx->deg, &ctx->deg_raw, &ctx->row_perm, &ctx->cmask, &ctx->cmask_tn, &ctx->hits, &ctx->hits_d2, &ctx->hits_n, &ctx->hit_thr, &ctx->lb_status, &ctx->row_ptr, &ctx->seo, &ctx->dst, &ctx->image, &ctx->dist, &ctx->rbf,
                      &ctx->tmp_in, &ctx->tmp_out, &ctx->stats, &ctx->col_x, &ctx->col_ei, &ctx->col_batch, &ctx->col_keys,
                      &ctx->col_lut, &ctx->col_table, &ctx->col_err, &ctx->view_a, &ctx->view_b, &ctx->tab_cnt, &ctx->tab_ptr,
                      &ctx->tab_val};
    for (DevBuf* b : bufs) release(*b);
    if (ctx->h_stats) cudaFreeHost(ctx->h_stats);
    for (auto& e : ctx->ev.e)
        if (e) cudaEventDestroy(e);
    for (auto& e : ctx->cev)
        if (e) cudaEventDestroy(e);
    for (auto& e : ctx->copy_ev)
        if (e) cudaEventDestroy(e);
    if (ctx->copy_meta_ev) cudaEventDestroy(ctx->copy_meta_ev);
    if (ctx->aux_fork) cudaEventDestroy(ctx->aux_fork);
    if (ctx->aux_join) cudaEventDestroy(ctx->aux_join);
    if (ctx->aux_stream) cudaStreamDestroy(ctx->aux_stream);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

void* dgnn_stream(dgnn_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

uint64_t dgnn_build_serial(const dgnn_ctx* ctx) { return ctx ? ctx->build_serial : 0; }

int dgnn_host_alloc(size_t bytes, void** out) {
    if (!out) return DGNN_ERR_INVALID_ARG;
    *out = nullptr;
    if (cudaMallocHost(out, bytes > 0 ? bytes : 1) != cudaSuccess) {
        cudaGetLastError();
        *out = nullptr;
        return DGNN_ERR_OOM;
    }
    return DGNN_OK;
}
void dgnn_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

// (dgnn_stream.cu) second context on the same device with the same options: the other arena of the streamed build
dgnn_ctx* dgnn_internal_twin(dgnn_ctx* ctx) {
    if (!ctx) return nullptr;
    if (!ctx->twin) {
        dgnn_ctx* t = nullptr;
        if (dgnn_create(ctx->device, 0, &t) != DGNN_OK) return nullptr;
        for (const auto& kv : ctx->opt_log) dgnn_set_option(t, kv.first.c_str(), kv.second);
        ctx->twin = t;
    }
    return ctx->twin;
}
void dgnn_internal_set_error(dgnn_ctx* ctx, const char* msg) {
    if (ctx) ctx->err = msg ? msg : "";
}


int dgnn_set_offset_sinks(dgnn_ctx* ctx, int64_t* host_row_ptr, int64_t* host_struct_edge_offsets) {
    if (!ctx) return DGNN_ERR_INVALID_ARG;
    ctx->sink_row_ptr = host_row_ptr;
    ctx->sink_seo = host_struct_edge_offsets;
    return DGNN_OK;
}

int dgnn_last_build_info(const dgnn_ctx* ctx, dgnn_build_info* out) {
    if (!ctx || !out) return DGNN_ERR_INVALID_ARG;
    *out = ctx->info;
    return DGNN_OK;
}

int dgnn_build(dgnn_ctx* ctx, const dgnn_batch* b, dgnn_graphs* out) {
    if (!ctx) return DGNN_ERR_INVALID_ARG;
    if (!b || !out) return fail(ctx, DGNN_ERR_INVALID_ARG, "null batch / output");
    if (ctx->live_exports.load() > 0)
        return fail(ctx, DGNN_ERR_BUSY, "dgnn_build: %d DLPack export(s) still borrow this context's buffers", ctx->live_exports.load());
    memset(out, 0, sizeof(*out));
    ctx->err.clear();
    ctx->last.valid = false;
    const long long S = b->n_structs;
    if (S < 0) return fail(ctx, DGNN_ERR_INVALID_ARG, "n_structs < 0");
    if (!(b->r_cutoff > 0.0) || !std::isfinite(b->r_cutoff))
        return fail(ctx, DGNN_ERR_INVALID_ARG, "r_cutoff must be positive and finite");
    if (!b->atom_offsets) return fail(ctx, DGNN_ERR_INVALID_ARG, "atom_offsets is null");
    if (S >= 0x7fffffffLL) return fail(ctx, DGNN_ERR_TOO_LARGE, "too many structures");
    const bool want_rbf = !(b->flags & DGNN_SKIP_RBF) && b->dr > 0.0;
    int K = 0;
    if (want_rbf) {
        if (!(b->rbf_r_cutoff > 0.0) || !std::isfinite(b->rbf_r_cutoff))
            return fail(ctx, DGNN_ERR_INVALID_ARG, "rbf_r_cutoff must be positive and finite");
        const double kf = std::floor(b->rbf_r_cutoff / b->dr);  // edge_features.cpp:10
        if (!(kf >= 1.0 && kf <= 65536.0)) return fail(ctx, DGNN_ERR_INVALID_ARG, "RBF bin count out of range");
        K = (int)kf;
    }
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    dg_num_kernel_launches_reset();
    dg_rbf_launch_count_reset();
    dg_direct_launch_count_reset();
    memset(&ctx->info, 0, sizeof(ctx->info));

    // ---- inputs ----------------------------------------------------------------------------------------------
    const bool dev_in = (b->flags & DGNN_INPUT_DEVICE) != 0;
    long long N = 0;
    // chunked coordinate upload (host inputs): atoms [up_a[c], up_a[c+1]) arrive with copy_ev[c].  The guard keeps the
    // caller's buffers from being released under an in-flight copy when the build returns early.
    int n_up = 1;
    long long up_a[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    struct UploadGuard {
        cudaStream_t s;
        bool armed;
        ~UploadGuard() {
            if (armed) cudaStreamSynchronize(s);
        }
    } upload{ctx->copy_stream, false};
    CU(cudaEventRecord(ctx->ev.e[0], st));
    const long long* d_off;
    const double *d_x, *d_y, *d_z, *d_lat;
    if (dev_in) {
        // N and the structure-aligned part boundaries of the pipelined plan
        const int np = S >= ctx->upload_chunks ? ctx->upload_chunks : 1;
        for (int c = 0; c <= np; ++c)
            CU(cudaMemcpyAsync(&up_a[c], b->atom_offsets + S * c / np, sizeof(long long), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        N = up_a[np];
        if (N >= ctx->upload_min_atoms && np > 1) {
            n_up = np;
        } else {
            up_a[0] = 0;
            up_a[1] = N;
        }
        d_off = (const long long*)b->atom_offsets;
        d_x = b->x; d_y = b->y; d_z = b->z; d_lat = b->lattice;
    } else {
        if (b->atom_offsets[0] != 0) return fail(ctx, DGNN_ERR_INVALID_ARG, "atom_offsets[0] must be 0");
        for (long long s = 0; s < S; ++s)
            if (b->atom_offsets[s + 1] < b->atom_offsets[s])
                return fail(ctx, DGNN_ERR_INVALID_ARG, "atom_offsets not monotone at structure %lld", s);
        N = b->atom_offsets[S];
    }
    if (N < 0 || N >= 0x7fffffffLL) return fail(ctx, DGNN_ERR_TOO_LARGE, "atom count %lld out of range", N);
    if (n_up == 1) up_a[1] = N;
    for (int c = 0; c < n_up; ++c)
        if (up_a[c + 1] < up_a[c] || up_a[c] < 0 || up_a[c + 1] > N)
            return fail(ctx, DGNN_ERR_INVALID_ARG, "atom_offsets not monotone");
    if (N > 0 && (!b->x || !b->y || !b->z)) return fail(ctx, DGNN_ERR_INVALID_ARG, "coordinate pointer is null");
    if (S > 0 && !b->lattice) return fail(ctx, DGNN_ERR_INVALID_ARG, "lattice pointer is null");
    if (!dev_in) {
        ENSURE(ctx->in_off, (size_t)(S + 1) * 8);
        ENSURE(ctx->in_x, (size_t)N * 8 + 8);
        ENSURE(ctx->in_y, (size_t)N * 8 + 8);
        ENSURE(ctx->in_z, (size_t)N * 8 + 8);
        ENSURE(ctx->in_lat, (size_t)S * 72 + 8);
        // Everything goes up on ONE queue (the copy stream) in the order it is needed: offsets and lattices first (the
        // geometry kernel waits for copy_meta_ev), then the coordinates in structure-aligned chunks, which the prep +
        // count launches below wait for chunk by chunk.  (Small copies on the compute stream would queue behind the
        // coordinate transfers in the copy engine: measured 0.75 ms for 0.8 MB.)
        upload.armed = true;
        CU(cudaMemcpyAsync(ctx->in_off.p, b->atom_offsets, (size_t)(S + 1) * 8, cudaMemcpyHostToDevice, ctx->copy_stream));
        if (S > 0) CU(cudaMemcpyAsync(ctx->in_lat.p, b->lattice, (size_t)S * 72, cudaMemcpyHostToDevice, ctx->copy_stream));
        CU(cudaEventRecord(ctx->copy_meta_ev, ctx->copy_stream));
        CU(cudaStreamWaitEvent(st, ctx->copy_meta_ev, 0));
        if (N >= ctx->upload_min_atoms && S >= ctx->upload_chunks) n_up = ctx->upload_chunks;
        for (int c = 0; c <= n_up; ++c) up_a[c] = b->atom_offsets[S * c / n_up];
        for (int c = 0; c < n_up; ++c) {
            const long long a0 = up_a[c], n = up_a[c + 1] - up_a[c];
            if (n > 0) {
                upload.armed = true;
                CU(cudaMemcpyAsync((double*)ctx->in_x.p + a0, b->x + a0, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->copy_stream));
                CU(cudaMemcpyAsync((double*)ctx->in_y.p + a0, b->y + a0, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->copy_stream));
                CU(cudaMemcpyAsync((double*)ctx->in_z.p + a0, b->z + a0, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->copy_stream));
            }
            CU(cudaEventRecord(ctx->copy_ev[c], ctx->copy_stream));
        }
        d_off = (const long long*)ctx->in_off.p;
        d_x = (const double*)ctx->in_x.p; d_y = (const double*)ctx->in_y.p; d_z = (const double*)ctx->in_z.p;
        d_lat = (const double*)ctx->in_lat.p;
    }
    CU(cudaEventRecord(ctx->ev.e[1], st));

    // ---- geometry, classification, per-atom fractional coordinates ------------------------------------------
    ENSURE(ctx->small, (size_t)(S + 1) * sizeof(DgSmall));
    ENSURE(ctx->astruct, (size_t)(N + 1) * 4);
    ENSURE(ctx->fa, (size_t)(N + 1) * 8);
    ENSURE(ctx->fb, (size_t)(N + 1) * 8);
    ENSURE(ctx->fc, (size_t)(N + 1) * 8);
    ENSURE(ctx->row_ptr, (size_t)(N + 2) * 8);
    ENSURE(ctx->deg, (size_t)(N + 1) * 4);
    ENSURE(ctx->deg_raw, (size_t)(N + 1) * 4);
    ENSURE(ctx->seo, (size_t)(S + 2) * 8);
    {
        const long long mx = (N > S ? N : S) + 1;
        ENSURE(ctx->scratch, (size_t)(mx / 2048 + 4) * 8);
    }
    CU(cudaMemsetAsync(ctx->stats.p, 0, sizeof(DgStats), st));
    DgStats* d_stats = (DgStats*)ctx->stats.p;
    DgSmall* small = (DgSmall*)ctx->small.p;
    const int eps_mode = !(b->epsilon > 0.0) ? 0 : (b->epsilon <= DG_EPS_FAST ? 1 : 2);
    dg_launch_small_geom(st, S, d_off, d_lat, b->r_cutoff, ctx->direct_max_atoms, eps_mode == 1 ? b->epsilon : 0.0,
                         small, d_stats);
    // hit lists: the count pass records every accepted (atom, image) pair so that the fill pass does not repeat the
    // traversal.  Slots per row come from the density estimate of the degree (rows that outgrow them, and structures
    // whose image range does not fit the 5-bit fields, are filled by the traversal pass instead).
    int hit_cap = ctx->hit_cap_opt > 0 ? round_up(ctx->hit_cap_opt, 32) : 0;  // rows stay 16-byte aligned
    double hit_target = 0.0;  // > 0: lists of truncated rows hold ~this many pairs (per-structure d2 threshold)
    bool have_est = false;
    if ((ctx->hit_cap_opt < 0 || ctx->onepass_opt != 0) && N > 0) {
        int r = fetch_stats(ctx);  // sync: density estimate, classification, input errors
        if (r != DGNN_OK) return r;
        have_est = true;
    }
    if (ctx->hit_cap_opt < 0 && N > 0) {
        const long long est = ctx->h_stats->est_deg;
        // Worth it only where the fill pass is bound by the traversal, not by the bytes it writes: the lists cost the
        // count pass ~35 % more time and the fill pass 12 B of extra reads per accepted pair.  Measured on C2: uncapped
        // rows with K = 40 features (176 B written per pair) are HBM-bound either way (traversal fill 8.5 ms vs hit-list
        // fill 9.1 ms + 0.8 ms in the count pass); capped or feature-less rows are 3-5x faster from the lists.
        const double keep_est = (double)(est < max_nb_for_est(b->max_neighbors) ? est : max_nb_for_est(b->max_neighbors));
        const double bytes_per_pair = est > 0 ? keep_est * (16.0 + 4.0 * K) / (double)est : 1e9;
        // ... and always for cell-list structures, whose tile-staged traversal is the expensive part
        // (small direct-path batches take the single-pass build instead, see below)
        const bool onepass_likely = ctx->onepass_opt > 0 ||
                                    (ctx->onepass_opt < 0 && N <= (long long)ctx->sm_count * 2 * 8 * 2);
        // (the fp32-tile traversal of cell-list structures needs no lists; only the fp64 one, option cell32 = 0, does)
        if (est > 0 && ((ctx->h_stats->n_cell_structs > 0 && !ctx->cell32) || (bytes_per_pair < 64.0 && !onepass_likely)))
            hit_cap = round_up((int)(est + est / 4 + 16 < DGNN_MAX_DEGREE ? est + est / 4 + 16 : DGNN_MAX_DEGREE), 32);
        // Truncated rows (the reference's default max_neighbors = 20 of ~160-330 accepted pairs): the lists only need
        // the pairs that can reach the first max_neighbors + 1 ranks.  The count pass stores the pairs below a
        // per-structure d2 threshold set for ~target of them (mean + 4 sigma of a Poisson count, plus the margin the
        // exactness argument needs) and the slots shrink to match; rows the threshold fails go to the traversal fill.
        if (hit_cap > 0 && ctx->hit_threshold && (ctx->h_stats->n_cell_structs == 0 || ctx->cell32)) {
            const double need = (double)max_nb_for_est(b->max_neighbors) + 1.0;
            const double target = need + 4.0 * sqrt(need) + 8.0;
            const int slots = round_up((int)(target + 6.0 * sqrt(target) + 16.0), 32);
            if (slots < hit_cap && target < 0.4 * (double)est) {
                hit_target = target;
                hit_cap = slots;
            }
        }
    }
    if (hit_cap > 512) hit_cap = 512;  // the hit-list fill pass sorts in registers: rows of up to 512 entries
    if (hit_cap > 0 && (double)N * (double)hit_cap * 12.0 > ctx->hits_max_bytes) hit_cap = 0;
    if (hit_cap == 0) hit_target = 0.0;
    if (hit_cap > 0) {
        ENSURE(ctx->hits, (size_t)N * (size_t)hit_cap * 4 + 16);
        ENSURE(ctx->hits_d2, (size_t)N * (size_t)hit_cap * 8 + 16);
    }
    if (hit_target > 0.0) {
        ENSURE(ctx->hits_n, (size_t)(N + 1) * 4);
        ENSURE(ctx->hit_thr, (size_t)(S + 1) * 8);
        dg_launch_hit_thresholds(st, S, small, hit_target, b->r_cutoff * b->r_cutoff, (double*)ctx->hit_thr.p);
    }
    CU(cudaEventRecord(ctx->ev.e[2], st));

    // ---- common traversal parameters -------------------------------------------------------------------------
    const double r2 = b->r_cutoff * b->r_cutoff;  // neighbor_list.cpp:40
    int near_lo_hi, near_hi_hi;
    {
        const double lo = b->r_cutoff - 2e-6 > 0 ? b->r_cutoff - 2e-6 : 0.0, hi = b->r_cutoff + 2e-6;
        const double lo2 = lo * lo, hi2 = hi * hi;
        long long lb, hb;
        memcpy(&lb, &lo2, 8);
        memcpy(&hb, &hi2, 8);
        near_lo_hi = (int)(lb >> 32);
        near_hi_hi = (int)(hb >> 32);
    }
    const int max_nb = b->max_neighbors > 0x7fffffffULL ? 0x7fffffff : (int)b->max_neighbors;

    // ---- direct path: count pass -----------------------------------------------------------------------------
    DgDirectArgs D;
    memset(&D, 0, sizeof(D));
    D.small = small;
    D.astruct = (const int*)ctx->astruct.p;
    D.n_atoms = N;
    D.x = d_x; D.y = d_y; D.z = d_z;
    D.fa = (const double*)ctx->fa.p; D.fb = (const double*)ctx->fb.p; D.fc = (const double*)ctx->fc.p;
    D.r2 = r2;
    D.r_cutoff = b->r_cutoff;
    D.epsilon = b->epsilon;
    D.inv_rc = (float)(1.0 / b->r_cutoff);
    D.key_scale = (4294967296.0 - 4096.0) / b->r_cutoff;
    D.near_lo_hi = near_lo_hi;
    D.near_hi_hi = near_hi_hi;
    D.eps_mode = eps_mode;
    near_band_d2(b->r_cutoff, &D.near_d2_lo, &D.near_d2_hi);
    D.max_neighbors = max_nb;
    D.n_warps = 8;
    D.n_buckets = 32;
    D.deg_capped = (int*)ctx->deg.p;
    D.deg_raw = (int*)ctx->deg_raw.p;
    D.hits = hit_cap > 0 ? (uint32_t*)ctx->hits.p : nullptr;
    D.hits_d2 = hit_cap > 0 ? (double*)ctx->hits_d2.p : nullptr;
    D.hit_cap = hit_cap;
    D.hit_d2_max = hit_target > 0.0 ? (const double*)ctx->hit_thr.p : nullptr;
    D.hits_n = hit_target > 0.0 ? (int*)ctx->hits_n.p : nullptr;
    D.stats = d_stats;
    // small batches (fewer than two rounds of 64-row chunks over a full grid): rows go to warps a few at a time from a
    // device counter -- uneven rows (tiny triclinic cells: 2-8 atoms, 343-729 images per pair) otherwise leave most of
    // the grid waiting for the CTAs that drew the expensive structures
    const bool dyn_rows = ctx->dyn_rows && N > 0 && N < (long long)ctx->sm_count * 2 * 8 * 64 * 2;
    int dyn_launch = 0;
    auto dyn_arm = [&](DgDirectArgs& X) {
        if (dyn_rows && dyn_launch < 8) {
            X.dyn_unit = 4;
            X.unit_ctr = &((DgStats*)ctx->stats.p)->unit_ctr[dyn_launch++];
        } else {
            X.dyn_unit = 0;
            X.unit_ctr = nullptr;
        }
    };
    // ---- single-pass build (direct-path batches): traversal + look-back row offsets + sort + emit in one kernel --
    // The edge arena and the row buffers are sized from the density estimate; rows that outgrow the buffer, or an
    // arena that turns out too small, are completed by the regular fill pass afterwards (row_ptr is final either way).
    bool onepass = false, onepass_fused_rbf = false;
    int row_cap_1p = 0;
    // Auto policy: the look-back couples every CTA to its predecessors with one step of slack, which costs large batches
    // more than the count pass saves (C2 10 k structures: 11.8 ms vs 10.9 ms count + fill), but a batch of a few steps
    // per CTA -- the reference's one-structure-per-call API above all -- finishes in one kernel and one sync less
    // (9 shipped POSCARs: 0.31 ms vs 1.37 ms).
    const long long onepass_auto_rows = (long long)ctx->sm_count * 2 * 8 * 2;  // two steps of a full grid
    if (have_est && ctx->onepass_opt != 0 && !ctx->onepass_skip && N > 0 && ctx->h_stats->err_status == 0 &&
        ctx->h_stats->n_cell_structs == 0 &&
        (ctx->onepass_opt > 0 || (N <= onepass_auto_rows && (hit_cap == 0 || ctx->hit_cap_opt < 0)))) {
        const long long est = ctx->h_stats->est_deg;
        long long rc = est + est / 8 + 8;
        if (rc > DGNN_MAX_DEGREE) rc = DGNN_MAX_DEGREE;
        if (ctx->onepass_row_cap > 0) rc = ctx->onepass_row_cap;
        row_cap_1p = round_up((int)(rc > 32 ? rc : 32), 32);
        double est_e = (double)ctx->h_stats->est_edges;
        if (est_e > (double)N * (double)max_nb) est_e = (double)N * (double)max_nb;
        est_e = est_e * ctx->onepass_arena_scale + (ctx->onepass_arena_scale >= 1.0 ? 64.0 * 1024.0 : 0.0);
        DgDirectArgs O = D;
        O.hit_cap = 0;
        O.hits = nullptr;
        O.hits_d2 = nullptr;
        O.row_cap = row_cap_1p;
        O.n_buckets = ctx->sort_buckets > 0 ? ctx->sort_buckets : (row_cap_1p <= 256 ? 256 : (row_cap_1p <= 1024 ? 512 : 1024));
        O.n_warps = 8;
        while (O.n_warps > 2 && (int)dg_direct_onepass_smem_bytes(O) > ctx->smem_optin) O.n_warps >>= 1;
        // (an arena the device cannot hold fails in ENSURE with DGNN_ERR_OOM exactly as the two-pass build would: the
        // estimate is within a few percent of the edge total the fill pass needs anyway)
        if ((int)dg_direct_onepass_smem_bytes(O) <= ctx->smem_optin && est_e < 9.0e18) {
            const size_t e_est = (size_t)est_e;
            ENSURE(ctx->dst, (e_est + 1) * 4);
            ENSURE(ctx->image, (e_est + 1) * 4);
            ENSURE(ctx->dist, (e_est + 1) * 8);
            if (want_rbf) ENSURE(ctx->rbf, e_est * (size_t)K * 4 + 16);
            ENSURE(ctx->lb_status, (size_t)(N + 1) * 8);
            size_t cap_e = ctx->dst.cap / 4;
            if (ctx->image.cap / 4 < cap_e) cap_e = ctx->image.cap / 4;
            if (ctx->dist.cap / 8 < cap_e) cap_e = ctx->dist.cap / 8;
            if (want_rbf && (ctx->rbf.cap - 16) / ((size_t)K * 4) < cap_e) cap_e = (ctx->rbf.cap - 16) / ((size_t)K * 4);
            O.edge_cap = (long long)cap_e - 1;
            O.dst = (int*)ctx->dst.p;
            O.image = (uint32_t*)ctx->image.p;
            O.dist = (double*)ctx->dist.p;
            O.lb_status = (unsigned long long*)ctx->lb_status.p;
            O.lb_spin_limit = ctx->lb_spin_limit;
            O.row_ptr_out = (long long*)ctx->row_ptr.p;
            const int keep_max_1p = row_cap_1p < max_nb ? row_cap_1p : max_nb;
            onepass_fused_rbf = want_rbf && ctx->fuse_rbf && (K & 3) == 0 &&
                                (long long)keep_max_1p * (K / 4) < (1LL << 22);
            if (onepass_fused_rbf) {
                O.rbf = dg_make_rbf_epi(K, b->rbf_r_cutoff, b->dr, (float*)ctx->rbf.p);
                O.rbf.recur = ctx->rbf_recur;
            }
            CU(cudaMemsetAsync(ctx->lb_status.p, 0, (size_t)(N + 1) * 8, st));
            for (int c = 0; c < n_up; ++c) {
                if (!dev_in) CU(cudaStreamWaitEvent(st, ctx->copy_ev[c], 0));
                dg_launch_direct_prep(st, up_a[c], up_a[c + 1], S, d_off, d_x, d_y, d_z, small, (int*)ctx->astruct.p,
                                      (double*)ctx->fa.p, (double*)ctx->fb.p, (double*)ctx->fc.p, d_stats);
            }
            CU(cudaEventRecord(ctx->ev.e[3], st));
            CU(dg_launch_direct_onepass(st, O, ctx->sm_count, ctx->direct_min_blocks));
            onepass = true;
        }
    }
    // Skewed degrees: the fill pass sizes every warp's row buffer for the largest row, and a few dense rows (config 3:
    // mean degree 180, max 841) take it from 16 warps per SM to 4.  split_cap is the largest buffer that still gives
    // 2 CTAs x 8 warps; the count pass tallies the rows beyond it, and when they are few they get a launch of their own.
    int split_cap = 0, split_main_smem = 0;
    if (ctx->split_rows) {
        DgDirectArgs T = D;
        T.n_warps = 8;
        // (two main CTAs per SM, and one main CTA next to the large-row CTA of >= 4 warps of up to ~30 KB each)
        const int budget = (ctx->smem_per_sm - 3 * 1024) * 7 / 16;
        for (int rc = 512; rc >= 128; rc -= 32) {
            T.row_cap = rc;
            T.n_buckets = ctx->sort_buckets > 0 ? ctx->sort_buckets : (rc <= 256 ? 256 : 512);
            if ((int)dg_direct_smem_bytes(T, true) <= budget) {
                split_cap = rc;
                split_main_smem = (int)dg_direct_smem_bytes(T, true) + 2 * 1024;
                break;
            }
        }
    }
    // (a batch of cell-list structures only -- config 4 -- has no rows for the direct kernels: they are not launched)
    const bool any_direct = !have_est || ctx->h_stats->n_direct_structs > 0;
    if (!onepass && any_direct) {
        // per upload chunk (structure-aligned): fractional coordinates, then the count pass of its rows -- the next
        // chunk's coordinates cross PCIe meanwhile
        // without hit lists the count pass is symmetric: one sweep decides i -> j and j -> i (dgnn_direct.cu)
        const bool sym = ctx->sym_count && D.hit_cap == 0 && eps_mode <= 1;
        if (sym) {
            D.split_cap = split_cap;
            D.sym_band = (D.near_d2_hi - D.near_d2_lo) + 1.0e-9 * r2;
            CU(cudaMemsetAsync(ctx->deg_raw.p, 0, (size_t)(N + 1) * 4, st));
        }
        {
            bool probe_forked = false;
            for (int c = 0; c < n_up; ++c) {
                if (!dev_in) CU(cudaStreamWaitEvent(st, ctx->copy_ev[c], 0));
                dg_launch_direct_prep(st, up_a[c], up_a[c + 1], S, d_off, d_x, d_y, d_z, small, (int*)ctx->astruct.p,
                                      (double*)ctx->fa.p, (double*)ctx->fb.p, (double*)ctx->fc.p, d_stats);
                // rows that may go through the uncapped fill kernel: sample their sort keys for ties -- beside the
                // last chunk's count pass, on the auxiliary stream (the result arrives with the statistics of sync 1)
                if (c == n_up - 1 && ctx->tie_probe == 1 && max_nb >= 32) {
                    CU(cudaEventRecord(ctx->aux_fork, st));
                    CU(cudaStreamWaitEvent(ctx->aux_stream, ctx->aux_fork, 0));
                    CU(dg_launch_tie_probe(ctx->aux_stream, D));
                    CU(cudaEventRecord(ctx->aux_join, ctx->aux_stream));
                    probe_forked = true;
                }
                if (up_a[c + 1] > up_a[c]) {
                    D.row_begin = up_a[c];
                    D.row_end = up_a[c + 1];
                    if (sym) CU(dg_launch_direct_count_sym(st, D, ctx->sm_count, ctx->direct_min_blocks));
                    else {
                        dyn_arm(D);
                        CU(dg_launch_direct(st, D, false, ctx->sm_count, ctx->direct_min_blocks));
                    }
                }
            }
            D.row_begin = D.row_end = 0;
            if (sym) dg_launch_cap_degrees(st, D, ctx->sm_count);  // one launch for all chunks' rows
            dg_scan_i32_i64(st, (const int*)ctx->deg.p, (long long*)ctx->row_ptr.p, N, (long long*)ctx->scratch.p);
            if (split_cap > 0 && D.hit_cap == 0 && !sym)  // (the symmetric count's k_cap_degrees has tallied them)
                dg_launch_count_big_rows(st, N, (const int*)ctx->astruct.p, small, (const int*)ctx->deg_raw.p, split_cap,
                                         d_stats, ctx->sm_count);
            if (probe_forked) CU(cudaStreamWaitEvent(st, ctx->aux_join, 0));
        }
    } else if (!onepass) {
        // the cell-list kernels still need the uploaded coordinates, and empty batches a valid (all zero) row_ptr
        for (int c = 0; c < n_up; ++c)
            if (!dev_in) CU(cudaStreamWaitEvent(st, ctx->copy_ev[c], 0));
        CU(cudaMemsetAsync(ctx->deg.p, 0, (size_t)(N + 1) * 4, st));
        CU(cudaMemsetAsync(ctx->deg_raw.p, 0, (size_t)(N + 1) * 4, st));
        dg_scan_i32_i64(st, (const int*)ctx->deg.p, (long long*)ctx->row_ptr.p, N, (long long*)ctx->scratch.p);
    }
    dg_launch_finish_rows(st, N, S, d_off, (const long long*)ctx->row_ptr.p, (long long*)ctx->seo.p, d_stats);
    CU(cudaGetLastError());
    {
        int r = fetch_stats(ctx);  // sync 1: errors, cell-path structures, E and the largest row (direct-only case)
        if (r != DGNN_OK) return r;
    }
    if (ctx->h_stats->err_status != 0)
        return fail(ctx, status_to_code(ctx->h_stats->err_status), "structure %d: %s", ctx->h_stats->err_struct,
                    dgnn_strerror(status_to_code(ctx->h_stats->err_status)));
    if (ctx->h_stats->lb_timeout) {
        // The look-back of the single-pass build waits for predecessor CTAs; if something else holds SMs (another
        // stream of the process, an MPS client) the missing CTAs may never start within the spin bound and the row
        // offsets are then wrong.  Redo the batch with count + scan + fill, which needs no co-residency.
        if (ctx->onepass_skip) return fail(ctx, DGNN_ERR_CUDA, "internal: single-pass look-back timed out");
        if (upload.armed) CU(cudaStreamSynchronize(ctx->copy_stream));
        ctx->onepass_skip = true;
        const int r = dgnn_build(ctx, b, out);
        ctx->onepass_skip = false;
        if (r == DGNN_OK) {
            ctx->lb_fallbacks += 1;
            ctx->info.onepass = 3;
        }
        return r;
    }
    const int n_cell_structs = ctx->h_stats->n_cell_structs;
    int n_fallback_rows = ctx->h_stats->n_fallback_rows;
    const bool onepass_redo_all = onepass && ctx->h_stats->cap_overflow != 0;

    // ---- cell-list path (large structures): binning + count pass ---------------------------------------------
    DgTraverseArgs A;
    memset(&A, 0, sizeof(A));
    long long n_items = 0, P = 0;
    int grid = 1;
    bool cell32 = false;
    int tile_cap_report = 0;
    if (n_cell_structs > 0) {
        ENSURE(ctx->geoms, (size_t)(S + 1) * sizeof(DgGeom));
        ENSURE(ctx->acell, (size_t)(N + 1) * 12);
        ENSURE(ctx->ncells32, (size_t)(S + 1) * 4);
        ENSURE(ctx->nitems32, (size_t)(S + 1) * 4);
        ENSURE(ctx->cell_base, (size_t)(S + 2) * 8);
        ENSURE(ctx->item_base, (size_t)(S + 2) * 8);
        DgGeom* geoms = (DgGeom*)ctx->geoms.p;
        dg_launch_geom(st, S, d_off, d_lat, b->r_cutoff, ctx->cell_frac, small, geoms);
        dg_launch_atom_cells(st, N, S, d_off, d_x, d_y, d_z, geoms, (int*)ctx->astruct.p, (int*)ctx->acell.p);
        dg_launch_plan(st, S, geoms, ctx->tile_target, ctx->src_target, (int*)ctx->ncells32.p,
                       (int*)ctx->nitems32.p, d_stats);
        dg_scan_i32_i64(st, (const int*)ctx->ncells32.p, (long long*)ctx->cell_base.p, S,
                        (long long*)ctx->scratch.p);
        dg_scan_i32_i64(st, (const int*)ctx->nitems32.p, (long long*)ctx->item_base.p, S,
                        (long long*)ctx->scratch.p);
        dg_launch_plan_finish(st, S, geoms, (const long long*)ctx->cell_base.p, (const long long*)ctx->item_base.p,
                              d_stats);
        CU(cudaGetLastError());
        {
            int r = fetch_stats(ctx);  // sync: totals for the cell arrays
            if (r != DGNN_OK) return r;
        }
        if (ctx->h_stats->err_status != 0)
            return fail(ctx, status_to_code(ctx->h_stats->err_status), "structure %d: %s", ctx->h_stats->err_struct,
                        dgnn_strerror(status_to_code(ctx->h_stats->err_status)));
        const long long n_cells = ctx->h_stats->n_cells;
        n_items = ctx->h_stats->n_items;
        if (n_cells >= 0x7fffffffLL)
            return fail(ctx, DGNN_ERR_TOO_LARGE, "%lld grid cells: split the batch", n_cells);

        // K1: cell-list binning of the pruned image cloud
        ENSURE(ctx->cellcount, (size_t)(n_cells + 1) * 4);
        ENSURE(ctx->cellstart, (size_t)(n_cells + 2) * 8);
        {
            const size_t need = (size_t)(n_cells / 2048 + 4) * 8;
            ENSURE(ctx->scratch, need);
        }
        CU(cudaMemsetAsync(ctx->cellcount.p, 0, (size_t)(n_cells + 1) * 4, st));
        dg_launch_cloud_count(st, N, geoms, (const int*)ctx->astruct.p, (const int*)ctx->acell.p,
                              (int*)ctx->cellcount.p);
        dg_scan_i32_i64(st, (const int*)ctx->cellcount.p, (long long*)ctx->cellstart.p, n_cells,
                        (long long*)ctx->scratch.p);
        dg_launch_item_stats(st, n_items, S, geoms, (const long long*)ctx->item_base.p,
                             (const long long*)ctx->cellstart.p, n_cells, d_stats);
        CU(cudaGetLastError());
        {
            int r = fetch_stats(ctx);  // sync: cloud size, tile sizes
            if (r != DGNN_OK) return r;
        }
        P = ctx->h_stats->cloud_points;
        if (P >= 0x7fffffffLL) return fail(ctx, DGNN_ERR_TOO_LARGE, "%lld image-cloud points: split the batch", P);
        ENSURE(ctx->cx, (size_t)(P + 1) * 8);
        ENSURE(ctx->cy, (size_t)(P + 1) * 8);
        ENSURE(ctx->cz, (size_t)(P + 1) * 8);
        ENSURE(ctx->cj, (size_t)(P + 1) * 4);
        ENSURE(ctx->cn, (size_t)(P + 1) * 4);
        dg_launch_cloud_scatter(st, N, geoms, (const int*)ctx->astruct.p, (const int*)ctx->acell.p, d_x, d_y, d_z,
                                (const long long*)ctx->cellstart.p, (int*)ctx->cellcount.p, (double*)ctx->cx.p,
                                (double*)ctx->cy.p, (double*)ctx->cz.p, (int*)ctx->cj.p, (uint32_t*)ctx->cn.p);

        // K2: count pass
        A.geoms = geoms;
        A.item_base = (const long long*)ctx->item_base.p;
        A.n_items = n_items;
        A.n_structs = S;
        A.cellstart = (const long long*)ctx->cellstart.p;
        A.cx = (const double*)ctx->cx.p; A.cy = (const double*)ctx->cy.p; A.cz = (const double*)ctx->cz.p;
        A.cj = (const int*)ctx->cj.p; A.cn = (const uint32_t*)ctx->cn.p;
        A.r2 = r2;
        A.r_cutoff = b->r_cutoff;
        A.rcm = DG_RCM(b->r_cutoff);
        A.epsilon = b->epsilon;
        A.inv_rc = (float)(1.0 / b->r_cutoff);
        A.key_scale = D.key_scale;
        A.near_lo_hi = near_lo_hi;
        A.near_hi_hi = near_hi_hi;
        A.max_neighbors = max_nb;
        A.deg_capped = (int*)ctx->deg.p;
        A.deg_raw = (int*)ctx->deg_raw.p;
        A.hits = D.hits;
        A.hits_d2 = D.hits_d2;
        A.hit_cap = hit_cap;
        A.stats = d_stats;
        // shared-memory plan: tile sized to the largest actual tile (bounded), rest goes the unstaged way
        int tile_cap = round_up(ctx->h_stats->max_tile_pts > 32 ? ctx->h_stats->max_tile_pts : 32, 32);
        if (tile_cap > ctx->tile_limit) tile_cap = ctx->tile_limit;
        if (tile_cap < 32) tile_cap = 32;  // tile_limit 0: everything unstaged, keep a token tile
        int tcs_cap = round_up(ctx->h_stats->max_tile_tcs > 4 ? ctx->h_stats->max_tile_tcs : 4, 4);
        if (tcs_cap > 4096) tcs_cap = 4096;
        A.tile_cap = ctx->tile_limit == 0 ? 0 : tile_cap;
        A.tcs_cap = tcs_cap;
        A.src_cap = 1024;
        A.row_cap = 0;
        A.n_buckets = 32;
        A.n_warps = 8;
        grid = (int)((long long)ctx->sm_count * ctx->traverse_grid_mult < n_items
                         ? (long long)ctx->sm_count * ctx->traverse_grid_mult
                         : (n_items > 0 ? n_items : 1));
        // fp32-tile traversal: the whole tile of every item must fit (float4 per point, 16-bit slots)
        {
            DgTraverseArgs C = A;
            C.tile_cap = round_up(ctx->h_stats->max_tile_pts > 32 ? ctx->h_stats->max_tile_pts : 32, 32);
            C.row_cap = 0;
            C.src_cap = 1024;  // (fp32-tile traversal: slot list entries per warp)
            cell32 = ctx->cell32 && C.tile_cap <= 65504 && ctx->h_stats->max_tile_tcs <= 4096 &&
                     (int)dg_cell32_smem_bytes(C, false) <= ctx->smem_optin;
            if (cell32) {
                A.tile_cap = C.tile_cap;
                A.src_cap = 1024;
                A.hits = nullptr;
                A.hits_d2 = nullptr;
                A.hit_cap = 0;
            }
        }
        tile_cap_report = A.tile_cap;
        if (cell32) {
            // candidate words for the fill pass: 132 B per atom of the batch (skipped beyond 2 GB)
            A.cmask = nullptr;
            A.cmask_tn = nullptr;
            if (ctx->cell32_mask && (double)N * 132.0 <= 2.0e9) {
                ENSURE(ctx->cmask, (size_t)N * 128 + 16);
                ENSURE(ctx->cmask_tn, (size_t)N * 4 + 16);
                A.cmask = (uint32_t*)ctx->cmask.p;
                A.cmask_tn = (int*)ctx->cmask_tn.p;
            }
            CU(dg_launch_cell32(st, A, false, grid));
        } else {
            if ((int)dg_traverse_smem_bytes(A, false) > ctx->smem_optin)
                return fail(ctx, DGNN_ERR_UNSUPPORTED, "count-pass shared memory plan does not fit");
            CU(dg_launch_traverse(st, A, false, grid));
        }
        dg_scan_i32_i64(st, (const int*)ctx->deg.p, (long long*)ctx->row_ptr.p, N, (long long*)ctx->scratch.p);
        dg_launch_finish_rows(st, N, S, d_off, (const long long*)ctx->row_ptr.p, (long long*)ctx->seo.p, d_stats);
        CU(cudaGetLastError());
        {
            int r = fetch_stats(ctx);  // sync: E and the largest row
            if (r != DGNN_OK) return r;
        }
        n_fallback_rows = ctx->h_stats->n_fallback_rows;
    }
    if (!onepass) CU(cudaEventRecord(ctx->ev.e[3], st));
    // offset sinks: row_ptr and the per-structure offsets are final here for the two-pass plans (the stream has just
    // been synchronised), so they travel back on the copy stream while the fill pass runs
    const bool sinks = ctx->sink_row_ptr != nullptr || ctx->sink_seo != nullptr;
    const bool sinks_early = sinks && !onepass;
    if (sinks_early) {
        upload.armed = true;  // (the guard also keeps these copies from outliving an early return)
        if (ctx->sink_row_ptr)
            CU(cudaMemcpyAsync(ctx->sink_row_ptr, ctx->row_ptr.p, (size_t)(N + 1) * 8, cudaMemcpyDeviceToHost, ctx->copy_stream));
        if (ctx->sink_seo)
            CU(cudaMemcpyAsync(ctx->sink_seo, ctx->seo.p, (size_t)(S + 1) * 8, cudaMemcpyDeviceToHost, ctx->copy_stream));
    }
    const long long E = ctx->h_stats->n_edges;
    const int max_deg = ctx->h_stats->max_deg;
    if (max_deg > DGNN_MAX_DEGREE)
        return fail(ctx, DGNN_ERR_DEGREE, "row degree %d exceeds %d: lower r_cutoff", max_deg, DGNN_MAX_DEGREE);

    // ---- fill pass (rows -> sorted CSR) ----------------------------------------------------------------------
    ENSURE(ctx->dst, (size_t)(E + 1) * 4);
    ENSURE(ctx->image, (size_t)(E + 1) * 4);
    ENSURE(ctx->dist, (size_t)(E + 1) * 8);
    int row_cap = round_up(max_deg > 32 ? max_deg : 32, 32);
    int n_buckets = ctx->sort_buckets > 0 ? ctx->sort_buckets : (row_cap <= 256 ? 256 : (row_cap <= 1024 ? 512 : 1024));
    // rows beyond split_cap are few: the bulk of the rows runs with the small row buffer (full occupancy), the rest
    // in a second launch with the large one
    const int row_cap_big = row_cap, n_buckets_big = n_buckets;
    const bool split_fill = !onepass && split_cap > 0 && hit_cap == 0 && row_cap > split_cap && any_direct &&
                            (long long)ctx->h_stats->n_big_rows * 8 <= N;
    if (split_fill) {
        row_cap = split_cap;
        n_buckets = ctx->sort_buckets > 0 ? ctx->sort_buckets : (row_cap <= 256 ? 256 : 512);
    }
    D.row_ptr = (const long long*)ctx->row_ptr.p;
    D.dst = (int*)ctx->dst.p;
    D.image = (uint32_t*)ctx->image.p;
    D.dist = (double*)ctx->dist.p;
    D.row_cap = row_cap;
    D.n_buckets = n_buckets;
    D.n_warps = 8;
    // mean number of entries sharing a truncated sort key over the probed rows: 1.0 for thermal / relaxed structures,
    // 10-30 on the coordination shells of an ideal lattice, where the compact ranking is quadratic
    const double tie_run_mean = ctx->h_stats->tie_n > 0 ? (double)ctx->h_stats->tie_sum / (double)ctx->h_stats->tie_n : 0.0;
    D.full_tie_sort = ctx->tie_probe == 2 || (ctx->tie_probe == 1 && tie_run_mean >= 3.0);
    // fused RBF epilogue: rows stream their K features right after the sort (needs K % 4 == 0 for the float4 stores)
    const int keep_max = max_deg < max_nb ? max_deg : max_nb;
    const bool fuse_rbf = onepass ? (onepass_fused_rbf && E > 0)
                                  : (want_rbf && E > 0 && ctx->fuse_rbf && (K & 3) == 0 &&
                                     (long long)keep_max * (K / 4) < (1LL << 22));
    if (want_rbf && E > 0) ENSURE(ctx->rbf, (size_t)E * (size_t)K * 4 + 16);
    if (fuse_rbf) {
        D.rbf = dg_make_rbf_epi(K, b->rbf_r_cutoff, b->dr, (float*)ctx->rbf.p);
        D.rbf.recur = ctx->rbf_recur;
    }
    while (D.n_warps > 1 && (int)dg_direct_smem_bytes(D, true) > ctx->smem_optin) D.n_warps >>= 1;
    if ((int)dg_direct_smem_bytes(D, true) > ctx->smem_optin)
        return fail(ctx, DGNN_ERR_DEGREE, "row buffer for degree %d does not fit in shared memory", max_deg);
    // rows with a stored hit list are filled from it; the traversal fill pass takes whatever is left
    DgDirectArgs H = D;
    H.fill_hits_bulk = ctx->fill_hits_bulk;
    H.fill_hits_warps = ctx->fill_hits_warps;
    if (n_cell_structs > 0 && !cell32) {
        H.cloud_j = (const int*)ctx->cj.p;
        H.cloud_n = (const uint32_t*)ctx->cn.p;
    }
    bool use_hits = hit_cap > 0 && !onepass;
    if (onepass) {
        // everything is already written, except rows that outgrew the estimated row buffer (or all rows, when the
        // estimated arena was too small and has just been regrown by the ENSUREs above)
        D.hit_cap = 0;
        D.skip_deg_le = onepass_redo_all ? 0 : row_cap_1p;
        if (E > 0 && (onepass_redo_all || n_fallback_rows > 0)) {
            if (onepass_redo_all) CU(cudaMemsetAsync(&((DgStats*)ctx->stats.p)->n_tie_rows, 0, sizeof(int), st));
            CU(dg_launch_direct(st, D, true, ctx->sm_count, ctx->direct_min_blocks));
        }
    }
    if (use_hits) {
        H.n_warps = 8;
        while (H.n_warps > 1 && (int)dg_fill_hits_smem_bytes(H) > ctx->smem_optin) H.n_warps >>= 1;
        if ((int)dg_fill_hits_smem_bytes(H) > ctx->smem_optin) use_hits = false;
    }
    if (!use_hits) D.hit_cap = 0;
    if (use_hits && hit_target > 0.0) {
        // thresholded lists: the stored count (INT_MAX: not served by the list) decides who fills a row
        H.deg_raw = (int*)ctx->hits_n.p;
        D.deg_raw = (int*)ctx->hits_n.p;
    }
    if (E > 0 && use_hits && (any_direct || (n_cell_structs > 0 && !cell32))) CU(dg_launch_fill_hits(st, H, ctx->sm_count));
    if (E > 0 && !onepass && any_direct && (!use_hits || n_fallback_rows > 0)) {
        // split fill: the few rows beyond split_cap go first, on the auxiliary stream, with a buffer for the largest row
        // (one CTA of few warps per SM); the main launch's CTAs fit beside it (split_cap leaves room), so the large
        // rows do not cost a low-occupancy tail of their own
        const bool aux = split_fill && ctx->h_stats->n_big_rows > 0;
        if (aux) {
            DgDirectArgs B2 = D;
            B2.row_cap = row_cap_big;
            B2.n_buckets = n_buckets_big;
            B2.skip_deg_le = split_cap;  // (rows up to split_cap entries are written by the main launch)
            B2.n_warps = 8;
            while (B2.n_warps > 1 && (int)dg_direct_smem_bytes(B2, true) > ctx->smem_optin - split_main_smem) B2.n_warps -= 1;
            if ((int)dg_direct_smem_bytes(B2, true) > ctx->smem_optin)
                return fail(ctx, DGNN_ERR_DEGREE, "row buffer for degree %d does not fit in shared memory", max_deg);
            dyn_arm(B2);
            CU(cudaEventRecord(ctx->aux_fork, st));
            CU(cudaStreamWaitEvent(ctx->aux_stream, ctx->aux_fork, 0));
            CU(dg_launch_direct(ctx->aux_stream, B2, true, ctx->sm_count, ctx->direct_min_blocks));
            CU(cudaEventRecord(ctx->aux_join, ctx->aux_stream));
        }
        dyn_arm(D);
        D.defer_big = split_fill ? 1 : 0;
        bool class_order = false;
        if (D.dyn_unit > 0 && hit_cap == 0 && split_cap > 0 && (split_fill || ctx->h_stats->n_big_rows == 0)) {
            // (worth it when at least 5 % of the rows are outside the dominant class)
            long long tot = 0, mx = 0;
            for (int k = 0; k < 4; ++k) {
                tot += ctx->h_stats->cls_n[k];
                if (ctx->h_stats->cls_n[k] > mx) mx = ctx->h_stats->cls_n[k];
            }
            class_order = (tot - mx) * 20 >= tot && tot > 0;
        }
        if (class_order) {
            // rows of very different lengths run different sort networks; in class order the warps of an SM execute
            // the same code at the same time (instruction-cache misses were the top stall of the config-3 fill pass)
            const long long n_perm = (long long)ctx->h_stats->cls_n[0] + ctx->h_stats->cls_n[1] + ctx->h_stats->cls_n[2] +
                                     ctx->h_stats->cls_n[3];
            ENSURE(ctx->row_perm, (size_t)(N + 32) * 4);
            dg_launch_class_scatter(st, N, (const int*)ctx->astruct.p, small, (const int*)ctx->deg_raw.p, split_cap,
                                    d_stats, (int*)ctx->row_perm.p, ctx->sm_count);
            D.row_perm = (const int*)ctx->row_perm.p;
            D.n_perm = n_perm;
        }
        CU(dg_launch_direct(st, D, true, ctx->sm_count, ctx->direct_min_blocks));
        D.defer_big = 0;
        D.row_perm = nullptr;
        if (aux) CU(cudaStreamWaitEvent(st, ctx->aux_join, 0));
    }
    bool cell32_rest = false;  // rows the fp32-tile fill left to the fp64 traversal (degree beyond the register sort)
    if (cell32 && E > 0) {
        A.row_ptr = (const long long*)ctx->row_ptr.p;
        A.dst = (int*)ctx->dst.p;
        A.image = (uint32_t*)ctx->image.p;
        A.dist = (double*)ctx->dist.p;
        A.row_cap = row_cap_big < 512 ? row_cap_big : 512;
        A.n_warps = ctx->cell32_fill_warps;  // the tile is staged in fp32 and in fp64: one CTA per SM shares it
        if (A.row_cap > 256 && A.n_warps > 16) A.n_warps = 16;  // (the 512-entry sort network needs 128 registers)
        A.rbf = D.rbf;
        // shared memory: first the slot list shrinks (cells whose stencil outgrows it take the range-by-range sweep),
        // then warps go
        if ((int)dg_cell32_smem_bytes(A, true) > ctx->smem_optin) A.src_cap = 768;
        while (A.n_warps > 8 && (int)dg_cell32_smem_bytes(A, true) > ctx->smem_optin) A.n_warps = A.n_warps > 16 ? A.n_warps - 4 : A.n_warps / 2;
        if ((int)dg_cell32_smem_bytes(A, true) <= ctx->smem_optin) {
            const int fgrid = (int)((long long)ctx->sm_count * 4 < n_items ? (long long)ctx->sm_count * 4 : n_items);
            CU(dg_launch_cell32(st, A, true, fgrid));
            A.n_warps = 8;
            A.src_cap = 1024;
            cell32_rest = max_deg > 512;
        } else {
            cell32_rest = true;
            A.row_cap = 0;  // (nothing written: the fp64 traversal takes every row)
        }
    }
    if (n_cell_structs > 0 && E > 0 && (cell32 ? cell32_rest : (!use_hits || n_fallback_rows > 0))) {
        A.hit_cap = cell32 ? A.row_cap : (use_hits ? hit_cap : 0);
        if (cell32) {
            A.tile_cap = 0;
            A.tcs_cap = 4;
        }
        A.row_ptr = (const long long*)ctx->row_ptr.p;
        A.dst = (int*)ctx->dst.p;
        A.image = (uint32_t*)ctx->image.p;
        A.dist = (double*)ctx->dist.p;
        A.row_cap = row_cap_big;
        A.n_buckets = n_buckets_big;
        A.n_warps = 8;
        A.rbf = D.rbf;
        while (A.n_warps > 1 && (int)dg_traverse_smem_bytes(A, true) > ctx->smem_optin) A.n_warps >>= 1;
        while (A.tile_cap > 0 && (int)dg_traverse_smem_bytes(A, true) > ctx->smem_optin) {
            A.tile_cap = A.tile_cap > 512 ? A.tile_cap / 2 : 0;  // shrink the tile: more items go unstaged
            if (A.tile_cap == 0) A.tcs_cap = 4;
        }
        if ((int)dg_traverse_smem_bytes(A, true) > ctx->smem_optin)
            return fail(ctx, DGNN_ERR_DEGREE, "row buffer for degree %d does not fit in shared memory", max_deg);
        CU(dg_launch_traverse(st, A, true, grid));
    }
    CU(cudaEventRecord(ctx->ev.e[4], st));

    // ---- K3: fused Gaussian RBF expansion ----------------------------------------------------------------------
    if (want_rbf && E > 0 && !fuse_rbf) {
        dg_launch_rbf_f32(st, (const double*)ctx->dist.p, E, K, b->rbf_r_cutoff, b->dr, (float*)ctx->rbf.p,
                          ctx->sm_count);
    }
    CU(cudaEventRecord(ctx->ev.e[5], st));
    CU(cudaGetLastError());
    if (sinks && !sinks_early) {
        if (ctx->sink_row_ptr)
            CU(cudaMemcpyAsync(ctx->sink_row_ptr, ctx->row_ptr.p, (size_t)(N + 1) * 8, cudaMemcpyDeviceToHost, st));
        if (ctx->sink_seo)
            CU(cudaMemcpyAsync(ctx->sink_seo, ctx->seo.p, (size_t)(S + 1) * 8, cudaMemcpyDeviceToHost, st));
    }
    if (sinks_early) CU(cudaStreamSynchronize(ctx->copy_stream));
    {
        int r = fetch_stats(ctx);  // final sync: completion + tie statistics
        if (r != DGNN_OK) return r;
    }
    if (ctx->h_stats->overflow) return fail(ctx, DGNN_ERR_DEGREE, "internal: row buffer overflow");

    out->n_structs = S;
    out->n_atoms = N;
    out->n_edges = E;
    out->K = want_rbf ? K : 0;
    out->max_degree = max_deg;
    out->row_ptr = (const int64_t*)ctx->row_ptr.p;
    out->dst = (const int32_t*)ctx->dst.p;
    out->image = (const int32_t*)ctx->image.p;
    out->dist = (const double*)ctx->dist.p;
    out->rbf = (want_rbf && E > 0) ? (const float*)ctx->rbf.p : nullptr;
    out->struct_edge_offsets = (const int64_t*)ctx->seo.p;
    out->n_near_cutoff = (int64_t)ctx->h_stats->n_near;
    out->n_tie_rows = ctx->h_stats->n_tie_rows;

    ctx->last.valid = true;
    ctx->build_serial += 1;
    ctx->last.S = S; ctx->last.N = N; ctx->last.E = E;
    ctx->last.max_deg = max_deg;
    ctx->last.off = d_off;
    ctx->last.x = d_x; ctx->last.y = d_y; ctx->last.z = d_z;

    dgnn_build_info& I = ctx->info;
    cudaEventElapsedTime(&I.h2d_ms, ctx->ev.e[0], ctx->ev.e[1]);
    cudaEventElapsedTime(&I.bin_ms, ctx->ev.e[1], ctx->ev.e[2]);
    cudaEventElapsedTime(&I.count_ms, ctx->ev.e[2], ctx->ev.e[3]);
    cudaEventElapsedTime(&I.fill_ms, ctx->ev.e[3], ctx->ev.e[4]);
    cudaEventElapsedTime(&I.rbf_ms, ctx->ev.e[4], ctx->ev.e[5]);
    cudaEventElapsedTime(&I.total_ms, ctx->ev.e[0], ctx->ev.e[5]);
    I.kernel_launches = dg_num_kernel_launches_reset() + dg_rbf_launch_count_reset() + dg_direct_launch_count_reset();
    I.n_items = (int32_t)n_items;
    I.cloud_points = P;
    I.staged_tile_cap = tile_cap_report;
    I.row_cap = row_cap;
    I.n_cell_structs = n_cell_structs;
    I.direct_warps = D.n_warps;
    I.rbf_fused = fuse_rbf ? 1 : 0;
    I.hit_cap = use_hits ? hit_cap : 0;
    I.onepass = onepass ? (onepass_redo_all ? 2 : 1) : 0;
    I.onepass_fallback_rows = onepass ? n_fallback_rows : 0;
    I.tie_run_mean = (float)tie_run_mean;
    I.full_tie_sort = (!onepass && max_nb >= row_cap && D.full_tie_sort) ? 1 : 0;
    I.split_big_rows = split_fill ? ctx->h_stats->n_big_rows : 0;
    return DGNN_OK;
}

}  // extern "C"
