// dgnn_consumers.cu -- C ABI, second half: host copies of a build, stand-alone Gaussian RBF expansion, the consumers of
// the device CSR (collation, viewer arrays, neighbourhood clouds, distance tables) and the DLPack export.
#include "dgnn_ctx.h"

extern "C" {

int dgnn_copy_edges_to_host(dgnn_ctx* ctx, const dgnn_graphs* g, int64_t e0, int64_t e1, int32_t* dst,
                            int32_t* image, double* dist, float* rbf) {
    if (!ctx || !g) return DGNN_ERR_INVALID_ARG;
    if (e0 < 0 || e1 < e0 || e1 > g->n_edges) return fail(ctx, DGNN_ERR_INVALID_ARG, "edge range out of bounds");
    CU(cudaSetDevice(ctx->device));
    const size_t n = (size_t)(e1 - e0);
    if (n == 0) return DGNN_OK;
    cudaStream_t st = ctx->stream;
    if (dst) CU(cudaMemcpyAsync(dst, g->dst + e0, n * 4, cudaMemcpyDeviceToHost, st));
    if (image) CU(cudaMemcpyAsync(image, g->image + e0, n * 4, cudaMemcpyDeviceToHost, st));
    if (dist) CU(cudaMemcpyAsync(dist, g->dist + e0, n * 8, cudaMemcpyDeviceToHost, st));
    if (rbf) {
        if (!g->rbf) return fail(ctx, DGNN_ERR_INVALID_ARG, "no RBF features in this build");
        CU(cudaMemcpyAsync(rbf, g->rbf + (size_t)e0 * g->K, n * (size_t)g->K * 4, cudaMemcpyDeviceToHost, st));
    }
    CU(cudaStreamSynchronize(st));
    return DGNN_OK;
}

int dgnn_copy_to_host(dgnn_ctx* ctx, const dgnn_graphs* g, int64_t* row_ptr, int32_t* dst, int32_t* image,
                      double* dist, float* rbf, int64_t* struct_edge_offsets) {
    if (!ctx || !g) return DGNN_ERR_INVALID_ARG;
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (row_ptr) CU(cudaMemcpyAsync(row_ptr, g->row_ptr, (size_t)(g->n_atoms + 1) * 8, cudaMemcpyDeviceToHost, st));
    if (struct_edge_offsets)
        CU(cudaMemcpyAsync(struct_edge_offsets, g->struct_edge_offsets, (size_t)(g->n_structs + 1) * 8,
                           cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return dgnn_copy_edges_to_host(ctx, g, 0, g->n_edges, dst, image, dist, rbf);
}

static int rbf_common(dgnn_ctx* ctx, const double* dist, int64_t n, double r_cutoff, double dr, void* out,
                      int32_t* K_out, bool f64) {
    if (!ctx) return DGNN_ERR_INVALID_ARG;
    if (n < 0 || (n > 0 && (!dist || !out))) return fail(ctx, DGNN_ERR_INVALID_ARG, "null pointer / negative n");
    if (!(r_cutoff > 0.0) || !(dr > 0.0)) return fail(ctx, DGNN_ERR_INVALID_ARG, "r_cutoff and dr must be positive");
    const double kf = std::floor(r_cutoff / dr);
    if (!(kf >= 1.0 && kf <= 65536.0)) return fail(ctx, DGNN_ERR_INVALID_ARG, "RBF bin count out of range");
    const int K = (int)kf;
    if (K_out) *K_out = K;
    if (n == 0) return DGNN_OK;
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const double* d_in = dist;
    void* d_out = out;
    const size_t out_bytes = (size_t)n * K * (f64 ? 8 : 4);
    const bool in_dev = is_device_ptr(dist), out_dev = is_device_ptr(out);
    if (!in_dev) {
        ENSURE(ctx->tmp_in, (size_t)n * 8);
        CU(cudaMemcpyAsync(ctx->tmp_in.p, dist, (size_t)n * 8, cudaMemcpyHostToDevice, st));
        d_in = (const double*)ctx->tmp_in.p;
    }
    if (!out_dev) {
        ENSURE(ctx->tmp_out, out_bytes + 16);
        d_out = ctx->tmp_out.p;
    }
    if (f64)
        dg_launch_rbf_f64_colmajor(st, d_in, n, K, r_cutoff, dr, (double*)d_out);
    else
        dg_launch_rbf_f32(st, d_in, n, K, r_cutoff, dr, (float*)d_out, ctx->sm_count);
    CU(cudaGetLastError());
    if (!out_dev) CU(cudaMemcpyAsync(out, d_out, out_bytes, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return DGNN_OK;
}

int dgnn_rbf_expand(dgnn_ctx* ctx, const double* dist, int64_t n, double r_cutoff, double dr, float* out,
                    int32_t* K_out) {
    return rbf_common(ctx, dist, n, r_cutoff, dr, out, K_out, false);
}

int dgnn_rbf_expand_f64(dgnn_ctx* ctx, const double* dist, int64_t n, double r_cutoff, double dr, double* out,
                        int32_t* K_out) {
    return rbf_common(ctx, dist, n, r_cutoff, dr, out, K_out, true);
}



// ---------------------------------------------------------------------------------------------------------------------
// consumers of the last build's CSR (dgnn_collate.cu)
// ---------------------------------------------------------------------------------------------------------------------
namespace {
int consumer_begin(dgnn_ctx* ctx, const dgnn_graphs* g, const char* who) {
    if (!g) return fail(ctx, DGNN_ERR_INVALID_ARG, "%s: null graphs", who);
    if (!ctx->last.valid || g->row_ptr != (const int64_t*)ctx->row_ptr.p || g->n_atoms != ctx->last.N ||
        g->n_edges != ctx->last.E || g->n_structs != ctx->last.S)
        return fail(ctx, DGNN_ERR_INVALID_ARG, "%s: graphs are not the result of the last dgnn_build on this context", who);
    CU(cudaSetDevice(ctx->device));
    for (auto& e : ctx->cev)
        if (!e) CU(cudaEventCreate(&e));
    memset(&ctx->cinfo, 0, sizeof(ctx->cinfo));
    dg_collate_launch_count_reset();
    return DGNN_OK;
}
int consumer_end(dgnn_ctx* ctx, long long bytes_read, long long bytes_written) {
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(ctx->stream));
    cudaEventElapsedTime(&ctx->cinfo.device_ms, ctx->cev[0], ctx->cev[1]);
    ctx->cinfo.kernel_launches = dg_collate_launch_count_reset();
    ctx->cinfo.bytes_read = bytes_read;
    ctx->cinfo.bytes_written = bytes_written;
    return DGNN_OK;
}
}  // namespace

int dgnn_last_consumer_info(const dgnn_ctx* ctx, dgnn_consumer_info* out) {
    if (!ctx || !out) return DGNN_ERR_INVALID_ARG;
    *out = ctx->cinfo;
    return DGNN_OK;
}

int dgnn_collate(dgnn_ctx* ctx, const dgnn_graphs* g, const int32_t* node_keys, const int32_t* emb_keys,
                 const double* emb_table, int32_t n_keys, int32_t F, uint32_t flags, dgnn_collated* out) {
    if (!ctx) return DGNN_ERR_INVALID_ARG;
    if (!out) return fail(ctx, DGNN_ERR_INVALID_ARG, "dgnn_collate: null output");
    if (ctx->live_exports.load() > 0)
        return fail(ctx, DGNN_ERR_BUSY, "dgnn_collate: %d DLPack export(s) still borrow this context's buffers", ctx->live_exports.load());
    memset(out, 0, sizeof(*out));
    {
        int r = consumer_begin(ctx, g, "dgnn_collate");
        if (r != DGNN_OK) return r;
    }
    const long long N = ctx->last.N, E = ctx->last.E;
    const bool want_x = node_keys != nullptr || emb_keys != nullptr || emb_table != nullptr || F > 0 || n_keys > 0;
    if (want_x) {
        if (!node_keys || !emb_keys || !emb_table) return fail(ctx, DGNN_ERR_INVALID_ARG, "dgnn_collate: node_keys, emb_keys and emb_table go together");
        if (F < 1 || F > 65536) return fail(ctx, DGNN_ERR_INVALID_ARG, "dgnn_collate: feature width out of range");
        if (n_keys < 1 || n_keys > 65536) return fail(ctx, DGNN_ERR_INVALID_ARG, "dgnn_collate: n_keys out of range");
    }
    cudaStream_t st = ctx->stream;
    int lut_size = 0;
    if (want_x) {
        // the std::map<int, VectorXd> as a dense key -> row table (keys are species indices / atomic numbers)
        int max_key = 0;
        for (int i = 0; i < n_keys; ++i) {
            if (emb_keys[i] < 0 || emb_keys[i] >= 65536) return fail(ctx, DGNN_ERR_INVALID_ARG, "dgnn_collate: embedding key %d out of range", emb_keys[i]);
            if (emb_keys[i] > max_key) max_key = emb_keys[i];
        }
        lut_size = max_key + 1;
        std::vector<int> lut((size_t)lut_size, -1);
        for (int i = 0; i < n_keys; ++i) {
            if (lut[emb_keys[i]] >= 0) return fail(ctx, DGNN_ERR_INVALID_ARG, "dgnn_collate: duplicate embedding key %d", emb_keys[i]);
            lut[emb_keys[i]] = i;
        }
        std::vector<float> tab((size_t)n_keys * F);
        for (size_t i = 0; i < tab.size(); ++i) tab[i] = (float)emb_table[i];  // torch.tensor(..., dtype=torch.float)
        ENSURE(ctx->col_lut, (size_t)lut_size * 4);
        ENSURE(ctx->col_table, tab.size() * 4);
        ENSURE(ctx->col_err, 16);
        ENSURE(ctx->col_x, (size_t)(N > 0 ? N : 1) * F * 4);
        CU(cudaMemcpyAsync(ctx->col_lut.p, lut.data(), (size_t)lut_size * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(ctx->col_table.p, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemsetAsync(ctx->col_err.p, 0xff, 4, st));  // -1: no missing key
        if (!(flags & DGNN_INPUT_DEVICE) && N > 0) {
            ENSURE(ctx->col_keys, (size_t)N * 4);
            CU(cudaMemcpyAsync(ctx->col_keys.p, node_keys, (size_t)N * 4, cudaMemcpyHostToDevice, st));
            node_keys = (const int32_t*)ctx->col_keys.p;
        }
        CU(cudaStreamSynchronize(st));  // lut / tab are stack-scoped host vectors
    }
    ENSURE(ctx->col_ei, (size_t)(E > 0 ? E : 1) * 16);
    ENSURE(ctx->col_batch, (size_t)(N > 0 ? N : 1) * 8);
    CU(cudaEventRecord(ctx->cev[0], st));
    dg_launch_collate_nodes(st, N, want_x ? F : 0, node_keys, (const int*)ctx->col_lut.p, lut_size,
                            (const float*)ctx->col_table.p, (const int*)ctx->astruct.p, want_x ? (float*)ctx->col_x.p : nullptr,
                            (long long*)ctx->col_batch.p, (int*)ctx->col_err.p, ctx->sm_count);
    dg_launch_collate_edges(st, N, E, (const long long*)ctx->row_ptr.p, (const int*)ctx->astruct.p, ctx->last.off,
                            (const int*)ctx->dst.p, (long long*)ctx->col_ei.p, ctx->sm_count);
    CU(cudaEventRecord(ctx->cev[1], st));
    if (want_x && N > 0) {
        int bad = -1;
        CU(cudaMemcpyAsync(&bad, ctx->col_err.p, 4, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        if (bad >= 0) return fail(ctx, DGNN_ERR_INVALID_ARG, "dgnn_collate: node %d: key not in the embedding table (std::map::at would throw)", bad);
    }
    {
        int r = consumer_end(ctx, 8 * (N + 1) + 4 * N + 4 * E + (want_x ? 4 * N : 0),
                             16 * E + 8 * N + (want_x ? 4LL * F * N : 0));
        if (r != DGNN_OK) return r;
    }
    out->n_nodes = N;
    out->n_edges = E;
    out->F = want_x ? F : 0;
    out->K = g->K;
    out->x = want_x ? (const float*)ctx->col_x.p : nullptr;
    out->edge_index = (const int64_t*)ctx->col_ei.p;
    out->batch = (const int64_t*)ctx->col_batch.p;
    out->edge_attr = g->rbf;
    return DGNN_OK;
}

int dgnn_copy_collated_to_host(dgnn_ctx* ctx, const dgnn_collated* c, float* x, int64_t* edge_index, int64_t* batch) {
    if (!ctx || !c) return DGNN_ERR_INVALID_ARG;
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (x) {
        if (!c->x) return fail(ctx, DGNN_ERR_INVALID_ARG, "no node features in this collation");
        if (c->n_nodes > 0) CU(cudaMemcpyAsync(x, c->x, (size_t)c->n_nodes * c->F * 4, cudaMemcpyDeviceToHost, st));
    }
    if (edge_index && c->n_edges > 0) CU(cudaMemcpyAsync(edge_index, c->edge_index, (size_t)c->n_edges * 16, cudaMemcpyDeviceToHost, st));
    if (batch && c->n_nodes > 0) CU(cudaMemcpyAsync(batch, c->batch, (size_t)c->n_nodes * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return DGNN_OK;
}

int dgnn_edge_views(dgnn_ctx* ctx, const dgnn_graphs* g, int32_t* sources, int32_t* targets, float* distances,
                    float* displacements) {
    if (!ctx) return DGNN_ERR_INVALID_ARG;
    {
        int r = consumer_begin(ctx, g, "dgnn_edge_views");
        if (r != DGNN_OK) return r;
    }
    const long long N = ctx->last.N, E = ctx->last.E;
    cudaStream_t st = ctx->stream;
    if (E == 0) return DGNN_OK;
    // staging: [sources | targets | distances | displacements], only what was asked for
    const size_t e4 = (size_t)E * 4;
    ENSURE(ctx->view_a, e4 * ((sources ? 1 : 0) + (targets ? 1 : 0) + (distances ? 1 : 0) + (displacements ? 3 : 0)) + 16);
    unsigned char* p = (unsigned char*)ctx->view_a.p;
    int* d_src = nullptr; int* d_tgt = nullptr; float* d_dist = nullptr; float* d_disp = nullptr;
    if (sources) { d_src = (int*)p; p += e4; }
    if (targets) { d_tgt = (int*)p; p += e4; }
    if (distances) { d_dist = (float*)p; p += e4; }
    if (displacements) { d_disp = (float*)p; p += 3 * e4; }
    CU(cudaEventRecord(ctx->cev[0], st));
    dg_launch_edge_views(st, N, (const long long*)ctx->row_ptr.p, (const int*)ctx->astruct.p, (const DgSmall*)ctx->small.p,
                         ctx->last.x, ctx->last.y, ctx->last.z, (const int*)ctx->dst.p, (const uint32_t*)ctx->image.p,
                         (const double*)ctx->dist.p, d_src, d_tgt, d_dist, d_disp, ctx->sm_count);
    CU(cudaEventRecord(ctx->cev[1], st));
    if (sources) CU(cudaMemcpyAsync(sources, d_src, e4, cudaMemcpyDeviceToHost, st));
    if (targets) CU(cudaMemcpyAsync(targets, d_tgt, e4, cudaMemcpyDeviceToHost, st));
    if (distances) CU(cudaMemcpyAsync(distances, d_dist, e4, cudaMemcpyDeviceToHost, st));
    if (displacements) CU(cudaMemcpyAsync(displacements, d_disp, 3 * e4, cudaMemcpyDeviceToHost, st));
    return consumer_end(ctx, 8 * (N + 1) + E * ((targets || displacements ? 4 : 0) + (distances ? 8 : 0) + (displacements ? 4 : 0)),
                        (long long)e4 * ((sources ? 1 : 0) + (targets ? 1 : 0) + (distances ? 1 : 0) + (displacements ? 3 : 0)));
}

int dgnn_neighborhood_clouds(dgnn_ctx* ctx, const dgnn_graphs* g, int64_t a0, int64_t a1, double* cloud,
                             int64_t* cloud_ptr) {
    if (!ctx) return DGNN_ERR_INVALID_ARG;
    {
        int r = consumer_begin(ctx, g, "dgnn_neighborhood_clouds");
        if (r != DGNN_OK) return r;
    }
    if (a0 < 0 || a1 < a0 || a1 > ctx->last.N) return fail(ctx, DGNN_ERR_INVALID_ARG, "dgnn_neighborhood_clouds: atom range out of bounds");
    if (!cloud || !cloud_ptr) return fail(ctx, DGNN_ERR_INVALID_ARG, "dgnn_neighborhood_clouds: null output");
    cudaStream_t st = ctx->stream;
    const long long n = a1 - a0;
    if (n == 0) {
        cloud_ptr[0] = 0;
        return DGNN_OK;
    }
    long long rp[2];
    CU(cudaMemcpyAsync(&rp[0], (const long long*)ctx->row_ptr.p + a0, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(&rp[1], (const long long*)ctx->row_ptr.p + a1, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const long long n_points = (rp[1] - rp[0]) + n;
    ENSURE(ctx->view_a, (size_t)n_points * 24 + 16);
    ENSURE(ctx->view_b, (size_t)(n + 1) * 8);
    CU(cudaEventRecord(ctx->cev[0], st));
    dg_launch_neighborhood_clouds(st, a0, a1, (const long long*)ctx->row_ptr.p, (const int*)ctx->astruct.p,
                                  (const DgSmall*)ctx->small.p, ctx->last.x, ctx->last.y, ctx->last.z,
                                  (const int*)ctx->dst.p, (const uint32_t*)ctx->image.p, (double*)ctx->view_a.p,
                                  (long long*)ctx->view_b.p, ctx->sm_count);
    CU(cudaEventRecord(ctx->cev[1], st));
    CU(cudaMemcpyAsync(cloud, ctx->view_a.p, (size_t)n_points * 24, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(cloud_ptr, ctx->view_b.p, (size_t)(n + 1) * 8, cudaMemcpyDeviceToHost, st));
    return consumer_end(ctx, 8 * (n + 1) + 24 * n + 8 * (rp[1] - rp[0]), 24 * n_points + 8 * (n + 1));
}

int dgnn_distance_tables(dgnn_ctx* ctx, const dgnn_graphs* g, int64_t a0, int64_t a1, int32_t mode,
                         dgnn_distance_tables_out* out) {
    if (!ctx) return DGNN_ERR_INVALID_ARG;
    {
        int r = consumer_begin(ctx, g, "dgnn_distance_tables");
        if (r != DGNN_OK) return r;
    }
    if (!out) return fail(ctx, DGNN_ERR_INVALID_ARG, "dgnn_distance_tables: null output");
    memset(out, 0, sizeof(*out));
    if (a0 < 0 || a1 < a0 || a1 > ctx->last.N) return fail(ctx, DGNN_ERR_INVALID_ARG, "dgnn_distance_tables: atom range out of bounds");
    if (mode != 0 && mode != 1) return fail(ctx, DGNN_ERR_INVALID_ARG, "dgnn_distance_tables: mode must be 0 or 1");
    cudaStream_t st = ctx->stream;
    const long long n = a1 - a0;
    out->mode = mode;
    out->n_atoms = n;
    ENSURE(ctx->tab_ptr, (size_t)(n + 2) * 8);
    out->table_ptr = (const int64_t*)ctx->tab_ptr.p;
    if (n == 0) {
        CU(cudaMemsetAsync(ctx->tab_ptr.p, 0, 8, st));
        CU(cudaStreamSynchronize(st));
        return DGNN_OK;
    }
    const size_t smem = (size_t)(ctx->last.max_deg + 1) * 32;
    if ((int)smem > ctx->smem_optin)
        return fail(ctx, DGNN_ERR_DEGREE, "dgnn_distance_tables: a cloud of %d points does not fit in shared memory", ctx->last.max_deg + 1);
    ENSURE(ctx->tab_cnt, (size_t)(n + 1) * 4);
    ENSURE(ctx->scratch, (size_t)(n / 2048 + 4) * 8);
    CU(cudaEventRecord(ctx->cev[0], st));
    dg_launch_table_counts(st, a0, a1, (const long long*)ctx->row_ptr.p, mode, (int*)ctx->tab_cnt.p);
    dg_scan_i32_i64(st, (const int*)ctx->tab_cnt.p, (long long*)ctx->tab_ptr.p, n, (long long*)ctx->scratch.p);
    long long total = 0;
    CU(cudaMemcpyAsync(&total, (const long long*)ctx->tab_ptr.p + n, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const size_t esz = mode == 0 ? 8 : 4;
    ENSURE(ctx->tab_val, (size_t)total * esz + 16);
    CU(dg_launch_distance_tables(st, a0, a1, (const long long*)ctx->row_ptr.p, (const int*)ctx->astruct.p,
                                 (const DgSmall*)ctx->small.p, ctx->last.x, ctx->last.y, ctx->last.z,
                                 (const int*)ctx->dst.p, (const uint32_t*)ctx->image.p,
                                 (const long long*)ctx->tab_ptr.p, ctx->tab_val.p, mode, ctx->last.max_deg,
                                 ctx->sm_count));
    CU(cudaEventRecord(ctx->cev[1], st));
    out->n_values = total;
    out->values = ctx->tab_val.p;
    long long rp[2];
    CU(cudaMemcpyAsync(&rp[0], (const long long*)ctx->row_ptr.p + a0, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(&rp[1], (const long long*)ctx->row_ptr.p + a1, 8, cudaMemcpyDeviceToHost, st));
    // read: row_ptr + per edge (dst, image) + the positions they point at; written: tables + offsets
    return consumer_end(ctx, 8 * (n + 1) + 8 * (rp[1] - rp[0]) + 24 * ((rp[1] - rp[0]) + n), (long long)(total * esz) + 8 * (n + 1));
}

int dgnn_copy_distance_tables_to_host(dgnn_ctx* ctx, const dgnn_distance_tables_out* t, void* values, int64_t* table_ptr) {
    if (!ctx || !t) return DGNN_ERR_INVALID_ARG;
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (values && t->n_values > 0)
        CU(cudaMemcpyAsync(values, t->values, (size_t)t->n_values * (t->mode == 0 ? 8 : 4), cudaMemcpyDeviceToHost, st));
    if (table_ptr) CU(cudaMemcpyAsync(table_ptr, t->table_ptr, (size_t)(t->n_atoms + 1) * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return DGNN_OK;
}

// ---- DLPack export of the collated batch (dlpack.h, the legacy "dltensor" capsule ABI) --------------------------------
namespace {
struct DgDLDevice { int32_t device_type, device_id; };
struct DgDLDataType { uint8_t code, bits; uint16_t lanes; };
struct DgDLTensor {
    void* data;
    DgDLDevice device;
    int32_t ndim;
    DgDLDataType dtype;
    int64_t* shape;
    int64_t* strides;
    uint64_t byte_offset;
};
struct DgDLManagedTensor {
    DgDLTensor dl_tensor;
    void* manager_ctx;
    void (*deleter)(DgDLManagedTensor*);
};
struct DgExport {
    DgDLManagedTensor mt;
    int64_t shape[2];
    dgnn_ctx* owner;
};
void dg_export_deleter(DgDLManagedTensor* t) {
    if (!t) return;
    DgExport* e = static_cast<DgExport*>(t->manager_ctx);
    dgnn_ctx* owner = e->owner;
    delete e;
    if (owner->live_exports.fetch_sub(1) == 1 && owner->zombie.load()) dgnn_destroy(owner);
}
}  // namespace

int dgnn_live_exports(const dgnn_ctx* ctx) { return ctx ? ctx->live_exports.load() : 0; }

int dgnn_collated_dlpack(dgnn_ctx* ctx, const dgnn_collated* c, int32_t which, void** managed_tensor) {
    if (!ctx || !c || !managed_tensor) return DGNN_ERR_INVALID_ARG;
    *managed_tensor = nullptr;
    if (c->edge_index != (const int64_t*)ctx->col_ei.p)
        return fail(ctx, DGNN_ERR_INVALID_ARG, "dgnn_collated_dlpack: not the result of the last dgnn_collate on this context");
    DgExport* e = new (std::nothrow) DgExport();
    if (!e) return DGNN_ERR_OOM;
    DgDLTensor& t = e->mt.dl_tensor;
    t.device = DgDLDevice{2 /* kDLCUDA */, ctx->device};
    t.strides = nullptr;
    t.byte_offset = 0;
    t.shape = e->shape;
    switch (which) {
        case 0: t.data = (void*)c->x; t.ndim = 2; e->shape[0] = c->n_nodes; e->shape[1] = c->F; t.dtype = DgDLDataType{2, 32, 1}; break;
        case 1: t.data = (void*)c->edge_index; t.ndim = 2; e->shape[0] = 2; e->shape[1] = c->n_edges; t.dtype = DgDLDataType{0, 64, 1}; break;
        case 2: t.data = (void*)c->batch; t.ndim = 1; e->shape[0] = c->n_nodes; e->shape[1] = 1; t.dtype = DgDLDataType{0, 64, 1}; break;
        case 3: t.data = (void*)c->edge_attr; t.ndim = 2; e->shape[0] = c->n_edges; e->shape[1] = c->K; t.dtype = DgDLDataType{2, 32, 1}; break;
        default: delete e; return fail(ctx, DGNN_ERR_INVALID_ARG, "dgnn_collated_dlpack: which must be 0..3");
    }
    if (!t.data && !(e->shape[0] == 0 || (t.ndim == 2 && e->shape[1] == 0))) {
        delete e;
        return fail(ctx, DGNN_ERR_INVALID_ARG, "dgnn_collated_dlpack: this array was not produced by the collate call");
    }
    e->owner = ctx;
    e->mt.manager_ctx = e;
    e->mt.deleter = dg_export_deleter;
    ctx->live_exports.fetch_add(1);
    *managed_tensor = &e->mt;
    return DGNN_OK;
}

void dgnn_dlpack_release(void* managed_tensor) {
    DgDLManagedTensor* t = static_cast<DgDLManagedTensor*>(managed_tensor);
    if (t && t->deleter) t->deleter(t);
}

}  // extern "C"
