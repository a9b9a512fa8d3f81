// dgnn_options.cu -- dgnn_set_option: the public options and the "tune.<name>" knobs of a context.
#include <string.h>

#include "dgnn_ctx.h"

extern "C" {

// Public options (include/dgnn_cabi.h): cell_frac, direct_max_atoms, hit_cap, onepass, fuse_rbf, upload_chunks, cell32,
// sym_count.  Everything else is a tuning / testing knob and has to be asked for as "tune.<name>".
int dgnn_set_option(dgnn_ctx* ctx, const char* name, double value) {
    if (!ctx || !name) return DGNN_ERR_INVALID_ARG;
    const char* full_name = name;
    const bool tune = !strncmp(name, "tune.", 5);
    if (tune) {
        name += 5;
    } else {
        static const char* const kPublic[] = {"cell_frac", "direct_max_atoms", "hit_cap", "onepass",
                                              "fuse_rbf",  "upload_chunks",    "cell32",  "sym_count"};
        bool ok = false;
        for (const char* q : kPublic) ok = ok || !strcmp(name, q);
        if (!ok) return fail(ctx, DGNN_ERR_INVALID_ARG, "unknown option '%s' (tuning knobs are spelled tune.<name>)", name);
    }
    if (!strcmp(name, "cell_frac")) {
        if (!(value >= 0.05 && value <= 4.0)) return fail(ctx, DGNN_ERR_INVALID_ARG, "cell_frac out of range");
        ctx->cell_frac = value;
    } else if (!strcmp(name, "tile_target")) {
        if (!(value >= 64 && value <= 8192)) return fail(ctx, DGNN_ERR_INVALID_ARG, "tile_target out of range");
        ctx->tile_target = (int)value;
    } else if (!strcmp(name, "tile_limit")) {
        if (!(value >= 0 && value <= 8192)) return fail(ctx, DGNN_ERR_INVALID_ARG, "tile_limit out of range");
        ctx->tile_limit = (int)value;  // 0 forces the unstaged (global-memory) traversal everywhere
    } else if (!strcmp(name, "src_target")) {
        if (!(value >= 1 && value <= 1024)) return fail(ctx, DGNN_ERR_INVALID_ARG, "src_target out of range");
        ctx->src_target = (int)value;
    } else if (!strcmp(name, "traverse_grid_mult")) {
        if (!(value >= 1 && value <= 64)) return fail(ctx, DGNN_ERR_INVALID_ARG, "traverse_grid_mult out of range");
        ctx->traverse_grid_mult = (int)value;
    } else if (!strcmp(name, "sort_buckets")) {
        const int v = (int)value;
        if (!(v == 0 || (v >= 32 && v <= 2048 && (v & (v - 1)) == 0)))
            return fail(ctx, DGNN_ERR_INVALID_ARG, "sort_buckets must be 0 (auto) or a power of two in [32, 2048]");
        ctx->sort_buckets = v;
    } else if (!strcmp(name, "fuse_rbf")) {
        ctx->fuse_rbf = value != 0.0;  // 0: separate K3 launch after the fill pass (phase-separated timing)
    } else if (!strcmp(name, "hit_cap")) {
        if (!(value >= -1 && value <= DGNN_MAX_DEGREE)) return fail(ctx, DGNN_ERR_INVALID_ARG, "hit_cap out of range");
        ctx->hit_cap_opt = (int)value;  // per-row hit-list slots written by the count pass (-1 auto, 0 off)
    } else if (!strcmp(name, "hits_max_bytes")) {
        if (!(value >= 0)) return fail(ctx, DGNN_ERR_INVALID_ARG, "hits_max_bytes out of range");
        ctx->hits_max_bytes = value;
    } else if (!strcmp(name, "upload_min_atoms")) {
        if (!(value >= 0)) return fail(ctx, DGNN_ERR_INVALID_ARG, "upload_min_atoms out of range");
        ctx->upload_min_atoms = (long long)value;
    } else if (!strcmp(name, "upload_chunks")) {
        if (!(value >= 1 && value <= 8)) return fail(ctx, DGNN_ERR_INVALID_ARG, "upload_chunks must be in [1, 8]");
        ctx->upload_chunks = (int)value;
    } else if (!strcmp(name, "fill_hits_bulk")) {
        ctx->fill_hits_bulk = value != 0.0;
    } else if (!strcmp(name, "cell32_fill_warps")) {
        if (!(value == 8 || value == 16 || value == 20 || value == 24))
            return fail(ctx, DGNN_ERR_INVALID_ARG, "cell32_fill_warps must be 8, 16, 20 or 24");
        ctx->cell32_fill_warps = (int)value;
    } else if (!strcmp(name, "cell32")) {
        ctx->cell32 = value != 0.0;
    } else if (!strcmp(name, "dyn_rows")) {
        ctx->dyn_rows = value != 0.0;
    } else if (!strcmp(name, "sym_count")) {
        ctx->sym_count = value != 0.0;
    } else if (!strcmp(name, "rbf_recur")) {
        ctx->rbf_recur = value != 0.0;
    } else if (!strcmp(name, "cell32_mask")) {
        ctx->cell32_mask = value != 0.0;
    } else if (!strcmp(name, "split_rows")) {
        ctx->split_rows = value != 0.0;
    } else if (!strcmp(name, "hit_threshold")) {
        ctx->hit_threshold = value != 0.0;
    } else if (!strcmp(name, "fill_hits_warps")) {
        if (!(value >= 0 && value <= 8)) return fail(ctx, DGNN_ERR_INVALID_ARG, "fill_hits_warps must be 0 ... 8");
        ctx->fill_hits_warps = (int)value;
    } else if (!strcmp(name, "tie_probe")) {
        // 1: sample rows decide the tie ranking of the uncapped fill pass (dgnn_tie_probe.cu), 0: always the compact
        // ranking, 2: always the warp-cooperative one
        if (!(value >= 0 && value <= 2)) return fail(ctx, DGNN_ERR_INVALID_ARG, "tie_probe must be 0, 1 or 2");
        ctx->tie_probe = (int)value;
    } else if (!strcmp(name, "onepass")) {
        if (!(value >= -1 && value <= 1)) return fail(ctx, DGNN_ERR_INVALID_ARG, "onepass must be -1, 0 or 1");
        ctx->onepass_opt = (int)value;  // single-pass build (look-back row offsets) instead of count + fill
    } else if (!strcmp(name, "lb_spin_limit")) {
        if (!(value >= 0 && value <= 1e9)) return fail(ctx, DGNN_ERR_INVALID_ARG, "lb_spin_limit out of range");
        ctx->lb_spin_limit = (int)value;
    } else if (!strcmp(name, "onepass_row_cap")) {
        if (!(value >= 0 && value <= DGNN_MAX_DEGREE)) return fail(ctx, DGNN_ERR_INVALID_ARG, "onepass_row_cap out of range");
        ctx->onepass_row_cap = (int)value;
    } else if (!strcmp(name, "onepass_arena_scale")) {
        if (!(value > 0.0 && value <= 16.0)) return fail(ctx, DGNN_ERR_INVALID_ARG, "onepass_arena_scale out of range");
        ctx->onepass_arena_scale = value;
    } else if (!strcmp(name, "direct_max_atoms")) {
        if (!(value >= 0 && value <= 65535)) return fail(ctx, DGNN_ERR_INVALID_ARG, "direct_max_atoms out of range");
        ctx->direct_max_atoms = (int)value;  // 0 sends every structure to the cell-list path
    } else {
        return fail(ctx, DGNN_ERR_INVALID_ARG, "unknown option '%s'", name);
    }
    ctx->opt_log.emplace_back(full_name, value);
    if (ctx->twin) dgnn_set_option(ctx->twin, full_name, value);
    return DGNN_OK;
}

}  // extern "C"
