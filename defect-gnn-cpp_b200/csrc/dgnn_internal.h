// dgnn_internal.h -- shared declarations between the kernel translation units and the C-ABI host layer.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "geom.h"

// Opt-in dynamic shared memory of a kernel is always raised to the DEVICE'S maximum, never to the size of the launch at
// hand: the attribute is per function and device, and several contexts (host threads: graph::ShardedBuilder, the
// streamed build's twin) launch the same kernels with different sizes -- between one thread's set and its launch
// another could lower it and the launch failed with "invalid argument".  A constant value makes the race harmless.
template <auto Kernel>
inline cudaError_t dg_raise_dyn_smem() {
    static thread_local int done_dev = -1;  // (once per host thread and device: the value never changes)
    int dev = 0, optin = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev == done_dev) return cudaSuccess;
    e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (e != cudaSuccess) return e;
    cudaFuncAttributes fa;
    e = cudaFuncGetAttributes(&fa, Kernel);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin - (int)fa.sharedSizeBytes);
    if (e == cudaSuccess) done_dev = dev;
    return e;
}

// Device-resident statistics block, copied to the host at the (few) sync points of a build.
struct DgStats {
    unsigned long long n_near;   // candidate pairs with |d - r_c| <= 1e-6
    long long n_cells;           // total halo cells of the batch
    long long n_items;           // total work items
    long long cloud_points;      // pruned image cloud size
    long long n_edges;
    int max_tile_pts;            // largest staged-tile point count over items
    int max_tile_tcs;            // largest tile cell-start table (rows * (tdc + 1))
    int max_deg;                 // largest uncapped row
    int n_tie_rows;
    int err_status;              // first non-zero DgGeom::status seen
    int err_struct;              // and the structure it was seen on
    int overflow;                // a row did not fit the row buffer (must not happen)
    int n_cell_structs;          // non-empty structures routed to the cell-list path
    int n_direct_structs;        // ... and to the direct path
    int n_fallback_rows;         // direct-path rows whose hit list was not stored (degree > hit_cap or image range > 15)
    int est_deg;                 // density estimate of the largest degree (sizes the hit lists of a ctx's first build)
    int cap_overflow;            // single-pass build: a row did not fit the (estimated) edge arena -> rerun as two passes
    int lb_timeout;              // single-pass build: a look-back spin gave up (must not happen: all CTAs are resident)
    unsigned long long est_edges;  // density estimate of the edge total, sum over structures of N_s * est_deg_s
    unsigned long long unit_ctr[8];  // row-unit counters of dynamically scheduled launches (one per launch)
    int est_sparse;              // 1000000 - smallest per-structure degree estimate (0: no non-empty structure)
    unsigned long long tie_sum;  // dgnn_tie_probe.cu: sum over probed entries of the entries sharing their sort key ...
    unsigned long long tie_n;    // ... and the number of probed entries (ratio = mean run length; 1.0: no ties)
    int n_big_rows;              // direct-path rows whose degree exceeds the row buffer that still gives the fill pass
                                 // its full occupancy (k_count_big_rows): they get a fill launch of their own
    int cls_n[4];                // ... and the other direct-path rows by sort-network class (<= 64, 128, 256, more entries)
    int cls_fill[4];             // k_class_scatter: slots handed out per class
};

// Fused RBF epilogue of the fill passes (row_sort.cuh).  K4 = 0: no features in this launch.
struct DgRbfEpi {
    float dr;      // bin spacing
    float coef;    // -0.5 / sigma^2 * log2(e)
    float lognorm; // log2(1 / (sigma * sqrt(2 pi)))
    float inv_k4;  // 1 / (K/4)
    int K4;        // K / 4
    float4* out;   // E x K fp32 row-major
    // Gaussian recurrence across the 4 bins of a float4 (row_sort.cuh): g(k+1)/g(k) = r * q^(2i+1) with
    // r = 2^(c2dr * x0), q = 2^(coef * dr^2): two ex2 instead of four per float4
    float c2dr;    // 2 * coef * dr
    float q;       // 2^(coef * dr^2)
    float q2;      // q * q
    int recur;     // 1: use the recurrence in the fused epilogue
};
DgRbfEpi dg_make_rbf_epi(int K, double r_cutoff, double dr, float* out);  // dgnn_rbf.cu

struct DgTraverseArgs {
    // inputs
    const DgGeom* geoms;
    const long long* item_base;  // [S+1]
    long long n_items;
    long long n_structs;
    const long long* cellstart;  // [n_cells+1] offsets into the cloud
    const double* cx;            // cloud: image positions (reference arithmetic), cell-sorted
    const double* cy;
    const double* cz;
    const int* cj;               // cloud: local atom index
    const uint32_t* cn;          // cloud: packed image
    double r2;                   // r_c * r_c (fp64, computed once like neighbor_list.cpp:40)
    double r_cutoff;
    double rcm;                  // r_c * (1 + margin) for candidate generation
    double epsilon;
    float inv_rc;
    double key_scale;            // (2^32 - 4096) / r_c: 32-bit sort keys of the row sort
    int near_lo_hi, near_hi_hi;  // high words of (r_c -+ 2e-6)^2
    int max_neighbors;           // clamped to INT_MAX
    // capacities (shared memory carve-up)
    int tile_cap, tcs_cap, src_cap, row_cap, n_buckets, n_warps;
    // outputs
    int* deg_capped;             // [N] (count pass)
    int* deg_raw;                // [N] uncapped degrees
    uint32_t* hits;              // [N * hit_cap] per-row hit lists (cloud point index of every accepted pair), or null
    double* hits_d2;             // [N * hit_cap] their d2
    int hit_cap;
    const long long* row_ptr;    // [N+1] (fill pass)
    int* dst;
    uint32_t* image;
    double* dist;
    DgStats* stats;
    DgRbfEpi rbf;
    // fp32-tile traversal: the count pass records, per source atom, which slots of its flattened stencil list passed
    // the fp32 pre-filter (bit s of word w = list position 32 w + s; lists hold up to 1024 slots) and the list length;
    // the fill pass then compacts its candidates from the words instead of repeating the sweep.  Null: not recorded.
    uint32_t* cmask;             // [N * 32]
    int* cmask_tn;               // [N] padded list length the words belong to (-1: none, the fill pass sweeps)
};

#define DG_ROWS_CAP 1024  // tile rows (td_a * td_b) <= 32 * 32
#define DG_STN_CAP 961    // stencil rows (2R_a+1)(2R_b+1) <= 31 * 31

// ---- launch wrappers (dgnn_kernels.cu) ----
// small: classification of the direct path (may be null); structures it owns get status DG_ST_SKIP here
void dg_launch_geom(cudaStream_t st, long long S, const long long* atom_offsets, const double* lattice,
                    double r_cutoff, double cell_frac, const DgSmall* small, DgGeom* geoms);
void dg_launch_atom_cells(cudaStream_t st, long long N, long long S, const long long* atom_offsets,
                          const double* x, const double* y, const double* z, DgGeom* geoms, int* astruct,
                          int* acell);
void dg_launch_plan(cudaStream_t st, long long S, DgGeom* geoms, int tile_target, int src_target, int* ncells32,
                    int* nitems32, DgStats* stats);
void dg_launch_plan_finish(cudaStream_t st, long long S, DgGeom* geoms, const long long* cell_base,
                           const long long* item_base, DgStats* stats);
void dg_launch_cloud_count(cudaStream_t st, long long N, const DgGeom* geoms, const int* astruct,
                           const int* acell, int* cellcount);
void dg_launch_item_stats(cudaStream_t st, long long n_items, long long S, const DgGeom* geoms,
                          const long long* item_base, const long long* cellstart, long long n_cells,
                          DgStats* stats);
void dg_launch_cloud_scatter(cudaStream_t st, long long N, const DgGeom* geoms, const int* astruct,
                             const int* acell, const double* x, const double* y, const double* z,
                             const long long* cellstart, int* cellcount, double* cx, double* cy, double* cz,
                             int* cj, uint32_t* cn);
// fp32-tile traversal (dgnn_cell32.cu): rows of up to 512 entries; the fp64 traversal keeps the rest and oversize tiles
size_t dg_cell32_smem_bytes(const DgTraverseArgs& a, bool fill);
cudaError_t dg_launch_cell32(cudaStream_t st, const DgTraverseArgs& a, bool fill, int grid);
size_t dg_traverse_smem_bytes(const DgTraverseArgs& a, bool fill);
cudaError_t dg_launch_traverse(cudaStream_t st, const DgTraverseArgs& a, bool fill, int grid);
void dg_launch_finish_rows(cudaStream_t st, long long N, long long S, const long long* atom_offsets,
                           const long long* row_ptr, long long* struct_edge_offsets, DgStats* stats);
// exclusive scan int32 -> int64, out has n+1 entries (out[n] = total); scratch >= n/2048 + 2 entries
void dg_scan_i32_i64(cudaStream_t st, const int* in, long long* out, long long n, long long* scratch);
// same, offsets start at *base (device pointer, may be null = 0): chains the scans of consecutive parts
void dg_scan_i32_i64_based(cudaStream_t st, const int* in, long long* out, long long n, long long* scratch,
                           const long long* base);
int dg_num_kernel_launches_reset();  // returns launches since last reset

// ---- dgnn_direct.cu: direct (small-structure) path ----
struct DgDirectArgs {
    const DgSmall* small;        // [S]
    const int* astruct;          // [N] owning structure of every atom
    long long n_atoms;
    const double* x;             // Cartesian positions (reference arithmetic operands)
    const double* y;
    const double* z;
    const double* fa;            // fractional coordinates (candidate windows only)
    const double* fb;
    const double* fc;
    double r2;                   // r_c * r_c (fp64, computed once like neighbor_list.cpp:40)
    double r_cutoff;
    double epsilon;
    float inv_rc;
    double key_scale;            // (2^32 - 4096) / r_c: 32-bit sort keys of the row sort
    int near_lo_hi, near_hi_hi;  // high words of (r_c -+ 2e-6)^2
    int max_neighbors;           // clamped to INT_MAX
    int row_cap, n_buckets, n_warps;
    int eps_mode;                // self-exclusion rule: 0 none (eps <= 0), 1 fast (0 < eps <= DG_EPS_FAST), 2 generic
    double near_d2_lo, near_d2_hi;  // |sqrt(d2) - r_c| <= 1e-6  <=>  near_d2_lo <= d2 <= near_d2_hi
    int* deg_capped;             // [N] (count pass)
    int* deg_raw;                // [N] uncapped degrees (count pass writes, fill passes read)
    uint32_t* hits;              // [N * hit_cap] per-row hit lists written by the count pass (dg_pack_hit), or null
    double* hits_d2;             // [N * hit_cap] the accepted pairs' d2 (bit-exact reference arithmetic)
    const int* cloud_j;          // cell-list structures: hit word = cloud point index -> atom / packed image
    const uint32_t* cloud_n;
    int hit_cap;                 // slots per row (0: no hit lists, the fill pass repeats the traversal)
    const long long* row_ptr;    // [N+1] (fill pass)
    int* dst;
    uint32_t* image;
    double* dist;
    DgStats* stats;
    DgRbfEpi rbf;
    // single-pass build (k_direct<ONE>): traversal, row offsets by decoupled look-back over per-row status words, sort,
    // emit -- no count pass.  Rows with more than row_cap hits only reserve their range (a fill pass with
    // skip_deg_le = that row_cap completes them); rows ending beyond edge_cap set stats->cap_overflow.
    unsigned long long* lb_status;  // [N] zeroed before the launch: flag (2 bits) | value (62 bits)
    long long* row_ptr_out;         // [N+1]
    long long edge_cap;             // edges the output arrays can hold
    int skip_deg_le;                // fill pass: skip rows with deg_raw <= this (0: off), they are already written
    long long row_begin, row_end;   // k_direct: rows [row_begin, row_end) only (row_end = 0: all rows)
    int fill_hits_bulk;             // k_fill_hits: features through the bulk-copy engine (1) or st.global.cs (0)
    int lb_spin_limit;              // single-pass build: look-back spins before giving up (0: default 2^22)
    int dyn_unit;                   // > 0: rows are handed out to warps in units of this many rows from *unit_ctr (zeroed)
    unsigned long long* unit_ctr;
    double sym_band;                // symmetric count pass: |d2 - r_c^2| <= sym_band recomputes the mirror edge exactly
    // truncated rows: the count pass stores only accepted pairs with d2 below the structure's threshold (direct_sweep.cuh,
    // HitTally) and writes the stored count -- or INT_MAX for rows the list cannot serve -- to hits_n; the fill passes
    // are then given hits_n as their deg_raw.  Null: everything is stored and deg_raw decides.
    const double* hit_d2_max;       // [S]
    int* hits_n;                    // [N]
    int fill_hits_warps;            // k_fill_hits: warps per CTA (0: the count that puts the most warps on an SM)
    int full_tie_sort;              // uncapped fill pass: 1 = warp-cooperative tie ranking (batches the probe found tie-heavy)
    int defer_big;                  // fill pass: rows with more than row_cap pairs are left to a later launch (no overflow flag)
    // dynamically scheduled fill pass: rows in the order of row_perm[0, n_perm) (grouped by sort-network class, so that
    // the warps of an SM run the same code at the same time: the kernel is ~190 KB of SASS); null = rows in order
    const int* row_perm;
    long long n_perm;
    int split_cap;                  // > 0: k_cap_degrees also tallies stats->n_big_rows / cls_n (as k_count_big_rows does)
};
#define DG_EPS_FAST 1.0e-3
// hit word: neighbour atom (17 bits) | n_a + 16 (5 bits) << 17 | n_b + 16 << 22 | n_c + 16 << 27
#define DG_HIT_NMAX 15
#define DG_HIT_JMAX 0x1ffff
// eps_fast > 0: the traversal will use the fast self-exclusion (eps_mode 1); cells thinner than 2 eps are rejected
void dg_launch_small_geom(cudaStream_t st, long long S, const long long* atom_offsets, const double* lattice,
                          double r_cutoff, int direct_max_atoms, double eps_fast, DgSmall* small, DgStats* stats);
void dg_launch_direct_prep(cudaStream_t st, long long a_begin, long long a_end, long long S,
                           const long long* atom_offsets, const double* x, const double* y, const double* z,
                           const DgSmall* small, int* astruct, double* fa, double* fb, double* fc, DgStats* stats);
size_t dg_direct_smem_bytes(const DgDirectArgs& a, bool fill);
cudaError_t dg_launch_direct(cudaStream_t st, const DgDirectArgs& a, bool fill, int sm_count, int min_blocks);
void dg_launch_cap_degrees(cudaStream_t st, const DgDirectArgs& a, int sm_count);
// stats->n_big_rows += direct-path rows with deg_raw > cap
void dg_launch_count_big_rows(cudaStream_t st, long long N, const int* astruct, const DgSmall* small, const int* deg_raw,
                              int cap, DgStats* stats, int sm_count);
// perm[0, sum of stats->cls_n) = the direct-path rows with deg_raw <= cap grouped by class (after dg_launch_count_big_rows)
void dg_launch_class_scatter(cudaStream_t st, long long N, const int* astruct, const DgSmall* small, const int* deg_raw,
                             int cap, DgStats* stats, int* perm, int sm_count);
cudaError_t dg_launch_tie_probe(cudaStream_t st, const DgDirectArgs& a);  // dgnn_tie_probe.cu
cudaError_t dg_launch_direct_count_sym(cudaStream_t st, const DgDirectArgs& a, int sm_count, int min_blocks);
size_t dg_direct_onepass_smem_bytes(const DgDirectArgs& a);
cudaError_t dg_launch_direct_onepass(cudaStream_t st, const DgDirectArgs& a, int sm_count, int min_blocks);
// fill pass from the count pass's hit lists (rows with degree <= hit_cap and image range <= DG_HIT_NMAX)
size_t dg_fill_hits_smem_bytes(const DgDirectArgs& a);
cudaError_t dg_launch_fill_hits(cudaStream_t st, const DgDirectArgs& a, int sm_count);
// per-structure d2 below which about `target` neighbours are expected (+inf where that is most of the cutoff sphere)
void dg_launch_hit_thresholds(cudaStream_t st, long long S, const DgSmall* small, double target, double r2, double* out);
int dg_direct_launch_count_reset();

// ---- dgnn_rbf.cu ----
void dg_launch_rbf_f32(cudaStream_t st, const double* dist, long long n_edges, int K, double r_cutoff, double dr,
                       float* out, int sm_count);
void dg_launch_rbf_f64_colmajor(cudaStream_t st, const double* dist, long long n_edges, int K, double r_cutoff,
                                double dr, double* out);
int dg_rbf_launch_count_reset();

// ---- dgnn_collate.cu: consumers of the CSR (collation, viewer arrays, neighbourhood clouds) ----
void dg_launch_collate_edges(cudaStream_t st, long long N, long long E, const long long* row_ptr, const int* astruct,
                             const long long* atom_offsets, const int* dst, long long* edge_index, int sm_count);
void dg_launch_collate_nodes(cudaStream_t st, long long N, int F, const int* node_keys, const int* lut, int lut_size,
                             const float* table, const int* astruct, float* x, long long* batch, int* err,
                             int sm_count);
void dg_launch_edge_views(cudaStream_t st, long long N, const long long* row_ptr, const int* astruct,
                          const DgSmall* small, const double* x, const double* y, const double* z, const int* dst,
                          const uint32_t* image, const double* dist, int* sources, int* targets, float* distances,
                          float* displacements, int sm_count);
void dg_launch_neighborhood_clouds(cudaStream_t st, long long a0, long long a1, const long long* row_ptr,
                                   const int* astruct, const DgSmall* small, const double* x, const double* y,
                                   const double* z, const int* dst, const uint32_t* image, double* cloud,
                                   long long* cloud_ptr, int sm_count);
void dg_launch_table_counts(cudaStream_t st, long long a0, long long a1, const long long* row_ptr, int mode, int* counts);
cudaError_t dg_launch_distance_tables(cudaStream_t st, long long a0, long long a1, const long long* row_ptr,
                                      const int* astruct, const DgSmall* small, const double* x, const double* y,
                                      const double* z, const int* dst, const uint32_t* image, const long long* table_ptr,
                                      void* out, int mode, int max_degree, int sm_count);
int dg_collate_launch_count_reset();
