// dgnn_kernels.cu -- sm_100a kernels of the crystal-graph builder: geometry/planning, uniform cell-list binning of
// the pruned periodic image cloud (K1), tile-staged pairwise distance traversal with warp ballot/popc compaction
// and in-warp row sort into CSR (K2).  Compile with -fmad=false: the accept test must round like the reference's
// FMA-free build (CMakeLists.txt:33-37); the critical expressions additionally use explicit __d*_rn intrinsics.
//
// Reference behaviour reproduced (paths relative to the reference repo):
//   image cloud point  = pos_j + ((n_a*row0 + n_b*row1) + n_c*row2)          src/graph/neighbor_list.cpp:82-87
//   accept             : ((dx*dx)+(dy*dy))+(dz*dz) < r_c*r_c, dx = q - p      :40 + nanoflann L2_Simple_Adaptor
//   self exclusion     : orig == i && sqrt(d2) < epsilon                      :47-49
//   distance           = sqrt(d2)                                              :53
//   row order          : ascending distance, canonical tie-break, truncate     :56-62
#include <cuda_runtime.h>
#include <stdint.h>

#include "dgnn_internal.h"
#include "row_sort.cuh"
#include "cell_items.cuh"

static thread_local int g_launches = 0;  // per calling thread: a build runs wholly on one thread
int dg_num_kernel_launches_reset() {
    int v = g_launches;
    g_launches = 0;
    return v;
}
#define DG_COUNT_LAUNCH() (++g_launches)
void dg_kernel_launch_bump(int n) { g_launches += n; }

#ifndef FULL_MASK
#define FULL_MASK 0xffffffffu
#endif

// =====================================================================================================================
// Exclusive scan int32 -> int64 (three small kernels; inputs are a few million elements at most)
// =====================================================================================================================
#define SCAN_THREADS 256
#define SCAN_ITEMS 8
#define SCAN_TILE (SCAN_THREADS * SCAN_ITEMS)

__device__ __forceinline__ long long warp_incl_scan_ll(long long v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        long long t = __shfl_up_sync(FULL_MASK, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

// inclusive block scan of one value per thread (blockDim.x <= 1024); returns inclusive value, total in *total
__device__ __forceinline__ long long block_incl_scan_ll(long long v, long long* total, long long* s_warp) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    long long inc = warp_incl_scan_ll(v, lane);
    if (lane == 31) s_warp[w] = inc;
    __syncthreads();
    if (w == 0) {
        long long t = lane < nw ? s_warp[lane] : 0;
        long long ti = warp_incl_scan_ll(t, lane);
        if (lane < nw) s_warp[lane] = ti - t;  // exclusive warp offsets
        if (lane == 31) s_warp[32] = ti;
    }
    __syncthreads();
    inc += s_warp[w];
    *total = s_warp[32];
    __syncthreads();
    return inc;
}

__global__ void __launch_bounds__(SCAN_THREADS) k_scan_reduce(const int* __restrict__ in, long long n,
                                                              long long* __restrict__ bsum) {
    __shared__ long long s_warp[33];
    const long long base = (long long)blockIdx.x * SCAN_TILE;
    long long acc = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        const long long idx = base + (long long)i * SCAN_THREADS + threadIdx.x;
        if (idx < n) acc += in[idx];
    }
    long long total;
    block_incl_scan_ll(acc, &total, s_warp);
    if (threadIdx.x == 0) bsum[blockIdx.x] = total;
}

// single block: exclusive scan of bsum[0..nb) in place, bsum[nb] = grand total
__global__ void __launch_bounds__(1024) k_scan_bsums(long long* bsum, long long nb, const long long* base) {
    __shared__ long long s_warp[33];
    __shared__ long long s_carry;
    if (threadIdx.x == 0) s_carry = base ? *base : 0;
    __syncthreads();
    for (long long base = 0; base < nb; base += blockDim.x) {
        const long long idx = base + threadIdx.x;
        const long long v = idx < nb ? bsum[idx] : 0;
        long long total;
        const long long inc = block_incl_scan_ll(v, &total, s_warp);
        const long long carry = s_carry;
        if (idx < nb) bsum[idx] = carry + inc - v;
        __syncthreads();
        if (threadIdx.x == 0) s_carry = carry + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) bsum[nb] = s_carry;
}

__global__ void __launch_bounds__(SCAN_THREADS) k_scan_apply(const int* __restrict__ in, long long n,
                                                             const long long* __restrict__ bsum,
                                                             long long* __restrict__ out, long long nb) {
    __shared__ long long s_warp[33];
    const long long base = (long long)blockIdx.x * SCAN_TILE;
    // thread-contiguous layout so the per-thread partial scan is a simple serial prefix
    int v[SCAN_ITEMS];
    long long acc = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        const long long idx = base + (long long)threadIdx.x * SCAN_ITEMS + i;
        v[i] = idx < n ? in[idx] : 0;
        acc += v[i];
    }
    long long total;
    const long long inc = block_incl_scan_ll(acc, &total, s_warp);
    long long run = bsum[blockIdx.x] + inc - acc;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        const long long idx = base + (long long)threadIdx.x * SCAN_ITEMS + i;
        if (idx < n) out[idx] = run;
        run += v[i];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = bsum[nb];
}

__global__ void k_scan_empty(long long* out, const long long* base) { out[0] = base ? *base : 0; }

void dg_scan_i32_i64_based(cudaStream_t st, const int* in, long long* out, long long n, long long* scratch,
                           const long long* base) {
    if (n <= 0) {
        k_scan_empty<<<1, 1, 0, st>>>(out, base);
        DG_COUNT_LAUNCH();
        return;
    }
    const long long nb = (n + SCAN_TILE - 1) / SCAN_TILE;
    k_scan_reduce<<<(unsigned)nb, SCAN_THREADS, 0, st>>>(in, n, scratch);
    k_scan_bsums<<<1, 1024, 0, st>>>(scratch, nb, base);
    k_scan_apply<<<(unsigned)nb, SCAN_THREADS, 0, st>>>(in, n, scratch, out, nb);
    g_launches += 3;
}
void dg_scan_i32_i64(cudaStream_t st, const int* in, long long* out, long long n, long long* scratch) {
    dg_scan_i32_i64_based(st, in, out, n, scratch, nullptr);
}

// =====================================================================================================================
// K0: per-structure geometry, per-atom cell coordinates, planning
// =====================================================================================================================
__global__ void k_geom(long long S, const long long* __restrict__ atom_offsets, const double* __restrict__ lattice,
                       double r_cutoff, double cell_frac, const DgSmall* __restrict__ small,
                       DgGeom* __restrict__ geoms) {
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    DgGeom G;
    double L[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) L[i] = lattice[9 * s + i];
    const long long a0 = atom_offsets[s], a1 = atom_offsets[s + 1];
    dg_geom_init(&G, L, (int)(a1 - a0), r_cutoff, cell_frac);
    if (small && small[s].mode != DG_MODE_CELL) G.status = DG_ST_SKIP;
    G.atom_base = a0;
    G.cell_base = 0;
    G.item_base = 0;
    geoms[s] = G;
}

void dg_launch_geom(cudaStream_t st, long long S, const long long* atom_offsets, const double* lattice,
                    double r_cutoff, double cell_frac, const DgSmall* small, DgGeom* geoms) {
    if (S <= 0) return;
    k_geom<<<(unsigned)((S + 127) / 128), 128, 0, st>>>(S, atom_offsets, lattice, r_cutoff, cell_frac, small, geoms);
    DG_COUNT_LAUNCH();
}

__global__ void __launch_bounds__(256) k_atom_cells(long long N, long long S,
                                                    const long long* __restrict__ atom_offsets,
                                                    const double* __restrict__ x, const double* __restrict__ y,
                                                    const double* __restrict__ z, DgGeom* __restrict__ geoms,
                                                    int* __restrict__ astruct, int* __restrict__ acell) {
    const long long a = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = a < N;
    int s = -1, c0 = 0, c1 = 0, c2 = 0;
    bool ok = false;
    if (valid) {
        long long lo = 0, hi = S;  // last s with atom_offsets[s] <= a
        while (hi - lo > 1) {
            const long long mid = (lo + hi) >> 1;
            if (atom_offsets[mid] <= a) lo = mid; else hi = mid;
        }
        s = (int)lo;
        astruct[a] = s;
        const DgGeom* G = geoms + s;
        if (G->status == DG_ST_OK) {
            ok = true;
            const double px = x[a], py = y[a], pz = z[a];
            c0 = dg_cell_coord(G, 0, px, py, pz);
            c1 = dg_cell_coord(G, 1, px, py, pz);
            c2 = dg_cell_coord(G, 2, px, py, pz);
        }
        acell[3 * a + 0] = c0;
        acell[3 * a + 1] = c1;
        acell[3 * a + 2] = c2;
    }
    // warp-aggregated min/max when the whole warp sits in one structure (the common case)
    const int s0 = __shfl_sync(FULL_MASK, s, 0);
    const bool uniform = __all_sync(FULL_MASK, valid && ok && s == s0);
    if (uniform) {
        const int mn0 = __reduce_min_sync(FULL_MASK, c0), mx0 = __reduce_max_sync(FULL_MASK, c0);
        const int mn1 = __reduce_min_sync(FULL_MASK, c1), mx1 = __reduce_max_sync(FULL_MASK, c1);
        const int mn2 = __reduce_min_sync(FULL_MASK, c2), mx2 = __reduce_max_sync(FULL_MASK, c2);
        if ((threadIdx.x & 31) == 0) {
            DgGeom* G = geoms + s0;
            atomicMin(&G->cmin[0], mn0); atomicMax(&G->cmax[0], mx0);
            atomicMin(&G->cmin[1], mn1); atomicMax(&G->cmax[1], mx1);
            atomicMin(&G->cmin[2], mn2); atomicMax(&G->cmax[2], mx2);
        }
    } else if (valid && ok) {
        DgGeom* G = geoms + s;
        atomicMin(&G->cmin[0], c0); atomicMax(&G->cmax[0], c0);
        atomicMin(&G->cmin[1], c1); atomicMax(&G->cmax[1], c1);
        atomicMin(&G->cmin[2], c2); atomicMax(&G->cmax[2], c2);
    }
}

void dg_launch_atom_cells(cudaStream_t st, long long N, long long S, const long long* atom_offsets,
                          const double* x, const double* y, const double* z, DgGeom* geoms, int* astruct,
                          int* acell) {
    if (N <= 0) return;
    k_atom_cells<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(N, S, atom_offsets, x, y, z, geoms, astruct, acell);
    DG_COUNT_LAUNCH();
}

__global__ void k_plan(long long S, DgGeom* __restrict__ geoms, int tile_target, int src_target,
                       int* __restrict__ ncells32, int* __restrict__ nitems32, DgStats* __restrict__ stats) {
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    DgGeom* G = geoms + s;
    dg_geom_plan(G, tile_target, src_target);
    if (G->status != DG_ST_OK) {
        if (G->status != DG_ST_SKIP && atomicCAS(&stats->err_status, 0, G->status) == 0) stats->err_struct = (int)s;
        G->ncells = 0;
        G->nitems = 0;
    }
    ncells32[s] = (int)G->ncells;
    nitems32[s] = G->nitems;
}

void dg_launch_plan(cudaStream_t st, long long S, DgGeom* geoms, int tile_target, int src_target, int* ncells32,
                    int* nitems32, DgStats* stats) {
    if (S <= 0) return;
    k_plan<<<(unsigned)((S + 127) / 128), 128, 0, st>>>(S, geoms, tile_target, src_target, ncells32, nitems32,
                                                         stats);
    DG_COUNT_LAUNCH();
}

__global__ void k_plan_finish(long long S, DgGeom* __restrict__ geoms, const long long* __restrict__ cell_base,
                              const long long* __restrict__ item_base, DgStats* __restrict__ stats) {
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s == 0) {
        stats->n_cells = cell_base[S];
        stats->n_items = item_base[S];
    }
    if (s >= S) return;
    geoms[s].cell_base = cell_base[s];
    geoms[s].item_base = item_base[s];
}

void dg_launch_plan_finish(cudaStream_t st, long long S, DgGeom* geoms, const long long* cell_base,
                           const long long* item_base, DgStats* stats) {
    k_plan_finish<<<(unsigned)((S + 128) / 128), 128, 0, st>>>(S, geoms, cell_base, item_base, stats);
    DG_COUNT_LAUNCH();
}

// =====================================================================================================================
// K1: uniform cell-list binning of the pruned periodic image cloud (count -> scan -> scatter), one warp per atom
// =====================================================================================================================
struct ImageBox {
    int nlo[3], cnt[3], total;
};

__device__ __forceinline__ ImageBox image_box(const DgGeom* G, const int* c) {
    ImageBox b;
    b.cnt[0] = dg_image_range(G, 0, c[0], &b.nlo[0]);
    b.cnt[1] = dg_image_range(G, 1, c[1], &b.nlo[1]);
    b.cnt[2] = dg_image_range(G, 2, c[2], &b.nlo[2]);
    b.total = b.cnt[0] * b.cnt[1] * b.cnt[2];
    return b;
}

__device__ __forceinline__ long long halo_cell_id(const DgGeom* G, int X0, int X1, int X2) {
    return G->cell_base +
           ((long long)(X0 - G->hmin[0]) * G->hd[1] + (X1 - G->hmin[1])) * G->hd[2] + (X2 - G->hmin[2]);
}

__global__ void __launch_bounds__(256) k_cloud_count(long long N, const DgGeom* __restrict__ geoms,
                                                     const int* __restrict__ astruct, const int* __restrict__ acell,
                                                     int* __restrict__ cellcount) {
    // one THREAD per atom: in a large cell almost every atom contributes exactly one cloud point (itself); the few
    // images of boundary atoms -- or the ~100 of an atom in a tiny cell -- are a short serial loop
    const long long a = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= N) return;
    const DgGeom* G = geoms + astruct[a];
    if (G->status != DG_ST_OK) return;
    const int c[3] = {acell[3 * a], acell[3 * a + 1], acell[3 * a + 2]};
    const ImageBox b = image_box(G, c);
    for (int ia = 0; ia < b.cnt[0]; ++ia)
        for (int ib = 0; ib < b.cnt[1]; ++ib)
            for (int ic = 0; ic < b.cnt[2]; ++ic) {
                const int X0 = c[0] + (b.nlo[0] + ia) * G->g[0];
                const int X1 = c[1] + (b.nlo[1] + ib) * G->g[1];
                const int X2 = c[2] + (b.nlo[2] + ic) * G->g[2];
                atomicAdd(&cellcount[halo_cell_id(G, X0, X1, X2)], 1);
            }
}

void dg_launch_cloud_count(cudaStream_t st, long long N, const DgGeom* geoms, const int* astruct,
                           const int* acell, int* cellcount) {
    if (N <= 0) return;
    k_cloud_count<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(N, geoms, astruct, acell, cellcount);
    DG_COUNT_LAUNCH();
}

__global__ void __launch_bounds__(256) k_cloud_scatter(long long N, const DgGeom* __restrict__ geoms,
                                                       const int* __restrict__ astruct,
                                                       const int* __restrict__ acell, const double* __restrict__ x,
                                                       const double* __restrict__ y, const double* __restrict__ z,
                                                       const long long* __restrict__ cellstart,
                                                       int* __restrict__ cellcount, double* __restrict__ cx,
                                                       double* __restrict__ cy, double* __restrict__ cz,
                                                       int* __restrict__ cj, uint32_t* __restrict__ cn) {
    const long long a = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // one thread per atom (see k_cloud_count)
    if (a >= N) return;
    const DgGeom* G = geoms + astruct[a];
    if (G->status != DG_ST_OK) return;
    const int c[3] = {acell[3 * a], acell[3 * a + 1], acell[3 * a + 2]};
    const ImageBox b = image_box(G, c);
    const double px = x[a], py = y[a], pz = z[a];
    const int jloc = (int)(a - G->atom_base);
    for (int ia = 0; ia < b.cnt[0]; ++ia)
        for (int ib = 0; ib < b.cnt[1]; ++ib)
            for (int ic = 0; ic < b.cnt[2]; ++ic) {
                const int na = b.nlo[0] + ia, nb = b.nlo[1] + ib, nc = b.nlo[2] + ic;
                const long long cid = halo_cell_id(G, c[0] + na * G->g[0], c[1] + nb * G->g[1], c[2] + nc * G->g[2]);
                const int slot = atomicSub(&cellcount[cid], 1) - 1;  // counts down to zero: cell fill order is free
                const long long p = cellstart[cid] + slot;
                // reference cloud point: position + offset, offset = (n_a*row0 + n_b*row1) + n_c*row2
                cx[p] = DG_ADD(px, dg_image_offset(G->L, 0, na, nb, nc));
                cy[p] = DG_ADD(py, dg_image_offset(G->L, 1, na, nb, nc));
                cz[p] = DG_ADD(pz, dg_image_offset(G->L, 2, na, nb, nc));
                cj[p] = jloc;
                cn[p] = dg_pack_image(na, nb, nc);
            }
}

void dg_launch_cloud_scatter(cudaStream_t st, long long N, const DgGeom* geoms, const int* astruct,
                             const int* acell, const double* x, const double* y, const double* z,
                             const long long* cellstart, int* cellcount, double* cx, double* cy, double* cz,
                             int* cj, uint32_t* cn) {
    if (N <= 0) return;
    k_cloud_scatter<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(N, geoms, astruct, acell, x, y, z, cellstart,
                                                                       cellcount, cx, cy, cz, cj, cn);
    DG_COUNT_LAUNCH();
}

// ---------------------------------------------------------------------------------------------------------------------
// work items: (structure, source-cell block).  Shared decode used by the stats kernel and the traversal.
// ---------------------------------------------------------------------------------------------------------------------
// (ItemBox / find_struct / item_box: cell_items.cuh)
__global__ void k_item_stats(long long n_items, long long S, const DgGeom* __restrict__ geoms,
                             const long long* __restrict__ item_base, const long long* __restrict__ cellstart,
                             long long n_cells, DgStats* __restrict__ stats) {
    const long long item = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (item == 0) stats->cloud_points = cellstart[n_cells];
    if (item >= n_items) return;
    const long long s = find_struct(item_base, S, item);
    const DgGeom* G = geoms + s;
    const ItemBox B = item_box(G, (int)(item - G->item_base));
    long long pts = 0;
    for (int ra = 0; ra < B.td[0]; ++ra)
        for (int rb = 0; rb < B.td[1]; ++rb) {
            const long long g0 = G->cell_base + ((long long)(B.t0[0] + ra) * G->hd[1] + (B.t0[1] + rb)) * G->hd[2] + B.t0[2];
            pts += cellstart[g0 + B.td[2]] - cellstart[g0];
        }
    const int ipts = pts > 0x7fffffffLL ? 0x7fffffff : (int)pts;
    atomicMax(&stats->max_tile_pts, ipts);
    atomicMax(&stats->max_tile_tcs, B.td[0] * B.td[1] * (B.td[2] + 1));
}

void dg_launch_item_stats(cudaStream_t st, long long n_items, long long S, const DgGeom* geoms,
                          const long long* item_base, const long long* cellstart, long long n_cells,
                          DgStats* stats) {
    const long long n = n_items > 0 ? n_items : 1;
    k_item_stats<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(n_items, S, geoms, item_base, cellstart, n_cells, stats);
    DG_COUNT_LAUNCH();
}

// =====================================================================================================================
// K2: tile-staged pairwise distance traversal.  One CTA per work item: the item's neighbour tile (the cell-sorted
// image points of its source block plus the stencil halo) is staged in shared memory.  A warp takes one source CELL at
// a time: the cell's stencil (contiguous point ranges, one per stencil row) is flattened once into a per-warp candidate
// list, and every source atom of the cell then sweeps that list 32 candidates per step at full lane occupancy,
// compacting in-cutoff pairs with ballot/popc.
//   COUNT pass: row degrees (+ the hit lists the fill pass sorts).     FILL pass: row buffer -> sort -> CSR arrays.
// =====================================================================================================================
#define DG_IDX_CAP 768   // candidate list entries per warp (a 5x5x5 stencil of r_c/2 cells holds ~500 points)
#define DG_SRC_GROUP 64  // source points of one cell handled per sweep of the stencil

struct Smem {
    double *tx, *ty, *tz;
    int* tj;
    uint32_t* tn;
    uint32_t* tg;  // global cloud index of every tile point (hit lists of the count pass)
    int* tcs;
    int* rowbase;
    signed char* stn;
    uint32_t* idx;  // per-warp candidate list
    int* msrc;      // per-warp running degrees of the current source group
    RowBuf rb;      // per-warp (fill)
};

__host__ __device__ inline size_t align16(size_t v) { return (v + 15) & ~(size_t)15; }

__host__ __device__ inline size_t warp_smem_bytes(int row_cap, int n_buckets, bool fill) {
    return align16((size_t)DG_IDX_CAP * 4) + align16((size_t)DG_SRC_GROUP * 4) +
           (fill ? dg_rowbuf_bytes(row_cap, n_buckets) : 0);
}

__host__ __device__ inline size_t shared_part_bytes(int tile_cap, int tcs_cap) {
    return align16((size_t)tile_cap * 8) * 3 + align16((size_t)tile_cap * 4) * 3 + align16((size_t)tcs_cap * 4) +
           align16((size_t)(DG_ROWS_CAP + 1) * 4) + align16((size_t)DG_STN_CAP * 2);
}

size_t dg_traverse_smem_bytes(const DgTraverseArgs& a, bool fill) {
    return shared_part_bytes(a.tile_cap, a.tcs_cap) + (size_t)a.n_warps * warp_smem_bytes(a.row_cap, a.n_buckets, fill);
}

__device__ __forceinline__ Smem carve(unsigned char* base, const DgTraverseArgs& A, bool fill, int warp) {
    Smem s;
    unsigned char* p = base;
    s.tx = (double*)p; p += align16((size_t)A.tile_cap * 8);
    s.ty = (double*)p; p += align16((size_t)A.tile_cap * 8);
    s.tz = (double*)p; p += align16((size_t)A.tile_cap * 8);
    s.tj = (int*)p; p += align16((size_t)A.tile_cap * 4);
    s.tn = (uint32_t*)p; p += align16((size_t)A.tile_cap * 4);
    s.tg = (uint32_t*)p; p += align16((size_t)A.tile_cap * 4);
    s.tcs = (int*)p; p += align16((size_t)A.tcs_cap * 4);
    s.rowbase = (int*)p; p += align16((size_t)(DG_ROWS_CAP + 1) * 4);
    s.stn = (signed char*)p; p += align16((size_t)DG_STN_CAP * 2);
    p += (size_t)warp * warp_smem_bytes(A.row_cap, A.n_buckets, fill);
    s.idx = (uint32_t*)p; p += align16((size_t)DG_IDX_CAP * 4);
    s.msrc = (int*)p; p += align16((size_t)DG_SRC_GROUP * 4);
    s.rb = RowBuf{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    if (fill) s.rb = dg_rowbuf_carve(p, A.row_cap);
    return s;
}

// Candidate point accessors: STAGED reads the shared-memory tile, otherwise the global cloud.
template <bool STAGED>
struct PointView {
    const double *x, *y, *z;
    const int* j;
    const uint32_t* n;
    const uint32_t* g;  // STAGED: tile point -> global cloud index
};

struct WarpAcc {
    unsigned near;  // per-lane count of near-cutoff candidates
    int maxdeg;     // warp-uniform
};

// point range [s, e) of cells [z0, z1] of tile row (ra, rb) in view coordinates
template <bool STAGED>
__device__ __forceinline__ void row_range(const DgTraverseArgs& A, const DgGeom& G, const ItemBox& B, const Smem& sm,
                                          int ra, int rb, int z0, int z1, long long& s, long long& e) {
    if (STAGED) {
        const int base = (ra * B.td[1] + rb) * (B.td[2] + 1);
        s = sm.tcs[base + z0];
        e = sm.tcs[base + z1 + 1];
    } else {
        const long long g0 = G.cell_base + ((long long)(B.t0[0] + ra) * G.hd[1] + (B.t0[1] + rb)) * G.hd[2] + B.t0[2];
        s = A.cellstart[g0 + z0];
        e = A.cellstart[g0 + z1 + 1];
    }
}

// stencil cursor: row q of the (2R_a+1) x (2R_b+1) stencil, roff points of it already consumed
struct StencilCursor {
    int q;
    long long roff;
};

// Append stencil ranges of the source cell (ta, tb, tc) to the warp's candidate list until it is full or the stencil
// is exhausted.  Returns the number of candidates (0: done).
template <bool STAGED>
__device__ __forceinline__ int build_batch(const DgTraverseArgs& A, const DgGeom& G, const ItemBox& B, const Smem& sm,
                                           int ta, int tb, int tc, int lane, StencilCursor& cur) {
    const int NBr = 2 * G.R[1] + 1;
    const int nst = (2 * G.R[0] + 1) * NBr;
    int T = 0;
    int qa = cur.q / NBr, qb = cur.q - qa * NBr;
    while (cur.q < nst) {
        const int zlo = sm.stn[2 * cur.q], zhi = sm.stn[2 * cur.q + 1];
        if (zlo <= zhi) {
            long long s, e;
            row_range<STAGED>(A, G, B, sm, ta + qa - G.R[0], tb + qb - G.R[1], tc + zlo, tc + zhi, s, e);
            s += cur.roff;
            const long long len = e - s;
            const int room = DG_IDX_CAP - T;
            const int take = len < (long long)room ? (int)len : room;
            for (int t = lane; t < take; t += 32) sm.idx[T + t] = (uint32_t)(s + t);
            T += take;
            if ((long long)take < len) {  // list full: resume inside this row
                cur.roff += take;
                break;
            }
        }
        ++cur.q;
        cur.roff = 0;
        if (++qb == NBr) {
            qb = 0;
            ++qa;
        }
    }
    __syncwarp();
    return T;
}

// One source atom against the warp's candidate list idx[0, T).  m: running (uncapped) degree of the row.
template <bool STAGED, bool FILL>
__device__ __forceinline__ void sweep_list(const DgTraverseArgs& A, const Smem& sm, const PointView<STAGED>& pv, int T,
                                           double xi, double yi, double zi, int iloc, uint32_t* __restrict__ hrow,
                                           double* __restrict__ hd2, int lane, unsigned lt_mask, int& m,
                                           WarpAcc& acc) {
    for (int f0 = 0; f0 < T; f0 += 32) {
        const int f = f0 + lane;
        const bool valid = f < T;
        const long long pc = sm.idx[valid ? f : 0];
        const double dx = DG_SUB(xi, pv.x[pc]);
        const double dy = DG_SUB(yi, pv.y[pc]);
        const double dz = DG_SUB(zi, pv.z[pc]);
        // nanoflann evalMetric: result = 0; result += dx*dx; += dy*dy; += dz*dz
        const double d2 = DG_ADD(DG_ADD(DG_MUL(dx, dx), DG_MUL(dy, dy)), DG_MUL(dz, dz));
        bool hit = valid && (d2 < A.r2);
        if (hit && pv.j[pc] == iloc) {
            if (__dsqrt_rn(d2) < A.epsilon) hit = false;  // neighbor_list.cpp:47-49
        }
        if (!FILL) {
            const int hi = __double2hiint(d2);
            if (valid && hi >= A.near_lo_hi && hi <= A.near_hi_hi) {
                if (fabs(__dsqrt_rn(d2) - A.r_cutoff) <= 1e-6) acc.near++;
            }
        }
        const unsigned mask = __ballot_sync(FULL_MASK, hit);
        if (hit) {
            const int pos = m + __popc(mask & lt_mask);
            if (FILL) {
                if (pos < A.row_cap) {
                    sm.rb.rd[pos] = d2;
                    sm.rb.rj[pos] = pv.j[pc];
                    sm.rb.rn[pos] = pv.n[pc];
                }
            } else if (hrow != nullptr && pos < A.hit_cap) {  // hit list for the fill pass: cloud index + d2
                hrow[pos] = STAGED ? pv.g[pc] : (uint32_t)pc;
                hd2[pos] = d2;
            }
        }
        m += __popc(mask);
    }
}

// All source atoms of the block cell (ta, tb, tc) (tile coordinates).
template <bool STAGED, bool FILL>
__device__ __forceinline__ void process_cell(const DgTraverseArgs& A, const DgGeom& G, const ItemBox& B,
                                             const Smem& sm, const PointView<STAGED>& pv, int ta, int tb, int tc,
                                             int lane, WarpAcc& acc) {
    long long p0, p1;
    row_range<STAGED>(A, G, B, sm, ta, tb, tc, tc, p0, p1);
    if (p0 >= p1) return;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int nst = (2 * G.R[0] + 1) * (2 * G.R[1] + 1);
    if (FILL) {
        // rows left to this pass are rare (hit-list overflow): one source at a time, rebuilding the list per batch
        for (long long ps = p0; ps < p1; ++ps) {
            if (pv.n[ps] != 0u) continue;  // a periodic image that shares the cell: not a source
            const int iloc = pv.j[ps];
            const long long iglob = G.atom_base + iloc;
            if (A.hit_cap > 0 && A.deg_raw[iglob] <= A.hit_cap) continue;  // the hit-list fill pass owns this row
            const double xi = pv.x[ps], yi = pv.y[ps], zi = pv.z[ps];
            StencilCursor cur{0, 0};
            int m = 0;
            while (cur.q < nst) {
                const int T = build_batch<STAGED>(A, G, B, sm, ta, tb, tc, lane, cur);
                if (T == 0) break;
                sweep_list<STAGED, true>(A, sm, pv, T, xi, yi, zi, iloc, nullptr, nullptr, lane, lt_mask, m, acc);
                __syncwarp();
            }
            __syncwarp();
            acc.maxdeg = m > acc.maxdeg ? m : acc.maxdeg;
            const int keep = m < A.max_neighbors ? m : A.max_neighbors;
            if (m > A.row_cap) {  // cannot happen: row_cap >= max degree of the count pass
                if (lane == 0) atomicExch(&A.stats->overflow, 1);
                continue;
            }
            dg_row_sort_emit(sm.rb, m, keep, A.n_buckets, A.inv_rc, A.key_scale, A.row_ptr[iglob], A.dst, A.image,
                             A.dist, A.stats, lane, A.rbf);
        }
        return;
    }
    // COUNT: the stencil is flattened once per source group and shared by the group's atoms
    for (long long g0 = p0; g0 < p1; g0 += DG_SRC_GROUP) {
        const long long g1 = g0 + DG_SRC_GROUP < p1 ? g0 + DG_SRC_GROUP : p1;
        for (int k = lane; k < (int)(g1 - g0); k += 32) sm.msrc[k] = 0;
        __syncwarp();
        StencilCursor cur{0, 0};
        while (cur.q < nst) {
            const int T = build_batch<STAGED>(A, G, B, sm, ta, tb, tc, lane, cur);
            if (T == 0) break;
            for (long long ps = g0; ps < g1; ++ps) {
                if (pv.n[ps] != 0u) continue;
                const int iloc = pv.j[ps];
                const long long iglob = G.atom_base + iloc;
                uint32_t* hrow = A.hits ? A.hits + (size_t)iglob * (size_t)A.hit_cap : nullptr;
                double* hd2 = A.hits ? A.hits_d2 + (size_t)iglob * (size_t)A.hit_cap : nullptr;
                int m = sm.msrc[ps - g0];
                sweep_list<STAGED, false>(A, sm, pv, T, pv.x[ps], pv.y[ps], pv.z[ps], iloc, hrow, hd2, lane, lt_mask,
                                          m, acc);
                __syncwarp();
                if (lane == 0) sm.msrc[ps - g0] = m;
            }
            __syncwarp();
        }
        for (long long ps = g0 + lane; ps < g1; ps += 32) {
            if (pv.n[ps] != 0u) continue;
            const long long iglob = G.atom_base + pv.j[ps];
            const int m = sm.msrc[ps - g0];
            A.deg_capped[iglob] = m < A.max_neighbors ? m : A.max_neighbors;
            A.deg_raw[iglob] = m;
            if (A.hits && m > A.hit_cap) atomicAdd(&A.stats->n_fallback_rows, 1);
            acc.maxdeg = m > acc.maxdeg ? m : acc.maxdeg;  // per lane here; reduced at the end of the kernel
        }
        __syncwarp();
    }
}

template <bool FILL>
__global__ void __launch_bounds__(256, 2) k_traverse(const DgTraverseArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ DgGeom sG;
    __shared__ long long s_struct;
    __shared__ int s_staged, s_maxdeg;
    __shared__ unsigned s_near;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nthreads = blockDim.x, nwarps = nthreads >> 5;
    const Smem sm = carve(smem_raw, A, FILL, warp);
    WarpAcc acc;
    acc.near = 0;
    acc.maxdeg = 0;
    if (tid == 0) {
        s_maxdeg = 0;
        s_near = 0;
    }

    for (long long item = blockIdx.x; item < A.n_items; item += gridDim.x) {
        __syncthreads();  // previous item's shared state is dead
        if (tid == 0) s_struct = find_struct(A.item_base, A.n_structs, item);
        __syncthreads();
        {
            const int* src = reinterpret_cast<const int*>(A.geoms + s_struct);
            int* dstw = reinterpret_cast<int*>(&sG);
            for (int i = tid; i < (int)(sizeof(DgGeom) / 4); i += nthreads) dstw[i] = src[i];
        }
        __syncthreads();
        const DgGeom& G = sG;
        const ItemBox B = item_box(&G, (int)(item - G.item_base));
        const int nrows = B.td[0] * B.td[1];
        const int tcs_need = nrows * (B.td[2] + 1);

        // ---- stage 1: per-row point counts -> exclusive row bases ----
        for (int r = tid; r < nrows; r += nthreads) {
            const int ra = r / B.td[1], rb = r % B.td[1];
            const long long g0 = G.cell_base + ((long long)(B.t0[0] + ra) * G.hd[1] + (B.t0[1] + rb)) * G.hd[2] + B.t0[2];
            sm.rowbase[r] = (int)(A.cellstart[g0 + B.td[2]] - A.cellstart[g0]);
        }
        __syncthreads();
        if (warp == 0) {
            const int per = (nrows + 31) >> 5;  // <= 32
            const int b0 = lane * per;
            int sum = 0;
            for (int i = 0; i < per; ++i) {
                const int r = b0 + i;
                if (r < nrows) sum += sm.rowbase[r];
            }
            int inc = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(FULL_MASK, inc, o);
                if (lane >= o) inc += t;
            }
            int run = inc - sum;
            for (int i = 0; i < per; ++i) {
                const int r = b0 + i;
                if (r < nrows) {
                    const int c = sm.rowbase[r];
                    sm.rowbase[r] = run;
                    run += c;
                }
            }
            const int total = __shfl_sync(FULL_MASK, inc, 31);
            if (lane == 0) {
                sm.rowbase[nrows] = total;
                s_staged = (total <= A.tile_cap && tcs_need <= A.tcs_cap) ? 1 : 0;
            }
        }
        // ---- (independent) stencil row table ----
        {
            const int NBr = 2 * G.R[1] + 1;
            const int nst = (2 * G.R[0] + 1) * NBr;
            for (int q = tid; q < nst; q += nthreads) {
                int zlo = 1, zhi = 0;
                if (!dg_stencil_row(&G, q / NBr - G.R[0], q % NBr - G.R[1], A.rcm, &zlo, &zhi)) {
                    zlo = 1;
                    zhi = 0;
                }
                sm.stn[2 * q] = (signed char)zlo;
                sm.stn[2 * q + 1] = (signed char)zhi;
            }
        }
        __syncthreads();

        const int nbc0 = B.c1[0] - B.c0[0], nbc1 = B.c1[1] - B.c0[1], nbc2 = B.c1[2] - B.c0[2];
        const int nbcells = nbc0 * nbc1 * nbc2;
        const bool staged = s_staged != 0;
        if (staged) {
            // ---- stage 2: copy the tile (cell-sorted image points) and its cell-start table ----
            for (int r = warp; r < nrows; r += nwarps) {
                const int ra = r / B.td[1], rb = r % B.td[1];
                const long long g0 = G.cell_base + ((long long)(B.t0[0] + ra) * G.hd[1] + (B.t0[1] + rb)) * G.hd[2] + B.t0[2];
                const long long gs = A.cellstart[g0];
                const int lb = sm.rowbase[r], cnt = sm.rowbase[r + 1] - lb;
                for (int p = lane; p < cnt; p += 32) {
                    sm.tx[lb + p] = A.cx[gs + p];
                    sm.ty[lb + p] = A.cy[gs + p];
                    sm.tz[lb + p] = A.cz[gs + p];
                    sm.tj[lb + p] = A.cj[gs + p];
                    sm.tn[lb + p] = A.cn[gs + p];
                    sm.tg[lb + p] = (uint32_t)(gs + p);
                }
                for (int c = lane; c <= B.td[2]; c += 32)
                    sm.tcs[r * (B.td[2] + 1) + c] = lb + (int)(A.cellstart[g0 + c] - gs);
            }
            __syncthreads();
            PointView<true> pv{sm.tx, sm.ty, sm.tz, sm.tj, sm.tn, sm.tg};
            for (int bc = warp; bc < nbcells; bc += nwarps) {
                const int c2 = bc % nbc2, c1 = (bc / nbc2) % nbc1, c0 = bc / (nbc2 * nbc1);
                process_cell<true, FILL>(A, G, B, sm, pv, c0 + G.R[0], c1 + G.R[1], c2 + G.R[2], lane, acc);
            }
        } else {
            // ---- unstaged fallback: same traversal straight from the global cloud (L1/L2 cached) ----
            PointView<false> pv{A.cx, A.cy, A.cz, A.cj, A.cn, nullptr};
            for (int bc = warp; bc < nbcells; bc += nwarps) {
                const int c2 = bc % nbc2, c1 = (bc / nbc2) % nbc1, c0 = bc / (nbc2 * nbc1);
                process_cell<false, FILL>(A, G, B, sm, pv, c0 + G.R[0], c1 + G.R[1], c2 + G.R[2], lane, acc);
            }
        }
    }
    // ---- per-CTA statistics ----
    if (!FILL) {
        unsigned near = acc.near;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) near += __shfl_xor_sync(FULL_MASK, near, o);
        const int maxdeg = __reduce_max_sync(FULL_MASK, acc.maxdeg);
        __syncthreads();
        if (lane == 0) {
            if (near) atomicAdd(&s_near, near);
            atomicMax(&s_maxdeg, maxdeg);
        }
        __syncthreads();
        if (tid == 0) {
            if (s_near) atomicAdd(&A.stats->n_near, (unsigned long long)s_near);
            atomicMax(&A.stats->max_deg, s_maxdeg);
        }
    }
}

cudaError_t dg_launch_traverse(cudaStream_t st, const DgTraverseArgs& a, bool fill, int grid) {
    if (a.n_items <= 0) return cudaSuccess;
    const size_t smem = dg_traverse_smem_bytes(a, fill);
    cudaError_t e;
    const int threads = a.n_warps * 32;
    if (fill) {
        e = dg_raise_dyn_smem<k_traverse<true>>();
        if (e != cudaSuccess) return e;
        k_traverse<true><<<grid, threads, smem, st>>>(a);
    } else {
        e = dg_raise_dyn_smem<k_traverse<false>>();
        if (e != cudaSuccess) return e;
        k_traverse<false><<<grid, threads, smem, st>>>(a);
    }
    DG_COUNT_LAUNCH();
    return cudaGetLastError();
}

// struct_edge_offsets[s] = row_ptr[atom_offsets[s]];  stats->n_edges = row_ptr[N]
__global__ void k_finish_rows(long long N, long long S, const long long* __restrict__ atom_offsets,
                              const long long* __restrict__ row_ptr, long long* __restrict__ seo,
                              DgStats* __restrict__ stats) {
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s == 0) stats->n_edges = row_ptr[N];
    if (s <= S) seo[s] = row_ptr[atom_offsets[s]];
}

void dg_launch_finish_rows(cudaStream_t st, long long N, long long S, const long long* atom_offsets,
                           const long long* row_ptr, long long* struct_edge_offsets, DgStats* stats) {
    k_finish_rows<<<(unsigned)((S + 1 + 127) / 128), 128, 0, st>>>(N, S, atom_offsets, row_ptr, struct_edge_offsets,
                                                                    stats);
    DG_COUNT_LAUNCH();
}
