// dgnn_direct.cu -- direct (small-structure) traversal: the hot path for batches of defect supercells (BASELINE
// configs 2, 3, 5).  A few hundred atoms per structure make a cell list degenerate (2-4 cells per axis), so no grid
// is built at all: one warp owns one source atom, sweeps the structure's atoms from L1 (a structure is a few KB),
// and for each candidate atom tests only the periodic images inside its fractional-coordinate window (geom.h,
// dg_axis_window) -- ~2 images per pair for a 16 x 16 x 8 A cell at r_c = 8 A instead of the reference's 125.
// In-cutoff pairs are compacted with ballot/popc into the warp's shared-memory row, sorted in-warp and written as
// one contiguous CSR row.  Two traversals: lanes over atoms (ATOMPAR, supercells) or lanes over the images of one
// pair (IMGPAR, tiny cells with many images per pair).  Compile with -fmad=false.
//
// Reference behaviour reproduced (paths relative to the reference repo):
//   image cloud point  = pos_j + ((n_a*row0 + n_b*row1) + n_c*row2)          src/graph/neighbor_list.cpp:82-87
//   accept             : ((dx*dx)+(dy*dy))+(dz*dz) < r_c*r_c, dx = q - p      :40 + nanoflann L2_Simple_Adaptor
//   image range        : |n_k| <= ceil(r_c / min|row|) + 1                     :68-72, 79-81
//   self exclusion     : orig == i && sqrt(d2) < epsilon                      :47-49
//   row order / cap    : ascending distance, canonical tie-break, truncate     :56-62
#include <cuda_runtime.h>
#include <stdint.h>

#include "dgnn_internal.h"
#include "row_sort.cuh"
#include "direct_sweep.cuh"

static thread_local int g_direct_launches = 0;
void dg_direct_launch_bump(int n) { g_direct_launches += n; }
int dg_direct_launch_count_reset() {
    int v = g_direct_launches;
    g_direct_launches = 0;
    return v;
}

// ---------------------------------------------------------------------------------------------------------------------
// per-structure geometry + classification
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_small_geom(long long S, const long long* __restrict__ atom_offsets,
                                                    const double* __restrict__ lattice, double r_cutoff,
                                                    int direct_max_atoms, double eps_fast,
                                                    DgSmall* __restrict__ small, DgStats* __restrict__ stats) {
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    double L[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) L[i] = lattice[9 * s + i];
    const long long a0 = atom_offsets[s], a1 = atom_offsets[s + 1];
    DgSmall G;
    dg_small_init(&G, L, a0, (int)(a1 - a0), r_cutoff, direct_max_atoms);
    if (G.status == DG_ST_OK && G.mode != DG_MODE_CELL && eps_fast > 0.0) {
        // fast self-exclusion (EPS = 1) assumes no periodic self-image is closer than epsilon: every lattice
        // translation is at least one perpendicular height h_k = rcm / w_k long
        const double rcm = DG_RCM(r_cutoff);
        const double wmax = fmax(G.w[0], fmax(G.w[1], G.w[2]));
        if (!(rcm > 2.0 * eps_fast * wmax)) G.status = DG_ST_GEOMETRY;
    }
    small[s] = G;
    if (G.status != DG_ST_OK) {
        if (atomicCAS(&stats->err_status, 0, G.status) == 0) stats->err_struct = (int)s;
    } else if (G.natoms > 0) {
        if (G.mode == DG_MODE_CELL) atomicAdd(&stats->n_cell_structs, 1);
        else atomicAdd(&stats->n_direct_structs, 1);
        // density estimate of the degree: N * (4/3) pi r^3 / V, V = h_0 h_1 h_2 / (the sines) >= rcm^3 / (w0 w1 w2)
        const double rcm = DG_RCM(r_cutoff);
        const double det = fabs(L[0] * (L[4] * L[8] - L[5] * L[7]) + L[1] * (L[5] * L[6] - L[3] * L[8]) +
                                L[2] * (L[3] * L[7] - L[4] * L[6]));
        const double est = (double)G.natoms * 4.18879020478639 * rcm * rcm * rcm / det;
        atomicMax(&stats->est_deg, est < 1.0e6 ? (int)est + 1 : 1000000);
        atomicMax(&stats->est_sparse, 1000000 - (est < 999999.0 ? (int)est : 999999));
        if (G.mode != DG_MODE_CELL)
            atomicAdd(&stats->est_edges, (unsigned long long)((double)G.natoms * (est < 1.0e6 ? est + 1.0 : 1.0e6)));
    }
}

void dg_launch_small_geom(cudaStream_t st, long long S, const long long* atom_offsets, const double* lattice,
                          double r_cutoff, int direct_max_atoms, double eps_fast, DgSmall* small, DgStats* stats) {
    if (S <= 0) return;
    k_small_geom<<<(unsigned)((S + 127) / 128), 128, 0, st>>>(S, atom_offsets, lattice, r_cutoff, direct_max_atoms,
                                                               eps_fast, small, stats);
    dg_direct_launch_bump(1);
}

// per-atom: owning structure and fractional coordinates
__global__ void __launch_bounds__(256) k_direct_prep(long long a_begin, long long N, long long S,
                                                     const long long* __restrict__ atom_offsets,
                                                     const double* __restrict__ x, const double* __restrict__ y,
                                                     const double* __restrict__ z,
                                                     const DgSmall* __restrict__ small, int* __restrict__ astruct,
                                                     double* __restrict__ fa, double* __restrict__ fb,
                                                     double* __restrict__ fc, DgStats* __restrict__ stats) {
    const long long a = a_begin + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= N) return;
    long long lo = 0, hi = S;  // last s with atom_offsets[s] <= a
    while (hi - lo > 1) {
        const long long mid = (lo + hi) >> 1;
        if (atom_offsets[mid] <= a) lo = mid; else hi = mid;
    }
    astruct[a] = (int)lo;
    const DgSmall* G = small + lo;
    const double px = x[a], py = y[a], pz = z[a];
    const double f0 = dg_frac_coord(G->Linv, 0, px, py, pz);
    const double f1 = dg_frac_coord(G->Linv, 1, px, py, pz);
    const double f2 = dg_frac_coord(G->Linv, 2, px, py, pz);
    fa[a] = f0;
    fb[a] = f1;
    fc[a] = f2;
    // the traversal's integer windows need bounded fractional coordinates (also rejects NaN / Inf positions)
    if (G->status == DG_ST_OK && G->mode != DG_MODE_CELL &&
        !(fabs(f0) <= 1.0e6 && fabs(f1) <= 1.0e6 && fabs(f2) <= 1.0e6)) {
        if (atomicCAS(&stats->err_status, 0, DG_ST_UNWRAPPED) == 0) stats->err_struct = (int)lo;
    }
}

// atoms [a_begin, a_end)
void dg_launch_direct_prep(cudaStream_t st, long long a_begin, long long a_end, long long S,
                           const long long* atom_offsets, const double* x, const double* y, const double* z,
                           const DgSmall* small, int* astruct, double* fa, double* fb, double* fc, DgStats* stats) {
    if (a_end <= a_begin) return;
    k_direct_prep<<<(unsigned)((a_end - a_begin + 255) / 256), 256, 0, st>>>(a_begin, a_end, S, atom_offsets, x, y, z,
                                                                              small, astruct, fa, fb, fc, stats);
    dg_direct_launch_bump(1);
}

// DYN: rows are handed out to WARPS in units of A.dyn_unit consecutive rows from a device counter instead of the static
// round-robin of 64-row chunks over CTAs -- for small batches of uneven rows (config 3: 2-8 atoms per cell, 343-729
// images per pair) the static schedule leaves most of the grid waiting for the CTAs that drew the expensive cells.
template <bool FILL, int MINB, int EPS, bool HITS, bool LEAN = false, bool DYN = false>
__global__ void __launch_bounds__(256, MINB) k_direct(const DgDirectArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    RowBuf rb = RowBuf{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    if (FILL) rb = dg_rowbuf_carve(smem_raw + (size_t)warp * dg_rowbuf_bytes(A.row_cap, A.n_buckets), A.row_cap);
    unsigned near = 0;
    int maxdeg = 0;
    // rows [row_begin, row_end) (row_end = 0: all rows); the count pass runs range by range behind the chunked upload
    const long long r0 = A.row_begin, r1 = A.row_end > 0 ? A.row_end : A.n_atoms;
    const long long n_chunks = (r1 - r0 + DG_ROW_CHUNK - 1) / DG_ROW_CHUNK;

    // (the row body is a macro shared by the two schedules: the static loop nest must stay exactly as it was -- routing
    // both through one generic loop cost the headline kernel 4.7 %, same-box A/B 8.72 against 8.32 ms)
#define DG_DIRECT_ROW_BODY                                                                                               \
            const DgSmall* __restrict__ G = A.small + A.astruct[a];                                                      \
            const int mode = G->mode;                                                                                    \
            if (mode == DG_MODE_CELL || G->status != DG_ST_OK) continue; /* the cell-list path owns this structure */    \
            const int nimg = G->nimg;                                                                                    \
            /* rows the hit-list fill pass (k_fill_hits) owns are skipped here; this pass takes the rest */              \
            if (FILL && A.hit_cap > 0 && nimg <= DG_HIT_NMAX && A.deg_raw[a] <= A.hit_cap) continue;                     \
            /* ... and so are rows the single-pass build has already written */                                          \
            if (FILL && A.skip_deg_le > 0 && A.deg_raw[a] <= A.skip_deg_le) continue;                                    \
            const bool store_hits = HITS && nimg <= DG_HIT_NMAX;                                                         \
            uint32_t* __restrict__ hrow = HITS ? A.hits + (size_t)a * (size_t)A.hit_cap : nullptr;                       \
            double* __restrict__ hd2 = HITS ? A.hits_d2 + (size_t)a * (size_t)A.hit_cap : nullptr;                       \
            HitTally ht;                                                                                                 \
            if (HITS) {                                                                                                  \
                ht.tot = ht.lo = 0;                                                                                      \
                ht.d2_max = A.hit_d2_max ? A.hit_d2_max[A.astruct[a]] : DG_INF;                                          \
                ht.d2_lo = ht.d2_max * (1.0 - 3.814697265625e-6);                                                        \
            }                                                                                                            \
            int m = HITS ? traverse_row<FILL, EPS, HITS, false, DYN>(A, G, a, mode, rb, hrow, hd2, lane, lt_mask, near, ht) \
                         : traverse_row<FILL, EPS, HITS, false, DYN>(A, G, a, mode, rb, hrow, hd2, lane, lt_mask, near);  \
            if (!FILL && !HITS) m = __reduce_add_sync(FULL_MASK, m);                                                     \
            int mh = 0, mlo = 0; /* HITS: stored pairs, pairs safely below the storing threshold */                      \
            if (HITS) {                                                                                                  \
                mh = m;                                                                                                  \
                m = __reduce_add_sync(FULL_MASK, ht.tot);                                                                \
                mlo = __reduce_add_sync(FULL_MASK, ht.lo);                                                               \
            }                                                                                                            \
            maxdeg = m > maxdeg ? m : maxdeg;                                                                            \
            const int keep = m < A.max_neighbors ? m : A.max_neighbors;                                                  \
            if (!FILL) {                                                                                                 \
                if (lane == 0) {                                                                                         \
                    A.deg_capped[a] = keep;                                                                              \
                    A.deg_raw[a] = m;                                                                                    \
                    if (HITS) {                                                                                          \
                        /* the list serves the fill pass if it is complete, or holds the first max_neighbors + 1 ranks */  \
                        const bool ok = store_hits && mh <= A.hit_cap && (mh == m || mlo > A.max_neighbors);             \
                        if (A.hits_n) A.hits_n[a] = ok ? mh : 0x7fffffff;                                                \
                        if (!ok) atomicAdd(&A.stats->n_fallback_rows, 1);                                                \
                    }                                                                                                    \
                }                                                                                                        \
            } else if (m > A.row_cap) { /* defer_big: a later launch with a larger row buffer owns the row; otherwise */ \
                /* it cannot happen (row_cap >= max degree of the count pass) */                                        \
                if (lane == 0 && !A.defer_big) atomicExch(&A.stats->overflow, 1);                                        \
            } else {                                                                                                     \
                dg_row_sort_emit<LEAN, true>(rb, m, keep, A.n_buckets, A.inv_rc, A.key_scale, A.row_ptr[a], A.dst, A.image,    \
                                       A.dist, A.stats, lane, A.rbf);                                                    \
            }
    if (DYN) {
        const int* __restrict__ perm = A.row_perm;
        const long long n_rows = perm ? A.n_perm : r1 - r0;
        for (;;) {
            unsigned long long u = 0;
            if (lane == 0) u = atomicAdd(A.unit_ctr, 1ULL);
            u = __shfl_sync(FULL_MASK, u, 0);
            const long long i_first = (long long)u * A.dyn_unit;
            if (i_first >= n_rows) break;
            const long long i_end = i_first + A.dyn_unit < n_rows ? i_first + A.dyn_unit : n_rows;
            for (long long i = i_first; i < i_end; ++i) {
                const long long a = perm ? (long long)perm[i] : r0 + i;
                DG_DIRECT_ROW_BODY
            }
        }
    } else {
        for (long long chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
            const long long a_end = r0 + (chunk + 1) * DG_ROW_CHUNK < r1 ? r0 + (chunk + 1) * DG_ROW_CHUNK : r1;
            for (long long a = r0 + chunk * DG_ROW_CHUNK + warp; a < a_end; a += nwarps) {
                DG_DIRECT_ROW_BODY
            }
        }
    }
#undef DG_DIRECT_ROW_BODY
    if (!FILL) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) near += __shfl_xor_sync(FULL_MASK, near, o);
        if (lane == 0) {
            if (near) atomicAdd(&A.stats->n_near, (unsigned long long)near);
            if (maxdeg) atomicMax(&A.stats->max_deg, maxdeg);
        }
    }
}

// Rows whose degree exceeds `cap` (direct-path structures only): a handful of dense rows would otherwise size every
// warp's row buffer and take the whole fill pass down to a quarter of its occupancy (config 3: mean degree 180, max 841).
__device__ __forceinline__ int dg_row_class(int m) { return m <= 64 ? 0 : (m <= 128 ? 1 : (m <= 256 ? 2 : 3)); }
__global__ void __launch_bounds__(256) k_count_big_rows(long long N, const int* __restrict__ astruct,
                                                        const DgSmall* __restrict__ small,
                                                        const int* __restrict__ deg_raw, int cap,
                                                        DgStats* __restrict__ stats) {
    int n = 0, c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    for (long long a = (long long)blockIdx.x * blockDim.x + threadIdx.x; a < N; a += (long long)gridDim.x * blockDim.x) {
        const DgSmall* __restrict__ G = small + astruct[a];
        if (G->mode == DG_MODE_CELL || G->status != DG_ST_OK) continue;
        const int m = deg_raw[a];
        if (m > cap) {
            ++n;
        } else {
            const int c = dg_row_class(m);
            c0 += c == 0; c1 += c == 1; c2 += c == 2; c3 += c == 3;
        }
    }
    n = __reduce_add_sync(FULL_MASK, n);
    c0 = __reduce_add_sync(FULL_MASK, c0);
    c1 = __reduce_add_sync(FULL_MASK, c1);
    c2 = __reduce_add_sync(FULL_MASK, c2);
    c3 = __reduce_add_sync(FULL_MASK, c3);
    if ((threadIdx.x & 31) == 0) {
        if (n) atomicAdd(&stats->n_big_rows, n);
        if (c0) atomicAdd(&stats->cls_n[0], c0);
        if (c1) atomicAdd(&stats->cls_n[1], c1);
        if (c2) atomicAdd(&stats->cls_n[2], c2);
        if (c3) atomicAdd(&stats->cls_n[3], c3);
    }
}
void dg_launch_count_big_rows(cudaStream_t st, long long N, const int* astruct, const DgSmall* small, const int* deg_raw,
                              int cap, DgStats* stats, int sm_count) {
    if (N <= 0) return;
    long long g = (N + 255) / 256;
    if (g > (long long)sm_count * 8) g = (long long)sm_count * 8;
    k_count_big_rows<<<(unsigned)g, 256, 0, st>>>(N, astruct, small, deg_raw, cap, stats);
    dg_direct_launch_bump(1);
}
// Rows grouped by class, runs of consecutive rows of a class kept together (a warp places its 32 rows class by class
// with one atomic per class): the atoms of a structure mostly stay adjacent, so its lines stay in L1.
__global__ void __launch_bounds__(256) k_class_scatter(long long N, const int* __restrict__ astruct,
                                                       const DgSmall* __restrict__ small,
                                                       const int* __restrict__ deg_raw, int cap,
                                                       DgStats* __restrict__ stats, int* __restrict__ perm) {
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int b1 = stats->cls_n[0], b2 = b1 + stats->cls_n[1], b3 = b2 + stats->cls_n[2];
    const long long n32 = (N + 31) & ~31LL;
    for (long long a = (long long)blockIdx.x * blockDim.x + threadIdx.x; a < n32; a += (long long)gridDim.x * blockDim.x) {
        int c = -1;
        if (a < N) {
            const DgSmall* __restrict__ G = small + astruct[a];
            if (G->mode != DG_MODE_CELL && G->status == DG_ST_OK) {
                const int m = deg_raw[a];
                if (m <= cap) c = dg_row_class(m);
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const unsigned mask = __ballot_sync(FULL_MASK, c == k);
            if (mask == 0u) continue;
            int base = 0;
            if (lane == __ffs(mask) - 1) base = atomicAdd(&stats->cls_fill[k], __popc(mask));
            base = __shfl_sync(FULL_MASK, base, __ffs(mask) - 1);
            if (c == k) perm[(k == 0 ? 0 : (k == 1 ? b1 : (k == 2 ? b2 : b3))) + base + __popc(mask & lt_mask)] = (int)a;
        }
    }
}
void dg_launch_class_scatter(cudaStream_t st, long long N, const int* astruct, const DgSmall* small, const int* deg_raw,
                             int cap, DgStats* stats, int* perm, int sm_count) {
    if (N <= 0) return;
    long long g = (N + 255) / 256;
    if (g > (long long)sm_count * 8) g = (long long)sm_count * 8;
    k_class_scatter<<<(unsigned)g, 256, 0, st>>>(N, astruct, small, deg_raw, cap, stats, perm);
    dg_direct_launch_bump(1);
}

size_t dg_direct_smem_bytes(const DgDirectArgs& a, bool fill) {
    return fill ? (size_t)a.n_warps * dg_rowbuf_bytes(a.row_cap, a.n_buckets) : 0;
}
template <bool FILL, int EPS, bool HITS, bool LEAN, bool DYN>
static cudaError_t launch_direct_t(cudaStream_t st, const DgDirectArgs& a, int sm_count, size_t smem, int threads) {
    cudaError_t e = dg_raise_dyn_smem<k_direct<FILL, 2, EPS, HITS, LEAN, DYN>>();
    if (e != cudaSuccess) return e;
    int occ = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_direct<FILL, 2, EPS, HITS, LEAN, DYN>, threads, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
    // persistent grid: every SM full; static schedule = chunks of DG_ROW_CHUNK consecutive rows round-robin over CTAs
    const long long rows = (DYN && a.row_perm) ? a.n_perm : (a.row_end > 0 ? a.row_end : a.n_atoms) - a.row_begin;
    const long long want = DYN ? (rows + (long long)a.dyn_unit * (threads / 32) - 1) / ((long long)a.dyn_unit * (threads / 32))
                               : (rows + DG_ROW_CHUNK - 1) / DG_ROW_CHUNK;
    long long grid = (long long)sm_count * occ;
    if (grid > want) grid = want;
    if (grid < 1) return cudaSuccess;
    k_direct<FILL, 2, EPS, HITS, LEAN, DYN><<<(unsigned)grid, threads, smem, st>>>(a);
    return cudaGetLastError();
}

template <bool FILL, bool HITS, bool LEAN>
static cudaError_t launch_direct_e(cudaStream_t st, const DgDirectArgs& a, int sm_count, size_t smem, int threads) {
    const bool dyn = a.dyn_unit > 0 && a.unit_ctr != nullptr;
#define DG_LD(E)                                                                              \
    return dyn ? launch_direct_t<FILL, E, HITS, LEAN, true>(st, a, sm_count, smem, threads)   \
               : launch_direct_t<FILL, E, HITS, LEAN, false>(st, a, sm_count, smem, threads);
    switch (a.eps_mode) {
        case 0: DG_LD(0)
        case 1: DG_LD(1)
        default: DG_LD(2)
    }
#undef DG_LD
}

// (a.dyn_unit > 0 with a zeroed a.unit_ctr: dynamic per-warp schedule, see k_direct)
cudaError_t dg_launch_direct(cudaStream_t st, const DgDirectArgs& a, bool fill, int sm_count, int min_blocks) {
    (void)min_blocks;  // (3 CTAs per SM at 80 registers measured slower, round 1: the variant is no longer compiled)
    if (a.n_atoms <= 0) return cudaSuccess;
    const size_t smem = dg_direct_smem_bytes(a, fill);
    const int threads = a.n_warps * 32;
    dg_direct_launch_bump(1);
    if (fill && a.max_neighbors >= a.row_cap && !a.full_tie_sort)  // no row is truncated, keys distinct: the lean sort
        return launch_direct_e<true, false, true>(st, a, sm_count, smem, threads);
    if (fill) return launch_direct_e<true, false, false>(st, a, sm_count, smem, threads);
    if (a.hit_cap > 0 && a.hits) return launch_direct_e<false, true, false>(st, a, sm_count, smem, threads);
    return launch_direct_e<false, false, false>(st, a, sm_count, smem, threads);
}

