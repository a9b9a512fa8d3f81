// dgnn_direct_count.cu -- symmetric count pass of the direct path (one sweep decides i -> j and j -> i).
// Compile with -fmad=false.
#include <cuda_runtime.h>
#include <stdint.h>

#include "dgnn_internal.h"
#include "row_sort.cuh"
#include "direct_sweep.cuh"

// ---------------------------------------------------------------------------------------------------------------------
// Symmetric count pass.  The edge i -> (j, n) and its mirror j -> (i, -n) have the same length in exact arithmetic; in the
// reference's rounded arithmetic (each direction computed on its own, neighbor_list.cpp:82-87 + nanoflann's metric) their
// d2 differ by a few ulp at most.  So a row only sweeps the candidates j >= i: a pair whose d2 is clear of the cutoff AND
// of the near-cutoff band by A.sym_band (>> the rounding difference, see dgnn_api.cu) decides both directions at once,
// the rare pair inside the band recomputes the mirror's d2 with the mirror's own operands
//   query = pos_j, cloud point = pos_i + ((-n_a)*row0 + (-n_b)*row1) + (-n_c)*row2 = pos_i - offset(n)   (negation is exact)
// so degrees and the near-cutoff statistic are bit-identical to the two-sided sweep at roughly half the work.  Degrees
// are accumulated with integer atomics into deg_raw (zeroed before the launch).  The atom's own images (j == i) and
// IMGPAR structures (tiny cells, lanes over images) take the two-sided sweep.  No hit lists in this mode.
// ---------------------------------------------------------------------------------------------------------------------
template <int EPS>
__device__ __forceinline__ void sweep_row_sym(const DgDirectArgs& A, const DgSmall* __restrict__ G, long long a,
                                              int lane, unsigned& near) {
    const long long base = G->atom_base;
    const int N = G->natoms, nimg = G->nimg;
    const double sx = A.x[a], sy = A.y[a], sz = A.z[a];
    const int iloc = (int)(a - base);
    const double fi0 = A.fa[a], fi1 = A.fb[a], fi2 = A.fc[a];
    const double w0 = G->w[0], w1 = G->w[1], w2 = G->w[2];
    const double L00 = G->L[0], L01 = G->L[1], L02 = G->L[2];
    const double L10 = G->L[3], L11 = G->L[4], L12 = G->L[5];
    const double L20 = G->L[6], L21 = G->L[7], L22 = G->L[8];
    // the two directions' d2 differ by rounding only: a few ulp of (|operands| * r_c), operands bounded by the source
    // position, the cutoff and the largest image offset of the reference cube -- 1e-14 * mag^2 is > 40 ulp of that
    const double omax = (double)nimg * (fabs(L00) + fabs(L01) + fabs(L02) + fabs(L10) + fabs(L11) + fabs(L12) +
                                         fabs(L20) + fabs(L21) + fabs(L22));
    const double mag = fabs(sx) + fabs(sy) + fabs(sz) + A.r_cutoff + omax;
    const double r2 = A.r2, band = A.sym_band + 1.0e-14 * mag * mag;
    int own = 0;
    long long gj = base + (iloc + lane < N ? iloc + lane : N - 1);
    double nx = A.x[gj], ny = A.y[gj], nz = A.z[gj];
    double nfa = A.fa[gj], nfb = A.fb[gj], nfc = A.fc[gj];
    for (int jb = iloc; jb < N; jb += 32) {
        const int j = jb + lane;
        const bool valid = j < N;
        const int jj = valid ? j : N - 1;
        const double xj = nx, yj = ny, zj = nz;
        const bool self = valid && jj == iloc;  // lane 0 of the first step
        const double t0 = fi0 - nfa, t1 = fi1 - nfb, t2 = fi2 - nfc;
        if (jb + 32 < N) {
            const int jn = j + 32;
            gj = base + (jn < N ? jn : N - 1);
            nx = A.x[gj]; ny = A.y[gj]; nz = A.z[gj];
            nfa = A.fa[gj]; nfb = A.fb[gj]; nfc = A.fc[gj];
        }
        int lo0, lo1, lo2;
        int c0 = axis_window_fast(t0, w0, nimg, &lo0);
        const int c1 = axis_window_fast(t1, w1, nimg, &lo1);
        const int c2 = axis_window_fast(t2, w2, nimg, &lo2);
        if (!valid) c0 = 0;
        const int mc0 = __reduce_max_sync(FULL_MASK, c0);
        const int mc1 = __reduce_max_sync(FULL_MASK, c0 ? c1 : 0);
        const int mc2 = __reduce_max_sync(FULL_MASK, c0 ? c2 : 0);
        int theirs = 0;
        // (innermost body: this direction's test and the mirror's)
#define DG_SYM_BODY                                                                                                  \
                    const bool act = act_ab && sc < c2;                                                              \
                    const double ox = DG_ADD(ob0, DG_MUL(ncd, L20)); /* offset(n), neighbor_list.cpp:82-84 */        \
                    const double oy = DG_ADD(ob1, DG_MUL(ncd, L21));                                                 \
                    const double oz = DG_ADD(ob2, DG_MUL(ncd, L22));                                                 \
                    const double dx = DG_SUB(sx, DG_ADD(xj, ox));                                                    \
                    const double dy = DG_SUB(sy, DG_ADD(yj, oy));                                                    \
                    const double dz = DG_SUB(sz, DG_ADD(zj, oz));                                                    \
                    const double d2 = DG_ADD(DG_ADD(DG_MUL(dx, dx), DG_MUL(dy, dy)), DG_MUL(dz, dz));                \
                    bool hit = act && d2 < r2;                                                                       \
                    if (EPS == 1) hit = hit && !(self && d2 == 0.0);                                                 \
                    bool hit_m = hit && !self; /* the mirror of a self image is another image of this very row */    \
                    if (act && fabs(d2 - r2) <= band) {                                                              \
                        near += (d2 >= A.near_d2_lo && d2 <= A.near_d2_hi) ? 1u : 0u;                                \
                        if (!self) {                                                                                 \
                            const double ex = DG_SUB(xj, DG_SUB(sx, ox)); /* pos_i + (-offset) == pos_i - offset */  \
                            const double ey = DG_SUB(yj, DG_SUB(sy, oy));                                            \
                            const double ez = DG_SUB(zj, DG_SUB(sz, oz));                                            \
                            const double e2 = DG_ADD(DG_ADD(DG_MUL(ex, ex), DG_MUL(ey, ey)), DG_MUL(ez, ez));        \
                            hit_m = e2 < r2;                                                                         \
                            near += (e2 >= A.near_d2_lo && e2 <= A.near_d2_hi) ? 1u : 0u;                            \
                        }                                                                                            \
                    }                                                                                                \
                    own += hit ? 1 : 0;                                                                              \
                    theirs += hit_m ? 1 : 0;
        if (mc0 == 1 && mc1 == 1) {
            // one image along a and b for every lane of the chunk (cells at least 2 r_c wide along those axes): only the
            // loop over n_c remains
            const double nad = (double)lo0, nbd = (double)lo1;
            const double ob0 = DG_ADD(DG_MUL(nad, L00), DG_MUL(nbd, L10));
            const double ob1 = DG_ADD(DG_MUL(nad, L01), DG_MUL(nbd, L11));
            const double ob2 = DG_ADD(DG_MUL(nad, L02), DG_MUL(nbd, L12));
            const bool act_ab = c0 > 0 && c1 > 0;
            double ncd = (double)lo2;
            for (int sc = 0; sc < mc2; ++sc, ncd += 1.0) {
                DG_SYM_BODY
            }
        } else {
        double nad = (double)lo0;
        for (int sa = 0; sa < mc0; ++sa, nad += 1.0) {
            const double oa0 = DG_MUL(nad, L00), oa1 = DG_MUL(nad, L01), oa2 = DG_MUL(nad, L02);
            double nbd = (double)lo1;
            for (int sb = 0; sb < mc1; ++sb, nbd += 1.0) {
                const double ob0 = DG_ADD(oa0, DG_MUL(nbd, L10));
                const double ob1 = DG_ADD(oa1, DG_MUL(nbd, L11));
                const double ob2 = DG_ADD(oa2, DG_MUL(nbd, L12));
                const bool act_ab = sa < c0 && sb < c1;
                double ncd = (double)lo2;
                for (int sc = 0; sc < mc2; ++sc, ncd += 1.0) {
                    DG_SYM_BODY
                }
            }
        }
        }
#undef DG_SYM_BODY
        if (theirs) atomicAdd(A.deg_raw + (base + jj), theirs);
    }
    own = __reduce_add_sync(FULL_MASK, own);
    if (lane == 0 && own) atomicAdd(A.deg_raw + a, own);
}

// IMGPAR structures (tiny cells, lanes over the images of one pair) take the generic two-sided sweep: out of line, so
// that its fp64 state does not count against the registers of the fp32 loop (A is a __grid_constant__ parameter: its
// address can be taken without a local copy)
template <int EPS>
__device__ __noinline__ void count_row_imgpar(const DgDirectArgs& A, const DgSmall* __restrict__ G, long long a, int mode,
                                              int lane, unsigned* near_io) {
    const RowBuf rb = RowBuf{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    unsigned near = 0;
    int m = traverse_row<false, EPS, false, false, true>(A, G, a, mode, rb, nullptr, nullptr, lane, (1u << lane) - 1u, near);
    m = __reduce_add_sync(FULL_MASK, m);
    if (lane == 0 && m) atomicAdd(A.deg_raw + a, m);
    *near_io += near;
}

template <int MINB, int EPS>
__global__ void __launch_bounds__(256, MINB) k_direct_count_sym(const __grid_constant__ DgDirectArgs A) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    unsigned near = 0;
    const long long r0 = A.row_begin, r1 = A.row_end > 0 ? A.row_end : A.n_atoms;
    const long long n_chunks = (r1 - r0 + DG_ROW_CHUNK - 1) / DG_ROW_CHUNK;
    for (long long chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
        const long long a_end = r0 + (chunk + 1) * DG_ROW_CHUNK < r1 ? r0 + (chunk + 1) * DG_ROW_CHUNK : r1;
        for (long long a = r0 + chunk * DG_ROW_CHUNK + warp; a < a_end; a += nwarps) {
            const DgSmall* __restrict__ G = A.small + A.astruct[a];
            const int mode = G->mode;
            if (mode == DG_MODE_CELL || G->status != DG_ST_OK) continue;
            if (mode == DG_MODE_ATOMPAR) {
                sweep_row_sym<EPS>(A, G, a, lane, near);
            } else {
                count_row_imgpar<EPS>(A, G, a, mode, lane, &near);
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) near += __shfl_xor_sync(FULL_MASK, near, o);
    if (lane == 0 && near) atomicAdd(&A.stats->n_near, (unsigned long long)near);
}


// deg_capped = min(deg_raw, max_neighbors) and the largest row, after the symmetric count pass (direct-path rows only)
__global__ void __launch_bounds__(256) k_cap_degrees(const DgDirectArgs A) {
    const long long r0 = A.row_begin, r1 = A.row_end > 0 ? A.row_end : A.n_atoms;
    int mx = 0;
    // (split_cap > 0: rows beyond the full-occupancy row buffer and the others by sort-network class, see dgnn_direct.cu)
    int nb = 0, c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    const int cap = A.split_cap;
    for (long long a = r0 + (long long)blockIdx.x * blockDim.x + threadIdx.x; a < r1; a += (long long)gridDim.x * blockDim.x) {
        const DgSmall* __restrict__ G = A.small + A.astruct[a];
        if (G->mode == DG_MODE_CELL || G->status != DG_ST_OK) continue;
        const int m = A.deg_raw[a];
        A.deg_capped[a] = m < A.max_neighbors ? m : A.max_neighbors;
        mx = m > mx ? m : mx;
        if (cap > 0) {
            const bool small_row = m <= cap;  // (classes as dg_row_class, dgnn_direct.cu)
            nb += !small_row;
            c0 += small_row && m <= 64;
            c1 += small_row && m > 64 && m <= 128;
            c2 += small_row && m > 128 && m <= 256;
            c3 += small_row && m > 256;
        }
    }
    mx = __reduce_max_sync(FULL_MASK, mx);
    if ((threadIdx.x & 31) == 0 && mx) atomicMax(&A.stats->max_deg, mx);
    if (cap > 0) {
        nb = __reduce_add_sync(FULL_MASK, nb);
        c0 = __reduce_add_sync(FULL_MASK, c0);
        c1 = __reduce_add_sync(FULL_MASK, c1);
        c2 = __reduce_add_sync(FULL_MASK, c2);
        c3 = __reduce_add_sync(FULL_MASK, c3);
        if ((threadIdx.x & 31) == 0) {
            if (nb) atomicAdd(&A.stats->n_big_rows, nb);
            if (c0) atomicAdd(&A.stats->cls_n[0], c0);
            if (c1) atomicAdd(&A.stats->cls_n[1], c1);
            if (c2) atomicAdd(&A.stats->cls_n[2], c2);
            if (c3) atomicAdd(&A.stats->cls_n[3], c3);
        }
    }
}

template <int MINB>
static cudaError_t launch_count_sym_m(cudaStream_t st, const DgDirectArgs& a, int sm_count) {
    const long long rows = (a.row_end > 0 ? a.row_end : a.n_atoms) - a.row_begin;
    if (rows <= 0) return cudaSuccess;
    const long long want = (rows + DG_ROW_CHUNK - 1) / DG_ROW_CHUNK;
    int occ = 1;
    cudaError_t e;
    long long grid;
#define DG_LAUNCH_SYM(E)                                                                                   \
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_direct_count_sym<MINB, E>, 256, 0);           \
    if (e != cudaSuccess) return e;                                                                         \
    grid = (long long)sm_count * (occ < 1 ? 1 : occ);                                                       \
    if (grid > want) grid = want;                                                                           \
    k_direct_count_sym<MINB, E><<<(unsigned)grid, 256, 0, st>>>(a);
    if (a.eps_mode == 0) { DG_LAUNCH_SYM(0) } else { DG_LAUNCH_SYM(1) }
#undef DG_LAUNCH_SYM
    dg_direct_launch_bump(1);
    return cudaGetLastError();
}
// (once per build, after the count pass of every upload chunk: the caller resets the row range)
void dg_launch_cap_degrees(cudaStream_t st, const DgDirectArgs& a, int sm_count) {
    const long long rows = (a.row_end > 0 ? a.row_end : a.n_atoms) - a.row_begin;
    if (rows <= 0) return;
    long long g2 = (rows + 255) / 256;
    if (g2 > (long long)sm_count * 8) g2 = (long long)sm_count * 8;
    k_cap_degrees<<<(unsigned)g2, 256, 0, st>>>(a);
    dg_direct_launch_bump(1);
}
// symmetric count pass of rows [row_begin, row_end): deg_raw must be zero there; eps_mode 0 or 1 only.  Leaves raw degrees;
// dg_launch_cap_degrees turns them into deg_capped / max_deg (and the row-class tallies) afterwards.
cudaError_t dg_launch_direct_count_sym(cudaStream_t st, const DgDirectArgs& a, int sm_count, int min_blocks) {
    if (a.n_atoms <= 0) return cudaSuccess;
    if (a.eps_mode > 1) return cudaErrorInvalidValue;
    (void)min_blocks;  // (3 CTAs per SM at 80 registers spills: not compiled)
    return launch_count_sym_m<2>(st, a, sm_count);
}

