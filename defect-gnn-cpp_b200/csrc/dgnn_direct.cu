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

static int g_direct_launches = 0;
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
    ++g_direct_launches;
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
    ++g_direct_launches;
}

// ---------------------------------------------------------------------------------------------------------------------
// traversal
// ---------------------------------------------------------------------------------------------------------------------
#define DG_ROW_CHUNK 64  // consecutive source atoms owned by one CTA at a time (keeps the structure's lines in L1)

// floor / ceil of |x| < 2^31 as int32 through the 2^52 + 2^51 trick: one FP64-pipe add instead of FRND + F2I on the
// quarter-rate conversion unit.  The prep kernel rejects fractional coordinates beyond +-1e6, so |x| is bounded.
#define DG_MAGIC 6755399441055744.0
__device__ __forceinline__ int dg_floor_i(double x) { return __double2loint(__dadd_rd(x, DG_MAGIC)); }
__device__ __forceinline__ int dg_ceil_i(double x) { return __double2loint(__dadd_ru(x, DG_MAGIC)); }

// images n of a candidate along one axis: integers in [t - w, t + w] clipped to [-nimg, nimg]  (= dg_axis_window)
__device__ __forceinline__ int axis_window_fast(double t, double w, int nimg, int* lo) {
    int a = dg_ceil_i(t - w), b = dg_floor_i(t + w);
    a = a < -nimg ? -nimg : a;
    b = b > nimg ? nimg : b;
    *lo = a;
    const int c = b - a + 1;
    return c > 0 ? c : 0;
}

struct Src {
    double x, y, z;
    int iloc;
};

// One candidate image point (px,py,pz) of atom jloc per lane; compacts hits into the row buffer.
// EPS selects the self-exclusion rule `orig == i && sqrt(d2) < epsilon` (neighbor_list.cpp:47-49):
//   0: epsilon <= 0, nothing is excluded;  1: 0 < epsilon <= DG_EPS_FAST: only the atom's own n = (0,0,0) image can
//   qualify (its d2 is exactly 0; k_small_geom rejects cells whose perpendicular height is below 2 epsilon), so the
//   test is the branch-free `self && d2 == 0`;  2: any epsilon, exact sqrt test behind a (rare) branch.
__device__ __forceinline__ uint32_t dg_pack_hit(int j, int na, int nb, int nc) {
    return (uint32_t)j | ((uint32_t)(na + 16) << 17) | ((uint32_t)(nb + 16) << 22) | ((uint32_t)(nc + 16) << 27);
}

// HITS (count pass only): in-cutoff candidates are also compacted (ballot/popc) into the row's hit list in global
// memory, so that the fill pass recomputes only the ~deg accepted pairs instead of repeating the traversal.
template <bool FILL, int EPS, bool HITS, bool ONE>
__device__ __forceinline__ void test_point(const DgDirectArgs& A, const Src& S, const RowBuf& rb,
                                           uint32_t* __restrict__ hrow, double* __restrict__ hd2, bool active,
                                           bool self, int jloc, double px, double py, double pz, int na, int nb, int nc,
                                           unsigned lt_mask, int& m, unsigned& near) {
    const double dx = DG_SUB(S.x, px);
    const double dy = DG_SUB(S.y, py);
    const double dz = DG_SUB(S.z, pz);
    // nanoflann evalMetric: result = 0; result += dx*dx; += dy*dy; += dz*dz
    const double d2 = DG_ADD(DG_ADD(DG_MUL(dx, dx), DG_MUL(dy, dy)), DG_MUL(dz, dz));
    bool hit = active && (d2 < A.r2);
    if (EPS == 1) hit = hit && !(self && d2 == 0.0);
    if (EPS == 2) {
        if (hit && self) {
            if (__dsqrt_rn(d2) < A.epsilon) hit = false;
        }
    }
    // near-cutoff band |sqrt(d2) - r_c| <= 1e-6 as two thresholds on d2 (sqrt is monotone; thresholds bisected on the
    // host); tallied by whichever pass sees every candidate once (count pass, or the single-pass build)
    if (!FILL || ONE) near += (active && d2 >= A.near_d2_lo && d2 <= A.near_d2_hi) ? 1u : 0u;
    if (!FILL) {
        // count pass: per-lane tallies, reduced once per row (no ballot per candidate)
        if (HITS) {
            const unsigned mask = __ballot_sync(FULL_MASK, hit);
            if (hit) {
                const int pos = m + __popc(mask & lt_mask);
                if (pos < A.hit_cap) {
                    hrow[pos] = dg_pack_hit(jloc, na, nb, nc);
                    hd2[pos] = d2;
                }
            }
            m += __popc(mask);  // warp-uniform in this mode
        } else {
            m += hit ? 1 : 0;
        }
        return;
    }
    const unsigned mask = __ballot_sync(FULL_MASK, hit);
    if (hit) {
        const int pos = m + __popc(mask & lt_mask);
        if (pos < A.row_cap) {
            rb.rd[pos] = d2;
            rb.rj[pos] = jloc;
            rb.rn[pos] = dg_pack_image(na, nb, nc);
        }
    }
    m += __popc(mask);
}

// Candidate sweep of one source atom a of structure G (mode: DG_MODE_ATOMPAR / DG_MODE_IMGPAR); returns the number of
// accepted pairs (per lane when !FILL && !HITS, warp-uniform otherwise).  All 32 lanes must call.
template <bool FILL, int EPS, bool HITS, bool ONE>
__device__ __forceinline__ int traverse_row(const DgDirectArgs& A, const DgSmall* __restrict__ G, long long a, int mode,
                                            const RowBuf& rb, uint32_t* __restrict__ hrow, double* __restrict__ hd2,
                                            int lane, unsigned lt_mask, unsigned& near) {
    const long long base = G->atom_base;
    const int N = G->natoms, nimg = G->nimg;
    Src S;
    S.x = A.x[a]; S.y = A.y[a]; S.z = A.z[a];
    S.iloc = (int)(a - base);
    const double fi0 = A.fa[a], fi1 = A.fb[a], fi2 = A.fc[a];
    const double w0 = G->w[0], w1 = G->w[1], w2 = G->w[2];
    const double L20 = G->L[6], L21 = G->L[7], L22 = G->L[8];
    int m = 0;
    if (mode == DG_MODE_ATOMPAR) {
        const double L00 = G->L[0], L01 = G->L[1], L02 = G->L[2];
        const double L10 = G->L[3], L11 = G->L[4], L12 = G->L[5];
        // lanes over candidate atoms (next chunk's loads in flight while this one is tested); nested loops
        // over the warp-wide image window, per-lane windows masked
        long long gj = base + (lane < N ? lane : N - 1);
        double nx = A.x[gj], ny = A.y[gj], nz = A.z[gj];
        double nfa = A.fa[gj], nfb = A.fb[gj], nfc = A.fc[gj];
        for (int jb = 0; jb < N; jb += 32) {
            const int j = jb + lane;
            const bool valid = j < N;
            const int jj = valid ? j : N - 1;
            const double xj = nx, yj = ny, zj = nz;
            const bool self = jj == S.iloc;
            const double t0 = fi0 - nfa, t1 = fi1 - nfb, t2 = fi2 - nfc;
            if (jb + 32 < N) {
                const int jn = j + 32;
                gj = base + (jn < N ? jn : N - 1);
                nx = A.x[gj]; ny = A.y[gj]; nz = A.z[gj];
                nfa = A.fa[gj]; nfb = A.fb[gj]; nfc = A.fc[gj];
            }
            int lo0, lo1, lo2;
            int c0 = axis_window_fast(t0, w0, nimg, &lo0);
            int c1 = axis_window_fast(t1, w1, nimg, &lo1);
            int c2 = axis_window_fast(t2, w2, nimg, &lo2);
            if (!valid) c0 = 0;
            const int mc0 = __reduce_max_sync(FULL_MASK, c0);
            const int mc1 = __reduce_max_sync(FULL_MASK, c0 ? c1 : 0);
            const int mc2 = __reduce_max_sync(FULL_MASK, c0 ? c2 : 0);
            double nad = (double)lo0;
            for (int sa = 0; sa < mc0; ++sa, nad += 1.0) {
                const double oa0 = DG_MUL(nad, L00), oa1 = DG_MUL(nad, L01), oa2 = DG_MUL(nad, L02);
                double nbd = (double)lo1;
                for (int sb = 0; sb < mc1; ++sb, nbd += 1.0) {
                    // offset = (n_a*row0 + n_b*row1) + n_c*row2, neighbor_list.cpp:82-84
                    const double ob0 = DG_ADD(oa0, DG_MUL(nbd, L10));
                    const double ob1 = DG_ADD(oa1, DG_MUL(nbd, L11));
                    const double ob2 = DG_ADD(oa2, DG_MUL(nbd, L12));
                    const bool act_ab = sa < c0 && sb < c1;
                    double ncd = (double)lo2;
                    for (int sc = 0; sc < mc2; ++sc, ncd += 1.0) {
                        const double px = DG_ADD(xj, DG_ADD(ob0, DG_MUL(ncd, L20)));
                        const double py = DG_ADD(yj, DG_ADD(ob1, DG_MUL(ncd, L21)));
                        const double pz = DG_ADD(zj, DG_ADD(ob2, DG_MUL(ncd, L22)));
                        test_point<FILL, EPS, HITS, ONE>(A, S, rb, hrow, hd2, act_ab && sc < c2, self, jj, px, py, pz,
                                                    lo0 + sa, lo1 + sb, lo2 + sc, lt_mask, m, near);
                    }
                }
            }
        }
    } else if (mode == DG_MODE_IMGPAR) {
        // lanes over the image window of one candidate atom at a time
        for (int j = 0; j < N; ++j) {
            const long long gj = base + j;
            const double xj = A.x[gj], yj = A.y[gj], zj = A.z[gj];
            int lo0, lo1, lo2;
            const int c0 = axis_window_fast(fi0 - A.fa[gj], w0, nimg, &lo0);
            const int c1 = axis_window_fast(fi1 - A.fb[gj], w1, nimg, &lo1);
            const int c2 = axis_window_fast(fi2 - A.fc[gj], w2, nimg, &lo2);
            const int T = c0 * c1 * c2;
            for (int tb = 0; tb < T; tb += 32) {
                const int t = tb + lane;
                const bool active = t < T;
                const int tt = active ? t : 0;
                const int ic = tt % c2, tq = tt / c2;
                const int na = lo0 + tq / c1, nb = lo1 + tq % c1, nc = lo2 + ic;
                const double px = DG_ADD(xj, dg_image_offset(G->L, 0, na, nb, nc));
                const double py = DG_ADD(yj, dg_image_offset(G->L, 1, na, nb, nc));
                const double pz = DG_ADD(zj, dg_image_offset(G->L, 2, na, nb, nc));
                test_point<FILL, EPS, HITS, ONE>(A, S, rb, hrow, hd2, active, j == S.iloc, j, px, py, pz, na, nb, nc,
                                            lt_mask, m, near);
            }
        }
    }
    return m;
}

// ---------------------------------------------------------------------------------------------------------------------
// Single-pass build: row offsets by decoupled look-back.  Every row owns one 64-bit status word, flag << 62 | value:
// AGG = the row's own (capped) degree, published right after its traversal; INC = the inclusive prefix of all rows up
// to and including it, published once the look-back has resolved.  A row walks back 32 predecessors per step summing
// AGG values until it meets an INC.  Flag and value share the word, so relaxed gpu-scope accesses are enough.  Chunks
// are handed out statically to a grid sized by the occupancy query (every CTA resident), so each predecessor row is
// either finished or being worked on: the spin always ends; a bounded spin count guards the box regardless.
// ---------------------------------------------------------------------------------------------------------------------
#define DG_LB_AGG (1ull << 62)
#define DG_LB_INC (2ull << 62)
#define DG_LB_VAL ((1ull << 62) - 1ull)
#define DG_LB_SPIN_LIMIT (1u << 22)
__device__ __forceinline__ unsigned long long dg_lb_load(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void dg_lb_store(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// exclusive prefix (sum of the values of entries 0 .. a-1); all 32 lanes must call.  A round trip covers DG_LB_WIDE * 32
// predecessors (the loads of a round are independent): with a few hundred chunks in flight the walk back to the last
// resolved prefix takes one to three L2 latencies.
#define DG_LB_WIDE 2
__device__ __forceinline__ long long dg_lookback(const unsigned long long* __restrict__ status, long long a, int lane,
                                                 DgStats* __restrict__ stats) {
    long long excl = 0;
    long long pos = a - 1;
    unsigned spins = 0;
    while (pos >= 0) {
        unsigned long long w[DG_LB_WIDE];
#pragma unroll
        for (int k = 0; k < DG_LB_WIDE; ++k) {
            const long long idx = pos - 32 * k - lane;
            w[k] = idx >= 0 ? dg_lb_load(status + idx) : DG_LB_INC;  // entry -1: inclusive prefix 0
        }
        long long sum = 0;
        bool done = false, retry = false;
#pragma unroll
        for (int k = 0; k < DG_LB_WIDE; ++k) {
            if (done || retry) continue;  // warp-uniform
            const unsigned flag = (unsigned)(w[k] >> 62);
            const unsigned inc = __ballot_sync(FULL_MASK, flag == 2u);
            const unsigned emp = __ballot_sync(FULL_MASK, flag == 0u);
            const int first = inc ? __ffs(inc) - 1 : 32;                   // nearest predecessor with a resolved prefix
            const unsigned need = first >= 31 ? 0xffffffffu : ((2u << first) - 1u);
            if (emp & need) {  // a predecessor in the window has not published its aggregate yet
                retry = true;
                continue;
            }
            // AGG values are sums of a few row degrees (< 2^20 each): 32 of them fit an unsigned sum
            sum += (long long)__reduce_add_sync(FULL_MASK, lane < first ? (unsigned)(w[k] & 0xffffffffull) : 0u);
            if (first < 32) {
                const unsigned long long v = w[k] & DG_LB_VAL;
                const unsigned lo = __shfl_sync(FULL_MASK, (unsigned)v, first);
                const unsigned hi = __shfl_sync(FULL_MASK, (unsigned)(v >> 32), first);
                sum += (long long)(((unsigned long long)hi << 32) | lo);
                done = true;
            }
        }
        if (retry) {
            if (++spins > DG_LB_SPIN_LIMIT) {
                if (lane == 0) atomicExch(&stats->lb_timeout, 1);
                break;
            }
            __nanosleep(20);
            continue;
        }
        excl += sum;
        if (done) break;
        pos -= 32 * DG_LB_WIDE;
    }
    return excl;
}

// LEAN (fill pass of batches no row of which is truncated, max_neighbors >= row_cap): the lean row sort (row_sort.cuh)
template <bool FILL, int MINB, int EPS, bool HITS, bool LEAN = false>
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

    for (long long chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
        const long long a_end = r0 + (chunk + 1) * DG_ROW_CHUNK < r1 ? r0 + (chunk + 1) * DG_ROW_CHUNK : r1;
        for (long long a = r0 + chunk * DG_ROW_CHUNK + warp; a < a_end; a += nwarps) {
            const DgSmall* __restrict__ G = A.small + A.astruct[a];
            const int mode = G->mode;
            if (mode == DG_MODE_CELL || G->status != DG_ST_OK) continue;  // the cell-list path owns this structure
            const int nimg = G->nimg;
            // rows the hit-list fill pass (k_fill_hits) owns are skipped here; this pass takes the rest
            if (FILL && A.hit_cap > 0 && nimg <= DG_HIT_NMAX && A.deg_raw[a] <= A.hit_cap) continue;  // (row_cap >= deg)
            // ... and so are rows the single-pass build has already written
            if (FILL && A.skip_deg_le > 0 && A.deg_raw[a] <= A.skip_deg_le) continue;
            const bool store_hits = HITS && nimg <= DG_HIT_NMAX;
            uint32_t* __restrict__ hrow = HITS ? A.hits + (size_t)a * (size_t)A.hit_cap : nullptr;
            double* __restrict__ hd2 = HITS ? A.hits_d2 + (size_t)a * (size_t)A.hit_cap : nullptr;
            int m = traverse_row<FILL, EPS, HITS, false>(A, G, a, mode, rb, hrow, hd2, lane, lt_mask, near);
            if (!FILL && !HITS) m = __reduce_add_sync(FULL_MASK, m);
            maxdeg = m > maxdeg ? m : maxdeg;
            const int keep = m < A.max_neighbors ? m : A.max_neighbors;
            if (!FILL) {
                if (lane == 0) {
                    A.deg_capped[a] = keep;
                    A.deg_raw[a] = m;
                    if (HITS && (m > A.hit_cap || !store_hits)) atomicAdd(&A.stats->n_fallback_rows, 1);
                }
            } else if (m > A.row_cap) {  // cannot happen: row_cap >= max degree of the count pass
                if (lane == 0) atomicExch(&A.stats->overflow, 1);
            } else {
                dg_row_sort_emit<LEAN>(rb, m, keep, A.n_buckets, A.inv_rc, A.key_scale, A.row_ptr[a], A.dst, A.image, A.dist,
                                       A.stats, lane, A.rbf);
            }
        }
    }
    if (!FILL) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) near += __shfl_xor_sync(FULL_MASK, near, o);
        if (lane == 0) {
            if (near) atomicAdd(&A.stats->n_near, (unsigned long long)near);
            if (maxdeg) atomicMax(&A.stats->max_deg, maxdeg);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Single-pass build.  A step of a CTA = nwarps consecutive rows (one per warp), step t of CTA b is chunk t * grid + b:
// at any moment the resident CTAs work on one contiguous window of rows, so every predecessor of a chunk is finished
// or in flight.  Per step: every warp sweeps its row into its shared-memory row buffer and posts the (capped) degree;
// the LAST warp to finish its sweep adds the degrees up, publishes the chunk aggregate, resolves the chunk's base by
// look-back over the chunk status words and hands the per-row offsets to its siblings -- which meanwhile have been
// sorting their rows in registers and only wait (named barrier) when they are ready to emit.  No count pass, no scan
// kernel, the sweep runs once.
// ---------------------------------------------------------------------------------------------------------------------
#define DG_ONE_MAX_WARPS 8
struct DgOneShared {
    long long base[2][DG_ONE_MAX_WARPS];  // per-row CSR offsets of the step (by step parity)
    int keep[2][DG_ONE_MAX_WARPS];
    int arrived[2];
};

// offset provider of the row sort (row_sort.cuh): the warp that ran the look-back already knows its offset, the others
// wait for the step's barrier.  Rows that must not be emitted (buffer overflow, arena overflow) yield -1.
struct DgOneOffset {
    const DgOneShared* sh;
    int par, warp, nthreads;
    long long o;        // valid when ready
    bool ready;
    long long edge_cap;
    DgStats* stats;
    int lane;
    __device__ __forceinline__ long long wait() {
        if (!ready) {
            // barrier id alternates with the step parity: a warp can be at most one step ahead of its siblings, so the
            // arrivals of consecutive steps never mix
            if (par) asm volatile("bar.sync 2, %0;" ::"r"(nthreads) : "memory");
            else asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory");
            o = sh->base[par][warp];
            ready = true;
        }
        return o;
    }
    __device__ __forceinline__ long long get(int keep) {
        const long long v = wait();
        if (v + keep > edge_cap) {  // the estimated arena is too small: the host regrows it and reruns the fill
            if (lane == 0) atomicExch(&stats->cap_overflow, 1);
            return -1;
        }
        return v;
    }
};

// The rows of two consecutive steps are alive at once (see k_direct_one): the entry arrays the sweep fills (d2, atom,
// image) exist twice per warp, the sort scratch once.
__host__ __device__ inline size_t dg_one_extra_bytes(int row_cap) {
    return dg_align16((size_t)row_cap * 8) + 2 * dg_align16((size_t)row_cap * 4);
}
__host__ __device__ inline size_t dg_one_warp_bytes(int row_cap, int n_buckets) {
    return dg_rowbuf_bytes(row_cap, n_buckets) + dg_one_extra_bytes(row_cap);
}
__device__ __forceinline__ RowBuf dg_one_rowbuf(unsigned char* wp, int row_cap, int n_buckets, int par) {
    RowBuf r = dg_rowbuf_carve(wp, row_cap);
    if (par) {
        unsigned char* q = wp + dg_rowbuf_bytes(row_cap, n_buckets);
        r.rd = (double*)q; q += dg_align16((size_t)row_cap * 8);
        r.rj = (int*)q; q += dg_align16((size_t)row_cap * 4);
        r.rn = (uint32_t*)q;
    }
    return r;
}

// a swept row waiting for its offset
struct DgOnePending {
    long long a, o;
    int m, keep;
    bool ready;
};

__device__ __forceinline__ void dg_one_finish_row(const DgDirectArgs& A, const DgOneShared* sh, const RowBuf& rb,
                                                  const DgOnePending& P, int par, int warp, int lane) {
    DgOneOffset op{sh, par, warp, (int)blockDim.x, P.o, P.ready, A.edge_cap, A.stats, lane};
    const bool valid = P.a < A.n_atoms;
    if (!valid || P.m > A.row_cap) {
        // nothing to emit here: rows that outgrew the estimated row buffer only reserve their range (a fill pass
        // completes them afterwards)
        op.wait();
        if (valid && lane == 0) atomicAdd(&A.stats->n_fallback_rows, 1);
    } else {
        dg_row_sort_emit_op(rb, P.m, P.keep, A.n_buckets, A.inv_rc, A.key_scale, op, A.dst, A.image, A.dist, A.stats,
                            lane, A.rbf);
    }
    if (valid && lane == 0) {
        A.deg_raw[P.a] = P.m;
        A.row_ptr_out[P.a] = op.o;
        if (P.a == A.n_atoms - 1) A.row_ptr_out[P.a + 1] = op.o + P.keep;
    }
}

template <int MINB, int EPS>
__global__ void __launch_bounds__(256, MINB) k_direct_one(const DgDirectArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ DgOneShared sh;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    unsigned char* wp = smem_raw + (size_t)warp * dg_one_warp_bytes(A.row_cap, A.n_buckets);
    if (threadIdx.x < 2) sh.arrived[threadIdx.x] = 0;
    __syncthreads();
    unsigned near = 0;
    int maxdeg = 0;
    const long long n_chunks = (A.n_atoms + nwarps - 1) / nwarps;
    int par = 0;
    bool have_pending = false;
    DgOnePending P{0, 0, 0, 0, false};
    for (long long chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x, par ^= 1) {
        // ---- sweep the row of this step into the entry arrays of its parity ----
        const long long a = chunk * nwarps + warp;
        int m = 0;
        if (a < A.n_atoms) {
            const DgSmall* __restrict__ G = A.small + A.astruct[a];
            const int mode = G->mode;
            // (never launched on batches with cell-list or invalid structures; such a row would count as empty)
            if (mode != DG_MODE_CELL && G->status == DG_ST_OK) {
                const RowBuf rb = dg_one_rowbuf(wp, A.row_cap, A.n_buckets, par);
                m = traverse_row<true, EPS, false, true>(A, G, a, mode, rb, nullptr, nullptr, lane, lt_mask, near);
            }
        }
        maxdeg = m > maxdeg ? m : maxdeg;
        const int keep = m < A.max_neighbors ? m : A.max_neighbors;
        int ticket = 0;
        if (lane == 0) {
            sh.keep[par][warp] = keep;
            __threadfence_block();
            ticket = atomicAdd(&sh.arrived[par], 1);
        }
        ticket = __shfl_sync(FULL_MASK, ticket, 0);
        DgOnePending C{a, 0, m, keep, false};
        if (ticket == nwarps - 1) {
            // last sweep of the step to finish: chunk aggregate, look-back, per-row offsets for everybody
            __threadfence_block();
            const int k = lane < nwarps ? sh.keep[par][lane] : 0;
            int incl = k;
#pragma unroll
            for (int off = 1; off < DG_ONE_MAX_WARPS; off <<= 1) {
                const int t = __shfl_up_sync(FULL_MASK, incl, off);
                if (lane >= off) incl += t;
            }
            const int total = __shfl_sync(FULL_MASK, incl, nwarps - 1);
            if (lane == 0) dg_lb_store(A.lb_status + chunk, DG_LB_AGG | (unsigned long long)total);
            const long long excl = dg_lookback(A.lb_status, chunk, lane, A.stats);
            if (lane == 0) {
                dg_lb_store(A.lb_status + chunk, DG_LB_INC | (unsigned long long)(excl + total));
                sh.arrived[par] = 0;  // next used two steps from now
            }
            const long long mine = excl + (long long)(incl - k);
            if (lane < nwarps) sh.base[par][lane] = mine;
            if (par) asm volatile("bar.arrive 2, %0;" ::"r"((int)blockDim.x) : "memory");
            else asm volatile("bar.arrive 1, %0;" ::"r"((int)blockDim.x) : "memory");
            C.o = __shfl_sync(FULL_MASK, mine, warp);
            C.ready = true;
        }
        // ---- sort and emit the row of the previous step: its offsets have had a whole sweep to arrive ----
        if (have_pending) {
            const RowBuf rbp = dg_one_rowbuf(wp, A.row_cap, A.n_buckets, par ^ 1);
            dg_one_finish_row(A, &sh, rbp, P, par ^ 1, warp, lane);
        }
        P = C;
        have_pending = true;
    }
    if (have_pending) {
        const RowBuf rbp = dg_one_rowbuf(wp, A.row_cap, A.n_buckets, par ^ 1);
        dg_one_finish_row(A, &sh, rbp, P, par ^ 1, warp, lane);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) near += __shfl_xor_sync(FULL_MASK, near, o);
    if (lane == 0) {
        if (near) atomicAdd(&A.stats->n_near, (unsigned long long)near);
        if (maxdeg) atomicMax(&A.stats->max_deg, maxdeg);
    }
}

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
                    const bool act = act_ab && sc < c2;
                    const double ox = DG_ADD(ob0, DG_MUL(ncd, L20));  // offset(n), neighbor_list.cpp:82-84
                    const double oy = DG_ADD(ob1, DG_MUL(ncd, L21));
                    const double oz = DG_ADD(ob2, DG_MUL(ncd, L22));
                    const double dx = DG_SUB(sx, DG_ADD(xj, ox));
                    const double dy = DG_SUB(sy, DG_ADD(yj, oy));
                    const double dz = DG_SUB(sz, DG_ADD(zj, oz));
                    const double d2 = DG_ADD(DG_ADD(DG_MUL(dx, dx), DG_MUL(dy, dy)), DG_MUL(dz, dz));
                    bool hit = act && d2 < r2;
                    if (EPS == 1) hit = hit && !(self && d2 == 0.0);
                    bool hit_m = hit && !self;  // the mirror of a self image is another image of this very row
                    if (act && fabs(d2 - r2) <= band) {
                        near += (d2 >= A.near_d2_lo && d2 <= A.near_d2_hi) ? 1u : 0u;
                        if (!self) {
                            const double ex = DG_SUB(xj, DG_SUB(sx, ox));  // pos_i + (-offset) == pos_i - offset
                            const double ey = DG_SUB(yj, DG_SUB(sy, oy));
                            const double ez = DG_SUB(zj, DG_SUB(sz, oz));
                            const double e2 = DG_ADD(DG_ADD(DG_MUL(ex, ex), DG_MUL(ey, ey)), DG_MUL(ez, ez));
                            hit_m = e2 < r2;
                            near += (e2 >= A.near_d2_lo && e2 <= A.near_d2_hi) ? 1u : 0u;
                        }
                    }
                    own += hit ? 1 : 0;
                    theirs += hit_m ? 1 : 0;
                }
            }
        }
        if (theirs) atomicAdd(A.deg_raw + (base + jj), theirs);
    }
    own = __reduce_add_sync(FULL_MASK, own);
    if (lane == 0 && own) atomicAdd(A.deg_raw + a, own);
}

template <int MINB, int EPS>
__global__ void __launch_bounds__(256, MINB) k_direct_count_sym(const DgDirectArgs A) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    const RowBuf rb = RowBuf{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
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
                int m = traverse_row<false, EPS, false, false>(A, G, a, mode, rb, nullptr, nullptr, lane, lt_mask, near);
                m = __reduce_add_sync(FULL_MASK, m);
                if (lane == 0 && m) atomicAdd(A.deg_raw + a, m);
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) near += __shfl_xor_sync(FULL_MASK, near, o);
    if (lane == 0 && near) atomicAdd(&A.stats->n_near, (unsigned long long)near);
}

// ---------------------------------------------------------------------------------------------------------------------
// Pipelined plan: the batch is cut into structure-aligned parts; this kernel FILLS part k (rows [row_begin, row_end),
// offsets from the scan of the part's degrees) and runs the symmetric COUNT of part k+1 (rows [cnt_begin, cnt_end)) in
// the same warps, row by row.  The fill is bound by its stores, the count by instruction issue: interleaved they share
// the SM instead of running back to back.  Arena and row buffer come from the density estimate: rows that would not
// fit the arena raise cap_overflow (the host regrows and refills), rows longer than the row buffer are left to a
// regular fill pass (skip_deg_le).
// ---------------------------------------------------------------------------------------------------------------------
template <int EPS>
__global__ void __launch_bounds__(256, 2) k_direct_fill_cnt(const DgDirectArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    const RowBuf rb = dg_rowbuf_carve(smem_raw + (size_t)warp * dg_rowbuf_bytes(A.row_cap, A.n_buckets), A.row_cap);
    const RowBuf none = RowBuf{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    unsigned near = 0, dummy = 0;
    const long long r0 = A.row_begin, r1 = A.row_end, c0 = A.cnt_begin, c1 = A.cnt_end;
    const long long nrows = (r1 - r0) > (c1 - c0) ? (r1 - r0) : (c1 - c0);
    const long long n_chunks = (nrows + DG_ROW_CHUNK - 1) / DG_ROW_CHUNK;
    for (long long chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
        const long long k_end = (chunk + 1) * DG_ROW_CHUNK < nrows ? (chunk + 1) * DG_ROW_CHUNK : nrows;
        for (long long k = chunk * DG_ROW_CHUNK + warp; k < k_end; k += nwarps) {
            const long long a = r0 + k;
            if (a < r1) {  // ---- fill one row of part k ----
                const DgSmall* __restrict__ G = A.small + A.astruct[a];
                const int mode = G->mode;
                const int deg = A.deg_raw[a];
                if (mode != DG_MODE_CELL && G->status == DG_ST_OK && deg <= A.row_cap) {
                    const long long o = A.row_ptr[a];
                    const int keep = deg < A.max_neighbors ? deg : A.max_neighbors;
                    if (o + keep > A.edge_cap) {
                        if (lane == 0) atomicExch(&A.stats->cap_overflow, 1);
                    } else {
                        const int m = traverse_row<true, EPS, false, false>(A, G, a, mode, rb, nullptr, nullptr, lane,
                                                                            lt_mask, dummy);
                        if (m != deg) {  // cannot happen: both passes evaluate the same predicate
                            if (lane == 0) atomicExch(&A.stats->overflow, 1);
                        } else {
                            dg_row_sort_emit(rb, m, keep, A.n_buckets, A.inv_rc, A.key_scale, o, A.dst, A.image, A.dist,
                                             A.stats, lane, A.rbf);
                        }
                    }
                }
            }
            const long long a2 = c0 + k;
            if (a2 < c1) {  // ---- count one row of part k+1 ----
                const DgSmall* __restrict__ G = A.small + A.astruct[a2];
                const int mode = G->mode;
                if (mode != DG_MODE_CELL && G->status == DG_ST_OK) {
                    if (mode == DG_MODE_ATOMPAR) {
                        sweep_row_sym<EPS>(A, G, a2, lane, near);
                    } else {
                        int m = traverse_row<false, EPS, false, false>(A, G, a2, mode, none, nullptr, nullptr, lane,
                                                                       lt_mask, near);
                        m = __reduce_add_sync(FULL_MASK, m);
                        if (lane == 0 && m) atomicAdd(A.deg_raw + a2, m);
                    }
                }
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) near += __shfl_xor_sync(FULL_MASK, near, o);
    if (lane == 0 && near) atomicAdd(&A.stats->n_near, (unsigned long long)near);
}

template <int EPS>
static cudaError_t launch_fill_cnt_t(cudaStream_t st, const DgDirectArgs& a, int sm_count, size_t smem) {
    cudaError_t e = cudaFuncSetAttribute(k_direct_fill_cnt<EPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int occ = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_direct_fill_cnt<EPS>, a.n_warps * 32, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
    const long long fr = a.row_end - a.row_begin, cr = a.cnt_end - a.cnt_begin;
    const long long want = ((fr > cr ? fr : cr) + DG_ROW_CHUNK - 1) / DG_ROW_CHUNK;
    long long grid = (long long)sm_count * occ;
    if (grid > want) grid = want;
    if (grid < 1) return cudaSuccess;
    k_direct_fill_cnt<EPS><<<(unsigned)grid, a.n_warps * 32, smem, st>>>(a);
    return cudaGetLastError();
}
// fill rows [row_begin, row_end) + symmetric count of rows [cnt_begin, cnt_end) (may be empty); eps_mode 0 or 1
cudaError_t dg_launch_direct_fill_cnt(cudaStream_t st, const DgDirectArgs& a, int sm_count) {
    if (a.eps_mode > 1) return cudaErrorInvalidValue;
    const size_t smem = dg_direct_smem_bytes(a, true);
    ++g_direct_launches;
    return a.eps_mode == 0 ? launch_fill_cnt_t<0>(st, a, sm_count, smem) : launch_fill_cnt_t<1>(st, a, sm_count, smem);
}
void dg_launch_cap_degrees(cudaStream_t st, const DgDirectArgs& a, int sm_count);

// deg_capped = min(deg_raw, max_neighbors) and the largest row, after the symmetric count pass (direct-path rows only)
__global__ void __launch_bounds__(256) k_cap_degrees(const DgDirectArgs A) {
    const long long r0 = A.row_begin, r1 = A.row_end > 0 ? A.row_end : A.n_atoms;
    int mx = 0;
    for (long long a = r0 + (long long)blockIdx.x * blockDim.x + threadIdx.x; a < r1; a += (long long)gridDim.x * blockDim.x) {
        const DgSmall* __restrict__ G = A.small + A.astruct[a];
        if (G->mode == DG_MODE_CELL || G->status != DG_ST_OK) continue;
        const int m = A.deg_raw[a];
        A.deg_capped[a] = m < A.max_neighbors ? m : A.max_neighbors;
        mx = m > mx ? m : mx;
    }
    mx = __reduce_max_sync(FULL_MASK, mx);
    if ((threadIdx.x & 31) == 0 && mx) atomicMax(&A.stats->max_deg, mx);
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
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    long long g2 = (rows + 255) / 256;
    if (g2 > (long long)sm_count * 8) g2 = (long long)sm_count * 8;
    k_cap_degrees<<<(unsigned)g2, 256, 0, st>>>(a);
    g_direct_launches += 2;
    return cudaGetLastError();
}
void dg_launch_cap_degrees(cudaStream_t st, const DgDirectArgs& a, int sm_count) {
    const long long rows = (a.row_end > 0 ? a.row_end : a.n_atoms) - a.row_begin;
    if (rows <= 0) return;
    long long g2 = (rows + 255) / 256;
    if (g2 > (long long)sm_count * 8) g2 = (long long)sm_count * 8;
    k_cap_degrees<<<(unsigned)g2, 256, 0, st>>>(a);
    ++g_direct_launches;
}
// symmetric count pass of rows [row_begin, row_end): deg_raw must be zero there; eps_mode 0 or 1 only
cudaError_t dg_launch_direct_count_sym(cudaStream_t st, const DgDirectArgs& a, int sm_count, int min_blocks) {
    if (a.n_atoms <= 0) return cudaSuccess;
    if (a.eps_mode > 1) return cudaErrorInvalidValue;
    return min_blocks <= 2 ? launch_count_sym_m<2>(st, a, sm_count) : launch_count_sym_m<3>(st, a, sm_count);
}

size_t dg_direct_smem_bytes(const DgDirectArgs& a, bool fill) {
    return fill ? (size_t)a.n_warps * dg_rowbuf_bytes(a.row_cap, a.n_buckets) : 0;
}
size_t dg_direct_onepass_smem_bytes(const DgDirectArgs& a) {
    return (size_t)a.n_warps * dg_one_warp_bytes(a.row_cap, a.n_buckets);
}

template <bool FILL, int MINB, int EPS, bool HITS, bool LEAN = false>
static cudaError_t launch_direct_t(cudaStream_t st, const DgDirectArgs& a, int sm_count, size_t smem, int threads) {
    cudaError_t e = cudaFuncSetAttribute(k_direct<FILL, MINB, EPS, HITS, LEAN>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int occ = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_direct<FILL, MINB, EPS, HITS, LEAN>, threads, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
    // persistent grid: every SM full, chunks of DG_ROW_CHUNK consecutive rows handed out round-robin (the single-pass
    // look-back relies on the grid being fully resident: never more CTAs than the occupancy query allows)
    const long long rows = (a.row_end > 0 ? a.row_end : a.n_atoms) - a.row_begin;
    const long long want = (rows + DG_ROW_CHUNK - 1) / DG_ROW_CHUNK;
    long long grid = (long long)sm_count * occ;
    if (grid > want) grid = want;
    if (grid < 1) return cudaSuccess;
    k_direct<FILL, MINB, EPS, HITS, LEAN><<<(unsigned)grid, threads, smem, st>>>(a);
    return cudaGetLastError();
}

template <bool FILL, int MINB, bool HITS, bool LEAN = false>
static cudaError_t launch_direct_e(cudaStream_t st, const DgDirectArgs& a, int sm_count, size_t smem, int threads) {
    switch (a.eps_mode) {
        case 0: return launch_direct_t<FILL, MINB, 0, HITS, LEAN>(st, a, sm_count, smem, threads);
        case 1: return launch_direct_t<FILL, MINB, 1, HITS, LEAN>(st, a, sm_count, smem, threads);
        default: return launch_direct_t<FILL, MINB, 2, HITS, LEAN>(st, a, sm_count, smem, threads);
    }
}

cudaError_t dg_launch_direct(cudaStream_t st, const DgDirectArgs& a, bool fill, int sm_count, int min_blocks) {
    if (a.n_atoms <= 0) return cudaSuccess;
    const size_t smem = dg_direct_smem_bytes(a, fill);
    const int threads = a.n_warps * 32;
    ++g_direct_launches;
    if (fill && a.max_neighbors >= a.row_cap)  // no row is truncated: the lean sort
        return min_blocks <= 2 ? launch_direct_e<true, 2, false, true>(st, a, sm_count, smem, threads)
                               : launch_direct_e<true, 3, false, true>(st, a, sm_count, smem, threads);
    if (fill)
        return min_blocks <= 2 ? launch_direct_e<true, 2, false>(st, a, sm_count, smem, threads)
                               : launch_direct_e<true, 3, false>(st, a, sm_count, smem, threads);
    if (a.hit_cap > 0 && a.hits)
        return min_blocks <= 2 ? launch_direct_e<false, 2, true>(st, a, sm_count, smem, threads)
                               : launch_direct_e<false, 3, true>(st, a, sm_count, smem, threads);
    return min_blocks <= 2 ? launch_direct_e<false, 2, false>(st, a, sm_count, smem, threads)
                           : launch_direct_e<false, 3, false>(st, a, sm_count, smem, threads);
}

// ---------------------------------------------------------------------------------------------------------------------
// Speculative capped build (the reference's default call: max_neighbors = 20 at r_c = 10, every atom of a crystal has
// hundreds of candidates).  With every row truncated to the cap the CSR offsets are known before any distance is
// computed: row a starts at a * cap.  One kernel sweeps a row into its shared-memory buffer, sorts, keeps the first
// `cap` entries and emits them (+ features); a row that comes up short (or outgrows the buffer sized from the density
// estimate) raises spec_fail and the host redoes the batch with counted offsets.
// ---------------------------------------------------------------------------------------------------------------------
template <int EPS>
__global__ void __launch_bounds__(256, 2) k_direct_spec(const DgDirectArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    const RowBuf rb = dg_rowbuf_carve(smem_raw + (size_t)warp * dg_rowbuf_bytes(A.row_cap, A.n_buckets), A.row_cap);
    unsigned near = 0;
    int maxdeg = 0;
    bool fail = false;
    const int cap = A.max_neighbors;
    const long long n_chunks = (A.n_atoms + DG_ROW_CHUNK - 1) / DG_ROW_CHUNK;
    for (long long chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
        const long long a_end = (chunk + 1) * DG_ROW_CHUNK < A.n_atoms ? (chunk + 1) * DG_ROW_CHUNK : A.n_atoms;
        for (long long a = chunk * DG_ROW_CHUNK + warp; a < a_end; a += nwarps) {
            const DgSmall* __restrict__ G = A.small + A.astruct[a];
            const int mode = G->mode;
            if (mode == DG_MODE_CELL || G->status != DG_ST_OK) {  // (the host speculates on direct-path batches only)
                fail = true;
                continue;
            }
            const int m = traverse_row<true, EPS, false, true>(A, G, a, mode, rb, nullptr, nullptr, lane, lt_mask, near);
            maxdeg = m > maxdeg ? m : maxdeg;
            if (lane == 0) {
                A.deg_raw[a] = m;
                A.deg_capped[a] = m < cap ? m : cap;
            }
            if (m < cap || m > A.row_cap) {
                fail = true;
                continue;
            }
            dg_row_sort_emit(rb, m, cap, A.n_buckets, A.inv_rc, A.key_scale, a * (long long)cap, A.dst, A.image, A.dist,
                             A.stats, lane, A.rbf);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) near += __shfl_xor_sync(FULL_MASK, near, o);
    if (lane == 0) {
        if (near) atomicAdd(&A.stats->n_near, (unsigned long long)near);
        if (maxdeg) atomicMax(&A.stats->max_deg, maxdeg);
        if (fail) atomicExch(&A.stats->spec_fail, 1);
    }
}

__global__ void k_spec_offsets(long long N, long long cap, long long* __restrict__ row_ptr) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i <= N; i += stride) row_ptr[i] = i * cap;
}

template <int EPS>
static cudaError_t launch_spec_t(cudaStream_t st, const DgDirectArgs& a, int sm_count, size_t smem) {
    cudaError_t e = cudaFuncSetAttribute(k_direct_spec<EPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int occ = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_direct_spec<EPS>, a.n_warps * 32, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) return cudaErrorLaunchOutOfResources;
    const long long want = (a.n_atoms + DG_ROW_CHUNK - 1) / DG_ROW_CHUNK;
    long long grid = (long long)sm_count * occ;
    if (grid > want) grid = want;
    k_direct_spec<EPS><<<(unsigned)grid, a.n_warps * 32, smem, st>>>(a);
    return cudaGetLastError();
}

cudaError_t dg_launch_direct_spec(cudaStream_t st, const DgDirectArgs& a, int sm_count) {
    if (a.n_atoms <= 0) return cudaSuccess;
    long long g = (a.n_atoms + 1 + 255) / 256;
    if (g > (long long)sm_count * 8) g = (long long)sm_count * 8;
    k_spec_offsets<<<(unsigned)g, 256, 0, st>>>(a.n_atoms, (long long)a.max_neighbors, a.row_ptr_out);
    const size_t smem = dg_direct_smem_bytes(a, true);
    g_direct_launches += 2;
    switch (a.eps_mode) {
        case 0: return launch_spec_t<0>(st, a, sm_count, smem);
        case 1: return launch_spec_t<1>(st, a, sm_count, smem);
        default: return launch_spec_t<2>(st, a, sm_count, smem);
    }
}

// single-pass build: a.lb_status ([ceil(N / n_warps)] chunk status words) must be zeroed on the stream before the launch
template <int MINB, int EPS>
static cudaError_t launch_one_t(cudaStream_t st, const DgDirectArgs& a, int sm_count, size_t smem, int threads) {
    cudaError_t e = cudaFuncSetAttribute(k_direct_one<MINB, EPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int occ = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_direct_one<MINB, EPS>, threads, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) return cudaErrorLaunchOutOfResources;
    // the look-back needs every CTA resident: never more CTAs than the occupancy query allows
    const long long want = (a.n_atoms + a.n_warps - 1) / a.n_warps;
    long long grid = (long long)sm_count * occ;
    if (grid > want) grid = want;
    k_direct_one<MINB, EPS><<<(unsigned)grid, threads, smem, st>>>(a);
    return cudaGetLastError();
}
template <int MINB>
static cudaError_t launch_one_e(cudaStream_t st, const DgDirectArgs& a, int sm_count, size_t smem, int threads) {
    switch (a.eps_mode) {
        case 0: return launch_one_t<MINB, 0>(st, a, sm_count, smem, threads);
        case 1: return launch_one_t<MINB, 1>(st, a, sm_count, smem, threads);
        default: return launch_one_t<MINB, 2>(st, a, sm_count, smem, threads);
    }
}
cudaError_t dg_launch_direct_onepass(cudaStream_t st, const DgDirectArgs& a, int sm_count, int min_blocks) {
    if (a.n_atoms <= 0) return cudaSuccess;
    if (a.n_warps > DG_ONE_MAX_WARPS) return cudaErrorInvalidValue;
    const size_t smem = dg_direct_onepass_smem_bytes(a);
    const int threads = a.n_warps * 32;
    ++g_direct_launches;
    return min_blocks <= 2 ? launch_one_e<2>(st, a, sm_count, smem, threads) : launch_one_e<3>(st, a, sm_count, smem, threads);
}

// ---------------------------------------------------------------------------------------------------------------------
// fill pass from hit lists: one warp per row reads the row's accepted (atom, image) pairs written by the count pass,
// recomputes their d2 with the reference arithmetic (bit-identical to what the count pass tested), sorts and emits.
// No traversal state: few registers, more resident warps to overlap the sort with the feature stores.
// ---------------------------------------------------------------------------------------------------------------------
// Row schedule of one warp: rows warp, warp + nwarps, ... of chunk blockIdx.x, then of chunk blockIdx.x + gridDim.x ...
struct RowIter {
    long long chunk, a;
};
__device__ __forceinline__ bool row_iter_valid(const RowIter& it, long long n_chunks) { return it.chunk < n_chunks; }
__device__ __forceinline__ void row_iter_first(RowIter& it, long long n_atoms, int warp) {
    it.chunk = blockIdx.x;
    it.a = it.chunk * DG_ROW_CHUNK + warp;
    (void)n_atoms;
}
__device__ __forceinline__ void row_iter_next(RowIter& it, long long n_atoms, long long n_chunks, int warp, int nwarps) {
    it.a += nwarps;
    while (true) {
        const long long end = (it.chunk + 1) * DG_ROW_CHUNK < n_atoms ? (it.chunk + 1) * DG_ROW_CHUNK : n_atoms;
        if (it.a < end) return;
        it.chunk += gridDim.x;
        if (it.chunk >= n_chunks) return;
        it.a = it.chunk * DG_ROW_CHUNK + warp;
    }
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// per-row scalars fetched one row ahead
struct RowHead {
    int s, m;
    long long o;
};
__device__ __forceinline__ RowHead row_head_load(const DgDirectArgs& A, long long a) {
    RowHead h;
    h.s = A.astruct[a];
    h.m = A.deg_raw[a];
    h.o = A.row_ptr[a];
    return h;
}

// shared memory of one warp of k_fill_hits: d2 / hit words double buffered (the next row streams in with cp.async while
// this one is sorted and emitted), sorted words, fp32 distances in row order, tie scratch
__host__ __device__ inline size_t dg_fill_hits_sk_words(int W) { return ((size_t)DG_SK_WORDS(W) + 3) & ~(size_t)3; }
__host__ __device__ inline size_t dg_fill_hits_warp_bytes(int W) {
    return (size_t)2 * W * 8 + (size_t)2 * W * 4 + dg_fill_hits_sk_words(W) * 4 + (size_t)W * 4 + 16 +
           (size_t)2 * DG_STAGE_Q * 16;
}

__global__ void __launch_bounds__(256, 3) k_fill_hits(const DgDirectArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    // rows are hit_cap entries apart in global memory (a multiple of 32: 16-byte aligned); only the first row_cap of a
    // row can be in use
    const int W = A.hit_cap < A.row_cap ? A.hit_cap : A.row_cap;
    unsigned char* base = smem_raw + (size_t)warp * dg_fill_hits_warp_bytes(W);
    double* rd = reinterpret_cast<double*>(base);                      // [2][W]
    uint32_t* hw = reinterpret_cast<uint32_t*>(base + (size_t)2 * W * 8);  // [2][W]
    uint32_t* sk = hw + 2 * W;  // padded (DG_SKI)
    float* fd = reinterpret_cast<float*>(sk + dg_fill_hits_sk_words(W));
    int* tie = reinterpret_cast<int*>(fd + W);
    float4* stage = reinterpret_cast<float4*>(reinterpret_cast<unsigned char*>(tie) + 16);  // 2 x DG_STAGE_Q float4 (TMA source)
    unsigned seq = 0;
    const long long n_chunks = (A.n_atoms + DG_ROW_CHUNK - 1) / DG_ROW_CHUNK;

    RowIter it;
    row_iter_first(it, A.n_atoms, warp);
    if (!row_iter_valid(it, n_chunks) || it.a >= A.n_atoms) return;  // (only when the grid exceeds the chunk count)
    RowHead cur = row_head_load(A, it.a);
    {
        const uint32_t* src = A.hits + (size_t)it.a * (size_t)A.hit_cap;
        const double* srcd = A.hits_d2 + (size_t)it.a * (size_t)A.hit_cap;
        for (int c = lane * 4; c < W; c += 128) cp_async16(hw + c, src + c);
        for (int c = lane * 2; c < W; c += 64) cp_async16(rd + c, srcd + c);
        cp_async_commit();
    }
    int buf = 0;
    while (true) {
        RowIter nx = it;
        row_iter_next(nx, A.n_atoms, n_chunks, warp, nwarps);
        const bool have_next = row_iter_valid(nx, n_chunks);
        cp_async_wait_all();
        __syncwarp();
        RowHead nxt = cur;
        if (have_next) {
            nxt = row_head_load(A, nx.a);
            const uint32_t* src = A.hits + (size_t)nx.a * (size_t)A.hit_cap;
            const double* srcd = A.hits_d2 + (size_t)nx.a * (size_t)A.hit_cap;
            uint32_t* hwn = hw + (buf ^ 1) * W;
            double* rdn = rd + (buf ^ 1) * W;
            for (int c = lane * 4; c < W; c += 128) cp_async16(hwn + c, src + c);
            for (int c = lane * 2; c < W; c += 64) cp_async16(rdn + c, srcd + c);
            cp_async_commit();
        }
        const DgSmall* __restrict__ G = A.small + cur.s;
        const int m = cur.m;
        // rows this pass does not own (image range beyond the 5-bit fields, more hits than slots) are filled by the
        // traversal passes
        const bool cell = G->mode == DG_MODE_CELL;
        if (G->status == DG_ST_OK && m <= W && (cell ? A.cloud_j != nullptr : G->nimg <= DG_HIT_NMAX)) {
            const int keep = m < A.max_neighbors ? m : A.max_neighbors;
            if (cell) {
                const CloudPayload P{hw + buf * W, A.cloud_j, A.cloud_n};
                dg_row_sort_bitonic_any(rd + buf * W, sk, fd, tie, P, m, keep, A.key_scale, cur.o, A.dst, A.image,
                                        A.dist, A.stats, lane);
            } else {
                const HitWordPayload P{hw + buf * W};
                dg_row_sort_bitonic_any(rd + buf * W, sk, fd, tie, P, m, keep, A.key_scale, cur.o, A.dst, A.image,
                                        A.dist, A.stats, lane);
            }
            if (A.rbf.K4 > 0) {
                if (A.fill_hits_bulk) dg_rbf_row_bulk(fd, keep, cur.o, lane, A.rbf, stage, seq);
                else dg_rbf_row(fd, keep, cur.o, lane, A.rbf);
                __syncwarp();
            }
        }
        if (!have_next) break;
        it = nx;
        cur = nxt;
        buf ^= 1;
    }
    if (lane == 0) dg_bulk_wait_all();  // the stages must outlive the copies that read them
}

size_t dg_fill_hits_smem_bytes(const DgDirectArgs& a) {
    const int W = a.hit_cap < a.row_cap ? a.hit_cap : a.row_cap;
    return (size_t)a.n_warps * dg_fill_hits_warp_bytes(W);
}

cudaError_t dg_launch_fill_hits(cudaStream_t st, const DgDirectArgs& a, int sm_count) {
    if (a.n_atoms <= 0) return cudaSuccess;
    const size_t smem = dg_fill_hits_smem_bytes(a);
    const int threads = a.n_warps * 32;
    cudaError_t e = cudaFuncSetAttribute(k_fill_hits, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int occ = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_fill_hits, threads, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
    const long long want = (a.n_atoms + DG_ROW_CHUNK - 1) / DG_ROW_CHUNK;
    long long grid = (long long)sm_count * occ;
    if (grid > want) grid = want;
    ++g_direct_launches;
    k_fill_hits<<<(unsigned)grid, threads, smem, st>>>(a);
    return cudaGetLastError();
}
