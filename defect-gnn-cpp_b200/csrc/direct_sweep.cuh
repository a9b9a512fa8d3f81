// direct_sweep.cuh -- the candidate sweep of the direct (small-structure) path, shared by its kernels
// (dgnn_direct.cu count / fill passes, dgnn_direct_one.cu single-pass build, dgnn_direct_count.cu symmetric count).
// Compile the including translation units with -fmad=false.
//
// Reference behaviour reproduced (paths relative to the reference repo):
//   image cloud point  = pos_j + ((n_a*row0 + n_b*row1) + n_c*row2)          src/graph/neighbor_list.cpp:82-87
//   accept             : ((dx*dx)+(dy*dy))+(dz*dz) < r_c*r_c, dx = q - p      :40 + nanoflann L2_Simple_Adaptor
//   image range        : |n_k| <= ceil(r_c / min|row|) + 1                     :68-72, 79-81
//   self exclusion     : orig == i && sqrt(d2) < epsilon                      :47-49
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "dgnn_internal.h"
#include "row_sort.cuh"

// kernel launches of the direct path issued by the calling thread since the last reset (a build runs on one thread)
void dg_direct_launch_bump(int n);

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

// Count pass with hit lists of truncated rows (HITS): only accepted pairs below a per-structure d2 threshold are stored
// (a little more than max_neighbors + 1 of them, from the density) -- the fill pass reads and sorts ~50 instead of all
// ~160-330.  tot / lo are per-lane tallies of all accepted pairs and of those safely below the threshold
// (d2 < d2_max (1 - 2^-18)): a row whose lo reaches max_neighbors + 1 has every entry of its first max_neighbors + 1
// canonical ranks, and every entry that ties with one of them, in the list (ties differ by < 2^-50 in d2); the others
// are left to the traversal fill.  d2_max = +inf: everything is stored.
#define DG_INF __longlong_as_double(0x7ff0000000000000LL)
struct HitTally {
    int tot, lo;
    double d2_max, d2_lo;
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
                                           unsigned lt_mask, int& m, unsigned& near, HitTally& ht) {
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
            ht.tot += hit ? 1 : 0;
            ht.lo += (hit && d2 < ht.d2_lo) ? 1 : 0;
            hit = hit && d2 < ht.d2_max;
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
// accepted pairs (per lane when !FILL && !HITS, warp-uniform otherwise; HITS: the STORED pairs, the tallies are in ht).
// All 32 lanes must call.
// IMGSTEP: the IMGPAR sweep walks a lane's (n_a, n_b, n_c) with mixed-radix additions of 32 instead of two divisions
// and two remainders per candidate (same images, same order).  Only the dynamically scheduled instantiations use it:
// the headline kernel's code generation is kept exactly as it was.
template <bool FILL, int EPS, bool HITS, bool ONE, bool IMGSTEP = false>
__device__ __forceinline__ int traverse_row(const DgDirectArgs& A, const DgSmall* __restrict__ G, long long a, int mode,
                                            const RowBuf& rb, uint32_t* __restrict__ hrow, double* __restrict__ hd2,
                                            int lane, unsigned lt_mask, unsigned& near, HitTally& ht) {
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
            if (mc0 == 1 && mc1 == 1) {
                // one image along a and b for every lane of the chunk (cells at least 2 r_c wide along those axes: the C2
                // supercells): only the loop over n_c remains -- same images, same order as the nest below, without its
                // loop-carried moves (headline fill 8.33 -> 8.10 ms, capped count 3.15 -> 2.88 ms, same-box A/B)
                const double na_d = (double)lo0, nb_d = (double)lo1;
                const double ob0 = DG_ADD(DG_MUL(na_d, L00), DG_MUL(nb_d, L10));
                const double ob1 = DG_ADD(DG_MUL(na_d, L01), DG_MUL(nb_d, L11));
                const double ob2 = DG_ADD(DG_MUL(na_d, L02), DG_MUL(nb_d, L12));
                const bool act_ab = c0 > 0 && c1 > 0;
                double ncd = (double)lo2;
                for (int sc = 0; sc < mc2; ++sc, ncd += 1.0) {
                    const double px = DG_ADD(xj, DG_ADD(ob0, DG_MUL(ncd, L20)));
                    const double py = DG_ADD(yj, DG_ADD(ob1, DG_MUL(ncd, L21)));
                    const double pz = DG_ADD(zj, DG_ADD(ob2, DG_MUL(ncd, L22)));
                    test_point<FILL, EPS, HITS, ONE>(A, S, rb, hrow, hd2, act_ab && sc < c2, self, jj, px, py, pz, lo0, lo1,
                                                     lo2 + sc, lt_mask, m, near, ht);
                }
                continue;
            }
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
                                                    lo0 + sa, lo1 + sb, lo2 + sc, lt_mask, m, near, ht);
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
            if (IMGSTEP) {
                if (T == 0) continue;
                // t = (ia * c1 + ib) * c2 + ic, t = lane, lane + 32, ...: + 32 = + (q / c1, q % c1, 32 % c2), q = 32 / c2
                const int s2 = 32 % c2, q = 32 / c2, s1 = q % c1, s0 = q / c1;
                int ic = lane % c2, tq = lane / c2;
                int ib = tq % c1, ia = tq / c1;
                for (int tb = 0; tb < T; tb += 32) {
                    const bool active = tb + lane < T;
                    const int na = active ? lo0 + ia : lo0, nb = active ? lo1 + ib : lo1, nc = active ? lo2 + ic : lo2;
                    const double px = DG_ADD(xj, dg_image_offset(G->L, 0, na, nb, nc));
                    const double py = DG_ADD(yj, dg_image_offset(G->L, 1, na, nb, nc));
                    const double pz = DG_ADD(zj, dg_image_offset(G->L, 2, na, nb, nc));
                    test_point<FILL, EPS, HITS, ONE>(A, S, rb, hrow, hd2, active, j == S.iloc, j, px, py, pz, na, nb, nc,
                                                lt_mask, m, near, ht);
                    ic += s2;
                    int carry = ic >= c2 ? 1 : 0;
                    ic -= carry ? c2 : 0;
                    ib += s1 + carry;
                    carry = ib >= c1 ? 1 : 0;
                    ib -= carry ? c1 : 0;
                    ia += s0 + carry;
                }
                continue;
            }
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
                                            lt_mask, m, near, ht);
            }
        }
    }
    return m;
}
template <bool FILL, int EPS, bool HITS, bool ONE, bool IMGSTEP = false>
__device__ __forceinline__ int traverse_row(const DgDirectArgs& A, const DgSmall* __restrict__ G, long long a, int mode,
                                            const RowBuf& rb, uint32_t* __restrict__ hrow, double* __restrict__ hd2,
                                            int lane, unsigned lt_mask, unsigned& near) {
    HitTally ht;  // (dead without HITS)
    ht.tot = ht.lo = 0;
    ht.d2_max = ht.d2_lo = 0.0;
    return traverse_row<FILL, EPS, HITS, ONE, IMGSTEP>(A, G, a, mode, rb, hrow, hd2, lane, lt_mask, near, ht);
}
