// dgnn_cell32.cu -- cell-list traversal with fp32 tiles staged in shared memory (north_star kernel (2)): the pairwise
// distance pass of large structures (BASELINE config 4, one ~1M-atom cell).  Compile with -fmad=false.
//
// One CTA per work item (a block of source cells of one structure).  The item's neighbour tile -- the cell-sorted
// image points of the block plus its stencil halo -- is staged ONCE in shared memory as float4 {x, y, z relative to the
// tile centre in fp32, cloud index}.  A warp takes one source cell at a time, flattens the cell's pruned stencil
// (contiguous slot ranges of the tile, one per stencil row) into a per-warp list of tile slots, and every source atom
// of the cell sweeps that list 32 candidates per step at full lane occupancy: ONE LDS.128 + 9 fp32 operations per
// candidate, against three dependent 8-byte loads from L2 and 8 fp64 operations in the fp64 traversal it replaces.
//
// Exactness.  fp32 only PRE-FILTERS (hard part 1): with the rigorous per-tile bound `band` on
// |d2_32 - d2| (derivation at c32_band), d2_32 < r2 - band is an edge and d2_32 > r2 + band is not; only candidates in
// between (and anything close enough to the source for the epsilon rule) are decided by the reference's own fp64
// arithmetic on the reference's own operands, gathered from the cloud by index:
//   d2 = ((dx*dx)+(dy*dy))+(dz*dz), dx = q - p, accept iff d2 < r_c*r_c            src/graph/neighbor_list.cpp:40 + nanoflann
//   drop iff orig == i && sqrt(d2) < epsilon                                          :47-49
// The COUNT pass tallies sure hits per lane and drains the few in-band candidates lane-parallel; the FILL pass compacts
// every candidate (ballot/popc) into the warp's buffer, recomputes all of them in fp64 -- the accepted pairs need their
// exact d2 for Neighbor::distance anyway -- then sorts in registers and emits the CSR row with the fused Gaussian
// epilogue.  No hit lists travel between the passes (the fp64 traversal wrote and re-read 12 B per accepted pair).
#include <cuda_runtime.h>
#include <stdint.h>

#include "dgnn_internal.h"
#include "row_sort.cuh"
#include "cell_items.cuh"

void dg_kernel_launch_bump(int n);  // dgnn_kernels.cu

// (the flattened stencil of one source cell -- a 5x5x5 stencil of r_c/2 cells holds ~650 tile slots -- has a.src_cap
// entries per warp: the host picks 1024 or what shared memory allows)
#define C32_QCAP 96       // COUNT: in-band candidates waiting for the fp64 test (drained at > QCAP - 32)
#define C32_IMG_FLAG 0x80000000u

struct C32Shared {
    DgGeom G;
    long long s_struct;
    double org[3];      // tile centre (Cartesian)
    float r2m, r2p;     // r2 -+ band, rounded outwards
    float lowthr;       // candidates with d2_32 <= lowthr may fall under the epsilon rule: decided in fp64
    double eps_guard2;  // d2 <= this: the epsilon rule needs the exact sqrt test
    int staged_total;
    int maxdeg;
    unsigned near;
};

__host__ __device__ inline size_t c32_align16(size_t v) { return (v + 15) & ~(size_t)15; }
// tile: float4 per point (+ one sentinel slot that pads the slot lists); FILL also stages the fp64 operands (24 B/point)
// and the point's (atom, packed image) (8 B/point): the sorted row is emitted without gathers from the cloud arrays
__host__ __device__ inline size_t c32_shared_part(int tile_cap, int tcs_cap, bool fill) {
    return c32_align16((size_t)(tile_cap + 1) * 16) + (fill ? c32_align16((size_t)tile_cap * 8) * 4 : 0) +
           c32_align16((size_t)tcs_cap * 4) + c32_align16((size_t)(DG_ROWS_CAP + 1) * 4) +
           c32_align16((size_t)DG_STN_CAP * 2) * 2;
}
__host__ __device__ inline size_t c32_sk_words(int W) { return ((size_t)DG_SK_WORDS(W) + 3) & ~(size_t)3; }
// FILL: the candidate buffer (W + 32 slots) aliases the sort scratch + fp32 distances, which are only written after
// the last candidate has been decided
__host__ __device__ inline size_t c32_warp_bytes(int W, int idx_cap, bool fill) {
    size_t b = c32_align16((size_t)idx_cap * 2);
    if (!fill) return b + c32_align16((size_t)C32_QCAP * 4);
    return b + (size_t)W * 8 + (size_t)W * 4 + c32_sk_words(W) * 4 + (size_t)W * 4 + 16;
}
size_t dg_cell32_smem_bytes(const DgTraverseArgs& a, bool fill) {
    return c32_shared_part(a.tile_cap, a.tcs_cap, fill) + (size_t)a.n_warps * c32_warp_bytes(a.row_cap, a.src_cap, fill);
}

// Guard band of the fp32 pre-filter.  Tile coordinates are fp32 roundings of (cloud position - tile centre):
// |error| <= u M each (u = 2^-24, M = bound on |x|+|y|+|z| inside the tile); a coordinate difference adds one rounding
// u |dx|; so dx_32 = dx (1 + rho) + a with |rho| <= u, |a| <= 2 u M.  The three products and two sums of d2_32 add a
// relative 3u.  With beta = rho r_c + sqrt(3) a:  |sqrt(d2_32) - d| <= beta for d on either side of r_c, hence
// |d2_32 - d2| <= 2 r_c beta + beta^2 + 4 u r_c^2 whenever the exact d2 is on the other side of r_c^2.  Constants are
// taken 8x larger (rho = 8u, a = 16 u M) and the total doubled; the near-cutoff band (|d - r_c| <= 1e-6, tallied by
// the count pass on drained candidates) is added so that it lies inside.
__device__ __forceinline__ double c32_band(double r_cutoff, double r2, double M) {
    const double u = 5.9604644775390625e-8;
    const double beta = 8.0 * u * r_cutoff + 1.7320508075688774 * 16.0 * u * (M + r_cutoff);
    return 2.0 * (2.0 * r_cutoff * beta + beta * beta + 4.0 * u * r2) + (4.0e-6 * r_cutoff + 1.0e-11);
}

// the reference's metric on the reference's operands (q = source position, p = image-cloud point)
__device__ __forceinline__ double c32_exact_d2(double qx, double qy, double qz, double px, double py, double pz) {
    const double dx = DG_SUB(qx, px), dy = DG_SUB(qy, py), dz = DG_SUB(qz, pz);
    return DG_ADD(DG_ADD(DG_MUL(dx, dx), DG_MUL(dy, dy)), DG_MUL(dz, dz));
}

struct C32Warp {
    unsigned short* idx;  // flattened stencil (tile slots)
    uint32_t* q;          // COUNT: in-band queue / FILL: candidate buffer (tile slots)
    double* rd;           // FILL: accepted d2 by entry
    uint32_t* hw;         // FILL: accepted cloud indices by entry
    uint32_t* sk;         // FILL: sort scratch
    float* fd;            // FILL: distances in final order (fp32), input of the fused RBF epilogue
    int* tie;
};

struct C32Tile {
    const float4* T;
    const double *X, *Y, *Z;  // FILL: the reference's fp64 operands of the tile points
    const uint2* JN;          // FILL: (local atom index, packed image) of the tile points
    const int* tcs;
    const signed char* stn;   // stencil row q: z-range [stn[2q], stn[2q+1]] (empty if lo > hi)
    const short* srow;        // stencil row q: tile-row offset d_a * td_b + d_b
    int sentinel;             // slot of a far-away dummy point (pads the slot lists to whole warp steps)
};

// Flatten the pruned stencil of source cell (ta, tb, tc) -- one contiguous slot range of the tile per stencil row --
// into the warp's slot list: a lane owns a stencil row (its range comes from two table reads), a warp scan places the
// ranges, and the list is padded to a whole number of warp steps with the sentinel slot.  Returns the padded length,
// or -1 when the stencil does not fit the list (the caller then sweeps the ranges one by one).
__device__ __forceinline__ int c32_build_list(const DgGeom& G, const ItemBox& B, const C32Tile& tl, unsigned short* idx,
                                              int idx_cap, int ta, int tb, int tc, int lane) {
    const int nst = (2 * G.R[0] + 1) * (2 * G.R[1] + 1);
    const int zs = B.td[2] + 1;
    const int row0 = ta * B.td[1] + tb;
    int T = 0;
    for (int q0 = 0; q0 < nst; q0 += 32) {
        const int q = q0 + lane;
        int s = 0, len = 0;
        if (q < nst) {
            const int zlo = tl.stn[2 * q], zhi = tl.stn[2 * q + 1];
            if (zlo <= zhi) {
                const int base = (row0 + tl.srow[q]) * zs + tc;
                s = tl.tcs[base + zlo];
                len = tl.tcs[base + zhi + 1] - s;
            }
        }
        int inc = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(FULL_MASK, inc, o);
            if (lane >= o) inc += t;
        }
        const int total = __shfl_sync(FULL_MASK, inc, 31);
        if (T + total > idx_cap - 32) return -1;
        const int off = T + inc - len;
        unsigned mask = __ballot_sync(FULL_MASK, len > 0);
        while (mask) {
            const int r = __ffs(mask) - 1;
            mask &= mask - 1;
            const int sr = __shfl_sync(FULL_MASK, s, r), lr = __shfl_sync(FULL_MASK, len, r);
            const int orr = __shfl_sync(FULL_MASK, off, r);
            for (int t = lane; t < lr; t += 32) idx[orr + t] = (unsigned short)(sr + t);
        }
        T += total;
    }
    const int Tp = (T + 31) & ~31;
    if (T + lane < Tp) idx[T + lane] = (unsigned short)tl.sentinel;
    __syncwarp();
    return Tp;
}

struct C32TilePayload {  // FILL: hit word = tile slot; (atom, image) from the staged table
    static constexpr bool kSweepOrder = false;
    uint32_t* hw;
    const uint2* JN;
    __device__ __forceinline__ uint2 load_item(int e) const { return make_uint2(hw[e], 0u); }
    __device__ __forceinline__ void store_item(int e, uint2 v) const { hw[e] = v.x; }
    __device__ __forceinline__ void get(int e, int& j, uint32_t& img) const {
        const uint2 v = JN[hw[e]];
        j = (int)v.x;
        img = v.y;
    }
};

struct C32Src {
    float x, y, z;       // tile coordinates
    double qx, qy, qz;   // reference operands
    uint32_t pc;         // cloud index of the source point
    int iloc;
};

// fp64 decision of queued candidates q[0, n): returns (per lane) how many were accepted; COUNT tallies the near band.
template <bool FILL>
__device__ __forceinline__ int c32_drain(const DgTraverseArgs& A, const C32Shared& sh, const C32Tile& tl, const C32Warp& w,
                                         const C32Src& S, int n, int lane, unsigned lt_mask, int& m, unsigned& near) {
    int acc_cnt = 0;
    for (int e0 = 0; e0 < n; e0 += 32) {
        const int e = e0 + lane;
        const bool valid = e < n;
        const unsigned slot = w.q[valid ? e : 0];
        const uint32_t pc = __float_as_uint(tl.T[slot].w) & ~C32_IMG_FLAG;
        const double d2 = FILL ? c32_exact_d2(S.qx, S.qy, S.qz, tl.X[slot], tl.Y[slot], tl.Z[slot])
                               : c32_exact_d2(S.qx, S.qy, S.qz, A.cx[pc], A.cy[pc], A.cz[pc]);
        bool acc = valid && d2 < A.r2;
        if (acc && d2 <= sh.eps_guard2) {  // neighbor_list.cpp:47-49
            if ((FILL ? (int)tl.JN[slot].x : A.cj[pc]) == S.iloc && __dsqrt_rn(d2) < A.epsilon) acc = false;
        }
        if (!FILL) {
            if (valid && fabs(__dsqrt_rn(d2) - A.r_cutoff) <= 1e-6) near++;
            acc_cnt += acc ? 1 : 0;
        } else {
            const unsigned mask = __ballot_sync(FULL_MASK, acc);
            if (acc) {
                const int pos = m + __popc(mask & lt_mask);
                if (pos < A.row_cap) {
                    w.rd[pos] = d2;
                    w.hw[pos] = slot;
                }
            }
            m += __popc(mask);
        }
    }
    __syncwarp();
    return acc_cnt;
}

// One source atom against candidates.  LIST: the warp's padded slot list idx[0, T) (T a multiple of 32); otherwise the
// contiguous slot range [s0, s0 + T).  COUNT: per-lane sure hits in cnt, in-band candidates accumulate in w.q (qn);
// FILL: every candidate goes to w.q.
template <bool FILL, bool LIST>
__device__ __forceinline__ void c32_sweep(const DgTraverseArgs& A, const C32Shared& sh, const C32Tile& tl, const C32Warp& w,
                                          const C32Src& S, int s0, int T, int lane, unsigned lt_mask, int& cnt, int& qn,
                                          int& m, unsigned& near, int qcap, uint32_t& cand_word) {
    const float r2m = sh.r2m, r2p = sh.r2p, lowthr = sh.lowthr;
    for (int f0 = 0; f0 < T; f0 += 32) {
        unsigned slot;
        if (LIST) slot = w.idx[f0 + lane];
        else slot = (f0 + lane < T) ? (unsigned)(s0 + f0 + lane) : (unsigned)tl.sentinel;
        const float4 c = tl.T[slot];
        const float dx = S.x - c.x, dy = S.y - c.y, dz = S.z - c.z;
        const float d2f = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
        // the fill pass's candidates of this step (lane f0 / 32 keeps the word).  Unconditional: testing A.cmask here
        // cost the count kernel 0.15 ms on C4 (same-box A/B), the vote itself nothing measurable
        if (!FILL && LIST) {
            const unsigned cm = __ballot_sync(FULL_MASK, d2f <= r2p);
            if (lane == (f0 >> 5)) cand_word = cm;
        }
        if (!FILL) {
            // sure hits are tallied per lane; what lies in the guard band, or close enough to the source for the
            // epsilon rule (tallied above, taken back here), waits for the fp64 decision
            cnt += d2f < r2m ? 1 : 0;
            const bool low = d2f <= lowthr;
            const bool maybe = (!(d2f < r2m) && d2f <= r2p) || low;
            if (__any_sync(FULL_MASK, maybe)) {
                cnt -= (low && d2f < r2m) ? 1 : 0;
                const unsigned mask = __ballot_sync(FULL_MASK, maybe);
                if (maybe) w.q[qn + __popc(mask & lt_mask)] = slot;
                qn += __popc(mask);
                if (qn > qcap - 32) {
                    __syncwarp();
                    cnt += c32_drain<false>(A, sh, tl, w, S, qn, lane, lt_mask, m, near);
                    qn = 0;
                }
            }
        } else {
            const bool cand = d2f <= r2p;
            const unsigned mask = __ballot_sync(FULL_MASK, cand);
            if (cand) w.q[qn + __popc(mask & lt_mask)] = slot;
            qn += __popc(mask);
            if (qn > qcap - 32) {
                __syncwarp();
                c32_drain<true>(A, sh, tl, w, S, qn, lane, lt_mask, m, near);
                qn = 0;
            }
        }
    }
}

// COUNT: 8 warps per CTA, three CTAs per SM.  FILL: 16 warps per CTA, one CTA per SM -- the tile is staged twice (fp32
// for the sweep, fp64 operands for the decision), so nothing in the hot loops waits on L2.
// NW: warps per CTA the kernel is compiled for (register budget 65536 / (32 NW)); MAXEPL: largest sort network (rows of up
// to 32 * MAXEPL entries)
template <bool FILL, int NW, int MAXEPL>
__global__ void __launch_bounds__(NW * 32, FILL ? 1 : 3) k_cell32(const __grid_constant__ DgTraverseArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ C32Shared sh;
    __shared__ int s_next_cell;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nthreads = blockDim.x, nwarps = nthreads >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    // carve
    unsigned char* p = smem_raw;
    float4* T = (float4*)p; p += c32_align16((size_t)(A.tile_cap + 1) * 16);
    double *X = nullptr, *Y = nullptr, *Z = nullptr;
    uint2* JN = nullptr;
    if (FILL) {
        X = (double*)p; p += c32_align16((size_t)A.tile_cap * 8);
        Y = (double*)p; p += c32_align16((size_t)A.tile_cap * 8);
        Z = (double*)p; p += c32_align16((size_t)A.tile_cap * 8);
        JN = (uint2*)p; p += c32_align16((size_t)A.tile_cap * 8);
    }
    int* tcs = (int*)p; p += c32_align16((size_t)A.tcs_cap * 4);
    int* rowbase = (int*)p; p += c32_align16((size_t)(DG_ROWS_CAP + 1) * 4);
    signed char* stn = (signed char*)p; p += c32_align16((size_t)DG_STN_CAP * 2);
    short* srow = (short*)p; p += c32_align16((size_t)DG_STN_CAP * 2);
    p += (size_t)warp * c32_warp_bytes(A.row_cap, A.src_cap, FILL);
    C32Warp w;
    w.idx = (unsigned short*)p; p += c32_align16((size_t)A.src_cap * 2);
    const int W = A.row_cap;
    const int qcap = FILL ? W + 32 : C32_QCAP;
    w.q = (uint32_t*)p;
    w.rd = nullptr; w.hw = nullptr; w.sk = nullptr; w.fd = nullptr; w.tie = nullptr;
    if (FILL) {
        w.rd = (double*)p; p += (size_t)W * 8;
        w.hw = (uint32_t*)p; p += (size_t)W * 4;
        w.sk = (uint32_t*)p; p += c32_sk_words(W) * 4;
        w.fd = (float*)p; p += (size_t)W * 4;
        w.tie = (int*)p;
        w.q = w.sk;  // candidates (dead before the sort starts) share the scratch
    }
    unsigned near = 0;
    int maxdeg = 0;
    if (tid == 0) {
        sh.maxdeg = 0;
        sh.near = 0;
    }

    for (long long item = blockIdx.x; item < A.n_items; item += gridDim.x) {
        __syncthreads();  // previous item's shared state is dead
        if (tid == 0) {
            sh.s_struct = find_struct(A.item_base, A.n_structs, item);
            s_next_cell = nwarps;
        }
        __syncthreads();
        {
            const int* src = reinterpret_cast<const int*>(A.geoms + sh.s_struct);
            int* dstw = reinterpret_cast<int*>(&sh.G);
            for (int i = tid; i < (int)(sizeof(DgGeom) / 4); i += nthreads) dstw[i] = src[i];
        }
        __syncthreads();
        const DgGeom& G = sh.G;
        const ItemBox B = item_box(&G, (int)(item - G.item_base));
        const int nrows = B.td[0] * B.td[1];
        const int zs = B.td[2] + 1;
        const int NBr = 2 * G.R[1] + 1;
        const int nst = (2 * G.R[0] + 1) * NBr;

        // ---- per-row point counts -> exclusive row bases; stencil row table; tile centre and guard band ----
        for (int r = tid; r < nrows; r += nthreads) {
            const int ra = r / B.td[1], rb = r % B.td[1];
            const long long g0 = G.cell_base + ((long long)(B.t0[0] + ra) * G.hd[1] + (B.t0[1] + rb)) * G.hd[2] + B.t0[2];
            rowbase[r] = (int)(A.cellstart[g0 + B.td[2]] - A.cellstart[g0]);
        }
        for (int q = tid; q < nst; q += nthreads) {
            const int da = q / NBr - G.R[0], db = q % NBr - G.R[1];
            int zlo = 1, zhi = 0;
            if (!dg_stencil_row(&G, da, db, A.rcm, &zlo, &zhi)) {
                zlo = 1;
                zhi = 0;
            }
            stn[2 * q] = (signed char)zlo;
            stn[2 * q + 1] = (signed char)zhi;
            srow[q] = (short)(da * B.td[1] + db);
        }
        if (tid == 32) {
            double M = 0.0, o[3] = {0.0, 0.0, 0.0};
            for (int k = 0; k < 3; ++k) {
                const double uc = (double)(G.hmin[k] + B.t0[k]) + 0.5 * (double)B.td[k];
                const double half = 0.5 * (double)B.td[k] + 1.0;  // + a cell for the rounding of the binning
                for (int c = 0; c < 3; ++c) {
                    o[c] += uc * G.cvec[3 * k + c];
                    M += half * fabs(G.cvec[3 * k + c]);
                }
            }
            sh.org[0] = o[0]; sh.org[1] = o[1]; sh.org[2] = o[2];
            const double band = c32_band(A.r_cutoff, A.r2, M);
            sh.r2m = __double2float_rd(A.r2 - band);
            sh.r2p = __double2float_ru(A.r2 + band);
            if (A.epsilon > 0.0) {
                const double el = A.epsilon * (1.0 + 1.0e-6) + 1.7320508075688774 * 16.0 * 5.9604644775390625e-8 * (M + A.r_cutoff);
                sh.lowthr = __double2float_ru(el * el * (1.0 + 1.0e-6));
                sh.eps_guard2 = A.epsilon * A.epsilon * (1.0 + 1.0e-9) + 1.0e-300;
            } else {
                sh.lowthr = -1.0f;
                sh.eps_guard2 = -1.0;
            }
        }
        __syncthreads();
        if (warp == 0) {
            const int per = (nrows + 31) >> 5;  // <= 32
            const int b0 = lane * per;
            int sum = 0;
            for (int i = 0; i < per; ++i) {
                const int r = b0 + i;
                if (r < nrows) sum += rowbase[r];
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
                    const int c = rowbase[r];
                    rowbase[r] = run;
                    run += c;
                }
            }
            const int total = __shfl_sync(FULL_MASK, inc, 31);
            if (lane == 0) {
                rowbase[nrows] = total;
                sh.staged_total = total;
            }
        }
        __syncthreads();
        const int total = sh.staged_total;
        if (total > A.tile_cap || nrows * zs > A.tcs_cap || total > 65535) {
            // (the host sizes the tile from the largest item: cannot happen)
            if (tid == 0) atomicExch(&A.stats->overflow, 1);
            continue;
        }
        // ---- stage the tile: fp32 coordinates relative to the tile centre + cloud index (| image flag) ----
        {
            const double ox = sh.org[0], oy = sh.org[1], oz = sh.org[2];
            for (int r = warp; r < nrows; r += nwarps) {
                const int ra = r / B.td[1], rb = r % B.td[1];
                const long long g0 = G.cell_base + ((long long)(B.t0[0] + ra) * G.hd[1] + (B.t0[1] + rb)) * G.hd[2] + B.t0[2];
                const long long gs = A.cellstart[g0];
                const int lb = rowbase[r], cnt = rowbase[r + 1] - lb;
                for (int k = lane; k < cnt; k += 32) {
                    const long long pc = gs + k;
                    const double px = A.cx[pc], py = A.cy[pc], pz = A.cz[pc];
                    const uint32_t pn = A.cn[pc];
                    const uint32_t wbits = (uint32_t)pc | (pn != 0u ? C32_IMG_FLAG : 0u);
                    T[lb + k] = make_float4((float)(px - ox), (float)(py - oy), (float)(pz - oz), __uint_as_float(wbits));
                    if (FILL) {
                        X[lb + k] = px;
                        Y[lb + k] = py;
                        Z[lb + k] = pz;
                        JN[lb + k] = make_uint2((uint32_t)A.cj[pc], pn);
                    }
                }
                for (int c = lane; c <= B.td[2]; c += 32) tcs[r * zs + c] = lb + (int)(A.cellstart[g0 + c] - gs);
            }
            if (tid == 0) T[total] = make_float4(3.0e18f, 3.0e18f, 3.0e18f, __uint_as_float(C32_IMG_FLAG));
        }
        __syncthreads();
        const C32Tile tl{T, X, Y, Z, JN, tcs, stn, srow, total};

        const int nbc0 = B.c1[0] - B.c0[0], nbc1 = B.c1[1] - B.c0[1], nbc2 = B.c1[2] - B.c0[2];
        const int nbcells = nbc0 * nbc1 * nbc2;
        // source cells are handed out dynamically (cells hold different numbers of atoms); the atoms of a cell share its
        // flattened stencil.  (Finer units -- a few atoms of a cell, enumerated into a shared list first -- were tried:
        // the extra barrier and list builds cost more than the balance gains, C4 10.25 against 9.92 ms.)
        int Tn = 0, ta = 0, tb = 0, tc = 0;
        for (int bc = warp; bc < nbcells;) {
            ta = bc / (nbc2 * nbc1) + G.R[0];
            tb = (bc / nbc2) % nbc1 + G.R[1];
            tc = bc % nbc2 + G.R[2];
            const int k0 = tcs[(ta * B.td[1] + tb) * zs + tc], k1 = tcs[(ta * B.td[1] + tb) * zs + tc + 1];
            if (k0 < k1) Tn = c32_build_list(G, B, tl, w.idx, A.src_cap, ta, tb, tc, lane);
            for (int pb = k0; pb < k1; pb += 32) {
            // FILL: the rows' degrees and CSR offsets of up to 32 source points are fetched lane-parallel, one round trip
            // per cell instead of two dependent ones per atom
            int pre_deg = 0, pre_tn = -1;
            long long pre_o = 0;
            if (FILL && pb + lane < k1) {
                const uint32_t wl = __float_as_uint(T[pb + lane].w);
                if (!(wl & C32_IMG_FLAG)) {
                    const long long ig = G.atom_base + (int)JN[pb + lane].x;
                    pre_deg = A.deg_raw[ig];
                    pre_o = A.row_ptr[ig];
                    if (A.cmask) pre_tn = A.cmask_tn[ig];
                }
            }
            const int pe = pb + 32 < k1 ? pb + 32 : k1;
            for (int ps = pb; ps < pe; ++ps) {
                const float4 sv = T[ps];
                const uint32_t wb = __float_as_uint(sv.w);
                if (wb & C32_IMG_FLAG) continue;  // a periodic image that shares the cell: not a source
                C32Src S;
                S.x = sv.x; S.y = sv.y; S.z = sv.z;
                S.pc = wb;
                if (FILL) {
                    S.iloc = (int)JN[ps].x;
                    S.qx = X[ps]; S.qy = Y[ps]; S.qz = Z[ps];
                } else {
                    S.iloc = A.cj[wb];
                    S.qx = A.cx[wb]; S.qy = A.cy[wb]; S.qz = A.cz[wb];
                }
                const long long iglob = G.atom_base + S.iloc;
                int deg_known = 0, pre_tn_s = -1;
                long long o = 0;
                if (FILL) {
                    deg_known = __shfl_sync(FULL_MASK, pre_deg, ps - pb);
                    o = __shfl_sync(FULL_MASK, pre_o, ps - pb);
                    pre_tn_s = __shfl_sync(FULL_MASK, pre_tn, ps - pb);
                    if (deg_known > W) continue;  // (rows beyond the register sort: the fp64 traversal's fill)
                }
                int cnt = 0, qn = 0, m = 0;
                uint32_t cand_word = 0u;
                if (FILL && Tn >= 0 && pre_tn_s == Tn) {
                    // candidates from the count pass's words: no tile reads, no arithmetic
                    const uint32_t myw = A.cmask[(size_t)iglob * 32 + lane];
                    for (int f0 = 0; f0 < Tn; f0 += 32) {
                        const unsigned cm = __shfl_sync(FULL_MASK, myw, f0 >> 5);
                        if (cm == 0u) continue;
                        if ((cm >> lane) & 1u) w.q[qn + __popc(cm & lt_mask)] = w.idx[f0 + lane];
                        qn += __popc(cm);
                        if (qn > qcap - 32) {
                            __syncwarp();
                            c32_drain<true>(A, sh, tl, w, S, qn, lane, lt_mask, m, near);
                            qn = 0;
                        }
                    }
                } else if (Tn >= 0) {
                    c32_sweep<FILL, true>(A, sh, tl, w, S, 0, Tn, lane, lt_mask, cnt, qn, m, near, qcap, cand_word);
                } else {
                    // the stencil does not fit the slot list (very dense cells): one contiguous range at a time
                    const int row0 = ta * B.td[1] + tb;
                    for (int q = 0; q < nst; ++q) {
                        const int zlo = stn[2 * q], zhi = stn[2 * q + 1];
                        if (zlo > zhi) continue;
                        const int base = (row0 + srow[q]) * zs + tc;
                        const int s0 = tcs[base + zlo];
                        c32_sweep<FILL, false>(A, sh, tl, w, S, s0, tcs[base + zhi + 1] - s0, lane, lt_mask, cnt, qn, m,
                                               near, qcap, cand_word);
                    }
                }
                __syncwarp();
                if (!FILL) {
                    if (qn > 0) cnt += c32_drain<false>(A, sh, tl, w, S, qn, lane, lt_mask, m, near);
                    const int deg = __reduce_add_sync(FULL_MASK, cnt);
                    if (lane == 0) {
                        A.deg_raw[iglob] = deg;
                        A.deg_capped[iglob] = deg < A.max_neighbors ? deg : A.max_neighbors;
                    }
                    if (A.cmask) {
                        if (Tn >= 0) A.cmask[(size_t)iglob * 32 + lane] = cand_word;
                        if (lane == 0) A.cmask_tn[iglob] = Tn;
                    }
                    maxdeg = deg > maxdeg ? deg : maxdeg;
                } else {
                    if (qn > 0) c32_drain<true>(A, sh, tl, w, S, qn, lane, lt_mask, m, near);
                    if (m != deg_known) {  // cannot happen: both passes decide every pair the same way
                        if (lane == 0) atomicExch(&A.stats->overflow, 1);
                        continue;
                    }
                    const int keep = m < A.max_neighbors ? m : A.max_neighbors;
                    const C32TilePayload P{w.hw, JN};
                    dg_row_sort_bitonic_upto<MAXEPL>(w.rd, w.sk, w.fd, w.tie, P, m, keep, A.key_scale, o, A.dst, A.image,
                                                     A.dist, A.stats, lane);
                    if (A.rbf.K4 > 0) {
                        dg_rbf_row(w.fd, keep, o, lane, A.rbf);
                        __syncwarp();
                    }
                }
            }
            }
            __syncwarp();
            int nxt = 0;
            if (lane == 0) nxt = atomicAdd(&s_next_cell, 1);
            bc = __shfl_sync(FULL_MASK, nxt, 0);
        }
    }
    if (!FILL) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) near += __shfl_xor_sync(FULL_MASK, near, o);
        __syncthreads();
        if (lane == 0) {
            if (near) atomicAdd(&sh.near, near);
            atomicMax(&sh.maxdeg, maxdeg);
        }
        __syncthreads();
        if (tid == 0) {
            if (sh.near) atomicAdd(&A.stats->n_near, (unsigned long long)sh.near);
            atomicMax(&A.stats->max_deg, sh.maxdeg);
        }
    }
}

template <bool FILL, int NW, int MAXEPL>
static cudaError_t c32_launch_t(cudaStream_t st, const DgTraverseArgs& a, int grid, size_t smem) {
    cudaError_t e = dg_raise_dyn_smem<k_cell32<FILL, NW, MAXEPL>>();
    if (e != cudaSuccess) return e;
    k_cell32<FILL, NW, MAXEPL><<<grid, a.n_warps * 32, smem, st>>>(a);
    return cudaGetLastError();
}

// fill: a.n_warps in {8, 16, 20, 24} (one CTA per SM shares the doubly staged tile); a.row_cap <= 512
cudaError_t dg_launch_cell32(cudaStream_t st, const DgTraverseArgs& a, bool fill, int grid) {
    if (a.n_items <= 0) return cudaSuccess;
    const size_t smem = dg_cell32_smem_bytes(a, fill);
    dg_kernel_launch_bump(1);
    if (!fill) return c32_launch_t<false, 8, 8>(st, a, grid, smem);
    if (a.row_cap <= 256) {
        if (a.n_warps > 20) return c32_launch_t<true, 24, 8>(st, a, grid, smem);
        if (a.n_warps > 16) return c32_launch_t<true, 20, 8>(st, a, grid, smem);
        return c32_launch_t<true, 16, 8>(st, a, grid, smem);
    }
    if (a.n_warps > 16) return cudaErrorInvalidValue;
    return c32_launch_t<true, 16, 16>(st, a, grid, smem);
}
