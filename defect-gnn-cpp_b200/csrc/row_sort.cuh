// row_sort.cuh -- per-warp row buffer in shared memory and the in-warp canonical row sort shared by both traversals
// (cell-list path in dgnn_kernels.cu, direct path in dgnn_direct.cu).
//
// Replaces the reference's per-row std::sort + resize (src/graph/neighbor_list.cpp:56-62).  Row order is the canonical
// key (distance, dst, n_a, n_b, n_c) -- the reference's std::sort is unstable, so its tie order is unspecified (SURVEY
// Appendix A, R8).
//
// Sort = counting sort on a monotone bucket of the distance ((d/r_c)^3: equal-volume shells) followed by a rank inside
// the bucket.  The in-bucket rank compares 32-bit fixed-point keys (d * 2^32 / r_c, resolution 2e-9 A) that were
// scattered into bucket order, so the inner loop is LDS.32 + two integer compares on contiguous words; only entries
// whose 32-bit keys collide (exact ties of symmetric crystals, or |delta d| < 2e-9 A) take the exact path that compares
// the fp64 distances and then (dst, image).  Both the bucket and the key are monotone non-decreasing functions of the
// fp64 distance, so the result is exactly the canonical order.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "dgnn_internal.h"
#include "rbf_math.cuh"

#ifndef FULL_MASK
#define FULL_MASK 0xffffffffu
#endif

__host__ __device__ inline size_t dg_align16(size_t v) { return (v + 15) & ~(size_t)15; }

struct RowBuf {
    double* rd;           // d2 while collecting, distance after the sort prologue            (by entry)
    int* rj;              // neighbour atom (local index)                                       (by entry)
    uint32_t* rn;         // packed image                                                       (by entry)
    uint32_t* bs;         // bucket | slot << 16                                                (by entry)
    uint32_t* sk;         // 32-bit distance keys                                               (bucket order)
    unsigned short* ord;  // entry ids                                                          (bucket order)
    unsigned short* sb;   // buckets                                                            (bucket order)
    float* fd;            // keys by entry during the prologue (as uint32); then the distances in final row order
                          // (fp32), input of the fused RBF epilogue
    int* cnt;             // bucket counters / starts, padded: index(b) = b + b / per  (+ 2 words for the tie check)
};

// counters live at b + (b >> log2(per)), per = n_buckets / 32: every lane's segment starts on its own bank
__host__ __device__ inline size_t dg_cnt_words(int n_buckets) { return (size_t)n_buckets + 32 + 4; }

__host__ __device__ inline size_t dg_rowbuf_bytes(int row_cap, int n_buckets) {
    return dg_align16((size_t)row_cap * 8) + dg_align16((size_t)row_cap * 4) * 5 + dg_align16((size_t)row_cap * 2) * 2 +
           dg_align16(dg_cnt_words(n_buckets) * 4);
}

__device__ __forceinline__ RowBuf dg_rowbuf_carve(unsigned char* p, int row_cap) {
    RowBuf r;
    r.rd = (double*)p; p += dg_align16((size_t)row_cap * 8);
    r.rj = (int*)p; p += dg_align16((size_t)row_cap * 4);
    r.rn = (uint32_t*)p; p += dg_align16((size_t)row_cap * 4);
    r.bs = (uint32_t*)p; p += dg_align16((size_t)row_cap * 4);
    r.sk = (uint32_t*)p; p += dg_align16((size_t)row_cap * 4);
    r.fd = (float*)p; p += dg_align16((size_t)row_cap * 4);
    r.ord = (unsigned short*)p; p += dg_align16((size_t)row_cap * 2);
    r.sb = (unsigned short*)p; p += dg_align16((size_t)row_cap * 2);
    r.cnt = (int*)p;
    return r;
}

// canonical tie-break key of an entry: (dst, n_a, n_b, n_c)
__device__ __forceinline__ unsigned long long dg_tie_key(const RowBuf& sm, int e) {
    return ((unsigned long long)(uint32_t)sm.rj[e] << 32) | dg_image_sort_key(sm.rn[e]);
}

// Fused Gaussian RBF epilogue (edge_features.cpp:7-24 applied per edge as crystal_graph.cpp:32-40 does): the row's
// keep x K fp32 features are one contiguous span of the E x K tensor starting at float4 index o * K4.
//  - K4 groups fit a warp (G = 32 / K4 edges per step, >= 28 lanes busy): every lane owns one fixed float4 column, so
//    the bin centres are loop constants and a warp store covers G whole feature rows (G * K * 4 contiguous bytes);
//  - otherwise flat float4 indexing (every warp store covers 512 contiguous bytes).
__device__ __forceinline__ void dg_rbf_row(const float* __restrict__ fd, int keep, long long o, int lane,
                                           const DgRbfEpi& rbf) {
    float4* __restrict__ o4 = rbf.out + o * (long long)rbf.K4;
    const int G = 32 / rbf.K4;  // 0 when K4 > 32
    if (G * rbf.K4 >= 28) {
        const int es = lane / rbf.K4, k4 = lane - es * rbf.K4;
        if (es < G) {
            const float c0 = (float)(4 * k4) * rbf.dr, c1 = (float)(4 * k4 + 1) * rbf.dr;
            const float c2 = (float)(4 * k4 + 2) * rbf.dr, c3 = (float)(4 * k4 + 3) * rbf.dr;
            float4* __restrict__ p = o4 + es * rbf.K4 + k4;
            const int step = G * rbf.K4;
#pragma unroll 2
            for (int e = es; e < keep; e += G, p += step) {
                const float d = fd[e];
                const float x0 = c0 - d, x1 = c1 - d, x2 = c2 - d, x3 = c3 - d;
                float4 v;
                v.x = dg_ex2(fmaf(rbf.coef * x0, x0, rbf.lognorm));
                v.y = dg_ex2(fmaf(rbf.coef * x1, x1, rbf.lognorm));
                v.z = dg_ex2(fmaf(rbf.coef * x2, x2, rbf.lognorm));
                v.w = dg_ex2(fmaf(rbf.coef * x3, x3, rbf.lognorm));
                __stcs(p, v);
            }
        }
    } else {
        const int nq = keep * rbf.K4;
        // (edge, bin group) of float4 q = lane, lane + 32, ... walked incrementally: no division in the loop
        const int de = 32 / rbf.K4, dk = 32 - de * rbf.K4;
        int e = lane / rbf.K4, k4 = lane - e * rbf.K4;
#pragma unroll 4
        for (int q = lane; q < nq; q += 32) {
            __stcs(o4 + q, dg_rbf4(fd[e], k4 << 2, rbf));
            k4 += dk;
            e += de;
            if (k4 >= rbf.K4) {
                k4 -= rbf.K4;
                ++e;
            }
        }
    }
}

// d * key_scale rounded to an integer in [0, 2^32): monotone non-decreasing in d (one rounding in the product, one in
// the magic-number add).  key_scale = (2^32 - 4096) / r_c and d <= r_c (1 + 2^-52), so the value stays below 2^32.
#define DG_KEY_MAGIC 4503599627370496.0  // 2^52
#ifndef DG_BITONIC_MAX
#define DG_BITONIC_MAX 512  // rows up to this many entries try the register bitonic sort first
#endif
__device__ __forceinline__ uint32_t dg_dist_key(double d, double key_scale) {
    return (uint32_t)__double2loint(__dadd_rn(__dmul_rn(d, key_scale), DG_KEY_MAGIC));
}

// ---------------------------------------------------------------------------------------------------------------------
// Fast path: bitonic sort in registers.  Entry e of the row becomes the word (key & ~mask) | e, key = 32-bit distance
// key, mask = 32 * EPL - 1 (the entry id needs log2(32 EPL) bits); EPL words per lane, 32 * EPL >= m, padding
// 0xffffffff.  The network is lane-major (position p = lane * EPL + r): strides below EPL are compare-exchanges between
// a lane's own registers, larger strides one shuffle + one min/max per word.  Afterwards neighbouring positions whose
// truncated keys are equal (exact ties of symmetric cells, or |delta d| < r_c 2^-(32 - log2(32 EPL))) cannot be ordered
// by the word alone: the row then takes the counting-sort path below, which resolves them exactly.  Returns false in
// that case (warp-uniform); rd[] already holds the distances either way.
// ---------------------------------------------------------------------------------------------------------------------
template <int EPL>
__device__ __forceinline__ bool dg_row_sort_bitonic(const RowBuf& sm, int m, int keep, double key_scale, long long o,
                                                    int* __restrict__ dst, uint32_t* __restrict__ image,
                                                    double* __restrict__ dist, int lane) {
    constexpr uint32_t IDMASK = 32u * EPL - 1u;
    uint32_t v[EPL];
#pragma unroll
    for (int t = 0; t < EPL; ++t) {
        v[t] = 0xffffffffu;
        if (32 * t < m) {  // warp-uniform
            const int e = lane + 32 * t;
            if (e < m) {
                const double d = __dsqrt_rn(sm.rd[e]);  // Neighbor::distance = sqrt(match.second), neighbor_list.cpp:53
                sm.rd[e] = d;
                v[t] = (dg_dist_key(d, key_scale) & ~IDMASK) | (uint32_t)e;
            }
        }
    }
    // all-ascending bitonic network: the first step of every merge pairs position i with its mirror i ^ (k - 1), the
    // remaining steps with i ^ j; the smaller word always goes to the lower position, so no direction bookkeeping
#pragma unroll
    for (int k = 2; k <= 32 * EPL; k <<= 1) {
        if (k <= EPL) {
#pragma unroll
            for (int r = 0; r < EPL; ++r) {
                const int q = r ^ (k - 1);
                if (r < q) {
                    const uint32_t lo = min(v[r], v[q]), hi = max(v[r], v[q]);
                    v[r] = lo;
                    v[q] = hi;
                }
            }
        } else {  // mirror across lanes: partner lane = lane ^ (k / EPL - 1), partner register = EPL - 1 - r
            const int lm = k / EPL - 1;
            const bool lower = (lane & (k / (2 * EPL))) == 0;
            uint32_t w[EPL];
#pragma unroll
            for (int r = 0; r < EPL; ++r) w[r] = __shfl_xor_sync(FULL_MASK, v[EPL - 1 - r], lm);
#pragma unroll
            for (int r = 0; r < EPL; ++r) v[r] = lower ? min(v[r], w[r]) : max(v[r], w[r]);
        }
#pragma unroll
        for (int j = k >> 2; j > 0; j >>= 1) {
            if (j >= EPL) {
                const int lj = j / EPL;
                const bool lower = (lane & lj) == 0;
#pragma unroll
                for (int r = 0; r < EPL; ++r) {
                    const uint32_t w = __shfl_xor_sync(FULL_MASK, v[r], lj);
                    v[r] = lower ? min(v[r], w) : max(v[r], w);
                }
            } else {
#pragma unroll
                for (int r = 0; r < EPL; ++r) {
                    if ((r & j) == 0) {
                        const uint32_t lo = min(v[r], v[r | j]), hi = max(v[r], v[r | j]);
                        v[r] = lo;
                        v[r | j] = hi;
                    }
                }
            }
        }
    }
    // ties the truncated key cannot order?  (positions >= m hold padding)
    bool clash = false;
    const uint32_t nxt = __shfl_down_sync(FULL_MASK, v[0], 1);
#pragma unroll
    for (int r = 0; r < EPL; ++r) {
        const uint32_t w = r + 1 < EPL ? v[r + 1] : nxt;
        const int p1 = lane * EPL + r + 1;
        clash |= p1 < m && ((v[r] ^ w) & ~IDMASK) == 0u;
    }
    if (__any_sync(FULL_MASK, clash)) return false;
    // transpose through shared memory so that the CSR stores are coalesced
#pragma unroll
    for (int r = 0; r < EPL; ++r) sm.ord[lane * EPL + r] = (unsigned short)(v[r] & IDMASK);
    __syncwarp();
    for (int p = lane; p < keep; p += 32) {
        const int e = sm.ord[p];
        const double d = sm.rd[e];
        dst[o + p] = sm.rj[e];
        image[o + p] = sm.rn[e];
        dist[o + p] = d;
        sm.fd[p] = (float)d;
    }
    return true;
}

// Sort the m collected entries of the warp's row buffer and write the first `keep` to the CSR arrays at offset o.
// NBk (number of buckets) is a power of two, 32 <= NBk <= 2048.  All 32 lanes must call.
// rbf.K4 > 0: fused Gaussian RBF epilogue straight from the sorted distances in shared memory.
__device__ __forceinline__ void dg_row_sort_emit(const RowBuf& sm, int m, int keep, int NBk, float inv_rc,
                                                 double key_scale, long long o, int* __restrict__ dst,
                                                 uint32_t* __restrict__ image, double* __restrict__ dist,
                                                 DgStats* __restrict__ stats, int lane, const DgRbfEpi& rbf) {
    bool have_dist = false;  // rd[] holds distances (the fast path converted them before giving up)
    if (m <= 32 * 16 && m <= DG_BITONIC_MAX) {
        bool done;
        if (m <= 128) done = dg_row_sort_bitonic<4>(sm, m, keep, key_scale, o, dst, image, dist, lane);
        else if (m <= 256) done = dg_row_sort_bitonic<8>(sm, m, keep, key_scale, o, dst, image, dist, lane);
        else done = dg_row_sort_bitonic<16>(sm, m, keep, key_scale, o, dst, image, dist, lane);
        if (done) {
            __syncwarp();
            if (rbf.K4 > 0) {
                dg_rbf_row(sm.fd, keep, o, lane, rbf);
                __syncwarp();
            }
            return;
        }
        have_dist = true;
        __syncwarp();
    }
    const int per = NBk >> 5;
    const int lper = 31 - __clz(per);
    const int nwords = NBk + 32 + 1;
    uint32_t* kk = reinterpret_cast<uint32_t*>(sm.fd);  // keys by entry (dead before fd is written)
    for (int b = lane; b < nwords; b += 32) sm.cnt[b] = 0;
    __syncwarp();
    // ---- A: distance, key, bucket histogram ----
    const float scale = (float)NBk;
    for (int e = lane; e < m; e += 32) {
        // Neighbor::distance = sqrt(match.second), neighbor_list.cpp:53
        const double d = have_dist ? sm.rd[e] : __dsqrt_rn(sm.rd[e]);
        sm.rd[e] = d;
        kk[e] = dg_dist_key(d, key_scale);
        const float t = (float)d * inv_rc;
        int b = (int)(t * t * t * scale);
        b = b < 0 ? 0 : (b >= NBk ? NBk - 1 : b);
        const int slot = atomicAdd(&sm.cnt[b + (b >> lper)], 1);
        sm.bs[e] = (uint32_t)b | ((uint32_t)slot << 16);
    }
    __syncwarp();
    {   // ---- B: exclusive scan of the counters in place (lane owns buckets [lane*per, (lane+1)*per)); index(NBk) = m
        const int base = lane * per + lane;
        int sum = 0;
        for (int i = 0; i < per; ++i) sum += sm.cnt[base + i];
        int inc = sum;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int t = __shfl_up_sync(FULL_MASK, inc, off);
            if (lane >= off) inc += t;
        }
        int run = inc - sum;
        for (int i = 0; i < per; ++i) {
            const int c = sm.cnt[base + i];
            sm.cnt[base + i] = run;
            run += c;
        }
        if (lane == 31) sm.cnt[NBk + 32] = run;  // index(NBk) = NBk + (NBk >> lper)
    }
    __syncwarp();
    // ---- C: scatter keys / entry ids / buckets into bucket order ----
    for (int e = lane; e < m; e += 32) {
        const uint32_t v = sm.bs[e];
        const int b = v & 0xffffu;
        const int p = sm.cnt[b + (b >> lper)] + (int)(v >> 16);
        sm.sk[p] = kk[e];
        sm.ord[p] = (unsigned short)e;
        sm.sb[p] = (unsigned short)b;
    }
    __syncwarp();
    // ---- D: rank inside the bucket, emit ----
    int* tie = sm.cnt + NBk + 33;  // two words: entry ids ranked keep-1 and keep
    for (int p = lane; p < m; p += 32) {
        const uint32_t kp = sm.sk[p];
        const int e = sm.ord[p];
        const int b = sm.sb[p], b1 = b + 1;
        const int q0 = sm.cnt[b + (b >> lper)], q1 = sm.cnt[b1 + (b1 >> lper)];
        int lt = 0, eq = 0;
#pragma unroll 4
        for (int q = q0; q < q1; ++q) {
            const uint32_t kq = sm.sk[q];
            lt += kq < kp ? 1 : 0;
            eq += kq == kp ? 1 : 0;
        }
        int rank = q0 + lt;
        const double de = sm.rd[e];
        if (eq > 1) {  // rare: 32-bit keys collide -> exact order on (distance, dst, n_a, n_b, n_c)
            const unsigned long long ke = dg_tie_key(sm, e);
            for (int q = q0; q < q1; ++q) {
                if (q == p || sm.sk[q] != kp) continue;
                const int f = sm.ord[q];
                const double df = sm.rd[f];
                if (df < de || (df == de && dg_tie_key(sm, f) < ke)) ++rank;
            }
        }
        if (rank < keep) {  // scattered inside the row's contiguous range; L2 merges the sectors
            dst[o + rank] = sm.rj[e];
            image[o + rank] = sm.rn[e];
            dist[o + rank] = de;
        }
        if (m > keep && (rank == keep - 1 || rank == keep)) tie[rank - keep + 1] = e;
        // fd aliases the by-entry keys, which phase C consumed; sk / ord / sb stay live for the other lanes
        if (rank < keep) sm.fd[rank] = (float)de;
    }
    if (m > keep && keep > 0) {
        __syncwarp();
        if (lane == 0 && sm.rd[tie[0]] == sm.rd[tie[1]]) atomicAdd(&stats->n_tie_rows, 1);
    }
    __syncwarp();
    if (rbf.K4 > 0) {
        dg_rbf_row(sm.fd, keep, o, lane, rbf);
        __syncwarp();
    }
}
