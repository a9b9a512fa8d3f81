// row_sort.cuh -- per-warp row buffer in shared memory and the in-warp canonical row sort shared by both traversals
// (cell-list path in dgnn_kernels.cu, direct path in dgnn_direct.cu).
//
// Replaces the reference's per-row std::sort + resize (src/graph/neighbor_list.cpp:56-62).  Row order is the canonical
// key (distance, dst, n_a, n_b, n_c) -- the reference's std::sort is unstable, so its tie order is unspecified (SURVEY
// Appendix A, R8).  The sort is a monotone bucket of the distance (d/r_c)^3 -> histogram -> exclusive scan -> rank
// inside the bucket; neighbour shells of crystals put ~1 entry in a bucket, so the quadratic in-bucket step is tiny.
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
    double* rd;           // d2 while collecting, distance after the sort prologue
    int* rj;              // neighbour atom (local index)
    uint32_t* rn;         // packed image
    uint32_t* bs;         // bucket | slot << 16
    unsigned short* ord;  // bucket-ordered entry ids
    float* fd;            // distances in final row order (fp32), input of the fused RBF epilogue
    int* cnt;             // bucket counters / starts, padded: index(b) = b + b / per  (+ 2 words for the tie check)
};

// counters live at b + (b >> log2(per)), per = n_buckets / 32: every lane's segment starts on its own bank
__host__ __device__ inline size_t dg_cnt_words(int n_buckets) { return (size_t)n_buckets + 32 + 4; }

__host__ __device__ inline size_t dg_rowbuf_bytes(int row_cap, int n_buckets) {
    return dg_align16((size_t)row_cap * 8) + dg_align16((size_t)row_cap * 4) * 4 + dg_align16((size_t)row_cap * 2) +
           dg_align16(dg_cnt_words(n_buckets) * 4);
}

__device__ __forceinline__ RowBuf dg_rowbuf_carve(unsigned char* p, int row_cap) {
    RowBuf r;
    r.rd = (double*)p; p += dg_align16((size_t)row_cap * 8);
    r.rj = (int*)p; p += dg_align16((size_t)row_cap * 4);
    r.rn = (uint32_t*)p; p += dg_align16((size_t)row_cap * 4);
    r.bs = (uint32_t*)p; p += dg_align16((size_t)row_cap * 4);
    r.ord = (unsigned short*)p; p += dg_align16((size_t)row_cap * 2);
    r.fd = (float*)p; p += dg_align16((size_t)row_cap * 4);
    r.cnt = (int*)p;
    return r;
}

// Sort the m collected entries of the warp's row buffer and write the first `keep` to the CSR arrays at offset o.
// NBk (number of buckets) is a power of two, 32 <= NBk <= 2048.  All 32 lanes must call.
// rbf.K4 > 0: fused Gaussian RBF epilogue (edge_features.cpp:7-24 applied per edge as crystal_graph.cpp:32-40 does):
// the row's keep x K fp32 features are one contiguous span of the E x K tensor, written with flat float4 indexing
// (every warp store covers 512 contiguous bytes) straight from the sorted distances in shared memory.
__device__ __forceinline__ void dg_row_sort_emit(const RowBuf& sm, int m, int keep, int NBk, float inv_rc,
                                                 long long o, int* __restrict__ dst, uint32_t* __restrict__ image,
                                                 double* __restrict__ dist, DgStats* __restrict__ stats, int lane,
                                                 const DgRbfEpi& rbf) {
    const int per = NBk >> 5;
    const int lper = 31 - __clz(per);
    const int nwords = NBk + 32 + 1;
    for (int b = lane; b < nwords; b += 32) sm.cnt[b] = 0;
    __syncwarp();
    const float scale = (float)NBk;
    for (int e = lane; e < m; e += 32) {
        const double d = __dsqrt_rn(sm.rd[e]);  // Neighbor::distance = sqrt(match.second), neighbor_list.cpp:53
        sm.rd[e] = d;
        const float t = (float)d * inv_rc;
        int b = (int)(t * t * t * scale);
        b = b < 0 ? 0 : (b >= NBk ? NBk - 1 : b);
        const int slot = atomicAdd(&sm.cnt[b + (b >> lper)], 1);
        sm.bs[e] = (uint32_t)b | ((uint32_t)slot << 16);
    }
    __syncwarp();
    {   // exclusive scan of the counters in place (lane owns buckets [lane*per, (lane+1)*per)); index(NBk) = m
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
    for (int e = lane; e < m; e += 32) {
        const uint32_t v = sm.bs[e];
        const int b = v & 0xffffu;
        sm.ord[sm.cnt[b + (b >> lper)] + (v >> 16)] = (unsigned short)e;
    }
    __syncwarp();
    int* tie = sm.cnt + NBk + 33;  // two words: entry ids ranked keep-1 and keep
    for (int e = lane; e < m; e += 32) {
        const int b = sm.bs[e] & 0xffffu, b1 = b + 1;
        const int q0 = sm.cnt[b + (b >> lper)], q1 = sm.cnt[b1 + (b1 >> lper)];
        const double de = sm.rd[e];
        int rank = q0;
        for (int q = q0; q < q1; ++q) {
            const int f = sm.ord[q];
            const double df = sm.rd[f];
            if (df < de) {
                ++rank;
            } else if (df == de && f != e) {  // rare: exact distance tie -> canonical (dst, n_a, n_b, n_c)
                const unsigned long long ke =
                    ((unsigned long long)(uint32_t)sm.rj[e] << 32) | dg_image_sort_key(sm.rn[e]);
                const unsigned long long kf =
                    ((unsigned long long)(uint32_t)sm.rj[f] << 32) | dg_image_sort_key(sm.rn[f]);
                if (kf < ke) ++rank;
            }
        }
        if (rank < keep) {  // scattered inside the row's contiguous range; L2 merges the sectors
            dst[o + rank] = sm.rj[e];
            image[o + rank] = sm.rn[e];
            dist[o + rank] = de;
            sm.fd[rank] = (float)de;
        }
        if (m > keep && (rank == keep - 1 || rank == keep)) tie[rank - keep + 1] = e;
    }
    if (m > keep && keep > 0) {
        __syncwarp();
        if (lane == 0 && sm.rd[tie[0]] == sm.rd[tie[1]]) atomicAdd(&stats->n_tie_rows, 1);
    }
    __syncwarp();
    if (rbf.K4 > 0) {
        float4* __restrict__ o4 = rbf.out + o * (long long)rbf.K4;
        const int nq = keep * rbf.K4;
        // (edge, bin group) of float4 q = lane, lane + 32, ... walked incrementally: no division in the loop
        const int de = 32 / rbf.K4, dk = 32 - de * rbf.K4;
        int e = lane / rbf.K4, k4 = lane - e * rbf.K4;
#pragma unroll 4
        for (int q = lane; q < nq; q += 32) {
            __stcs(o4 + q, dg_rbf4(sm.fd[e], k4 << 2, rbf));
            k4 += dk;
            e += de;
            if (k4 >= rbf.K4) {
                k4 -= rbf.K4;
                ++e;
            }
        }
        __syncwarp();
    }
}
