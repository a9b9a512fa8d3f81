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
    unsigned short* fin;  // final rank -> entry id
    int* cnt;             // [n_buckets + 1]
};

__host__ __device__ inline size_t dg_rowbuf_bytes(int row_cap, int n_buckets) {
    return dg_align16((size_t)row_cap * 8) + dg_align16((size_t)row_cap * 4) * 3 + dg_align16((size_t)row_cap * 2) * 2 +
           dg_align16((size_t)(n_buckets + 1) * 4);
}

__device__ __forceinline__ RowBuf dg_rowbuf_carve(unsigned char* p, int row_cap) {
    RowBuf r;
    r.rd = (double*)p; p += dg_align16((size_t)row_cap * 8);
    r.rj = (int*)p; p += dg_align16((size_t)row_cap * 4);
    r.rn = (uint32_t*)p; p += dg_align16((size_t)row_cap * 4);
    r.bs = (uint32_t*)p; p += dg_align16((size_t)row_cap * 4);
    r.ord = (unsigned short*)p; p += dg_align16((size_t)row_cap * 2);
    r.fin = (unsigned short*)p; p += dg_align16((size_t)row_cap * 2);
    r.cnt = (int*)p;
    return r;
}

// Sort the m collected entries of the warp's row buffer and write the first `keep` to the CSR arrays at offset o.
// NBk (number of buckets) is a multiple of 32 and <= 512.  All 32 lanes must call.
__device__ __forceinline__ void dg_row_sort_emit(const RowBuf& sm, int m, int keep, int NBk, float inv_rc,
                                                 long long o, int* __restrict__ dst, uint32_t* __restrict__ image,
                                                 double* __restrict__ dist, DgStats* __restrict__ stats, int lane) {
    for (int b = lane; b <= NBk; b += 32) sm.cnt[b] = 0;
    __syncwarp();
    const float scale = (float)NBk;
    for (int e = lane; e < m; e += 32) {
        const double d = __dsqrt_rn(sm.rd[e]);  // Neighbor::distance = sqrt(match.second), neighbor_list.cpp:53
        sm.rd[e] = d;
        const float t = (float)d * inv_rc;
        int b = (int)(t * t * t * scale);
        b = b < 0 ? 0 : (b >= NBk ? NBk - 1 : b);
        const int slot = atomicAdd(&sm.cnt[b], 1);
        sm.bs[e] = (uint32_t)b | ((uint32_t)slot << 16);
    }
    __syncwarp();
    {   // exclusive scan of cnt[0..NBk) in place, cnt[NBk] = m
        const int per = NBk >> 5;
        int sum = 0;
        for (int i = 0; i < per; ++i) sum += sm.cnt[lane * per + i];
        int inc = sum;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int t = __shfl_up_sync(FULL_MASK, inc, off);
            if (lane >= off) inc += t;
        }
        int run = inc - sum;
        for (int i = 0; i < per; ++i) {
            const int c = sm.cnt[lane * per + i];
            sm.cnt[lane * per + i] = run;
            run += c;
        }
        if (lane == 31) sm.cnt[NBk] = run;
    }
    __syncwarp();
    for (int e = lane; e < m; e += 32) {
        const uint32_t v = sm.bs[e];
        sm.ord[sm.cnt[v & 0xffffu] + (v >> 16)] = (unsigned short)e;
    }
    __syncwarp();
    for (int e = lane; e < m; e += 32) {
        const int b = sm.bs[e] & 0xffffu;
        const int q0 = sm.cnt[b], q1 = sm.cnt[b + 1];
        int rank = q0;
        if (q1 - q0 > 1) {
            const double de = sm.rd[e];
            const unsigned long long ke =
                ((unsigned long long)(uint32_t)sm.rj[e] << 32) | dg_image_sort_key(sm.rn[e]);
            for (int q = q0; q < q1; ++q) {
                const int f = sm.ord[q];
                if (f == e) continue;
                const double df = sm.rd[f];
                if (df < de) {
                    ++rank;
                } else if (df == de) {
                    const unsigned long long kf =
                        ((unsigned long long)(uint32_t)sm.rj[f] << 32) | dg_image_sort_key(sm.rn[f]);
                    if (kf < ke) ++rank;
                }
            }
        }
        sm.fin[rank] = (unsigned short)e;
    }
    __syncwarp();
    for (int r = lane; r < keep; r += 32) {
        const int e = sm.fin[r];
        dst[o + r] = sm.rj[e];
        image[o + r] = sm.rn[e];
        dist[o + r] = sm.rd[e];
    }
    if (m > keep && keep > 0 && lane == 0) {
        if (sm.rd[sm.fin[keep - 1]] == sm.rd[sm.fin[keep]]) atomicAdd(&stats->n_tie_rows, 1);
    }
    __syncwarp();
}
