// cell_items.cuh -- work items of the cell-list path: (structure, block of source cells).  Shared decode used by the
// planning statistics (dgnn_kernels.cu), the fp64 traversal (k_traverse) and the fp32-tile traversal (dgnn_cell32.cu).
#pragma once

#include <cuda_runtime.h>

#include "dgnn_internal.h"

struct ItemBox {
    int c0[3], c1[3];  // source cell box, halo-local coordinates [c0, c1)
    int t0[3], td[3];  // tile box origin (halo-local) and dims
};

__device__ __forceinline__ long long find_struct(const long long* __restrict__ item_base, long long S,
                                                 long long item) {
    long long lo = 0, hi = S;
    while (hi - lo > 1) {
        const long long mid = (lo + hi) >> 1;
        if (item_base[mid] <= item) lo = mid; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ ItemBox item_box(const DgGeom* G, int local) {
    ItemBox B;
    const int b2 = local % G->nb[2], b1 = (local / G->nb[2]) % G->nb[1], b0 = local / (G->nb[2] * G->nb[1]);
    const int bb[3] = {b0, b1, b2};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int D = G->hd[k] - 2 * G->R[k];
        B.c0[k] = G->R[k] + bb[k] * G->bd[k];
        const int e = B.c0[k] + G->bd[k];
        B.c1[k] = e < G->R[k] + D ? e : G->R[k] + D;
        B.t0[k] = B.c0[k] - G->R[k];
        B.td[k] = (B.c1[k] - B.c0[k]) + 2 * G->R[k];
    }
    return B;
}

