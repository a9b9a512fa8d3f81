// dgnn_tie_probe.cu -- which tie ranking should the uncapped sweep-fill kernel use?  Compile with -fmad=false.
//
// Rows leave the register sort ordered by a truncated 32-bit distance key; entries that share a key are ranked exactly
// (distance, dst, n_a, n_b, n_c -- the canonical order of the reference's unstable sort, neighbor_list.cpp:57-63).  The
// headline kernel instantiates the compact per-entry ranking (row_sort.cuh, LEAN): free when keys are distinct (thermal
// or relaxed structures: every run has one entry), quadratic in the run length when they are not -- the coordination
// shells of an ideal lattice put 6-48 entries on one key and the compact ranking then costs 3x.
// The warp-cooperative ranking handles those runs at a small constant price per row but costs the headline kernel 6 %.
//
// So the choice is made per batch, from the data: this kernel sweeps DG_PROBE_ROWS sample rows spread over the batch
// (same traversal, same keys as the fill pass) and accumulates  sum over entries of (entries sharing its key)  and the
// entry count into DgStats; the ratio is the mean run length an entry sits in (1.0 without ties).  It runs behind the
// count pass on the same stream and its two numbers come back with the statistics the build reads anyway -- no extra
// synchronisation, ~10 us of device time.
#include <cuda_runtime.h>
#include <stdint.h>

#include "dgnn_internal.h"
#include "row_sort.cuh"
#include "direct_sweep.cuh"

#define DG_PROBE_ROWS 64
#define DG_PROBE_CAP 512  // = DG_BITONIC_MAX: longer rows take the bucket sort, which has one ranking only

template <int EPS>
__global__ void __launch_bounds__(32) k_tie_probe(const DgDirectArgs A) {
    __shared__ double rd[DG_PROBE_CAP];
    __shared__ int rj[DG_PROBE_CAP];
    __shared__ uint32_t rn[DG_PROBE_CAP];
    __shared__ uint32_t key[DG_PROBE_CAP];
    const int lane = threadIdx.x;
    const unsigned lt_mask = (1u << lane) - 1u;
    // sample row: evenly spread over the atoms of the batch
    const long long a = (long long)(((unsigned long long)blockIdx.x * (unsigned long long)A.n_atoms) / gridDim.x);
    if (a >= A.n_atoms) return;
    const DgSmall* __restrict__ G = A.small + A.astruct[a];
    const int mode = G->mode;
    if (mode == DG_MODE_CELL || G->status != DG_ST_OK) return;
    RowBuf rb = RowBuf{rd, rj, rn, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    unsigned near = 0;
    const int m = traverse_row<true, EPS, false, false>(A, G, a, mode, rb, nullptr, nullptr, lane, lt_mask, near);
    if (m <= 1 || m > DG_PROBE_CAP) return;  // warp-uniform
    __syncwarp();
    // the fill pass's truncated key: the low bits of the 32-bit key carry the entry id (row_sort.cuh)
    const uint32_t idmask = m <= 128 ? 127u : (m <= 256 ? 255u : 511u);
    for (int e = lane; e < m; e += 32) key[e] = dg_dist_key(__dsqrt_rn(rd[e]), A.key_scale) & ~idmask;
    __syncwarp();
    unsigned sum = 0;
    for (int e = lane; e < m; e += 32) {
        const uint32_t k = key[e];
        for (int f = 0; f < m; ++f) sum += key[f] == k ? 1u : 0u;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(FULL_MASK, sum, o);
    if (lane == 0) {
        atomicAdd(&A.stats->tie_sum, (unsigned long long)sum);
        atomicAdd(&A.stats->tie_n, (unsigned long long)m);
    }
}

// D as for the fill pass of the same build (row_cap is the probe's own)
cudaError_t dg_launch_tie_probe(cudaStream_t st, const DgDirectArgs& D) {
    if (D.n_atoms <= 0) return cudaSuccess;
    DgDirectArgs P = D;
    P.row_cap = DG_PROBE_CAP;
    P.row_begin = P.row_end = 0;
    const unsigned grid = (unsigned)(D.n_atoms < DG_PROBE_ROWS ? D.n_atoms : DG_PROBE_ROWS);
    switch (P.eps_mode) {
        case 0: k_tie_probe<0><<<grid, 32, 0, st>>>(P); break;
        case 1: k_tie_probe<1><<<grid, 32, 0, st>>>(P); break;
        default: k_tie_probe<2><<<grid, 32, 0, st>>>(P); break;
    }
    dg_direct_launch_bump(1);
    return cudaGetLastError();
}
