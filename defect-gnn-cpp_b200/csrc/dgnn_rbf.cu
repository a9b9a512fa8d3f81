// dgnn_rbf.cu -- K3: fused Gaussian RBF expansion (reference src/graph/edge_features.cpp:7-24 applied to every
// edge as in src/graph/crystal_graph.cpp:32-40), writing the E x K edge-feature tensor.
//
//   K = floor(r_c / dr);  sigma = r_c / 3;  g[e,k] = 1/(sigma*sqrt(2*pi)) * exp(-0.5 * (k*dr - d_e)^2 / sigma^2)
//
// fp32 row-major output (the consumer layout, utilities.py:69-71).  Pure HBM-store stream: 8 B read + 4K B written
// per edge.  Work is assigned by flat float4 index over the E x K array so every warp store instruction covers 512
// contiguous bytes regardless of K (rows are K*4 bytes: 160 B at K = 40, not 128 B-line aligned per edge); distances
// of a tile are staged once in shared memory; stores are streaming (st.global.cs, written once, never re-read here).
// Persistent grid: a multiple of the SM count, tiles handed out grid-stride.
#include <cuda_runtime.h>
#include <stdint.h>

#include <cmath>

#include "dgnn_internal.h"
#include "rbf_math.cuh"

static thread_local int g_rbf_launches = 0;
int dg_rbf_launch_count_reset() {
    int v = g_rbf_launches;
    g_rbf_launches = 0;
    return v;
}

#ifndef RBF_THREADS
#define RBF_THREADS 256
#endif
#ifndef RBF_TILE_EDGES
#define RBF_TILE_EDGES 512
#endif
#ifndef RBF_CTAS_PER_SM
#define RBF_CTAS_PER_SM 8
#endif  // edges per tile: 512 * K/4 float4 per tile (5120 at K = 40 -> 20 per thread)

__global__ void __launch_bounds__(RBF_THREADS) k_rbf_f32_vec4(const double* __restrict__ dist, long long n_edges,
                                                              float4* __restrict__ out, DgRbfEpi P) {
    __shared__ float s_d[RBF_TILE_EDGES];
    const long long n_tiles = (n_edges + RBF_TILE_EDGES - 1) / RBF_TILE_EDGES;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long e0 = tile * RBF_TILE_EDGES;
        const int ne = (int)((n_edges - e0) < RBF_TILE_EDGES ? (n_edges - e0) : RBF_TILE_EDGES);
        __syncthreads();
        for (int i = threadIdx.x; i < ne; i += RBF_THREADS) s_d[i] = (float)__ldcs(dist + e0 + i);
        __syncthreads();
        const int nq = ne * P.K4;
        float4* __restrict__ o = out + e0 * P.K4;
#pragma unroll 4
        for (int q = threadIdx.x; q < nq; q += RBF_THREADS) {
            const int e = (int)(((float)q + 0.5f) * P.inv_k4);  // q / K4, exact for q < 2^22
            const int k = (q - e * P.K4) << 2;
            __stcs(o + q, dg_rbf4(s_d[e], k, P));
        }
    }
}

// K not a multiple of 4: rows are not 16-byte aligned, but a tile of RBF_TILE_EDGES (a multiple of 4) edges is -- the
// tile's K * RBF_TILE_EDGES floats are written as float4s that may straddle two edges (one 32-bit division per float4)
__global__ void __launch_bounds__(RBF_THREADS) k_rbf_f32_flat4(const double* __restrict__ dist, long long n_edges,
                                                               unsigned K, float* __restrict__ out, DgRbfEpi P) {
    __shared__ float s_d[RBF_TILE_EDGES];
    const long long n_tiles = (n_edges + RBF_TILE_EDGES - 1) / RBF_TILE_EDGES;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long e0 = tile * RBF_TILE_EDGES;
        const unsigned ne = (unsigned)((n_edges - e0) < RBF_TILE_EDGES ? (n_edges - e0) : RBF_TILE_EDGES);
        __syncthreads();
        for (unsigned i = threadIdx.x; i < ne; i += RBF_THREADS) s_d[i] = (float)__ldcs(dist + e0 + i);
        __syncthreads();
        const unsigned nf = ne * K;  // floats of this tile
        float* __restrict__ o = out + e0 * K;
#pragma unroll 2
        for (unsigned f = threadIdx.x * 4u; f < nf; f += RBF_THREADS * 4u) {
            unsigned e = f / K, k = f - e * K;
            float g[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                g[i] = e < ne ? dg_rbf1(s_d[e], (int)k, P) : 0.0f;
                if (++k == K) k = 0, ++e;
            }
            if (f + 4u <= nf) {
                __stcs(reinterpret_cast<float4*>(o + f), make_float4(g[0], g[1], g[2], g[3]));
            } else {  // last, partial float4 of the last tile
                for (unsigned i = 0; f + i < nf; ++i) o[f + i] = g[i];
            }
        }
    }
}

// fallback (unaligned output pointer, K beyond the 32-bit tile index): scalar stores, flat-indexed and coalesced
__global__ void __launch_bounds__(RBF_THREADS) k_rbf_f32_scalar(const double* __restrict__ dist, long long n_edges,
                                                                int K, float* __restrict__ out, DgRbfEpi P) {
    const long long total = n_edges * K;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += stride) {
        const long long e = q / K;
        const int k = (int)(q - e * K);
        out[q] = dg_rbf1((float)dist[e], k, P);
    }
}

DgRbfEpi dg_make_rbf_epi(int K, double r_cutoff, double dr, float* out) {
    DgRbfEpi P;
    const double sigma = r_cutoff / 3.0;
    P.dr = (float)dr;
    P.coef = (float)(-0.5 / (sigma * sigma) * 1.4426950408889634074);
    P.lognorm = (float)std::log2(1.0 / (sigma * 2.5066282746310005024));
    {
        const double coef = -0.5 / (sigma * sigma) * 1.4426950408889634074;
        P.c2dr = (float)(2.0 * coef * dr);
        P.q = (float)std::exp2(coef * dr * dr);
        P.q2 = (float)std::exp2(2.0 * coef * dr * dr);
        P.recur = 0;
    }
    P.K4 = K / 4;
    P.inv_k4 = P.K4 > 0 ? 1.0f / (float)P.K4 : 0.0f;
    P.out = reinterpret_cast<float4*>(out);
    return P;
}

void dg_launch_rbf_f32(cudaStream_t st, const double* dist, long long n_edges, int K, double r_cutoff, double dr,
                       float* out, int sm_count) {
    if (n_edges <= 0 || K <= 0) return;
    const DgRbfEpi P = dg_make_rbf_epi(K, r_cutoff, dr, out);
    if (sm_count <= 0) sm_count = 148;
    if ((K & 3) == 0 && (long long)RBF_TILE_EDGES * (K / 4) < (1 << 22) && ((uintptr_t)out & 15) == 0) {
        const long long n_tiles = (n_edges + RBF_TILE_EDGES - 1) / RBF_TILE_EDGES;
        long long grid = (long long)sm_count * RBF_CTAS_PER_SM;  // resident CTAs per SM
        if (grid > n_tiles) grid = n_tiles;
        k_rbf_f32_vec4<<<(unsigned)grid, RBF_THREADS, 0, st>>>(dist, n_edges, reinterpret_cast<float4*>(out), P);
    } else if (((uintptr_t)out & 15) == 0 && (long long)RBF_TILE_EDGES * K < (1LL << 31) && (RBF_TILE_EDGES & 3) == 0) {
        const long long n_tiles = (n_edges + RBF_TILE_EDGES - 1) / RBF_TILE_EDGES;
        long long grid = (long long)sm_count * RBF_CTAS_PER_SM;
        if (grid > n_tiles) grid = n_tiles;
        k_rbf_f32_flat4<<<(unsigned)grid, RBF_THREADS, 0, st>>>(dist, n_edges, (unsigned)K, out, P);
    } else {
        const long long total = n_edges * K;
        long long grid = (total + RBF_THREADS - 1) / RBF_THREADS;
        const long long cap = (long long)sm_count * 16;
        if (grid > cap) grid = cap;
        k_rbf_f32_scalar<<<(unsigned)grid, RBF_THREADS, 0, st>>>(dist, n_edges, K, out, P);
    }
    ++g_rbf_launches;
}

// fp64, column-major E x K: the reference's own Eigen::MatrixXd storage (crystal_graph.cpp:30,37), for the legacy
// CrystalGraph::edge_attr() accessor.  Consecutive threads walk consecutive edges of one bin -> coalesced.
__global__ void __launch_bounds__(RBF_THREADS) k_rbf_f64_colmajor(const double* __restrict__ dist, long long n_edges,
                                                                  int K, double dr, double inv_sigma2, double norm,
                                                                  double* __restrict__ out) {
    const long long total = n_edges * K;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += stride) {
        const long long k = q / n_edges;
        const long long e = q - k * n_edges;
        const double x = (double)k * dr - dist[e];
        out[q] = norm * exp(-0.5 * (x * x) * inv_sigma2);
    }
}

void dg_launch_rbf_f64_colmajor(cudaStream_t st, const double* dist, long long n_edges, int K, double r_cutoff,
                                double dr, double* out) {
    if (n_edges <= 0 || K <= 0) return;
    const double sigma = r_cutoff / 3.0;
    const long long total = n_edges * K;
    long long grid = (total + RBF_THREADS - 1) / RBF_THREADS;
    if (grid > 148 * 16) grid = 148 * 16;
    k_rbf_f64_colmajor<<<(unsigned)grid, RBF_THREADS, 0, st>>>(dist, n_edges, K, dr, 1.0 / (sigma * sigma),
                                                               1.0 / (sigma * 2.5066282746310005024), out);
    ++g_rbf_launches;
}
