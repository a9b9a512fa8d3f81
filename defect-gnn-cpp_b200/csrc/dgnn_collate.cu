// dgnn_collate.cu -- consumers of the CSR the build leaves in HBM (the callers on either side of the path):
//
//   f1  PyG-style batch collation, device-resident: node features gathered from the atom-embedding table
//       (crystal_graph.cpp:15-21), edge_index [2, E] int64 with per-structure node offsets (crystal_graph.cpp:32-37 per
//       structure + the planned train::BatchedGraph collation, PROJECT_ROADMAP.md:892-904), batch vector; dtypes of the
//       Python consumer (.reference_code/Defect_GNN/utilities.py:69-71: x float, edge_index long, edge_attr float).
//   f4  the viewer's flat getters (src/viz/wasm_bindings.cpp:206-294): edge sources / targets (int32), distances and
//       displacements (float32) -- CSR expansions served by one kernel instead of four host walks.
//   f2  the Betti caller's per-atom neighbourhood clouds (src/topology/betti_features.cpp:65-75):
//       cloud[0] = position_i, cloud[1 + e] = position_i + displacement_e, fp64.
//
// All three are HBM-bound streaming kernels: one warp per CSR row, lanes over the row's edges, coalesced stores.
// Compile with -fmad=false: displacements follow the reference's FMA-free operand order.
#include <cstdint>
#include <cuda_runtime.h>
#include <stdint.h>

#include "dgnn_internal.h"

static thread_local int g_collate_launches = 0;
int dg_collate_launch_count_reset() {
    int v = g_collate_launches;
    g_collate_launches = 0;
    return v;
}

#define DG_COLLATE_THREADS 256

// ---- f1: edge_index [2, E] int64 (row 0 = global source node, row 1 = global target node) ---------------------------
// One warp per row; each lane stores two edges as one 16-byte streaming store (512 B per warp instruction), pairs
// aligned separately for the two output rows (row 1 starts at element E, whose parity shifts the 16-byte grid).
__device__ __forceinline__ void dg_stcs_ll2(long long* p, long long v0, long long v1) {
    asm volatile("st.global.cs.v2.s64 [%0], {%1, %2};" ::"l"(p), "l"(v0), "l"(v1) : "memory");
}
__global__ void __launch_bounds__(DG_COLLATE_THREADS) k_collate_edges(long long N, long long E,
                                                                      const long long* __restrict__ row_ptr,
                                                                      const int* __restrict__ astruct,
                                                                      const long long* __restrict__ atom_offsets,
                                                                      const int* __restrict__ dst,
                                                                      long long* __restrict__ edge_index) {
    const int lane = threadIdx.x & 31;
    const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    long long* __restrict__ out_src = edge_index;
    long long* __restrict__ out_dst = edge_index + E;
    const long long par = E & 1;  // out_dst + e is 16-byte aligned when (e + E) is even (the allocation is)
    // the row header (row_ptr pair, structure base: three dependent loads) of the NEXT row is fetched before this row's
    // stores are issued, so the latency chain overlaps the store stream
    long long n0 = 0, n1 = 0, nbase = 0;
    if (warp0 < N) {
        n0 = row_ptr[warp0], n1 = row_ptr[warp0 + 1];
        nbase = atom_offsets[astruct[warp0]];
    }
    for (long long a = warp0; a < N; a += nwarps) {
        const long long e0 = n0, e1 = n1, base = nbase;
        const long long an = a + nwarps;
        if (an < N) {
            n0 = row_ptr[an], n1 = row_ptr[an + 1];
            nbase = atom_offsets[astruct[an]];
        }
        if (e1 <= e0) continue;
        // sources: head / aligned pairs / tail
        {
            const long long b0 = e0 + (e0 & 1), b1 = e1 & ~1LL;
            if (lane == 0 && b0 != e0) __stcs(out_src + e0, a);
            if (lane == 1 && b1 != e1 && b1 >= b0) __stcs(out_src + b1, a);
            for (long long e = b0 + 2 * lane; e < b1; e += 64) dg_stcs_ll2(out_src + e, a, a);
        }
        // targets
        {
            const long long b0 = e0 + ((e0 + par) & 1), b1 = e1 - ((e1 + par) & 1);
            if (lane == 0 && b0 != e0) __stcs(out_dst + e0, base + (long long)dst[e0]);
            if (lane == 1 && b1 != e1 && b1 >= b0) __stcs(out_dst + b1, base + (long long)dst[b1]);
            for (long long e = b0 + 2 * lane; e < b1; e += 64)
                dg_stcs_ll2(out_dst + e, base + (long long)__ldcs(dst + e), base + (long long)__ldcs(dst + e + 1));
        }
    }
}

// ---- f1: x [N, F] fp32 gathered from the embedding table, batch [N] int64 -------------------------------------------
// lut[key] = row of the table or -1 (key absent: std::map::at would throw, crystal_graph.cpp:20)
__global__ void __launch_bounds__(DG_COLLATE_THREADS) k_collate_nodes(long long N, int F,
                                                                      const int* __restrict__ node_keys,
                                                                      const int* __restrict__ lut, int lut_size,
                                                                      const float* __restrict__ table,
                                                                      const int* __restrict__ astruct,
                                                                      float* __restrict__ x,
                                                                      long long* __restrict__ batch,
                                                                      int* __restrict__ err) {
    const long long total = N * (long long)F;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const long long a = i / F;
        const int c = (int)(i - a * F);
        const int key = node_keys[a];
        const int row = (key >= 0 && key < lut_size) ? lut[key] : -1;
        if (row < 0) {
            if (c == 0) atomicCAS(err, -1, (int)(a < 0x7fffffffLL ? a : 0x7fffffffLL));
            continue;
        }
        x[i] = table[(long long)row * F + c];
        if (c == 0 && batch) batch[a] = (long long)astruct[a];
    }
}
// F % 4 == 0 and N * F / 4 < 2^31: one float4 per thread, 32-bit index arithmetic (the scalar kernel pays a 64-bit
// division per element)
__global__ void __launch_bounds__(DG_COLLATE_THREADS) k_collate_nodes_v4(unsigned N, unsigned F4,
                                                                         const int* __restrict__ node_keys,
                                                                         const int* __restrict__ lut, int lut_size,
                                                                         const float4* __restrict__ table,
                                                                         const int* __restrict__ astruct,
                                                                         float4* __restrict__ x,
                                                                         long long* __restrict__ batch,
                                                                         int* __restrict__ err) {
    const unsigned total = N * F4;
    const unsigned stride = gridDim.x * blockDim.x;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const unsigned a = i / F4;
        const unsigned c = i - a * F4;
        const int key = __ldg(node_keys + a);
        const int row = (key >= 0 && key < lut_size) ? __ldg(lut + key) : -1;
        if (row < 0) {
            if (c == 0) atomicCAS(err, -1, (int)(a < 0x7fffffffu ? a : 0x7fffffffu));
            continue;
        }
        __stcs(x + i, __ldg(table + (size_t)row * F4 + c));
        if (c == 0 && batch) batch[a] = (long long)astruct[a];
    }
}
__global__ void __launch_bounds__(DG_COLLATE_THREADS) k_collate_batch(long long N, const int* __restrict__ astruct,
                                                                      long long* __restrict__ batch) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long a = (long long)blockIdx.x * blockDim.x + threadIdx.x; a < N; a += stride)
        batch[a] = (long long)astruct[a];
}

// ---- f4 / f2: per-edge expansion with the reference's displacement ---------------------------------------------------
// displacement = cloud_pos - query_pos, cloud_pos = pos_j + ((n_a*row0 + n_b*row1) + n_c*row2)   neighbor_list.cpp:82-87,51
__device__ __forceinline__ void dg_edge_displacement(const DgSmall* __restrict__ G, const double* __restrict__ x,
                                                     const double* __restrict__ y, const double* __restrict__ z,
                                                     long long ai, long long aj, uint32_t packed, double& dx,
                                                     double& dy, double& dz) {
    const int na = (int)(int8_t)(packed & 0xffu), nb = (int)(int8_t)((packed >> 8) & 0xffu),
              nc = (int)(int8_t)((packed >> 16) & 0xffu);
    dx = DG_SUB(DG_ADD(x[aj], dg_image_offset(G->L, 0, na, nb, nc)), x[ai]);
    dy = DG_SUB(DG_ADD(y[aj], dg_image_offset(G->L, 1, na, nb, nc)), y[ai]);
    dz = DG_SUB(DG_ADD(z[aj], dg_image_offset(G->L, 2, na, nb, nc)), z[ai]);
}

__global__ void __launch_bounds__(DG_COLLATE_THREADS) k_edge_views(long long N, const long long* __restrict__ row_ptr,
                                                                   const int* __restrict__ astruct,
                                                                   const DgSmall* __restrict__ small,
                                                                   const double* __restrict__ x,
                                                                   const double* __restrict__ y,
                                                                   const double* __restrict__ z,
                                                                   const int* __restrict__ dst,
                                                                   const uint32_t* __restrict__ image,
                                                                   const double* __restrict__ dist,
                                                                   int* __restrict__ sources, int* __restrict__ targets,
                                                                   float* __restrict__ distances,
                                                                   float* __restrict__ displacements) {
    const int lane = threadIdx.x & 31;
    const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long a = warp0; a < N; a += nwarps) {
        const long long e0 = row_ptr[a], e1 = row_ptr[a + 1];
        const DgSmall* __restrict__ G = small + astruct[a];
        const long long base = G->atom_base;
        const int iloc = (int)(a - base);
        for (long long e = e0 + lane; e < e1; e += 32) {
            const int j = dst[e];
            if (sources) sources[e] = iloc;                    // static_cast<int>(i)
            if (targets) targets[e] = j;                       // static_cast<int>(neighbor.idx)
            if (distances) distances[e] = (float)dist[e];      // static_cast<float>(neighbor.distance)
            if (displacements) {
                double dx, dy, dz;
                dg_edge_displacement(G, x, y, z, a, base + j, image[e], dx, dy, dz);
                displacements[3 * e + 0] = (float)dx;
                displacements[3 * e + 1] = (float)dy;
                displacements[3 * e + 2] = (float)dz;
            }
        }
    }
}

// f2: clouds of the atoms [a0, a1): atom a owns rows [cloud_ptr(a), cloud_ptr(a) + deg + 1) of the (x, y, z) fp64 array,
// cloud_ptr(a) = (row_ptr[a] - row_ptr[a0]) + (a - a0)
__global__ void __launch_bounds__(DG_COLLATE_THREADS) k_neighborhood_clouds(long long a0, long long a1,
                                                                            const long long* __restrict__ row_ptr,
                                                                            const int* __restrict__ astruct,
                                                                            const DgSmall* __restrict__ small,
                                                                            const double* __restrict__ x,
                                                                            const double* __restrict__ y,
                                                                            const double* __restrict__ z,
                                                                            const int* __restrict__ dst,
                                                                            const uint32_t* __restrict__ image,
                                                                            double* __restrict__ cloud,
                                                                            long long* __restrict__ cloud_ptr) {
    const int lane = threadIdx.x & 31;
    const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const long long ebase = row_ptr[a0];
    for (long long a = a0 + warp0; a < a1; a += nwarps) {
        const long long e0 = row_ptr[a], e1 = row_ptr[a + 1];
        const DgSmall* __restrict__ G = small + astruct[a];
        const long long base = G->atom_base;
        const long long p0 = (e0 - ebase) + (a - a0);
        const double xi = x[a], yi = y[a], zi = z[a];
        if (lane == 0) {
            cloud_ptr[a - a0] = p0;
            if (a == a1 - 1) cloud_ptr[a1 - a0] = (e1 - ebase) + (a1 - a0);
            cloud[3 * p0 + 0] = xi;  // point_cloud.row(0) = center.position
            cloud[3 * p0 + 1] = yi;
            cloud[3 * p0 + 2] = zi;
        }
        for (long long e = e0 + lane; e < e1; e += 32) {
            double dx, dy, dz;
            dg_edge_displacement(G, x, y, z, a, base + dst[e], image[e], dx, dy, dz);
            const long long p = p0 + 1 + (e - e0);
            cloud[3 * p + 0] = DG_ADD(xi, dx);  // center.position + neighbors[i].displacement
            cloud[3 * p + 1] = DG_ADD(yi, dy);
            cloud[3 * p + 2] = DG_ADD(zi, dz);
        }
    }
}

// f2: the (deg + 1) x (deg + 1) pairwise-distance tables the Betti caller builds from every cloud
// (src/topology/ripser_wrapper.cpp:60-66, Eigen without FMA):
//   sq_norms = rowwise squaredNorm           s_i  = (x_i x_i + y_i y_i) + z_i z_i
//   sq_dist  = s_i + s_j - 2 P P^T           d2   = (s_i + s_j) - 2 ((x_i x_j + y_i y_j) + z_i z_j)   (GEMM, depth 3 in order)
//   dist     = sq_dist.cwiseMax(0).cwiseSqrt()
// One CTA per atom: the cloud (row 0 = the atom, row 1 + e = position + displacement of neighbour e,
// betti_features.cpp:67-74) and its squared norms are staged in shared memory, a warp owns a column, lanes the rows, so
// every store is a contiguous run.  MODE 0: full fp64 table, column-major (the MatrixXd `dist`); MODE 1: the compressed
// lower triangle in fp32, row i = entries (i, 0..i-1) -- the `distances` vector handed to Ripser (:18-24).
// No tensor cores: DMMA fuses multiply and add (one rounding), the reference rounds both -- the tables would differ in
// the last bit, and the Gram product is 5 of the ~14 fp64 operations per entry (the rest: norms, max, IEEE sqrt).
#define DG_TABLE_THREADS 256
template <int MODE>
__global__ void __launch_bounds__(DG_TABLE_THREADS) k_distance_tables(long long a0, long long a1,
                                                                      const long long* __restrict__ row_ptr,
                                                                      const int* __restrict__ astruct,
                                                                      const DgSmall* __restrict__ small,
                                                                      const double* __restrict__ x,
                                                                      const double* __restrict__ y,
                                                                      const double* __restrict__ z,
                                                                      const int* __restrict__ dst,
                                                                      const uint32_t* __restrict__ image,
                                                                      const long long* __restrict__ table_ptr,
                                                                      void* __restrict__ out, int m_cap) {
    extern __shared__ __align__(16) double tsm[];
    double* px = tsm;
    double* py = px + m_cap;
    double* pz = py + m_cap;
    double* ps = pz + m_cap;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    for (long long a = a0 + blockIdx.x; a < a1; a += gridDim.x) {
        const long long e0 = row_ptr[a], e1 = row_ptr[a + 1];
        const int m = (int)(e1 - e0) + 1;
        const DgSmall* __restrict__ G = small + astruct[a];
        const long long abase = G->atom_base;
        const double xi = x[a], yi = y[a], zi = z[a];
        __syncthreads();  // the previous atom's cloud is dead
        for (int t = tid; t < m; t += blockDim.x) {
            double cx = xi, cy = yi, cz = zi;
            if (t > 0) {
                double dx, dy, dz;
                const long long e = e0 + t - 1;
                dg_edge_displacement(G, x, y, z, a, abase + dst[e], image[e], dx, dy, dz);
                cx = DG_ADD(xi, dx);
                cy = DG_ADD(yi, dy);
                cz = DG_ADD(zi, dz);
            }
            px[t] = cx; py[t] = cy; pz[t] = cz;
            ps[t] = DG_ADD(DG_ADD(DG_MUL(cx, cx), DG_MUL(cy, cy)), DG_MUL(cz, cz));
        }
        __syncthreads();
        const long long base = table_ptr[a - a0];
        if (MODE == 0) {
            double* __restrict__ o = reinterpret_cast<double*>(out) + base;
            for (int j = warp; j < m; j += nwarps) {
                const double xj = px[j], yj = py[j], zj = pz[j], sj = ps[j];
                for (int i = lane; i < m; i += 32) {
                    const double dot = DG_ADD(DG_ADD(DG_MUL(px[i], xj), DG_MUL(py[i], yj)), DG_MUL(pz[i], zj));
                    const double d2 = DG_SUB(DG_ADD(ps[i], sj), DG_MUL(2.0, dot));
                    o[(long long)j * m + i] = __dsqrt_rn(d2 > 0.0 ? d2 : 0.0);
                }
            }
        } else {
            float* __restrict__ o = reinterpret_cast<float*>(out) + base;
            for (int i = 1 + warp; i < m; i += nwarps) {
                const double xr = px[i], yr = py[i], zr = pz[i], sr = ps[i];
                float* __restrict__ orow = o + (long long)i * (i - 1) / 2;
                for (int j = lane; j < i; j += 32) {
                    const double dot = DG_ADD(DG_ADD(DG_MUL(xr, px[j]), DG_MUL(yr, py[j])), DG_MUL(zr, pz[j]));
                    const double d2 = DG_SUB(DG_ADD(sr, ps[j]), DG_MUL(2.0, dot));
                    orow[j] = (float)__dsqrt_rn(d2 > 0.0 ? d2 : 0.0);
                }
            }
        }
    }
}

// element counts of the tables of atoms [a0, a1) (int32: (deg + 1)^2 <= 4097^2)
__global__ void k_table_counts(long long a0, long long a1, const long long* __restrict__ row_ptr, int mode,
                               int* __restrict__ counts) {
    const long long a = a0 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= a1) return;
    const long long m = row_ptr[a + 1] - row_ptr[a] + 1;
    counts[a - a0] = (int)(mode == 0 ? m * m : m * (m - 1) / 2);
}

static unsigned grid_for(long long work_items, int sm_count, int per_block) {
    long long g = (work_items + per_block - 1) / per_block;
    const long long cap = (long long)sm_count * 16;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (unsigned)g;
}

void dg_launch_collate_edges(cudaStream_t st, long long N, long long E, const long long* row_ptr, const int* astruct,
                             const long long* atom_offsets, const int* dst, long long* edge_index, int sm_count) {
    if (N <= 0 || E <= 0) return;
    k_collate_edges<<<grid_for(N, sm_count, DG_COLLATE_THREADS / 32), DG_COLLATE_THREADS, 0, st>>>(
        N, E, row_ptr, astruct, atom_offsets, dst, edge_index);
    ++g_collate_launches;
}

void dg_launch_collate_nodes(cudaStream_t st, long long N, int F, const int* node_keys, const int* lut, int lut_size,
                             const float* table, const int* astruct, float* x, long long* batch, int* err,
                             int sm_count) {
    if (N <= 0) return;
    if (F > 0 && x && F % 4 == 0 && N * (long long)(F / 4) < 0x7fffffffLL && ((uintptr_t)table & 15) == 0 &&
        ((uintptr_t)x & 15) == 0) {
        k_collate_nodes_v4<<<grid_for(N * (long long)(F / 4), sm_count, DG_COLLATE_THREADS), DG_COLLATE_THREADS, 0, st>>>(
            (unsigned)N, (unsigned)(F / 4), node_keys, lut, lut_size, (const float4*)table, astruct, (float4*)x, batch,
            err);
    } else if (F > 0 && x) {
        k_collate_nodes<<<grid_for(N * (long long)F, sm_count, DG_COLLATE_THREADS), DG_COLLATE_THREADS, 0, st>>>(
            N, F, node_keys, lut, lut_size, table, astruct, x, batch, err);
    } else if (batch) {
        k_collate_batch<<<grid_for(N, sm_count, DG_COLLATE_THREADS), DG_COLLATE_THREADS, 0, st>>>(N, astruct, batch);
    }
    ++g_collate_launches;
}

void dg_launch_edge_views(cudaStream_t st, long long N, const long long* row_ptr, const int* astruct,
                          const DgSmall* small, const double* x, const double* y, const double* z, const int* dst,
                          const uint32_t* image, const double* dist, int* sources, int* targets, float* distances,
                          float* displacements, int sm_count) {
    if (N <= 0) return;
    k_edge_views<<<grid_for(N, sm_count, DG_COLLATE_THREADS / 32), DG_COLLATE_THREADS, 0, st>>>(
        N, row_ptr, astruct, small, x, y, z, dst, image, dist, sources, targets, distances, displacements);
    ++g_collate_launches;
}

void dg_launch_neighborhood_clouds(cudaStream_t st, long long a0, long long a1, const long long* row_ptr,
                                   const int* astruct, const DgSmall* small, const double* x, const double* y,
                                   const double* z, const int* dst, const uint32_t* image, double* cloud,
                                   long long* cloud_ptr, int sm_count) {
    if (a1 <= a0) return;
    k_neighborhood_clouds<<<grid_for(a1 - a0, sm_count, DG_COLLATE_THREADS / 32), DG_COLLATE_THREADS, 0, st>>>(
        a0, a1, row_ptr, astruct, small, x, y, z, dst, image, cloud, cloud_ptr);
    ++g_collate_launches;
}

void dg_launch_table_counts(cudaStream_t st, long long a0, long long a1, const long long* row_ptr, int mode, int* counts) {
    if (a1 <= a0) return;
    k_table_counts<<<(unsigned)((a1 - a0 + 255) / 256), 256, 0, st>>>(a0, a1, row_ptr, mode, counts);
    ++g_collate_launches;
}

cudaError_t dg_launch_distance_tables(cudaStream_t st, long long a0, long long a1, const long long* row_ptr,
                                      const int* astruct, const DgSmall* small, const double* x, const double* y,
                                      const double* z, const int* dst, const uint32_t* image, const long long* table_ptr,
                                      void* out, int mode, int max_degree, int sm_count) {
    if (a1 <= a0) return cudaSuccess;
    const int m_cap = max_degree + 1;
    const size_t smem = (size_t)m_cap * 4 * sizeof(double);
    long long grid = a1 - a0;
    if (grid > (long long)sm_count * 16) grid = (long long)sm_count * 16;
    cudaError_t e;
    if (mode == 0) {
        e = dg_raise_dyn_smem<k_distance_tables<0>>();
        if (e != cudaSuccess) return e;
        k_distance_tables<0><<<(unsigned)grid, DG_TABLE_THREADS, smem, st>>>(a0, a1, row_ptr, astruct, small, x, y, z, dst,
                                                                              image, table_ptr, out, m_cap);
    } else {
        e = dg_raise_dyn_smem<k_distance_tables<1>>();
        if (e != cudaSuccess) return e;
        k_distance_tables<1><<<(unsigned)grid, DG_TABLE_THREADS, smem, st>>>(a0, a1, row_ptr, astruct, small, x, y, z, dst,
                                                                              image, table_ptr, out, m_cap);
    }
    ++g_collate_launches;
    return cudaGetLastError();
}
