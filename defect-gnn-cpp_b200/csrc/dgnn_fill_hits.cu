// dgnn_fill_hits.cu -- fill pass from the count pass's hit lists.  Compile with -fmad=false.
#include <cuda_runtime.h>
#include <stdint.h>

#include "dgnn_internal.h"
#include "row_sort.cuh"
#include "direct_sweep.cuh"

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

// Storing threshold of the count pass's hit lists for truncated rows: the d2 below which a structure of N atoms in volume
// V is expected to hold `target` neighbours, r^3 = target V / (N 4/3 pi).  Rows whose neighbourhood is sparser than the
// structure's mean (next to a vacancy cluster, a surface) may fall short; the count pass notices and leaves them to the
// traversal fill (direct_sweep.cuh, HitTally).
__global__ void __launch_bounds__(128) k_hit_thresholds(long long S, const DgSmall* __restrict__ small, double target,
                                                        double r2, double* __restrict__ out) {
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    const DgSmall* G = small + s;
    double thr = DG_INF;
    if (G->status == DG_ST_OK && G->natoms > 0 && G->mode != DG_MODE_CELL) {
        const double* L = G->L;
        const double det = fabs(L[0] * (L[4] * L[8] - L[5] * L[7]) + L[1] * (L[5] * L[6] - L[3] * L[8]) +
                                L[2] * (L[3] * L[7] - L[4] * L[6]));
        const double r3 = target * det / ((double)G->natoms * 4.18879020478639);
        const double t = cbrt(r3 * r3);
        if (t < 0.8 * r2) thr = t;  // (otherwise nearly everything would be stored anyway)
    }
    out[s] = thr;
}
void dg_launch_hit_thresholds(cudaStream_t st, long long S, const DgSmall* small, double target, double r2, double* out) {
    if (S <= 0) return;
    k_hit_thresholds<<<(unsigned)((S + 127) / 128), 128, 0, st>>>(S, small, target, r2, out);
    dg_direct_launch_bump(1);
}

// Warps per CTA: the kernel is bound by latency at the occupancy its shared memory allows (10 KB per warp at 192-entry
// rows: 2 CTAs of 8 warps), so the CTA size is the one that puts the most warps on an SM -- 3 CTAs of 7 warps there.
static int pick_fill_hits_warps(const DgDirectArgs& a, int max_warps) {
    static thread_local size_t cached_bytes = 0;  // (a context is driven by one host thread at a time)
    static thread_local int cached_max = 0, cached_pick = 0;
    const int W = a.hit_cap < a.row_cap ? a.hit_cap : a.row_cap;
    const size_t per_warp = dg_fill_hits_warp_bytes(W);
    if (per_warp == cached_bytes && max_warps == cached_max) return cached_pick;
    int best = max_warps, best_warps = 0;
    for (int n = max_warps; n >= 4 && n >= max_warps - 4; --n) {
        const size_t smem = (size_t)n * per_warp;
        if (dg_raise_dyn_smem<k_fill_hits>() != cudaSuccess) continue;
        int occ = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_fill_hits, n * 32, smem) != cudaSuccess) continue;
        if (occ * n > best_warps) {
            best_warps = occ * n;
            best = n;
        }
    }
    (void)cudaGetLastError();
    cached_bytes = per_warp;
    cached_max = max_warps;
    cached_pick = best;
    return best;
}

cudaError_t dg_launch_fill_hits(cudaStream_t st, const DgDirectArgs& a_in, int sm_count) {
    if (a_in.n_atoms <= 0) return cudaSuccess;
    DgDirectArgs a = a_in;
    if (a.fill_hits_warps > 0) a.n_warps = a.fill_hits_warps < a.n_warps ? a.fill_hits_warps : a.n_warps;
    else a.n_warps = pick_fill_hits_warps(a, a.n_warps);
    const size_t smem = dg_fill_hits_smem_bytes(a);
    const int threads = a.n_warps * 32;
    cudaError_t e = dg_raise_dyn_smem<k_fill_hits>();
    if (e != cudaSuccess) return e;
    int occ = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_fill_hits, threads, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
    const long long want = (a.n_atoms + DG_ROW_CHUNK - 1) / DG_ROW_CHUNK;
    long long grid = (long long)sm_count * occ;
    if (grid > want) grid = want;
    dg_direct_launch_bump(1);
    k_fill_hits<<<(unsigned)grid, threads, smem, st>>>(a);
    return cudaGetLastError();
}
