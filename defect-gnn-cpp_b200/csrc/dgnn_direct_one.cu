// dgnn_direct_one.cu -- single-pass build of small direct-path batches (the reference's one-structure-per-call API):
// sweep, row offsets by decoupled look-back, sort and emit in one kernel.  Compile with -fmad=false.
#include <cuda_runtime.h>
#include <stdint.h>

#include "dgnn_internal.h"
#include "row_sort.cuh"
#include "direct_sweep.cuh"

// ---------------------------------------------------------------------------------------------------------------------
// Single-pass build: row offsets by decoupled look-back.  Every row owns one 64-bit status word, flag << 62 | value:
// AGG = the row's own (capped) degree, published right after its traversal; INC = the inclusive prefix of all rows up
// to and including it, published once the look-back has resolved.  A row walks back 32 predecessors per step summing
// AGG values until it meets an INC.  Flag and value share the word, so relaxed gpu-scope accesses are enough.  Chunks
// are handed out statically to a grid sized by the occupancy query (every CTA resident), so each predecessor row is
// either finished or being worked on: the spin always ends; a bounded spin count guards the box regardless.
// ---------------------------------------------------------------------------------------------------------------------
#define DG_LB_AGG (1ull << 62)
#define DG_LB_INC (2ull << 62)
#define DG_LB_VAL ((1ull << 62) - 1ull)
#define DG_LB_SPIN_LIMIT (1u << 22)
__device__ __forceinline__ unsigned long long dg_lb_load(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void dg_lb_store(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// exclusive prefix (sum of the values of entries 0 .. a-1); all 32 lanes must call.  A round trip covers DG_LB_WIDE * 32
// predecessors (the loads of a round are independent): with a few hundred chunks in flight the walk back to the last
// resolved prefix takes one to three L2 latencies.
#define DG_LB_WIDE 2
__device__ __forceinline__ long long dg_lookback(const unsigned long long* __restrict__ status, long long a, int lane,
                                                 DgStats* __restrict__ stats, unsigned spin_limit) {
    long long excl = 0;
    long long pos = a - 1;
    unsigned spins = 0;
    while (pos >= 0) {
        unsigned long long w[DG_LB_WIDE];
#pragma unroll
        for (int k = 0; k < DG_LB_WIDE; ++k) {
            const long long idx = pos - 32 * k - lane;
            w[k] = idx >= 0 ? dg_lb_load(status + idx) : DG_LB_INC;  // entry -1: inclusive prefix 0
        }
        long long sum = 0;
        bool done = false, retry = false;
#pragma unroll
        for (int k = 0; k < DG_LB_WIDE; ++k) {
            if (done || retry) continue;  // warp-uniform
            const unsigned flag = (unsigned)(w[k] >> 62);
            const unsigned inc = __ballot_sync(FULL_MASK, flag == 2u);
            const unsigned emp = __ballot_sync(FULL_MASK, flag == 0u);
            const int first = inc ? __ffs(inc) - 1 : 32;                   // nearest predecessor with a resolved prefix
            const unsigned need = first >= 31 ? 0xffffffffu : ((2u << first) - 1u);
            if (emp & need) {  // a predecessor in the window has not published its aggregate yet
                retry = true;
                continue;
            }
            // AGG values are sums of a few row degrees (< 2^20 each): 32 of them fit an unsigned sum
            sum += (long long)__reduce_add_sync(FULL_MASK, lane < first ? (unsigned)(w[k] & 0xffffffffull) : 0u);
            if (first < 32) {
                const unsigned long long v = w[k] & DG_LB_VAL;
                const unsigned lo = __shfl_sync(FULL_MASK, (unsigned)v, first);
                const unsigned hi = __shfl_sync(FULL_MASK, (unsigned)(v >> 32), first);
                sum += (long long)(((unsigned long long)hi << 32) | lo);
                done = true;
            }
        }
        if (retry) {
            if (++spins > spin_limit) {
                if (lane == 0) atomicExch(&stats->lb_timeout, 1);
                break;
            }
            __nanosleep(20);
            continue;
        }
        excl += sum;
        if (done) break;
        pos -= 32 * DG_LB_WIDE;
    }
    return excl;
}

// LEAN (fill pass of batches no row of which is truncated, max_neighbors >= row_cap): the lean row sort (row_sort.cuh)
// ---------------------------------------------------------------------------------------------------------------------
// Single-pass build.  A step of a CTA = nwarps consecutive rows (one per warp), step t of CTA b is chunk t * grid + b:
// at any moment the resident CTAs work on one contiguous window of rows, so every predecessor of a chunk is finished
// or in flight.  Per step: every warp sweeps its row into its shared-memory row buffer and posts the (capped) degree;
// the LAST warp to finish its sweep adds the degrees up, publishes the chunk aggregate, resolves the chunk's base by
// look-back over the chunk status words and hands the per-row offsets to its siblings -- which meanwhile have been
// sorting their rows in registers and only wait (named barrier) when they are ready to emit.  No count pass, no scan
// kernel, the sweep runs once.
// ---------------------------------------------------------------------------------------------------------------------
#define DG_ONE_MAX_WARPS 8
struct DgOneShared {
    long long base[2][DG_ONE_MAX_WARPS];  // per-row CSR offsets of the step (by step parity)
    int keep[2][DG_ONE_MAX_WARPS];
    int arrived[2];
};

// offset provider of the row sort (row_sort.cuh): the warp that ran the look-back already knows its offset, the others
// wait for the step's barrier.  Rows that must not be emitted (buffer overflow, arena overflow) yield -1.
struct DgOneOffset {
    const DgOneShared* sh;
    int par, warp, nthreads;
    long long o;        // valid when ready
    bool ready;
    long long edge_cap;
    DgStats* stats;
    int lane;
    __device__ __forceinline__ long long wait() {
        if (!ready) {
            // barrier id alternates with the step parity: a warp can be at most one step ahead of its siblings, so the
            // arrivals of consecutive steps never mix
            if (par) asm volatile("bar.sync 2, %0;" ::"r"(nthreads) : "memory");
            else asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory");
            o = sh->base[par][warp];
            ready = true;
        }
        return o;
    }
    __device__ __forceinline__ long long get(int keep) {
        const long long v = wait();
        if (v + keep > edge_cap) {  // the estimated arena is too small: the host regrows it and reruns the fill
            if (lane == 0) atomicExch(&stats->cap_overflow, 1);
            return -1;
        }
        return v;
    }
};

// The rows of two consecutive steps are alive at once (see k_direct_one): the entry arrays the sweep fills (d2, atom,
// image) exist twice per warp, the sort scratch once.
__host__ __device__ inline size_t dg_one_extra_bytes(int row_cap) {
    return dg_align16((size_t)row_cap * 8) + 2 * dg_align16((size_t)row_cap * 4);
}
__host__ __device__ inline size_t dg_one_warp_bytes(int row_cap, int n_buckets) {
    return dg_rowbuf_bytes(row_cap, n_buckets) + dg_one_extra_bytes(row_cap);
}
__device__ __forceinline__ RowBuf dg_one_rowbuf(unsigned char* wp, int row_cap, int n_buckets, int par) {
    RowBuf r = dg_rowbuf_carve(wp, row_cap);
    if (par) {
        unsigned char* q = wp + dg_rowbuf_bytes(row_cap, n_buckets);
        r.rd = (double*)q; q += dg_align16((size_t)row_cap * 8);
        r.rj = (int*)q; q += dg_align16((size_t)row_cap * 4);
        r.rn = (uint32_t*)q;
    }
    return r;
}

// a swept row waiting for its offset
struct DgOnePending {
    long long a, o;
    int m, keep;
    bool ready;
};

__device__ __forceinline__ void dg_one_finish_row(const DgDirectArgs& A, const DgOneShared* sh, const RowBuf& rb,
                                                  const DgOnePending& P, int par, int warp, int lane) {
    DgOneOffset op{sh, par, warp, (int)blockDim.x, P.o, P.ready, A.edge_cap, A.stats, lane};
    const bool valid = P.a < A.n_atoms;
    if (!valid || P.m > A.row_cap) {
        // nothing to emit here: rows that outgrew the estimated row buffer only reserve their range (a fill pass
        // completes them afterwards)
        op.wait();
        if (valid && lane == 0) atomicAdd(&A.stats->n_fallback_rows, 1);
    } else {
        dg_row_sort_emit_op<DgOneOffset, false, true>(rb, P.m, P.keep, A.n_buckets, A.inv_rc, A.key_scale, op, A.dst, A.image,
                                                      A.dist, A.stats, lane, A.rbf);
    }
    if (valid && lane == 0) {
        A.deg_raw[P.a] = P.m;
        A.row_ptr_out[P.a] = op.o;
        if (P.a == A.n_atoms - 1) A.row_ptr_out[P.a + 1] = op.o + P.keep;
    }
}

template <int MINB, int EPS>
__global__ void __launch_bounds__(256, MINB) k_direct_one(const DgDirectArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ DgOneShared sh;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    unsigned char* wp = smem_raw + (size_t)warp * dg_one_warp_bytes(A.row_cap, A.n_buckets);
    if (threadIdx.x < 2) sh.arrived[threadIdx.x] = 0;
    __syncthreads();
    unsigned near = 0;
    int maxdeg = 0;
    const long long n_chunks = (A.n_atoms + nwarps - 1) / nwarps;
    int par = 0;
    bool have_pending = false;
    DgOnePending P{0, 0, 0, 0, false};
    for (long long chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x, par ^= 1) {
        // ---- sweep the row of this step into the entry arrays of its parity ----
        const long long a = chunk * nwarps + warp;
        int m = 0;
        if (a < A.n_atoms) {
            const DgSmall* __restrict__ G = A.small + A.astruct[a];
            const int mode = G->mode;
            // (never launched on batches with cell-list or invalid structures; such a row would count as empty)
            if (mode != DG_MODE_CELL && G->status == DG_ST_OK) {
                const RowBuf rb = dg_one_rowbuf(wp, A.row_cap, A.n_buckets, par);
                m = traverse_row<true, EPS, false, true>(A, G, a, mode, rb, nullptr, nullptr, lane, lt_mask, near);
            }
        }
        maxdeg = m > maxdeg ? m : maxdeg;
        const int keep = m < A.max_neighbors ? m : A.max_neighbors;
        int ticket = 0;
        if (lane == 0) {
            sh.keep[par][warp] = keep;
            __threadfence_block();
            ticket = atomicAdd(&sh.arrived[par], 1);
        }
        ticket = __shfl_sync(FULL_MASK, ticket, 0);
        DgOnePending C{a, 0, m, keep, false};
        if (ticket == nwarps - 1) {
            // last sweep of the step to finish: chunk aggregate, look-back, per-row offsets for everybody
            __threadfence_block();
            const int k = lane < nwarps ? sh.keep[par][lane] : 0;
            int incl = k;
#pragma unroll
            for (int off = 1; off < DG_ONE_MAX_WARPS; off <<= 1) {
                const int t = __shfl_up_sync(FULL_MASK, incl, off);
                if (lane >= off) incl += t;
            }
            const int total = __shfl_sync(FULL_MASK, incl, nwarps - 1);
            if (lane == 0) dg_lb_store(A.lb_status + chunk, DG_LB_AGG | (unsigned long long)total);
            const long long excl = dg_lookback(A.lb_status, chunk, lane, A.stats, A.lb_spin_limit > 0 ? (unsigned)A.lb_spin_limit : DG_LB_SPIN_LIMIT);
            if (lane == 0) {
                dg_lb_store(A.lb_status + chunk, DG_LB_INC | (unsigned long long)(excl + total));
                sh.arrived[par] = 0;  // next used two steps from now
            }
            const long long mine = excl + (long long)(incl - k);
            if (lane < nwarps) sh.base[par][lane] = mine;
            if (par) asm volatile("bar.arrive 2, %0;" ::"r"((int)blockDim.x) : "memory");
            else asm volatile("bar.arrive 1, %0;" ::"r"((int)blockDim.x) : "memory");
            C.o = __shfl_sync(FULL_MASK, mine, warp);
            C.ready = true;
        }
        // ---- sort and emit the row of the previous step: its offsets have had a whole sweep to arrive ----
        if (have_pending) {
            const RowBuf rbp = dg_one_rowbuf(wp, A.row_cap, A.n_buckets, par ^ 1);
            dg_one_finish_row(A, &sh, rbp, P, par ^ 1, warp, lane);
        }
        P = C;
        have_pending = true;
    }
    if (have_pending) {
        const RowBuf rbp = dg_one_rowbuf(wp, A.row_cap, A.n_buckets, par ^ 1);
        dg_one_finish_row(A, &sh, rbp, P, par ^ 1, warp, lane);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) near += __shfl_xor_sync(FULL_MASK, near, o);
    if (lane == 0) {
        if (near) atomicAdd(&A.stats->n_near, (unsigned long long)near);
        if (maxdeg) atomicMax(&A.stats->max_deg, maxdeg);
    }
}

size_t dg_direct_onepass_smem_bytes(const DgDirectArgs& a) {
    return (size_t)a.n_warps * dg_one_warp_bytes(a.row_cap, a.n_buckets);
}

// single-pass build: a.lb_status ([ceil(N / n_warps)] chunk status words) must be zeroed on the stream before the launch
template <int MINB, int EPS>
static cudaError_t launch_one_t(cudaStream_t st, const DgDirectArgs& a, int sm_count, size_t smem, int threads) {
    cudaError_t e = dg_raise_dyn_smem<k_direct_one<MINB, EPS>>();
    if (e != cudaSuccess) return e;
    int occ = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_direct_one<MINB, EPS>, threads, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) return cudaErrorLaunchOutOfResources;
    // the look-back needs every CTA resident: never more CTAs than the occupancy query allows
    const long long want = (a.n_atoms + a.n_warps - 1) / a.n_warps;
    long long grid = (long long)sm_count * occ;
    if (grid > want) grid = want;
    k_direct_one<MINB, EPS><<<(unsigned)grid, threads, smem, st>>>(a);
    return cudaGetLastError();
}
template <int MINB>
static cudaError_t launch_one_e(cudaStream_t st, const DgDirectArgs& a, int sm_count, size_t smem, int threads) {
    switch (a.eps_mode) {
        case 0: return launch_one_t<MINB, 0>(st, a, sm_count, smem, threads);
        case 1: return launch_one_t<MINB, 1>(st, a, sm_count, smem, threads);
        default: return launch_one_t<MINB, 2>(st, a, sm_count, smem, threads);
    }
}
cudaError_t dg_launch_direct_onepass(cudaStream_t st, const DgDirectArgs& a, int sm_count, int min_blocks) {
    if (a.n_atoms <= 0) return cudaSuccess;
    if (a.n_warps > DG_ONE_MAX_WARPS) return cudaErrorInvalidValue;
    const size_t smem = dg_direct_onepass_smem_bytes(a);
    const int threads = a.n_warps * 32;
    dg_direct_launch_bump(1);
    (void)min_blocks;
    return launch_one_e<2>(st, a, sm_count, smem, threads);
}

