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
#ifndef DG_STORE_MODE
#define DG_STORE_MODE 0
#endif
#ifndef DG_RBF_UNROLL
#define DG_RBF_UNROLL 4  // independent float4 stores in flight per lane
#endif
#ifndef DG_RBF_SCHEME
#define DG_RBF_SCHEME 0  // 0: fixed float4 column per lane; 1: flat float4 index, warp stores aligned to 128-byte lines
#endif
__device__ __forceinline__ void dg_store_f4(float4* p, const float4& v) {
#if DG_STORE_MODE == 0
    __stcs(p, v);
#elif DG_STORE_MODE == 1
    *p = v;
#elif DG_STORE_MODE == 2
    __stcg(p, v);
#else
    __stwt(p, v);
#endif
}

__device__ __forceinline__ void dg_rbf_row(const float* __restrict__ fd, int keep, long long o, int lane,
                                           const DgRbfEpi& rbf) {
    float4* __restrict__ o4 = rbf.out + o * (long long)rbf.K4;
    const int G = 32 / rbf.K4;  // 0 when K4 > 32
    if (DG_RBF_SCHEME == 0 && G * rbf.K4 >= 28) {
        const int es = lane / rbf.K4, k4 = lane - es * rbf.K4;
        if (es < G) {
            const float c0 = (float)(4 * k4) * rbf.dr, c1 = (float)(4 * k4 + 1) * rbf.dr;
            const float c2 = (float)(4 * k4 + 2) * rbf.dr, c3 = (float)(4 * k4 + 3) * rbf.dr;
            float4* __restrict__ p = o4 + es * rbf.K4 + k4;
            const int step = G * rbf.K4;
            constexpr int kUnroll = DG_RBF_UNROLL;
            if (rbf.recur) {
                // (x0 + i dr)^2 = x0^2 + 2 i dr x0 + i^2 dr^2  =>  g_{i+1} = g_i * r * q^(2i+1): 2 MUFU + 10 FP32 ops
                // per float4 instead of 4 MUFU + 12; the three chained products add < 4 ulp (tolerance 1e-5 relative)
                const float q = rbf.q, q2 = rbf.q2;
#pragma unroll kUnroll
                for (int e = es; e < keep; e += G, p += step) {
                    const float x0 = c0 - fd[e];
                    const float g0 = dg_ex2(fmaf(rbf.coef * x0, x0, rbf.lognorm));
                    const float r0 = dg_ex2(rbf.c2dr * x0) * q;
                    const float r1 = r0 * q2, r2 = r1 * q2;
                    float4 v;
                    v.x = g0;
                    v.y = g0 * r0;
                    v.z = v.y * r1;
                    v.w = v.z * r2;
                    dg_store_f4(p, v);
                }
            } else {
#pragma unroll kUnroll
                for (int e = es; e < keep; e += G, p += step) {
                    const float d = fd[e];
                    const float x0 = c0 - d, x1 = c1 - d, x2 = c2 - d, x3 = c3 - d;
                    float4 v;
                    v.x = dg_ex2(fmaf(rbf.coef * x0, x0, rbf.lognorm));
                    v.y = dg_ex2(fmaf(rbf.coef * x1, x1, rbf.lognorm));
                    v.z = dg_ex2(fmaf(rbf.coef * x2, x2, rbf.lognorm));
                    v.w = dg_ex2(fmaf(rbf.coef * x3, x3, rbf.lognorm));
                    dg_store_f4(p, v);
                }
            }
        }
    } else {
        const int nq = keep * rbf.K4;
        // flat float4 index q = lane - shift, + 32, ...: the shift puts lane 0 on a 128-byte line boundary, so every
        // warp store covers four whole lines; (edge, bin group) walked incrementally: no division in the loop
        const int shift = DG_RBF_SCHEME == 1 ? (int)((o * (long long)rbf.K4) & 7) : 0;
        const int de = 32 / rbf.K4, dk = 32 - de * rbf.K4;
        int q = lane - shift;
        int e = 0, k4 = q;
        if (q >= 0) {
            e = q / rbf.K4;
            k4 = q - e * rbf.K4;
        } else {  // first step is a partial one for the lanes before the row start: they begin one round later
            q += 32;
            e = q / rbf.K4;
            k4 = q - e * rbf.K4;
        }
#pragma unroll 4
        for (; q < nq; q += 32) {
            dg_store_f4(o4 + q, dg_rbf4(fd[e], k4 << 2, rbf));
            k4 += dk;
            e += de;
            if (k4 >= rbf.K4) {
                k4 -= rbf.K4;
                ++e;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Same epilogue through the TMA: features are produced into a small shared-memory stage (the row-major layout of the
// E x K tensor, so a stage is one contiguous span of it) and an elected lane hands every full stage to the bulk-copy
// engine (cp.async.bulk shared -> global).  The feature stream -- 91 % of all bytes the path writes -- then never sits
// in the LSU / MIO queues, so back-pressure from HBM does not block the shuffles and shared-memory traffic of the warps
// that are sorting their next row.  Two stages per warp, used alternately across rows (`seq` persists).
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void dg_bulk_store(void* gdst, const void* ssrc, unsigned bytes) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(ssrc);
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(s), "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void dg_bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void dg_bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void dg_fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// stage: 2 x DG_STAGE_Q float4 per warp (16-byte aligned).  A stage takes DG_STAGE_STEPS warp steps of `step` float4
// (step = G * K4 <= 32 whole feature rows in the fixed-column scheme, else 32).  All 32 lanes must call.
#define DG_STAGE_STEPS 4
#define DG_STAGE_Q (32 * DG_STAGE_STEPS)
__device__ __forceinline__ void dg_rbf_row_bulk(const float* __restrict__ fd, int keep, long long o, int lane,
                                                const DgRbfEpi& rbf, float4* __restrict__ stage, unsigned& seq) {
    float4* __restrict__ o4 = rbf.out + o * (long long)rbf.K4;
    const int nq = keep * rbf.K4;
    const int G = 32 / rbf.K4;  // 0 when K4 > 32
    if (G * rbf.K4 >= 28) {
        // every lane owns one fixed float4 column: bin centres are loop constants, a step is G whole feature rows
        const int step = G * rbf.K4, cq = DG_STAGE_STEPS * step;
        const int es = lane / rbf.K4, kc = lane - es * rbf.K4;
        const float c0 = (float)(4 * kc) * rbf.dr, c1 = (float)(4 * kc + 1) * rbf.dr;
        const float c2 = (float)(4 * kc + 2) * rbf.dr, c3 = (float)(4 * kc + 3) * rbf.dr;
        int e = es;
        for (int q0 = 0; q0 < nq; q0 += cq) {
            float4* __restrict__ sb = stage + (seq & 1u) * DG_STAGE_Q + lane;
            if (lane == 0) dg_bulk_wait_read<1>();  // the copy that last read this stage is two groups back
            __syncwarp();
#pragma unroll
            for (int it = 0; it < DG_STAGE_STEPS; ++it, e += G, sb += step) {
                // branch-free: lanes past the row end (and the 32 - step idle lanes) recompute the last edge; what
                // they write lies beyond the bytes this stage copies or is overwritten by the next step (same warp,
                // program order)
                const float d = fd[e < keep ? e : keep - 1];
                const float x0 = c0 - d, x1 = c1 - d, x2 = c2 - d, x3 = c3 - d;
                float4 v;
                v.x = dg_ex2(fmaf(rbf.coef * x0, x0, rbf.lognorm));
                v.y = dg_ex2(fmaf(rbf.coef * x1, x1, rbf.lognorm));
                v.z = dg_ex2(fmaf(rbf.coef * x2, x2, rbf.lognorm));
                v.w = dg_ex2(fmaf(rbf.coef * x3, x3, rbf.lognorm));
                *sb = v;
            }
            dg_fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
                const int nq_c = nq - q0 < cq ? nq - q0 : cq;
                dg_bulk_store(o4 + q0, stage + (seq & 1u) * DG_STAGE_Q, (unsigned)nq_c * 16u);
            }
            ++seq;
        }
    } else {
        // flat float4 index; (edge, bin group) walked incrementally: no division in the loop
        const int de = 32 / rbf.K4, dk = 32 - de * rbf.K4;
        int e = lane / rbf.K4, k4 = lane - e * rbf.K4;
        for (int q0 = 0; q0 < nq; q0 += DG_STAGE_Q) {
            float4* __restrict__ sb = stage + (seq & 1u) * DG_STAGE_Q + lane;
            if (lane == 0) dg_bulk_wait_read<1>();
            __syncwarp();
#pragma unroll
            for (int it = 0; it < DG_STAGE_STEPS; ++it, sb += 32) {
                if (q0 + it * 32 + lane < nq) *sb = dg_rbf4(fd[e], k4 << 2, rbf);
                k4 += dk;
                e += de;
                if (k4 >= rbf.K4) {
                    k4 -= rbf.K4;
                    ++e;
                }
            }
            dg_fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
                const int nq_c = nq - q0 < DG_STAGE_Q ? nq - q0 : DG_STAGE_Q;
                dg_bulk_store(o4 + q0, stage + (seq & 1u) * DG_STAGE_Q, (unsigned)nq_c * 16u);
            }
            ++seq;
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
// Register bitonic sort (rows of up to 512 entries).  Entry e of the row becomes the word (key & ~mask) | e, key =
// 32-bit distance key, mask = 32 * EPL - 1 (the entry id needs log2(32 EPL) bits); EPL words per lane, 32 * EPL >= m,
// padding 0xffffffff.  The network is lane-major (position p = lane * EPL + r): strides below EPL are compare-exchanges
// between a lane's own registers, larger strides one shuffle + one min/max per word.  Positions whose truncated keys are
// equal (exact ties of symmetric cells, or |delta d| < r_c 2^-(32 - log2(32 EPL))) form contiguous runs the word alone
// cannot order: inside such runs the rank is recomputed with the exact canonical comparison (fp64 distance, then
// (dst, n_a, n_b, n_c)).
//
// Payload policies give the sort access to an entry's (dst, packed image):
// ---------------------------------------------------------------------------------------------------------------------
// (load_item / store_item move an entry inside the row buffer: the top-k prefilter compacts the survivors in place)
// kSweepOrder: entries of one neighbour atom appear in the row buffer in ascending (n_a, n_b, n_c) -- true for rows the
// direct-path sweeps compact (a lane owns an atom and walks its images in order; the top-k prefilter's compaction is
// stable); the tie ranking then orders a group of exactly equal distances by dst alone (dg_row_sort_bitonic).
struct RowBufPayload {  // traversal kernels: rj / rn arrays filled by the ballot compaction
    static constexpr bool kSweepOrder = false;
    int* rj;
    uint32_t* rn;
    __device__ __forceinline__ void get(int e, int& j, uint32_t& img) const {
        j = rj[e];
        img = rn[e];
    }
    __device__ __forceinline__ uint2 load_item(int e) const { return make_uint2((uint32_t)rj[e], rn[e]); }
    __device__ __forceinline__ void store_item(int e, uint2 v) const {
        rj[e] = (int)v.x;
        rn[e] = v.y;
    }
};
struct SweepRowPayload : RowBufPayload {  // ... of the direct-path sweeps (k_direct, k_direct_one)
    static constexpr bool kSweepOrder = true;
};
template <bool SWEEP> struct DgSelectPayload { typedef RowBufPayload type; };
template <> struct DgSelectPayload<true> { typedef SweepRowPayload type; };
struct HitWordPayload {  // hit-list fill pass: the count pass's packed words (dg_pack_hit in dgnn_direct.cu)
    static constexpr bool kSweepOrder = true;  // (the lists are compacted by the same sweep, the threshold only filters)
    uint32_t* hw;
    __device__ __forceinline__ uint2 load_item(int e) const { return make_uint2(hw[e], 0u); }
    __device__ __forceinline__ void store_item(int e, uint2 v) const { hw[e] = v.x; }
    __device__ __forceinline__ void get(int e, int& j, uint32_t& img) const {
        const uint32_t v = hw[e];
        j = (int)(v & 0x1ffffu);
        img = dg_pack_image((int)((v >> 17) & 31u) - 16, (int)((v >> 22) & 31u) - 16, (int)(v >> 27) - 16);
    }
};
struct CloudPayload {  // hit-list fill pass, cell-list structures: hit word = index into the binned image cloud
    static constexpr bool kSweepOrder = false;
    uint32_t* hw;
    __device__ __forceinline__ uint2 load_item(int e) const { return make_uint2(hw[e], 0u); }
    __device__ __forceinline__ void store_item(int e, uint2 v) const { hw[e] = v.x; }
    const int* cj;
    const uint32_t* cn;
    __device__ __forceinline__ void get(int e, int& j, uint32_t& img) const {
        const uint32_t p = hw[e];
        j = cj[p];
        img = cn[p];
    }
};
template <class Payload>
__device__ __forceinline__ int dg_payload_dst(const Payload& P, int e) {
    int j;
    uint32_t img;
    P.get(e, j, img);
    return j;
}
template <class Payload>
__device__ __forceinline__ unsigned long long dg_payload_tie_key(const Payload& P, int e) {
    int j;
    uint32_t img;
    P.get(e, j, img);
    return ((unsigned long long)(uint32_t)j << 32) | dg_image_sort_key(img);
}

// rd: d2 by entry (replaced by the distance), sk: scratch for the sorted words, fd: distances in final order (fp32),
// tie: two ints of scratch.  All 32 lanes must call; m <= 32 * EPL.
// Where a row's edges go.  The two-pass builds know the offset up front (row_ptr from the scan); the single-pass build
// learns it from the look-back while the row is being sorted, so the sort asks for it only when it starts to emit.
// get() is called exactly once per row by all 32 lanes; a negative offset means "do not emit this row".
struct DgFixedOffset {
    long long o;
    __device__ __forceinline__ long long get(int) { return o; }
};

// The sorted words leave the registers lane-major (position p = lane * EPL + r): with EPL = 4 / 8 / 16 a plain sk[p]
// store puts the 32 lanes on 8 / 4 / 2 banks.  One pad word per 32 positions (DG_SK_PAD = 1) makes the store and the
// later p = lane + 32 k reads conflict-free -- and the fill kernel 4 % SLOWER (same-box A/B: 8.69 vs 8.35 ms): the
// kernel is bound by instruction issue, the extra index arithmetic costs more than the conflicts.  Kept off.
#ifndef DG_SK_PAD
#define DG_SK_PAD 0
#endif
#if DG_SK_PAD
#define DG_SKI(p) ((p) + ((p) >> 5))
#else
#define DG_SKI(p) (p)
#endif
#define DG_SK_WORDS(n) ((n) + (((n) >> 5) + 4))

// LEAN: the variant the uncapped sweep-fill kernel (the headline kernel, 128 registers, bound by instruction issue)
// instantiates -- no top-k prefilter (its rows are never truncated) and the compact per-entry tie ranking instead of
// the warp-cooperative one.  Same results; measured on the same box, the full variant costs that kernel 6 % (8.82
// against 8.33 ms: two registers spill), and calling either piece out of line costs more (8.96 - 9.0 ms).
// Tie runs longer than the 32-position window of the ranking below (sweep-ordered rows only, see SweepRowPayload): the
// part of the run that sits in another window is brought into the lanes 32 entries at a time (vS: lane holds an entry of
// the run, dS its distance, jS its dst) and every lane of the run in the CURRENT window (mine: distance de, dst dj) adds
// the number of those entries ordered before it -- groups of a smaller distance by their sizes, its own distance's group
// by a radix walk over the bits of dst.  inclusive: the block lies BEFORE the window (its entries have the smaller entry
// ids: with the same dst they come first); otherwise after it.  All 32 lanes must call.
__device__ __forceinline__ int dg_tie_block_count(bool vS, double dS, int jS, bool mine, double de, int dj, bool inclusive,
                                                  int lane) {
    const unsigned vmask = __ballot_sync(FULL_MASK, vS);
    const long long pat = vS ? __double_as_longlong(dS) : (long long)(0xfff8000000000000ULL | (unsigned long long)lane);
    const unsigned eqS = __match_any_sync(FULL_MASK, pat) & vmask;
    const unsigned leadS = __ballot_sync(FULL_MASK, vS && lane == __ffs(eqS) - 1);
    int cnt = 0;
    unsigned mygrp = 0u;
    for (unsigned lm = leadS; lm; lm &= lm - 1) {
        const int r = __ffs(lm) - 1;
        const double d_r = __shfl_sync(FULL_MASK, dS, r);
        const unsigned g_r = __shfl_sync(FULL_MASK, eqS, r);
        if (mine) {
            if (d_r < de) cnt += __popc(g_r);
            else if (d_r == de) mygrp = g_r;
        }
    }
    const unsigned ord = __reduce_or_sync(FULL_MASK, (unsigned)(vS ? jS : 0) | (unsigned)(mine ? dj : 0));
    unsigned ltm = 0u, eq = mygrp;
    for (int b = 31 - __clz(ord | 1u); b >= 0; --b) {
        const unsigned B = __ballot_sync(FULL_MASK, vS && ((jS >> b) & 1) != 0);
        if ((dj >> b) & 1) {
            ltm |= eq & ~B;
            eq &= B;
        } else {
            eq &= ~B;
        }
    }
    if (mine) cnt += __popc(ltm) + (inclusive ? __popc(eq) : 0);
    return cnt;
}

template <int EPL, class Payload, class OP, bool LEAN = false>
__device__ __forceinline__ long long dg_row_sort_bitonic(double* __restrict__ rd, uint32_t* __restrict__ sk,
                                                         float* __restrict__ fd, int* __restrict__ tie,
                                                         const Payload& P, int m, int keep, double key_scale, OP& op,
                                                         int* __restrict__ dst, uint32_t* __restrict__ image,
                                                         double* __restrict__ dist, DgStats* __restrict__ stats,
                                                         int lane) {
    constexpr uint32_t IDMASK = 32u * EPL - 1u;
    uint32_t v[EPL];
#pragma unroll
    for (int t = 0; t < EPL; ++t) {
        v[t] = 0xffffffffu;
        if (32 * t < m) {  // warp-uniform
            const int e = lane + 32 * t;
            if (e < m) {
                const double d = __dsqrt_rn(rd[e]);  // Neighbor::distance = sqrt(match.second), neighbor_list.cpp:53
                rd[e] = d;
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
    clash = __any_sync(FULL_MASK, clash);
    // transpose through shared memory: coalesced CSR stores, and the run scan below needs random access
#pragma unroll
    for (int r = 0; r < EPL; ++r)
        if (lane * EPL + r < m) sk[DG_SKI(lane * EPL + r)] = v[r];
    __syncwarp();
    const long long o = op.get(keep);
    if (o < 0) return o;
    if (!clash) {
        for (int p = lane; p < keep; p += 32) {
            const int e = (int)(sk[DG_SKI(p)] & IDMASK);
            const double d = rd[e];
            int j;
            uint32_t img;
            P.get(e, j, img);
            dst[o + p] = j;
            image[o + p] = img;
            dist[o + p] = d;
            fd[p] = (float)d;
        }
    } else {
        if (LEAN) {
        // truncated rows: a position beyond `keep` whose key differs from the key AT position keep belongs to a run that
        // starts after it -- its rank can neither be emitted (rank < keep) nor enter the tie statistic (rank <= keep),
        // so only the runs that reach into [0, keep] are resolved (capped rows of ideal lattices: 20 of ~300 entries)
        const uint32_t tk_edge = keep < m ? (sk[DG_SKI(keep)] & ~IDMASK) : 0u;
        for (int p = lane; p < m; p += 32) {
            const uint32_t w = sk[DG_SKI(p)];
            const uint32_t tk = w & ~IDMASK;
            if (p > keep && tk != tk_edge) continue;
            const int e = (int)(w & IDMASK);
            const double de = rd[e];
            int s = p, t = p + 1;
            while (s > 0 && (sk[DG_SKI(s - 1)] & ~IDMASK) == tk) --s;
            while (t < m && (sk[DG_SKI(t)] & ~IDMASK) == tk) ++t;
            int rank = p;
            if (t - s > 1) {  // run of equal truncated keys: exact canonical rank inside it
                const unsigned long long ke = dg_payload_tie_key(P, e);
                rank = s;
                for (int q = s; q < t; ++q) {
                    if (q == p) continue;
                    const int f = (int)(sk[DG_SKI(q)] & IDMASK);
                    const double df = rd[f];
                    if (df < de || (df == de && dg_payload_tie_key(P, f) < ke)) ++rank;
                }
            }
            if (rank < keep) {
                int j;
                uint32_t img;
                P.get(e, j, img);
                dst[o + rank] = j;
                image[o + rank] = img;
                dist[o + rank] = de;
                fd[rank] = (float)de;
            }
            if (m > keep && (rank == keep - 1 || rank == keep)) tie[rank - keep + 1] = e;
        }
        } else {
        // Runs of equal truncated keys (coordination shells of an ideal lattice: every shell is one run) get their
        // exact canonical order (distance, dst, image).  Truncated rows: a position beyond `keep` whose key differs
        // from the key AT position keep belongs to a run that starts after it -- its rank can neither be emitted
        // (rank < keep) nor enter the tie statistic (rank <= keep), so only positions up to the end of the run that
        // holds position keep are resolved (capped rows of ideal lattices: 20 of ~300 entries).
        int lim = m;
        if (keep < m) {
            const uint32_t tk_edge = sk[DG_SKI(keep)] & ~IDMASK;
            lim = m;
            for (int c0 = keep + 1; c0 < m; c0 += 32) {
                const int q = c0 + lane;
                const unsigned diff = __ballot_sync(FULL_MASK, q < m && (sk[DG_SKI(q)] & ~IDMASK) != tk_edge);
                if (diff) {
                    lim = c0 + __ffs(diff) - 1;
                    break;
                }
            }
        }
        // 32 positions at a time: the entries that sit in a run are broadcast one by one from registers and every lane
        // of the same run counts the ones ordered before it (no divergent per-entry loops, no shared-memory chase);
        // only runs that cross the 32-position window finish with a scalar walk over the part outside it
        // The window is cut back to the start of its last run when that run continues past it (and does not fill the
        // whole window), so runs of up to 32 entries are always ranked inside one window.
        for (int c0 = 0, adv = 32; c0 < lim; c0 += adv) {
            const int p = c0 + lane;
            adv = 32;
            uint32_t w = p < lim ? sk[DG_SKI(p)] : 0u;
            uint32_t tk = w & ~IDMASK;
            if (c0 + 32 < lim) {
                const uint32_t tk_next = sk[DG_SKI(c0 + 32)] & ~IDMASK;
                const unsigned same = __ballot_sync(FULL_MASK, tk == tk_next);  // a suffix of the lanes (sorted keys)
                if (same != 0u && same != FULL_MASK) adv = __ffs(same) - 1;
            }
            const bool valid = p < lim && lane < adv;
            if (!valid) w = 0u, tk = 0u;
            const int e = (int)(w & IDMASK);
            const double de = valid ? rd[e] : 0.0;
            const bool nb_l = valid && p > 0 && (sk[DG_SKI(p - 1)] & ~IDMASK) == tk;
            const bool nb_r = valid && p + 1 < m && (sk[DG_SKI(p + 1)] & ~IDMASK) == tk;
            const unsigned inrun = __ballot_sync(FULL_MASK, nb_l || nb_r);
            int rank = p;
            if (inrun) {
                // A shell of an ideal lattice is one run of ~26 entries but only a few DISTINCT distances (equal up to
                // the last bit or two): the lanes of a run are grouped by exact distance (match.any), a lane counts the
                // groups of its run with a smaller distance by their sizes -- one broadcast per group, not per entry --
                // and orders itself inside its own group by (dst, image), one broadcast per member of the largest group.
                const bool in = nb_l || nb_r;
                const unsigned long long ke = in ? dg_payload_tie_key(P, e) : 0ull;
                // (lanes outside every run get keys of their own: a truncated key has its id bits clear)
                const unsigned runmask = __match_any_sync(FULL_MASK, in ? tk : (1u | ((uint32_t)lane << 1)));
                const unsigned eqmask = __match_any_sync(FULL_MASK, __double_as_longlong(de)) & runmask;
                const int gsz = __popc(eqmask);
                const unsigned leaders = __ballot_sync(FULL_MASK, in && lane == __ffs(eqmask) - 1);
                int cnt = 0;
                const int first = __ffs(runmask) - 1;
                for (unsigned lm = leaders; lm; lm &= lm - 1) {
                    const int r = __ffs(lm) - 1;
                    const uint32_t tk_r = __shfl_sync(FULL_MASK, tk, r);
                    const double de_r = __shfl_sync(FULL_MASK, de, r);
                    const int sz_r = __shfl_sync(FULL_MASK, gsz, r);
                    if (tk_r == tk && de_r < de) cnt += sz_r;
                }
                if (Payload::kSweepOrder) {
                    // inside a group the order is (dst, image), and the images of one dst already sit in order (lane
                    // order = entry order): a radix walk over the bits of dst splits the group into smaller / equal dst
                    const int dj = in ? (int)(ke >> 32) : 0;
                    const unsigned ord = __reduce_or_sync(FULL_MASK, (unsigned)dj);
                    unsigned ltm = 0u, eq = in ? eqmask : 0u;
                    for (int b = 31 - __clz(ord | 1u); b >= 0; --b) {
                        const bool bit = ((dj >> b) & 1) != 0;
                        const unsigned B = __ballot_sync(FULL_MASK, bit);
                        if (bit) {
                            ltm |= eq & ~B;
                            eq &= B;
                        } else {
                            eq &= ~B;
                        }
                    }
                    cnt += __popc(ltm) + __popc(eq & ((1u << lane) - 1u));
                } else {
                    unsigned gm = in ? (eqmask & ~(1u << lane)) : 0u;
                    const int iters = __reduce_max_sync(FULL_MASK, __popc(gm));
                    for (int t = 0; t < iters; ++t) {
                        const int src = gm ? __ffs(gm) - 1 : lane;
                        gm &= gm - 1u;
                        const unsigned long long ke_s = __shfl_sync(FULL_MASK, ke, src);
                        cnt += (src != lane && ke_s < ke) ? 1 : 0;
                    }
                }
                int run_start = c0;
                if (Payload::kSweepOrder) {
                    // parts of a run outside this window, 32 entries at a time (dg_tie_block_count); only the run that
                    // holds position c0 can begin earlier, only a run that fills the whole window can go on after it
                    const int dj = in ? (int)(ke >> 32) : 0;
                    const uint32_t tk0 = __shfl_sync(FULL_MASK, tk, 0);
                    const bool in0 = __shfl_sync(FULL_MASK, in ? 1 : 0, 0) != 0;
                    const bool mine0 = in && in0 && tk == tk0;
                    bool more = in0 && c0 > 0 && (sk[DG_SKI(c0 - 1)] & ~IDMASK) == tk0;
                    for (int qend = c0; more; qend -= 32) {
                        const int q = qend - 32 + lane;
                        const uint32_t wq = q >= 0 ? sk[DG_SKI(q)] : 0u;
                        const bool vS = q >= 0 && (wq & ~IDMASK) == tk0;
                        const int f = (int)(wq & IDMASK);
                        cnt += dg_tie_block_count(vS, vS ? rd[f] : 0.0, vS ? dg_payload_dst(P, f) : 0, mine0, de, dj, true, lane);
                        const unsigned vm = __ballot_sync(FULL_MASK, vS);
                        run_start = qend - __popc(vm);
                        more = vm == FULL_MASK && qend - 32 > 0 && (sk[DG_SKI(qend - 33)] & ~IDMASK) == tk0;
                    }
                    const int last = adv - 1;  // (a run that goes on after the window fills all of it: adv == 32)
                    const uint32_t tkL = __shfl_sync(FULL_MASK, tk, last);
                    const bool inL = __shfl_sync(FULL_MASK, in ? 1 : 0, last) != 0;
                    const bool mineL = in && inL && tk == tkL;
                    more = inL && c0 + adv < lim && (sk[DG_SKI(c0 + adv)] & ~IDMASK) == tkL;
                    for (int qs = c0 + adv; more; qs += 32) {
                        const int q = qs + lane;
                        const uint32_t wq = q < lim ? sk[DG_SKI(q)] : 0u;
                        const bool vS = q < lim && (wq & ~IDMASK) == tkL;
                        const int f = (int)(wq & IDMASK);
                        cnt += dg_tie_block_count(vS, vS ? rd[f] : 0.0, vS ? dg_payload_dst(P, f) : 0, mineL, de, dj, false, lane);
                        const unsigned vm = __ballot_sync(FULL_MASK, vS);
                        more = vm == FULL_MASK && qs + 32 < lim && (sk[DG_SKI(qs + 32)] & ~IDMASK) == tkL;
                    }
                }
                if (in && Payload::kSweepOrder) {
                    rank = (first == 0 ? run_start : c0 + first) + cnt;
                } else if (in) {
                    int s0 = c0 + first;
                    if (first == 0) {  // the run may begin in an earlier window
                        for (int q = c0 - 1; q >= 0; --q) {
                            const uint32_t wq = sk[DG_SKI(q)];
                            if ((wq & ~IDMASK) != tk) break;
                            const int f = (int)(wq & IDMASK);
                            const double df = rd[f];
                            cnt += (df < de || (df == de && dg_payload_tie_key(P, f) < ke)) ? 1 : 0;
                            s0 = q;
                        }
                    }
                    // ... and may run on past this window (positions beyond lim share no key with positions before it)
                    for (int q = c0 + adv; q < lim; ++q) {
                        const uint32_t wq = sk[DG_SKI(q)];
                        if ((wq & ~IDMASK) != tk) break;
                        const int f = (int)(wq & IDMASK);
                        const double df = rd[f];
                        cnt += (df < de || (df == de && dg_payload_tie_key(P, f) < ke)) ? 1 : 0;
                    }
                    rank = s0 + cnt;
                }
            }
            if (valid) {
                if (rank < keep) {
                    int j;
                    uint32_t img;
                    P.get(e, j, img);
                    dst[o + rank] = j;
                    image[o + rank] = img;
                    dist[o + rank] = de;
                    fd[rank] = (float)de;
                }
                if (m > keep && (rank == keep - 1 || rank == keep)) tie[rank - keep + 1] = e;
            }
        }
        }
        if (m > keep && keep > 0) {  // truncated through a distance tie?  (statistic of rule R8)
            __syncwarp();
            if (lane == 0 && rd[tie[0]] == rd[tie[1]]) atomicAdd(&stats->n_tie_rows, 1);
        }
    }
    __syncwarp();
    return o;
}

// Top-k prefilter of truncated rows (max_neighbors = 20 of ~330 candidates at the reference's defaults): instead of
// sorting all m entries, find a d2 threshold that keeps a little more than keep + 1 of them, compact the survivors to
// the front of the row buffer and sort only those.  The threshold is searched on fp32 copies of d2 held in registers
// (the count of a cubic-law guess, then a bracketed search); exactness: the first keep + 1 entries of the canonical
// order all lie below T' (count(f < T') >= keep + 1), the survivors are everything below T = T' (1 + 2^-19), and two
// entries whose rounded square roots tie differ by < 2^-50 relative in d2 -- far less than the 2^-19 gap minus fp32
// rounding -- so every entry that ties with one of the first keep + 1 survives too (tie order and the tie statistic see
// the same candidates as a full sort).  Returns the new entry count (m if the row is left alone).
template <int EPL, class Payload>
__device__ __forceinline__ int dg_row_prefilter_t(double* __restrict__ rd, const Payload& P, int m, int keep, float r2,
                                                  int lane) {
    float f[EPL];
#pragma unroll
    for (int r = 0; r < EPL; ++r) {
        const int e = r * 32 + lane;
        f[r] = e < m ? (float)rd[e] : 3.0e38f;
    }
    const int need = keep + 1;
    // stop when need <= count <= good; small caps (the reference's default 20) aim at the 64-entry network
    const int good = need <= 32 ? 56 : (need + 48 > 2 * need ? need + 48 : 2 * need);
    const float x = (float)(need + 8) / (float)m;
    float T = r2 * cbrtf(x * x) * 1.1f;  // count ~ d^3
    float lo = 0.0f, hi = 0.0f;          // lo: count < need, hi: count >= need (0: not bracketed yet)
    int chi = m;
    for (int it = 0; it < 10; ++it) {
        int c = 0;
#pragma unroll
        for (int r = 0; r < EPL; ++r) c += f[r] < T ? 1 : 0;
        c = __reduce_add_sync(FULL_MASK, c);
        if (c < need) {
            lo = T;
        } else {
            hi = T;
            chi = c;
            if (c <= good) break;
        }
        if (hi == 0.0f) T = lo * 1.3f;
        else if (lo == 0.0f) T = hi * 0.75f;
        else T = 0.5f * (lo + hi);
        if (hi != 0.0f && !(T > lo && T < hi)) break;  // bracket closed (a large shell of equal distances)
    }
    if (hi == 0.0f || chi >= m - (m >> 2)) return m;  // nothing found worth a compaction pass
    const float Tk = hi * (1.0f + 1.9073486328125e-6f);
    const unsigned lt_mask = (1u << lane) - 1u;
    int base = 0;
    // (the decision is made on the fp32 copies already in registers -- the same values the search counted)
#pragma unroll
    for (int r = 0; r < EPL; ++r) {
        if (r * 32 < m) {  // warp-uniform
            const int e = r * 32 + lane;
            const bool pred = f[r] < Tk;  // (padding is 3e38: never)
            double d2 = 0.0;
            uint2 item = make_uint2(0u, 0u);
            if (pred) {
                d2 = rd[e];
                item = P.load_item(e);
            }
            const unsigned mask = __ballot_sync(FULL_MASK, pred);
            __syncwarp();
            if (pred) {
                const int pos = base + __popc(mask & lt_mask);  // pos <= e: never ahead of an unread entry
                rd[pos] = d2;
                P.store_item(pos, item);
            }
            base += __popc(mask);
            __syncwarp();
        }
    }
    return base;
}
template <class Payload, int MAXEPL = 16>
__device__ __forceinline__ int dg_row_prefilter(double* rd, const Payload& P, int m, int keep, double key_scale,
                                                int lane) {
    if (keep >= m || m <= 128 || 3 * (keep + 1) + 48 > m) return m;
    const float rc = (float)((4294967296.0 - 4096.0) / key_scale);  // key_scale = (2^32 - 4096) / r_c
    const float r2 = rc * rc;
    if (MAXEPL <= 8 || m <= 256) return dg_row_prefilter_t<8>(rd, P, m, keep, r2, lane);
    return dg_row_prefilter_t<16>(rd, P, m, keep, r2, lane);
}

// MAXEPL: largest network compiled in (rows of up to 32 * MAXEPL entries) -- a kernel that only ever sees rows of up to
// 256 entries saves the registers of the 512-entry network
template <class Payload, class OP, bool LEAN = false, int MAXEPL = 16>
__device__ __forceinline__ long long dg_row_sort_bitonic_any_op(double* rd, uint32_t* sk, float* fd, int* tie,
                                                                const Payload& P, int m, int keep, double key_scale,
                                                                OP& op, int* dst, uint32_t* image, double* dist,
                                                                DgStats* stats, int lane) {
    if (!LEAN) {
        const int m0 = m;
        m = dg_row_prefilter<Payload, MAXEPL>(rd, P, m, keep, key_scale, lane);
        if (m < m0 && m <= 64)  // what the prefilter of a small cap leaves
            return dg_row_sort_bitonic<2, Payload, OP, LEAN>(rd, sk, fd, tie, P, m, keep, key_scale, op, dst, image, dist, stats, lane);
    }
    if (m <= 128)
        return dg_row_sort_bitonic<4, Payload, OP, LEAN>(rd, sk, fd, tie, P, m, keep, key_scale, op, dst, image, dist, stats, lane);
    if (MAXEPL <= 8 || m <= 256)
        return dg_row_sort_bitonic<8, Payload, OP, LEAN>(rd, sk, fd, tie, P, m, keep, key_scale, op, dst, image, dist, stats, lane);
    return dg_row_sort_bitonic<16, Payload, OP, LEAN>(rd, sk, fd, tie, P, m, keep, key_scale, op, dst, image, dist, stats, lane);
}
template <int MAXEPL, class Payload>
__device__ __forceinline__ void dg_row_sort_bitonic_upto(double* rd, uint32_t* sk, float* fd, int* tie, const Payload& P,
                                                         int m, int keep, double key_scale, long long o, int* dst,
                                                         uint32_t* image, double* dist, DgStats* stats, int lane) {
    DgFixedOffset op{o};
    dg_row_sort_bitonic_any_op<Payload, DgFixedOffset, false, MAXEPL>(rd, sk, fd, tie, P, m, keep, key_scale, op, dst, image,
                                                                      dist, stats, lane);
}
template <class Payload>
__device__ __forceinline__ void dg_row_sort_bitonic_any(double* rd, uint32_t* sk, float* fd, int* tie, const Payload& P,
                                                        int m, int keep, double key_scale, long long o, int* dst,
                                                        uint32_t* image, double* dist, DgStats* stats, int lane) {
    DgFixedOffset op{o};
    dg_row_sort_bitonic_any_op(rd, sk, fd, tie, P, m, keep, key_scale, op, dst, image, dist, stats, lane);
}

// Sort the m collected entries of the warp's row buffer and write the first `keep` to the CSR arrays at offset o.
// NBk (number of buckets) is a power of two, 32 <= NBk <= 2048.  All 32 lanes must call.
// rbf.K4 > 0: fused Gaussian RBF epilogue straight from the sorted distances in shared memory.
// SWEEP: the row buffer was filled by a direct-path sweep (SweepRowPayload)
template <class OP, bool LEAN = false, bool SWEEP = false>
__device__ __forceinline__ void dg_row_sort_emit_op(const RowBuf& sm, int m, int keep, int NBk, float inv_rc,
                                                    double key_scale, OP& op, int* __restrict__ dst,
                                                    uint32_t* __restrict__ image, double* __restrict__ dist,
                                                    DgStats* __restrict__ stats, int lane, const DgRbfEpi& rbf) {
    if (m <= DG_BITONIC_MAX) {
        typedef typename DgSelectPayload<SWEEP>::type PL;
        PL P;
        P.rj = sm.rj;
        P.rn = sm.rn;
        // padded sorted words: bs (unused by this path) and sk are adjacent, 2 * row_cap words
        const long long o = dg_row_sort_bitonic_any_op<PL, OP, LEAN>(sm.rd, sm.bs, sm.fd, sm.cnt, P, m, keep, key_scale, op,
                                                                     dst, image, dist, stats, lane);
        if (o >= 0 && rbf.K4 > 0) {
            dg_rbf_row(sm.fd, keep, o, lane, rbf);
            __syncwarp();
        }
        return;
    }
    const long long o = op.get(keep);
    if (o < 0) return;
    const int per = NBk >> 5;
    const int lper = 31 - __clz(per);
    const int nwords = NBk + 32 + 1;
    uint32_t* kk = reinterpret_cast<uint32_t*>(sm.fd);  // keys by entry (dead before fd is written)
    for (int b = lane; b < nwords; b += 32) sm.cnt[b] = 0;
    __syncwarp();
    // ---- A: distance, key, bucket histogram ----
    const float scale = (float)NBk;
    for (int e = lane; e < m; e += 32) {
        const double d = __dsqrt_rn(sm.rd[e]);  // Neighbor::distance = sqrt(match.second), neighbor_list.cpp:53
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

template <bool LEAN = false, bool SWEEP = false>
__device__ __forceinline__ void dg_row_sort_emit(const RowBuf& sm, int m, int keep, int NBk, float inv_rc,
                                                 double key_scale, long long o, int* __restrict__ dst,
                                                 uint32_t* __restrict__ image, double* __restrict__ dist,
                                                 DgStats* __restrict__ stats, int lane, const DgRbfEpi& rbf) {
    DgFixedOffset op{o};
    dg_row_sort_emit_op<DgFixedOffset, LEAN, SWEEP>(sm, m, keep, NBk, inv_rc, key_scale, op, dst, image, dist, stats, lane,
                                                    rbf);
}
