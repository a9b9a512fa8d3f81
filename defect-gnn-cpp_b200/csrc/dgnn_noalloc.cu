// dgnn_noalloc.cu -- dgnn_row_offsets_noalloc: CSR row offsets from per-row degrees without any allocation.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is a variant of the degree scan
// that dgnn_build runs between its count and fill passes (dgnn_kernels.cu: k_scan_reduce / k_scan_bsums /
// k_scan_apply over ctx->scratch, a cudaMalloc'ed buffer): same three steps, but the per-tile sums live in a
// file-scope __device__ array of fixed size, so the call needs no context object and allocates nothing -- no
// cudaMalloc / cudaFree / cudaMallocHost, no library routine that allocates.
//
//   * Limit: DGNN_NOALLOC_MAX_ROWS = 2^24 rows per call (8192 tiles of 2048 rows); larger inputs are rejected with
//     DGNN_ERR_TOO_LARGE.  The array is 8193 x 8 B = 65 544 B (64 KB + 8 B) and every GPU that loads the library
//     holds it until the process exits.
//   * The array is shared by every call on a device and keeps the previous call's contents.  Nothing has to start at
//     zero: words [0, n_tiles) are written by step 1 before step 2 reads them, word n_tiles by step 2 before step 3
//     reads it; words beyond are never read.  (So there is no memset in the call.)
//   * Calls that overlap on one GPU (two streams, two host threads) RACE on the array: serialise them, e.g. by
//     issuing them on one stream.  Calls on different GPUs are independent (one array per device).
//   * In host code the array's name is a host-side stand-in, not a device address; the device address is looked up
//     once per device with cudaGetSymbolAddress (that device current) and passed to the kernels.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdarg.h>
#include <stdio.h>

#include <mutex>

#include "../../include/dgnn_cabi.h"

#define NA_THREADS 256
#define NA_ITEMS 8
#define NA_TILE (NA_THREADS * NA_ITEMS)
#define NA_MAX_TILES (DGNN_NOALLOC_MAX_ROWS / NA_TILE)  // 8192
#define NA_MAX_DEVICES 64

static __device__ __align__(256) long long g_na_tile_sums[NA_MAX_TILES + 1];

namespace {

__device__ __forceinline__ long long na_warp_incl(long long v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const long long t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

// inclusive scan of one value per thread over the CTA; *total = CTA sum.  s_warp: 33 words.
__device__ __forceinline__ long long na_block_incl(long long v, long long* total, long long* s_warp) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    long long inc = na_warp_incl(v, lane);
    if (lane == 31) s_warp[w] = inc;
    __syncthreads();
    if (w == 0) {
        const long long t = lane < nw ? s_warp[lane] : 0;
        const long long ti = na_warp_incl(t, lane);
        if (lane < nw) s_warp[lane] = ti - t;
        if (lane == 31) s_warp[32] = ti;
    }
    __syncthreads();
    inc += s_warp[w];
    *total = s_warp[32];
    __syncthreads();
    return inc;
}

// step 1: sums[t] = sum of the degrees of tile t
__global__ void __launch_bounds__(NA_THREADS) k_na_tile_sums(const int* __restrict__ deg, long long n,
                                                             long long* __restrict__ sums) {
    __shared__ long long s_warp[33];
    const long long base = (long long)blockIdx.x * NA_TILE;
    long long acc = 0;
#pragma unroll
    for (int i = 0; i < NA_ITEMS; ++i) {
        const long long idx = base + (long long)i * NA_THREADS + threadIdx.x;
        if (idx < n) acc += deg[idx];
    }
    long long total;
    na_block_incl(acc, &total, s_warp);
    if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

// step 2 (one CTA): exclusive scan of sums[0 .. nt) in place, sums[nt] = grand total
__global__ void __launch_bounds__(1024) k_na_scan_sums(long long* sums, long long nt) {
    __shared__ long long s_warp[33];
    __shared__ long long s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (long long b0 = 0; b0 < nt; b0 += blockDim.x) {
        const long long idx = b0 + threadIdx.x;
        const long long v = idx < nt ? sums[idx] : 0;
        long long total;
        const long long inc = na_block_incl(v, &total, s_warp);
        const long long carry = s_carry;
        if (idx < nt) sums[idx] = carry + inc - v;
        __syncthreads();
        if (threadIdx.x == 0) s_carry = carry + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) sums[nt] = s_carry;
}

// step 3: row_ptr[i] = sums[tile] + exclusive prefix inside the tile; row_ptr[n] = grand total
__global__ void __launch_bounds__(NA_THREADS) k_na_offsets(const int* __restrict__ deg, long long n,
                                                           const long long* __restrict__ sums,
                                                           long long* __restrict__ row_ptr, long long nt) {
    __shared__ long long s_warp[33];
    const long long base = (long long)blockIdx.x * NA_TILE;
    int v[NA_ITEMS];
    long long acc = 0;
#pragma unroll
    for (int i = 0; i < NA_ITEMS; ++i) {
        const long long idx = base + (long long)threadIdx.x * NA_ITEMS + i;
        v[i] = idx < n ? deg[idx] : 0;
        acc += v[i];
    }
    long long total;
    const long long inc = na_block_incl(acc, &total, s_warp);
    long long run = sums[blockIdx.x] + inc - acc;
#pragma unroll
    for (int i = 0; i < NA_ITEMS; ++i) {
        const long long idx = base + (long long)threadIdx.x * NA_ITEMS + i;
        if (idx < n) row_ptr[idx] = run;
        run += v[i];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) row_ptr[n] = sums[nt];
}

__global__ void k_na_empty(long long* row_ptr) { row_ptr[0] = 0; }

// device address of g_na_tile_sums, looked up once per device
std::mutex g_na_mutex;
long long* g_na_ptr[NA_MAX_DEVICES] = {nullptr};

int na_fail(char* err, size_t err_len, int code, const char* fmt, ...) {
    if (err && err_len > 0) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, err_len, fmt, ap);
        va_end(ap);
    }
    return code;
}

}  // namespace

extern "C" int dgnn_row_offsets_noalloc(int device, const int32_t* degrees, int64_t n_rows, int64_t* row_ptr,
                                        void* cuda_stream, char* err, size_t err_len) {
    if (err && err_len > 0) err[0] = 0;
    if (n_rows < 0 || !row_ptr || (n_rows > 0 && !degrees))
        return na_fail(err, err_len, DGNN_ERR_INVALID_ARG,
                       "dgnn_row_offsets_noalloc: null pointer or negative n_rows (%lld)", (long long)n_rows);
    if (n_rows > DGNN_NOALLOC_MAX_ROWS)
        return na_fail(err, err_len, DGNN_ERR_TOO_LARGE,
                       "dgnn_row_offsets_noalloc: n_rows = %lld exceeds the fixed scratch of %lld rows "
                       "(DGNN_NOALLOC_MAX_ROWS): split the input or use dgnn_build",
                       (long long)n_rows, (long long)DGNN_NOALLOC_MAX_ROWS);
    if (device < 0 || device >= NA_MAX_DEVICES)
        return na_fail(err, err_len, DGNN_ERR_INVALID_ARG, "dgnn_row_offsets_noalloc: device %d out of range [0, %d)",
                       device, NA_MAX_DEVICES);
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess)
        return na_fail(err, err_len, DGNN_ERR_CUDA, "dgnn_row_offsets_noalloc: cudaSetDevice(%d) failed: %s", device,
                       cudaGetErrorString(e));
    long long* sums = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_na_mutex);
        if (!g_na_ptr[device]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_na_tile_sums);  // (device `device` is current)
            if (e != cudaSuccess || !p) {
                (void)cudaGetLastError();
                return na_fail(err, err_len, DGNN_ERR_CUDA,
                               "dgnn_row_offsets_noalloc: cudaGetSymbolAddress failed on device %d: %s", device,
                               cudaGetErrorString(e));
            }
            g_na_ptr[device] = static_cast<long long*>(p);
        }
        sums = g_na_ptr[device];
    }
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const int* deg = reinterpret_cast<const int*>(degrees);
    long long* out = reinterpret_cast<long long*>(row_ptr);
    if (n_rows == 0) {
        k_na_empty<<<1, 1, 0, st>>>(out);
    } else {
        const long long nt = (n_rows + NA_TILE - 1) / NA_TILE;
        k_na_tile_sums<<<(unsigned)nt, NA_THREADS, 0, st>>>(deg, n_rows, sums);
        k_na_scan_sums<<<1, 1024, 0, st>>>(sums, nt);
        k_na_offsets<<<(unsigned)nt, NA_THREADS, 0, st>>>(deg, n_rows, sums, out, nt);
    }
    e = cudaGetLastError();
    if (e != cudaSuccess)
        return na_fail(err, err_len, DGNN_ERR_CUDA, "dgnn_row_offsets_noalloc: launch failed (%lld rows): %s", (long long)n_rows,
                       cudaGetErrorString(e));
    return DGNN_OK;
}
