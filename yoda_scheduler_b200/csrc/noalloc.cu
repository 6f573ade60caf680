// noalloc.cu -- yoda_normalize_scores_noalloc: NormalizeScore for callers that must not allocate.
//
// Who this is for.  yoda_normalize_scores() belongs to a context and grows that context's device
// buffer with cudaMalloc when a longer list arrives.  Some callers cannot afford that:
//   * a scheduler thread that renormalises a node list while a FULL-mode batch of the same process
//     owns most of HBM -- cudaMalloc may fail there, and cudaMalloc/cudaFree serialise with every
//     stream of the device, so they would stall the evaluation that is running;
//   * a host that gave the device a fixed memory budget up front (MPS memory limits, a sidecar
//     that shares the GPU with training jobs) and treats any later allocation as a fault;
//   * code that holds no yoda_ctx at all, e.g. a health probe or a unit test of the rescale.
// This entry point needs no context and allocates nothing, on the host or on the device: its only
// working memory is the fixed array below, which lives in the library image.
//
// The price.  The array is loaded once into every GPU that loads the library and stays there until
// the process exits (16 KiB per device).  Its size caps the list length at
// YODA_NOALLOC_MAX_SCORES.  All calls on one device share it, so two calls whose kernels overlap on
// the GPU (different streams, or different host threads) corrupt each other's lowest/highest:
// serialise them.  The array keeps whatever the previous call left in it; that is harmless because
// a call reads back only the slots it has just written itself, so nothing has to be cleared.
//
// One caveat on "allocates nothing": the first call on a device makes the CUDA runtime load this
// module there (code and the array itself are placed in device memory by the driver at that moment).
// Make that first call at start-up -- before HBM fills up and before any stream capture -- and every
// later call on that device touches no allocator at all.
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "../../include/yoda_b200.h"

// thread-local message buffer behind yoda_last_error(), defined in runtime.cu
extern thread_local char yoda_g_err[512];
static int na_fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(yoda_g_err, sizeof yoda_g_err, fmt, ap);
    va_end(ap);
    return code;
}

#define NA_THREADS 256
#define NA_PER_BLOCK 1024                                   // scores handled by one block
#define NA_MAX_BLOCKS (YODA_NOALLOC_MAX_SCORES / NA_PER_BLOCK)  // 1024

// Per-block partial (lowest, highest) pairs.  In host code this name is only a stand-in for the
// device object, never a device address: kernels get the address from cudaGetSymbolAddress below.
static __device__ __align__(256) long long g_na_partials[2 * NA_MAX_BLOCKS];

// scheduler.go:134-145, first half: every block reduces its slice to one (lowest, highest) pair
__global__ void __launch_bounds__(NA_THREADS) k_na_partials(const long long *__restrict__ in, uint32_t n, long long *partials) {
    __shared__ long long s_lo[NA_THREADS / 32], s_hi[NA_THREADS / 32];
    const uint32_t beg = blockIdx.x * NA_PER_BLOCK, end = min(n, beg + NA_PER_BLOCK);
    long long lo = INT64_MAX, hi = INT64_MIN;
    for (uint32_t i = beg + threadIdx.x; i < end; i += NA_THREADS) {
        const long long v = in[i];
        lo = min(lo, v);
        hi = max(hi, v);
    }
    for (int o = 16; o; o >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xFFFFFFFFu, lo, o));
        hi = max(hi, __shfl_xor_sync(0xFFFFFFFFu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) {
        s_lo[threadIdx.x >> 5] = lo;
        s_hi[threadIdx.x >> 5] = hi;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < NA_THREADS / 32; ++w) {
            lo = min(lo, s_lo[w]);
            hi = max(hi, s_hi[w]);
        }
        partials[2 * blockIdx.x] = lo;
        partials[2 * blockIdx.x + 1] = hi;
    }
}

// second half: fold the n_blocks partial pairs, then rescale this block's slice (scheduler.go:147-153)
// `in` and `out` may be the same buffer (the header allows it), so neither is __restrict__
__global__ void __launch_bounds__(NA_THREADS) k_na_apply(const long long *in, long long *out, uint32_t n,
                                                         const long long *partials, uint32_t n_blocks) {
    __shared__ long long s_lo[NA_THREADS / 32], s_hi[NA_THREADS / 32];
    long long lo = INT64_MAX, hi = 0;  // highest starts at 0 (scheduler.go:134); lowest at scores[0], i.e. the plain minimum
    for (uint32_t b = threadIdx.x; b < n_blocks; b += NA_THREADS) {
        lo = min(lo, partials[2 * b]);
        hi = max(hi, partials[2 * b + 1]);
    }
    for (int o = 16; o; o >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xFFFFFFFFu, lo, o));
        hi = max(hi, __shfl_xor_sync(0xFFFFFFFFu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) {
        s_lo[threadIdx.x >> 5] = lo;
        s_hi[threadIdx.x >> 5] = hi;
    }
    __syncthreads();
    lo = s_lo[0];
    hi = s_hi[0];
    for (int w = 1; w < NA_THREADS / 32; ++w) {
        lo = min(lo, s_lo[w]);
        hi = max(hi, s_hi[w]);
    }
    if (hi == lo) lo = (long long)((unsigned long long)lo - 1ull);  // scheduler.go:147-149
    const long long den = (long long)((unsigned long long)hi - (unsigned long long)lo);
    const uint32_t beg = blockIdx.x * NA_PER_BLOCK, end = min(n, beg + NA_PER_BLOCK);
    for (uint32_t i = beg + threadIdx.x; i < end; i += NA_THREADS) {
        // Go int64 arithmetic wraps; MinInt64 / -1 stays MinInt64; division truncates toward zero
        const long long num = (long long)(((unsigned long long)in[i] - (unsigned long long)lo) * 100ull);
        out[i] = (num == INT64_MIN && den == -1) ? num : num / den;
    }
}

// device address of g_na_partials, looked up once per device with that device current
#define NA_MAX_DEVICES 64
static long long *g_na_addr[NA_MAX_DEVICES];
static std::mutex g_na_mu;

extern "C" int yoda_normalize_scores_noalloc(const int64_t *d_scores_in, int64_t *d_scores_out, uint32_t n, void *cuda_stream) {
    if (!d_scores_in || !d_scores_out) return na_fail(YODA_E_ARG, "yoda_normalize_scores_noalloc: null device pointer");
    if (n == 0) return na_fail(YODA_E_ARG, "empty score list (the reference indexes scores[0], scheduler.go:135)");
    if (n > YODA_NOALLOC_MAX_SCORES)
        return na_fail(YODA_E_RANGE, "yoda_normalize_scores_noalloc: %u scores exceed the fixed scratch (at most %u); use yoda_normalize_scores",
                       n, (unsigned)YODA_NOALLOC_MAX_SCORES);
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return na_fail(YODA_E_CUDA, "cudaGetDevice failed: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= NA_MAX_DEVICES) return na_fail(YODA_E_RANGE, "device ordinal %d not supported", dev);
    long long *partials;
    {
        std::lock_guard<std::mutex> lk(g_na_mu);
        if (!g_na_addr[dev]) {  // set-up: first call on this device
            long long *p = nullptr;
            e = cudaGetSymbolAddress((void **)&p, g_na_partials);
            if (e != cudaSuccess || !p) return na_fail(YODA_E_CUDA, "cudaGetSymbolAddress(scratch) failed: %s", cudaGetErrorString(e));
            g_na_addr[dev] = p;
        }
        partials = g_na_addr[dev];
    }
    const uint32_t n_blocks = (n + NA_PER_BLOCK - 1) / NA_PER_BLOCK;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    k_na_partials<<<n_blocks, NA_THREADS, 0, st>>>((const long long *)d_scores_in, n, partials);
    k_na_apply<<<n_blocks, NA_THREADS, 0, st>>>((const long long *)d_scores_in, (long long *)d_scores_out, n, partials, n_blocks);
    e = cudaGetLastError();
    if (e != cudaSuccess) return na_fail(YODA_E_CUDA, "kernel launch failed: %s", cudaGetErrorString(e));
    return YODA_OK;
}
