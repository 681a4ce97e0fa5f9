// sg_reduce_stats_static -- the population aggregate for callers that must not allocate device memory.
//
// sg_reduce_stats keeps its per-block partial sums in a buffer that sg_create obtained from cudaMalloc.  Some hosts
// of this library cannot do that: a plugin that is only handed device pointers by its embedding application and is
// forbidden (by a memory budget, a pool allocator that must see every byte, or a real-time section where a driver
// allocation may block) to call cudaMalloc itself, and therefore cannot hold an sg_handle at all.  For them the
// same reduction runs here with its scratch in a file-scope __device__ array that is part of the library image:
// 592 rows x 16 statistics x 8 bytes = 75,776 bytes of every GPU's memory from the moment the library's module is
// loaded on that GPU until the process exits.  Nothing in this file calls cudaMalloc, cudaFree or any library
// routine that allocates.
//
// The two kernels restate reduce_partial_kernel / reduce_final_kernel of sg_kernels.cu operation for operation
// (same block size, same grid rule, same shuffle tree, same fixed order over the partial rows), so the result is
// bit-identical to sg_reduce_stats on the same statistics (tests/test_zz_no_alloc.py).
#include "../../include/simglucose_b200.h"

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
#include <mutex>

namespace sgk { void set_last_error(const char *what, cudaError_t ce); }   // sg_capi.cu: the text sg_last_error returns

namespace {

constexpr int NSTAT = SG_NSTAT;
constexpr int RBLOCK = 256;                                    // threads per block, as sg_kernels.cu
constexpr int MAX_BLOCKS = SG_REDUCE_STATIC_MAX_BLOCKS;        // 148 SMs x 4, the grid cap of sg_reduce_stats
constexpr int MAX_DEVICES = 64;
constexpr int STAT_FIRST_DONE = 13;

// The scratch itself.  In HOST code this name is only a stand-in the runtime uses to look the symbol up; it is not
// a device address and must never be handed to a kernel.  scratch_of() below asks the runtime for the real address
// on the current device and that is what the kernels receive.
static __device__ __align__(256) double g_reduce_scratch[MAX_BLOCKS * NSTAT];

__device__ __forceinline__ double combine(int i, double u, double v)
{
    if (i == 2) return fmin(u, v);
    if (i == 3) return fmax(u, v);
    return u + v;
}

__device__ __forceinline__ double identity(int i)
{
    return i == 2 ? INFINITY : (i == 3 ? -INFINITY : 0.0);
}

// One partial row per block: rows [0, gridDim.x) of `partials` are overwritten in full, nothing is accumulated
// into what an earlier call left there.
__global__ void __launch_bounds__(RBLOCK) reduce_partial_static_kernel(const double *__restrict__ stats, int64_t n,
                                                                       int64_t cap, double *__restrict__ partials)
{
    __shared__ double warp_part[RBLOCK / 32][NSTAT];
    double v[NSTAT];
#pragma unroll
    for (int i = 0; i < NSTAT; i++) v[i] = identity(i);
    for (int64_t e = (int64_t)blockIdx.x * RBLOCK + threadIdx.x; e < n; e += (int64_t)gridDim.x * RBLOCK) {
#pragma unroll
        for (int i = 0; i < NSTAT; i++) {
            double u = stats[(int64_t)i * cap + e];
            if (i == STAT_FIRST_DONE) u = (u >= 0.0) ? 1.0 : 0.0;     // counts the episodes that reported done
            v[i] = combine(i, v[i], u);
        }
    }
#pragma unroll
    for (int i = 0; i < NSTAT; i++) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1)
            v[i] = combine(i, v[i], __shfl_xor_sync(0xffffffffu, v[i], off));
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < NSTAT; i++) warp_part[warp][i] = v[i];
    }
    __syncthreads();
    if (threadIdx.x < NSTAT) {
        double acc = identity(threadIdx.x);
        for (int w = 0; w < RBLOCK / 32; w++) acc = combine(threadIdx.x, acc, warp_part[w][threadIdx.x]);
        partials[(int64_t)blockIdx.x * NSTAT + threadIdx.x] = acc;
    }
}

// Reads only the n_blocks rows the kernel above has just written, in row order.
__global__ void reduce_final_static_kernel(const double *__restrict__ partials, int n_blocks, double *__restrict__ out)
{
    if (threadIdx.x < NSTAT) {
        double acc = identity(threadIdx.x);
        for (int b = 0; b < n_blocks; b++) acc = combine(threadIdx.x, acc, partials[(int64_t)b * NSTAT + threadIdx.x]);
        out[threadIdx.x] = acc;
    }
}

int grid_rule(int64_t n)       // reduce_blocks() of sg_kernels.cu
{
    int64_t b = (n + RBLOCK - 1) / RBLOCK;
    if (b > MAX_BLOCKS) b = MAX_BLOCKS;
    if (b < 1) b = 1;
    return (int)b;
}

std::mutex g_setup_lock;
double *g_scratch_addr[MAX_DEVICES] = {};      // device address of g_reduce_scratch, per device ordinal

// Set-up, once per device, with that device current: resolve the array's address on this GPU.
cudaError_t scratch_of(int device, double **addr)
{
    std::lock_guard<std::mutex> hold(g_setup_lock);
    if (!g_scratch_addr[device]) {
        double *p = nullptr;
        cudaError_t ce = cudaGetSymbolAddress((void **)&p, g_reduce_scratch);
        if (ce != cudaSuccess) return ce;
        g_scratch_addr[device] = p;
    }
    *addr = g_scratch_addr[device];
    return cudaSuccess;
}

int fail(int code, const char *what, cudaError_t ce = cudaSuccess)
{
    sgk::set_last_error(what, ce);
    return code;
}

}  // namespace

extern "C" int sg_reduce_stats_static(const double *stats, int64_t n_episodes, int64_t capacity, double *agg_out,
                                      void *stream)
{
    if (!stats || !agg_out) return fail(SG_ERR_ARG, "sg_reduce_stats_static: null argument");
    if (n_episodes <= 0 || capacity < n_episodes)
        return fail(SG_ERR_ARG, "sg_reduce_stats_static: need 0 < n_episodes <= capacity");
    if (n_episodes > (int64_t)SG_REDUCE_STATIC_MAX_EPISODES) {
        char msg[200];
        snprintf(msg, sizeof msg, "sg_reduce_stats_static: %lld episodes exceed the limit of %lld of the built-in "
                 "scratch; use sg_reduce_stats (which allocates) or reduce in pieces",
                 (long long)n_episodes, (long long)SG_REDUCE_STATIC_MAX_EPISODES);
        return fail(SG_ERR_ARG, msg);
    }
    const int n_blocks = grid_rule(n_episodes);
    if (n_blocks > MAX_BLOCKS) return fail(SG_ERR_ARG, "sg_reduce_stats_static: grid exceeds the built-in scratch");
    int device = -1;
    cudaError_t ce = cudaGetDevice(&device);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_reduce_stats_static: cudaGetDevice", ce);
    if (device < 0 || device >= MAX_DEVICES) return fail(SG_ERR_ARG, "sg_reduce_stats_static: device ordinal beyond 63");
    double *scratch = nullptr;
    ce = scratch_of(device, &scratch);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_reduce_stats_static: cudaGetSymbolAddress", ce);
    cudaStream_t s = (cudaStream_t)stream;
    reduce_partial_static_kernel<<<n_blocks, RBLOCK, 0, s>>>(stats, n_episodes, capacity, scratch);
    ce = cudaGetLastError();
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_reduce_stats_static: launch (partial rows)", ce);
    reduce_final_static_kernel<<<1, 32, 0, s>>>(scratch, n_blocks, agg_out);
    ce = cudaGetLastError();
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_reduce_stats_static: launch (final pass)", ce);
    return SG_OK;
}
