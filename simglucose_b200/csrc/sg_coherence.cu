// Schedule-coherent launch order for populations whose episodes have meal schedules of their own
// (BASELINE.json configs[2] / configs[3]; SURVEY.md section 7 "warp divergence": 59.7 % lane efficiency).
//
// A minute with a meal onset / offset or a bolus edge costs its lane up to 36 DOPRI5 steps against 3 for an
// ordinary minute, and the other 31 lanes of the warp wait.  Nothing in the arithmetic can change (parity), but
// WHICH episodes share a warp is free: before a segment of env-steps the episodes are sorted by the minute of
// their next meal event (and the grams, which fix the end of the meal and the bolus), and thread i of the
// following launches owns episode perm[i].  Every array stays indexed by the episode, so per-episode results are
// bit-identical to the unsorted run; only the waiting changes.
//
// Keys (counting sort, n_keys = (window + 1) * 128 + 1 buckets):
//   [0, 128)                 the episode is in the middle of a meal: minutes of eating left (planned / EAT_RATE)
//   (dt + 1) * 128 + grams   its next announced meal starts in dt < window minutes (RandomScenario amounts are
//                            whole grams, scenario_gen.py:58; clamped to 127)
//   n_keys - 1               no meal inside the window
#include "sg_kernels.cuh"

namespace sgc {
using namespace sgk;

constexpr int SBLOCK = 256;

// The sort key of episode e (see the header comment and, for mode 2, the comment inside).
__device__ __forceinline__ int episode_key(const SortArgs &a, int64_t e)
{
    const int minute = a.si[(int64_t)SI_MINUTE * a.cap + e];
    int pos = a.si[(int64_t)SI_MEAL_POS * a.cap + e];
    const double planned = a.sd[(int64_t)SD_PLANNED * a.cap + e];
    int key = a.n_keys - 1;
    if (a.mode == 2) {
        // Adaptive key, for launch orders that are re-sorted every few env-steps (lockstep vector environments):
        //   [0, 64)                      DOPRI5 steps per env-step since the previous sort above the calm level of
        //                                3 per minute; 63 = an episode that was (re)started since then -- its first
        //                                minute costs 8 - 9 steps and the next twenty 4 - 5 (states that are exactly
        //                                zero are held to atol = 1e-12 until they have grown)
        //   64 + dt * 128 + grams        a meal announced in dt < window minutes (the onset minute costs up to 36)
        // Whatever makes an episode expensive -- a fresh start, eating, the hour after a meal, a jump of the insulin
        // rate -- shows in the steps it just took, and such episodes end up in the same warps.
        const double cum = a.stats[(int64_t)14 * a.cap + e];
        const double prev = a.prev_steps[e];
        a.prev_steps[e] = cum;
        const double per_step = (cum - prev) / (double)(a.steps_between > 0 ? a.steps_between : 1);
        int b = (cum < prev || cum == 0.0) ? 63 : (int)(per_step + 0.5) - 3 * a.st;
        key = b < 0 ? 0 : (b > 63 ? 63 : b);
        int mm = -1;
        while (pos < a.max_meals) {
            mm = a.meal_minute[(int64_t)pos * a.n + e];
            if (mm < 0 || mm >= minute) break;
            pos++;
        }
        if (planned <= 0.0 && pos < a.max_meals && mm >= minute && mm - minute < a.window) {
            int g = (int)a.meal_grams[(int64_t)pos * a.n + e];
            g = g < 0 ? 0 : (g > GRAM_BUCKETS - 1 ? GRAM_BUCKETS - 1 : g);
            key = 64 + (mm - minute) * GRAM_BUCKETS + g;
        }
    } else if (planned > 0.0) {
        const int left = (int)ceil(planned / 5.0);                    // EAT_RATE, t1dpatient.py:20
        key = left < GRAM_BUCKETS - 1 ? left : GRAM_BUCKETS - 1;
    } else {
        int mm = -1;
        while (pos < a.max_meals) {                                   // the table lookup of episode_steps
            mm = a.meal_minute[(int64_t)pos * a.n + e];
            if (mm < 0 || mm >= minute) break;
            pos++;
        }
        if (pos < a.max_meals && mm >= minute && mm - minute < a.window) {
            int g = (int)a.meal_grams[(int64_t)pos * a.n + e];
            g = g < 0 ? 0 : (g > GRAM_BUCKETS - 1 ? GRAM_BUCKETS - 1 : g);
            key = (mm - minute + 1) * GRAM_BUCKETS + g;
        }
    }
    return key;
}

// key of every episode + histogram (warp-aggregated atomics: most lanes of a warp share the "no meal" bucket)
__global__ void __launch_bounds__(SBLOCK) key_hist_kernel(const SortArgs a)
{
    const int64_t e = (int64_t)blockIdx.x * SBLOCK + threadIdx.x;
    const bool valid = e < a.n;
    const unsigned act = __ballot_sync(0xffffffffu, valid);
    if (!valid) return;
    const int key = episode_key(a, e);
    a.keys[e] = key;
    const unsigned peers = __match_any_sync(act, key);
    if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(a.hist + key, __popc(peers));
}

// The same two passes for SMALL key spaces (a launch order that is re-sorted every env-step tells only the next few
// minutes apart: 513 keys at a 3-minute segment): the histogram of a block lives in shared memory, so an episode costs
// a shared-memory atomic and a block a global one per key it holds -- nearly every episode of such a sort sits in the
// same handful of "calm" buckets, which serialised the warp-aggregated global atomics of the general path.
constexpr int SMALL_KEYS = 1024;

__global__ void __launch_bounds__(SBLOCK) key_hist_small_kernel(const SortArgs a)
{
    __shared__ int32_t sh[SMALL_KEYS];
    for (int i = threadIdx.x; i < a.n_keys; i += SBLOCK) sh[i] = 0;
    __syncthreads();
    const int64_t e = (int64_t)blockIdx.x * SBLOCK + threadIdx.x;
    if (e < a.n) {
        const int key = episode_key(a, e);
        a.keys[e] = key;
        atomicAdd(&sh[key], 1);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < a.n_keys; i += SBLOCK)
        if (sh[i]) atomicAdd(a.hist + i, sh[i]);
}

__global__ void __launch_bounds__(SBLOCK) scatter_small_kernel(const SortArgs a)
{
    __shared__ int32_t cnt[SMALL_KEYS], base[SMALL_KEYS];
    for (int i = threadIdx.x; i < a.n_keys; i += SBLOCK) cnt[i] = 0;
    __syncthreads();
    const int64_t e = (int64_t)blockIdx.x * SBLOCK + threadIdx.x;
    int key = 0, rank = 0;
    if (e < a.n) {
        key = a.keys[e];
        rank = atomicAdd(&cnt[key], 1);                  // arrival order inside the block: any order will do
    }
    __syncthreads();
    for (int i = threadIdx.x; i < a.n_keys; i += SBLOCK)
        if (cnt[i]) base[i] = atomicAdd(a.hist + i, cnt[i]);
    __syncthreads();
    if (e < a.n) a.perm[base[key] + rank] = (int32_t)e;
}

// exclusive prefix sum of the histogram in place: one block, every thread owns a run of consecutive buckets
__global__ void __launch_bounds__(1024) scan_kernel(int32_t *hist, int32_t n_keys)
{
    __shared__ int32_t warp_sum[32];
    const int t = threadIdx.x;
    const int chunk = (n_keys + 1023) / 1024;
    const int lo = t * chunk, hi = min(lo + chunk, n_keys);
    int32_t mine = 0;
    for (int i = lo; i < hi; i++) mine += hist[i];
    int32_t incl = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const int32_t v = __shfl_up_sync(0xffffffffu, incl, off);
        if ((t & 31) >= off) incl += v;
    }
    if ((t & 31) == 31) warp_sum[t >> 5] = incl;
    __syncthreads();
    if (t < 32) {
        int32_t w = warp_sum[t];
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int32_t v = __shfl_up_sync(0xffffffffu, w, off);
            if (t >= off) w += v;
        }
        warp_sum[t] = w;
    }
    __syncthreads();
    int32_t run = incl - mine + ((t >> 5) > 0 ? warp_sum[(t >> 5) - 1] : 0);
    for (int i = lo; i < hi; i++) {
        const int32_t c = hist[i];
        hist[i] = run;
        run += c;
    }
}

// perm[offset[key]++] = e, again warp-aggregated.  The order inside a bucket is whatever the atomics give: it
// cannot reach a result, only which equal-key episodes sit next to each other.
__global__ void __launch_bounds__(SBLOCK) scatter_kernel(const SortArgs a)
{
    const int64_t e = (int64_t)blockIdx.x * SBLOCK + threadIdx.x;
    const bool valid = e < a.n;
    const unsigned act = __ballot_sync(0xffffffffu, valid);
    if (!valid) return;
    const int key = a.keys[e];
    const unsigned peers = __match_any_sync(act, key);
    const int lane = threadIdx.x & 31, leader = __ffs(peers) - 1;
    int32_t base = 0;
    if (lane == leader) base = atomicAdd(a.hist + key, __popc(peers));
    base = __shfl_sync(peers, base, leader);
    a.perm[base + __popc(peers & ((1u << lane) - 1u))] = (int32_t)e;
}

cudaError_t launch_sort(const SortArgs &a, cudaStream_t s)
{
    cudaError_t e = cudaMemsetAsync(a.hist, 0, sizeof(int32_t) * a.n_keys, s);
    if (e != cudaSuccess) return e;
    const unsigned blocks = (unsigned)((a.n + SBLOCK - 1) / SBLOCK);
    const bool small = a.n_keys <= SMALL_KEYS;
    if (small) key_hist_small_kernel<<<blocks, SBLOCK, 0, s>>>(a);
    else key_hist_kernel<<<blocks, SBLOCK, 0, s>>>(a);
    scan_kernel<<<1, 1024, 0, s>>>(a.hist, a.n_keys);
    if (small) scatter_small_kernel<<<blocks, SBLOCK, 0, s>>>(a);
    else scatter_kernel<<<blocks, SBLOCK, 0, s>>>(a);
    return cudaGetLastError();
}

}  // namespace sgc
