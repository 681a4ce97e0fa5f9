// Device-side generators of the per-episode INPUT streams.
//
// The reference draws these from numpy's MT19937 one Python call at a time:
//   CGM noise       CGMNoise / noise15_iter            simglucose/sensor/noise_gen.py:15-97
//   meal schedules  RandomScenario.create_scenario     simglucose/simulation/scenario_gen.py:33-60
//   parameter jitter (BASELINE.json configs[3]: +-10 % per used parameter)
// For seeds that must replay the reference bit for bit the host generators in
// simglucose_b200/streams.py are used instead; the kernels here produce the same DISTRIBUTIONS from a
// counter-based Philox4x32-10 stream (key = seed, counter = (global episode index, draw index)), so
// any shard of a population can be generated independently on any GPU.
#include "../../include/simglucose_b200.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cuda_runtime.h>

#include "sg_philox.cuh"

namespace {

using sgp::Philox;

constexpr int MAX_W = 151 * 11;   // Navigator: 151 samples per window

// CGMNoise: AR(1)-type chain at 15-minute knots, Johnson SU transform, cubic interpolation to the
// sensor grid expressed as a constant matrix W [(nsample) x 11] (row 0 = the carried knot, dropped).
// Which stream an episode draws from, and whether it is written at all.  Positional: episodes (ep0 + e) / group share a
// stream (the 30 patients of one seed group).  Keyed: key[e] names the stream (a vector environment restarts its
// episodes one by one: key = env index + num_envs * episodes started so far) and only columns with mask[e] != 0 are
// rewritten (mask null = all).
struct StreamSel {
    int64_t ep0, group;
    const int64_t *key;
    const int32_t *mask;
    __device__ __forceinline__ bool skip(int64_t e) const { return mask && mask[e] == 0; }
    __device__ __forceinline__ uint64_t stream(int64_t e) const
    {
        return key ? (uint64_t)key[e] : (uint64_t)((ep0 + e) / group);
    }
};

// The arithmetic of one CGMNoise knot, written with rounding-explicit intrinsics: the file is compiled with FMA
// contraction, and a contraction across these statements (pacf * (ar + z) - gamma as one fma) would make the value
// depend on the shape of the surrounding code -- the two kernels below must agree bit for bit.  numpy evaluates the
// same expressions without contraction (noise_gen.py:24, 86-88).
__device__ __forceinline__ double ar_next(double pacf, double ar, double z) { return __dmul_rn(pacf, __dadd_rn(ar, z)); }
__device__ __forceinline__ double johnson(double ar, double gamma, double lambda, double delta, double xi)
{
    return __dadd_rn(xi, __dmul_rn(lambda, sinh(__ddiv_rn(__dsub_rn(ar, gamma), delta))));
}
__device__ __forceinline__ double interp11(const double *w, const double (&knots)[11])
{
    double v = 0.0;
#pragma unroll
    for (int k = 0; k < 11; k++) v = fma(w[k], knots[k], v);
    return v;
}

__global__ void __launch_bounds__(128) noise_kernel(double *out, int64_t n, int64_t n_noise, int nsample,
                                                   const double *W, double pacf, double gamma, double lambda,
                                                   double delta, double xi, uint64_t seed, const StreamSel sel)
{
    extern __shared__ double sW[];
    for (int i = threadIdx.x; i < nsample * 11; i += blockDim.x) sW[i] = W[i];
    __syncthreads();
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n || sel.skip(e)) return;
    Philox rng(seed, sel.stream(e), 0x6e6f6973u /* 'nois' */);
    double ar = rng.normal();                                   // noise_gen.py:84-85
    double knots[11];
    knots[10] = johnson(ar, gamma, lambda, delta, xi);          // _noise_init, noise_gen.py:24
    int64_t pos = 0;
    while (pos < n_noise) {
        knots[0] = knots[10];                                   // noise_gen.py:33-36
#pragma unroll
        for (int k = 1; k <= 10; k++) {
            ar = ar_next(pacf, ar, rng.normal());               // noise_gen.py:86-88
            knots[k] = johnson(ar, gamma, lambda, delta, xi);
        }
        for (int r = 1; r < nsample && pos < n_noise; r++, pos++)     // noise[1:], noise_gen.py:47
            out[pos * n + e] = interp11(sW + r * 11, knots);
    }
}

// The same streams for a FEW episodes of a large population (a vector environment restarting the episodes that ended
// in this env-step: mask[e] != 0 for well under 1 % of them).  One thread per episode would leave 31 lanes of a warp idle
// behind a lane that walks ~1000 normal deviates and writes ~480 samples 8 * n bytes apart.  Here a block of four warps
// looks at its 128 episodes, lists the masked ones, and works through the list one episode at a time with ALL its threads:
// thread w owns the 15-minute window w (128 windows per round):
//   1. thread w draws the ten deviates of window w by random access into the Philox stream (Philox::block_at) -- the
//      expensive part (inverse normal CDF), 128 windows at a time;
//   2. the AR chain ar <- pacf * (ar + z) is a serial recurrence: one thread walks the round's deviates in shared memory
//      (two dependent FP64 operations per deviate) and overwrites each with the chain value;
//   3. thread w transforms its eleven knots and interpolates the samples of its window.
// Same deviates, same chain, same expressions as noise_kernel: bit-identical streams (tests/test_gpu_generators.py).
constexpr int NM_THREADS = 128;

// One episode's stream by the whole block (see above): z [NM_THREADS * 10] and ar_carry live in shared memory.
__device__ __forceinline__ void noise_episode_by_block(double *out, int64_t n, int64_t n_noise, int nsample, const double *sW,
                                                       double pacf, double gamma, double lambda, double delta, double xi,
                                                       uint64_t seed, uint64_t stream, int64_t e, double *z, double *ar_carry)
{
    const int t = threadIdx.x;
    const int per_window = nsample - 1;
    const int64_t n_windows = (n_noise + per_window - 1) / per_window;
    const Philox rng(seed, stream, 0x6e6f6973u /* 'nois' */);
    uint32_t w4[4];
    rng.block_at(0, w4);
    double ar0 = Philox::normal_of(w4, 0);                  // deviate 0: _noise_init, noise_gen.py:84-85
    for (int64_t w0 = 0; w0 < n_windows; w0 += NM_THREADS) {
        const int64_t win = w0 + t;
        __syncthreads();                                    // z and ar_carry of the round before are consumed
        if (win < n_windows) {
            // deviates 1 + 10 win .. 10 + 10 win: an odd first index, i.e. the second half of one draw block, four
            // whole blocks, the first half of a sixth
            const uint32_t b0 = (uint32_t)((1 + 10 * win) >> 1);
            rng.block_at(b0, w4);
            z[t * 10 + 0] = Philox::normal_of(w4, 1);
#pragma unroll
            for (int b = 1; b <= 4; b++) {
                rng.block_at(b0 + b, w4);
                z[t * 10 + 2 * b - 1] = Philox::normal_of(w4, 0);
                z[t * 10 + 2 * b] = Philox::normal_of(w4, 1);
            }
            rng.block_at(b0 + 5, w4);
            z[t * 10 + 9] = Philox::normal_of(w4, 0);
        }
        __syncthreads();
        // the chain: ONE thread walks the round's deviates and leaves the chain value in each deviate's slot (all
        // 128 threads doing it redundantly would saturate the FP64 pipe of an SM that holds sixteen such blocks)
        const int n_win_round = (int)min((int64_t)NM_THREADS, n_windows - w0);
        if (t == 0) {
            // software-pipelined by hand: the ten deviates of the NEXT window are loaded before this window's chain
            // values are stored over their own deviates, so no load ever waits behind a store to the same array and
            // the loop runs at the pace of its dependent chain (20 FP64 operations per window)
            double ar = ar0, zn[10];
#pragma unroll
            for (int k = 0; k < 10; k++) zn[k] = z[k];
            for (int w = 0; w < n_win_round; w++) {
                double zz[10];
#pragma unroll
                for (int k = 0; k < 10; k++) zz[k] = zn[k];
                if (w + 1 < n_win_round) {
#pragma unroll
                    for (int k = 0; k < 10; k++) zn[k] = z[(w + 1) * 10 + k];
                }
#pragma unroll
                for (int k = 0; k < 10; k++) {
                    ar = ar_next(pacf, ar, zz[k]);          // noise_gen.py:86-88
                    z[w * 10 + k] = ar;
                }
            }
            *ar_carry = ar;
        }
        __syncthreads();
        const double before = t == 0 ? ar0 : z[t * 10 - 1];
        if (win < n_windows) {
            double knots[11];
            knots[0] = johnson(before, gamma, lambda, delta, xi);           // the knot carried over, noise_gen.py:33-36
#pragma unroll
            for (int k = 1; k <= 10; k++) knots[k] = johnson(z[t * 10 + k - 1], gamma, lambda, delta, xi);
            for (int r = 1; r < nsample; r++) {
                const int64_t pos = win * per_window + r - 1;
                if (pos >= n_noise) break;
                out[pos * n + e] = interp11(sW + r * 11, knots);
            }
        }
        __syncthreads();
        ar0 = *ar_carry;
    }
}

__global__ void __launch_bounds__(NM_THREADS) noise_masked_kernel(double *out, int64_t n, int64_t n_noise, int nsample,
                                                                 const double *W, double pacf, double gamma,
                                                                 double lambda, double delta, double xi, uint64_t seed,
                                                                 const StreamSel sel)
{
    extern __shared__ double sW[];                         // nsample * 11 interpolation weights
    __shared__ double z[NM_THREADS * 10];                  // deviates of a round
    __shared__ int list[NM_THREADS];
    __shared__ int n_list;
    __shared__ double ar_carry;                            // chain value at the end of a round
    const int t = threadIdx.x, lane = t & 31;
    const int64_t mine = (int64_t)blockIdx.x * blockDim.x + t;
    if (t == 0) n_list = 0;
    __syncthreads();
    const bool sel_me = mine < n && !sel.skip(mine);
    const unsigned ballot = __ballot_sync(0xffffffffu, sel_me);
    int base_slot = 0;
    if (lane == 0 && ballot) base_slot = atomicAdd(&n_list, __popc(ballot));
    base_slot = __shfl_sync(0xffffffffu, base_slot, 0);
    if (sel_me) list[base_slot + __popc(ballot & ((1u << lane) - 1u))] = t;
    __syncthreads();
    const int count = n_list;
    if (count == 0) return;
    for (int i = t; i < nsample * 11; i += blockDim.x) sW[i] = W[i];
    for (int li = 0; li < count; li++) {                   // (the order inside the list cannot reach a result)
        const int64_t e = (int64_t)blockIdx.x * blockDim.x + list[li];
        noise_episode_by_block(out, n, n_noise, nsample, sW, pacf, gamma, lambda, delta, xi, seed, sel.stream(e), e, z,
                               &ar_carry);
    }
}

// The same for a LIST of episodes gathered over the whole population (sg_vec_finish appends the environments that
// ended): every block takes listed episodes in turn, so the time of the launch is (listed / resident blocks) episodes
// instead of the largest number of masked episodes that happen to share a block of 128.
__global__ void __launch_bounds__(NM_THREADS) noise_listed_kernel(double *out, int64_t n, int64_t n_noise, int nsample,
                                                                 const double *W, double pacf, double gamma,
                                                                 double lambda, double delta, double xi, uint64_t seed,
                                                                 const int64_t *key, const int32_t *list,
                                                                 const int32_t *count)
{
    extern __shared__ double sW[];
    __shared__ double z[NM_THREADS * 10];
    __shared__ double ar_carry;
    const int n_listed = *count;
    if ((int)blockIdx.x >= n_listed) return;
    for (int i = threadIdx.x; i < nsample * 11; i += blockDim.x) sW[i] = W[i];
    for (int li = blockIdx.x; li < n_listed; li += gridDim.x) {
        const int64_t e = list[li];
        noise_episode_by_block(out, n, n_noise, nsample, sW, pacf, gamma, lambda, delta, xi, seed, (uint64_t)key[e], e, z,
                               &ar_carry);
    }
}

// RandomScenario: one schedule per calendar day, six candidate meals (scenario_gen.py:38-58).
__global__ void __launch_bounds__(128) scenario_kernel(int32_t *meal_minute, double *meal_grams, int64_t n,
                                                      int max_meals, int start_minute_all, const int32_t *start_minute_rows,
                                                      int horizon, uint64_t seed, const StreamSel sel)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n || sel.skip(e)) return;
    const int start_minute = start_minute_rows ? start_minute_rows[e] : start_minute_all;
    const double prob[6] = {0.95, 0.3, 0.95, 0.3, 0.95, 0.3};
    const double lb[6] = {300, 540, 600, 840, 960, 1200}, ub[6] = {540, 600, 840, 960, 1200, 1380};
    const double mu[6] = {420, 570, 720, 900, 1080, 1290}, sd[6] = {60, 30, 60, 30, 60, 30};
    const double amu[6] = {45, 10, 70, 10, 80, 10}, asd[6] = {10, 5, 10, 5, 10, 5};
    Philox rng(seed, sel.stream(e), 0x6d65616cu /* 'meal' */);
    int pos = 0, last = -1;
    const int n_days = (start_minute + horizon + 1439) / 1440;
    for (int day = 0; day < n_days; day++) {
        for (int i = 0; i < 6; i++) {
            const double u = rng.uniform(), ut = rng.uniform(), za = rng.normal();
            if (!(u < prob[i])) continue;
            const double Fa = normcdf((lb[i] - mu[i]) / sd[i]), Fb = normcdf((ub[i] - mu[i]) / sd[i]);
            // (explicit fma: the masked kernel below must round these three expressions the same way)
            double t = fma(sd[i], normcdfinv(fma(ut, Fb - Fa, Fa)), mu[i]);
            t = fmin(fmax(t, lb[i]), ub[i]);
            const int tmin = (int)rint(t);                               // np.round
            const double grams = fmax(rint(fma(asd[i], za, amu[i])), 0.0);  // max(round(normal), 0)
            const int m = day * 1440 + tmin - start_minute;              // episode minute
            if (m < 0 || m >= horizon || m <= last) continue;            // duplicates: first wins (list.index)
            if (grams == 0.0) { last = m; continue; }                    // Action(meal=0) is "no meal" (get_action)
            if (pos < max_meals) {
                meal_minute[(int64_t)pos * n + e] = m;
                meal_grams[(int64_t)pos * n + e] = grams;
                pos++; last = m;
            }
        }
    }
    for (; pos < max_meals; pos++) {
        meal_minute[(int64_t)pos * n + e] = -1;
        meal_grams[(int64_t)pos * n + e] = 0.0;
    }
}

// The same schedules for a FEW episodes of a large population (mask[e] != 0 for well under 1 %: the episodes a vector
// environment restarts in this env-step).  With one thread per episode a warp would wait for its one masked lane to walk
// up to 18 candidate meals (three inverse normal CDFs each); here the lanes of a warp take one candidate each -- candidate
// c = 6 * day + slot consumes the words 6 c .. 6 c + 5 of the Philox stream, so any lane can draw it (Philox::block_at) --
// and lane 0 then applies the sequential rules (in range, one meal per minute, first wins) in the order of the list.
// Bit-identical to scenario_kernel (tests/test_gpu_generators.py).
// One episode's schedule by one warp (see above); cand_m / cand_g: 32 slots of shared memory of this warp.
__device__ __forceinline__ void scenario_episode_by_warp(int32_t *meal_minute, double *meal_grams, int64_t n, int max_meals,
                                                         int start_minute, int horizon, uint64_t seed, uint64_t stream,
                                                         int64_t e, int *cand_m, double *cand_g)
{
    const double prob[6] = {0.95, 0.3, 0.95, 0.3, 0.95, 0.3};
    const double lb[6] = {300, 540, 600, 840, 960, 1200}, ub[6] = {540, 600, 840, 960, 1200, 1380};
    const double mu[6] = {420, 570, 720, 900, 1080, 1290}, sd[6] = {60, 30, 60, 30, 60, 30};
    const double amu[6] = {45, 10, 70, 10, 80, 10}, asd[6] = {10, 5, 10, 5, 10, 5};
    const int lane = threadIdx.x & 31;
    const Philox rng(seed, stream, 0x6d65616cu /* 'meal' */);
    const int n_cand = 6 * ((start_minute + horizon + 1439) / 1440);
    int pos = 0, last = -1;
    for (int c0 = 0; c0 < n_cand; c0 += 32) {
        const int c = c0 + lane;
        int m = INT_MIN;
        double grams = 0.0;
        if (c < n_cand) {
            const int day = c / 6, i = c - 6 * day;
            // words 6 c .. 6 c + 5 of the stream, handed out from the top of each draw block: (3, 2), (1, 0)
            uint32_t w0[4], w1[4];
            const uint32_t b = (uint32_t)((6 * c) >> 2);
            rng.block_at(b, w0);
            rng.block_at(b + 1, w1);
            double u, ut, za;
            if ((c & 1) == 0) {
                u = Philox::uniform_of(w0[3], w0[2]); ut = Philox::uniform_of(w0[1], w0[0]);
                za = normcdfinv(Philox::uniform_of(w1[3], w1[2]));
            } else {
                u = Philox::uniform_of(w0[1], w0[0]); ut = Philox::uniform_of(w1[3], w1[2]);
                za = normcdfinv(Philox::uniform_of(w1[1], w1[0]));
            }
            if (u < prob[i]) {
                const double Fa = normcdf((lb[i] - mu[i]) / sd[i]), Fb = normcdf((ub[i] - mu[i]) / sd[i]);
                double tt = fma(sd[i], normcdfinv(fma(ut, Fb - Fa, Fa)), mu[i]);
                tt = fmin(fmax(tt, lb[i]), ub[i]);
                grams = fmax(rint(fma(asd[i], za, amu[i])), 0.0);
                m = day * 1440 + (int)rint(tt) - start_minute;
            }
        }
        cand_m[lane] = m;
        cand_g[lane] = grams;
        __syncwarp();
        if (lane == 0) {
            const int n_here = min(32, n_cand - c0);
            for (int k = 0; k < n_here; k++) {
                const int mk = cand_m[k];
                if (mk == INT_MIN || mk < 0 || mk >= horizon || mk <= last) continue;   // duplicates: first wins
                const double g = cand_g[k];
                if (g == 0.0) { last = mk; continue; }                                  // Action(meal=0) is "no meal"
                if (pos < max_meals) {
                    meal_minute[(int64_t)pos * n + e] = mk;
                    meal_grams[(int64_t)pos * n + e] = g;
                    pos++; last = mk;
                }
            }
        }
        __syncwarp();
    }
    if (lane == 0)
        for (; pos < max_meals; pos++) {
            meal_minute[(int64_t)pos * n + e] = -1;
            meal_grams[(int64_t)pos * n + e] = 0.0;
        }
}

__global__ void __launch_bounds__(128) scenario_masked_kernel(int32_t *meal_minute, double *meal_grams, int64_t n,
                                                             int max_meals, const int32_t *start_minute_rows,
                                                             int horizon, uint64_t seed, const StreamSel sel)
{
    __shared__ int list[128];
    __shared__ int n_list;
    __shared__ int cand_m[4][32];                          // episode minute of the candidate, INT_MIN = not drawn
    __shared__ double cand_g[4][32];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int64_t mine = (int64_t)blockIdx.x * blockDim.x + t;
    if (t == 0) n_list = 0;
    __syncthreads();
    const bool sel_me = mine < n && !sel.skip(mine);
    const unsigned ballot = __ballot_sync(0xffffffffu, sel_me);
    int base_slot = 0;
    if (lane == 0 && ballot) base_slot = atomicAdd(&n_list, __popc(ballot));
    base_slot = __shfl_sync(0xffffffffu, base_slot, 0);
    if (sel_me) list[base_slot + __popc(ballot & ((1u << lane) - 1u))] = t;
    __syncthreads();
    const int count = n_list;
    for (int li = warp; li < count; li += 4) {
        const int64_t e = (int64_t)blockIdx.x * blockDim.x + list[li];
        scenario_episode_by_warp(meal_minute, meal_grams, n, max_meals, start_minute_rows[e], horizon, seed, sel.stream(e),
                                 e, cand_m[warp], cand_g[warp]);
    }
}

// The same for a LIST of episodes gathered over the whole population: one warp per listed episode.
__global__ void __launch_bounds__(128) scenario_listed_kernel(int32_t *meal_minute, double *meal_grams, int64_t n,
                                                             int max_meals, const int32_t *start_minute_rows,
                                                             int horizon, uint64_t seed, const int64_t *key,
                                                             const int32_t *list, const int32_t *count)
{
    __shared__ int cand_m[4][32];
    __shared__ double cand_g[4][32];
    const int warp = threadIdx.x >> 5;
    const int n_listed = *count;
    for (int li = blockIdx.x * 4 + warp; li < n_listed; li += gridDim.x * 4) {
        const int64_t e = list[li];
        scenario_episode_by_warp(meal_minute, meal_grams, n, max_meals, start_minute_rows[e], horizon, seed,
                                 (uint64_t)key[e], e, cand_m[warp], cand_g[warp]);
    }
}

// params_soa[i][e] = table[row[e]][i] * (1 + amp * u), u ~ U(-1, 1) i.i.d. per parameter.
// Where and how an episode starts, as the gym wrapper draws it (envs/simglucose_gym_env.py:58-73): start hour
// ~ randint(0, 24), and T1DPatient(random_init_bg=True): x[3], x[4], x[12] ~ N(mean, 0.1 * mean) independently
// (multivariate_normal with a diagonal covariance, t1dpatient.py:256-270).  n_draws successive randomisations, each
// around the result of the one before with the SAME three deviates: the wrapper resets the patient twice (constructor
// and env.reset()) with RandomState(seed) re-created each time and init_state mutated in place.
__global__ void __launch_bounds__(128) episode_start_kernel(int32_t *start_minute, double *x0, const double *table_x0,
                                                           const int32_t *row, int64_t n, int n_draws, uint64_t seed,
                                                           const StreamSel sel)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n || sel.skip(e)) return;
    Philox rng(seed, sel.stream(e), 0x73746172u /* 'star' */);
    const double uh = rng.uniform();
    const double z3 = rng.normal(), z4 = rng.normal(), z12 = rng.normal();
    if (start_minute) {
        int hour = (int)(uh * 24.0);
        start_minute[e] = 60 * (hour > 23 ? 23 : hour);
    }
    if (x0) {
        const double *t = table_x0 + (int64_t)row[e] * SG_NX;
        double v3 = t[3], v4 = t[4], v12 = t[12];
        for (int d = 0; d < n_draws; d++) {
            v3 = v3 + sqrt(0.1 * v3) * z3;
            v4 = v4 + sqrt(0.1 * v4) * z4;
            v12 = v12 + sqrt(0.1 * v12) * z12;
        }
#pragma unroll
        for (int i = 0; i < SG_NX; i++) x0[(int64_t)i * n + e] = (i == 3) ? v3 : (i == 4) ? v4 : (i == 12) ? v12 : t[i];
    }
}

__global__ void __launch_bounds__(128) perturb_kernel(double *params_soa, const double *table, const int32_t *row,
                                                     int64_t n, double amp, uint64_t seed, int64_t ep0, int64_t group)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    Philox rng(seed, (uint64_t)((ep0 + e) / group), 0x70617261u /* 'para' */);
    const double *t = table + (int64_t)row[e] * SG_NPARAM;
#pragma unroll 4
    for (int i = 0; i < SG_NPARAM; i++) {
        const double u = 2.0 * rng.uniform() - 1.0;
        params_soa[(int64_t)i * n + e] = t[i] * (1.0 + amp * u);
    }
}

__global__ void __launch_bounds__(256) uniform_kernel(double *out, int64_t n, double lo, double hi, uint64_t seed,
                                                     uint64_t offset)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Philox rng(seed, offset + (uint64_t)i, 0x756e6966u /* 'unif' */);
    out[i] = lo + (hi - lo) * rng.uniform();
}

// Book-keeping of a lockstep vector environment around its per-environment resets (envs/vector_env.py), one launch each
// instead of a dozen elementwise ones.  vec_finish: after an env-step, which environments ended their episode (the
// step reported done, env.py:103, or the episode reached the horizon) and the Philox key of the episode that replaces
// each one.  vec_restart: after the masked reset, count the new episodes and merge the first observations in.
__global__ void __launch_bounds__(256) vec_finish_kernel(const double *row, int64_t n, int done_row, int obs_row,
                                                        int32_t *episode_steps, int32_t horizon,
                                                        const int64_t *episodes_started, int64_t env0, int64_t group,
                                                        int key_shift, int32_t *mask, int64_t *key, double *terminal_obs,
                                                        uint8_t *done, uint8_t *truncated, int32_t *list, int32_t *count)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    const int32_t steps = episode_steps[e] + 1;
    episode_steps[e] = steps;
    const bool d = row[(int64_t)done_row * n + e] != 0.0;
    const bool fin = d || steps >= horizon;
    mask[e] = fin ? 1 : 0;
    key[e] = (((env0 + e) / group) << key_shift) | episodes_started[e];
    terminal_obs[e] = row[(int64_t)obs_row * n + e];
    done[e] = d ? 1 : 0;
    truncated[e] = (fin && !d) ? 1 : 0;
    if (fin && list) list[atomicAdd(count, 1)] = (int32_t)e;     // (the order of the list cannot reach a result)
}

__global__ void __launch_bounds__(256) vec_restart_kernel(const int32_t *mask, int64_t n, int64_t *episodes_started,
                                                         int32_t *episode_steps, const double *obs_new,
                                                         const double *obs_step, double *obs_out)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    const bool fin = mask[e] != 0;
    if (fin) {
        episodes_started[e] += 1;
        episode_steps[e] = 0;
    }
    obs_out[e] = fin ? obs_new[e] : obs_step[e];
}

constexpr int LISTED_BLOCKS = 148 * 16;    // one resident wave of 128-thread blocks: the listed kernels stride over their list

inline unsigned blocks(int64_t n, int b) { return (unsigned)((n + b - 1) / b); }

}  // namespace

extern "C" {

int sg_gen_noise(double *out, int64_t n_episodes, int64_t n_noise, int32_t nsample, const double *W_dev,
                 const double *model_host, uint64_t seed, int64_t episode_offset, int64_t group_size, void *stream)
{
    if (!out || !W_dev || !model_host || n_episodes <= 0 || n_noise <= 0 || nsample < 2 || nsample * 11 > MAX_W ||
        group_size < 1)
        return SG_ERR_ARG;
    noise_kernel<<<blocks(n_episodes, 128), 128, sizeof(double) * nsample * 11, (cudaStream_t)stream>>>(
        out, n_episodes, n_noise, nsample, W_dev, model_host[0], model_host[1], model_host[2], model_host[3],
        model_host[4], seed, StreamSel{episode_offset, group_size, nullptr, nullptr});
    return cudaGetLastError() == cudaSuccess ? SG_OK : SG_ERR_CUDA;
}

int sg_gen_noise_keyed(double *out, int64_t n_episodes, int64_t n_noise, int32_t nsample, const double *W_dev,
                       const double *model_host, uint64_t seed, const int64_t *key, const int32_t *mask, void *stream)
{
    if (!out || !W_dev || !model_host || !key || n_episodes <= 0 || n_noise <= 0 || nsample < 2 || nsample * 11 > MAX_W)
        return SG_ERR_ARG;
    const StreamSel sel{0, 1, key, mask};
    if (mask)       // a few columns of a resident buffer: warp-cooperative
        noise_masked_kernel<<<blocks(n_episodes, NM_THREADS), NM_THREADS, sizeof(double) * nsample * 11, (cudaStream_t)stream>>>(
            out, n_episodes, n_noise, nsample, W_dev, model_host[0], model_host[1], model_host[2], model_host[3],
            model_host[4], seed, sel);
    else
        noise_kernel<<<blocks(n_episodes, 128), 128, sizeof(double) * nsample * 11, (cudaStream_t)stream>>>(
            out, n_episodes, n_noise, nsample, W_dev, model_host[0], model_host[1], model_host[2], model_host[3],
            model_host[4], seed, sel);
    return cudaGetLastError() == cudaSuccess ? SG_OK : SG_ERR_CUDA;
}

int sg_gen_scenarios(int32_t *meal_minute, double *meal_grams, int64_t n_episodes, int32_t max_meals,
                     int32_t start_minute, int32_t horizon_minutes, uint64_t seed, int64_t episode_offset,
                     int64_t group_size, void *stream)
{
    if (!meal_minute || !meal_grams || n_episodes <= 0 || max_meals <= 0 || start_minute < 0 ||
        start_minute >= 1440 || horizon_minutes <= 0 || group_size < 1)
        return SG_ERR_ARG;
    scenario_kernel<<<blocks(n_episodes, 128), 128, 0, (cudaStream_t)stream>>>(
        meal_minute, meal_grams, n_episodes, max_meals, start_minute, nullptr, horizon_minutes, seed,
        StreamSel{episode_offset, group_size, nullptr, nullptr});
    return cudaGetLastError() == cudaSuccess ? SG_OK : SG_ERR_CUDA;
}

int sg_gen_scenarios_keyed(int32_t *meal_minute, double *meal_grams, int64_t n_episodes, int32_t max_meals,
                           const int32_t *start_minute, int32_t horizon_minutes, uint64_t seed, const int64_t *key,
                           const int32_t *mask, void *stream)
{
    if (!meal_minute || !meal_grams || !start_minute || !key || n_episodes <= 0 || max_meals <= 0 || horizon_minutes <= 0)
        return SG_ERR_ARG;
    if (mask)       // a few columns of resident tables: warp-cooperative
        scenario_masked_kernel<<<blocks(n_episodes, 128), 128, 0, (cudaStream_t)stream>>>(
            meal_minute, meal_grams, n_episodes, max_meals, start_minute, horizon_minutes, seed, StreamSel{0, 1, key, mask});
    else
        scenario_kernel<<<blocks(n_episodes, 128), 128, 0, (cudaStream_t)stream>>>(
            meal_minute, meal_grams, n_episodes, max_meals, 0, start_minute, horizon_minutes, seed, StreamSel{0, 1, key, mask});
    return cudaGetLastError() == cudaSuccess ? SG_OK : SG_ERR_CUDA;
}

int sg_gen_noise_listed(double *out, int64_t n_episodes, int64_t n_noise, int32_t nsample, const double *W_dev,
                        const double *model_host, uint64_t seed, const int64_t *key, const int32_t *list,
                        const int32_t *count, void *stream)
{
    if (!out || !W_dev || !model_host || !key || !list || !count || n_episodes <= 0 || n_noise <= 0 || nsample < 2 ||
        nsample * 11 > MAX_W)
        return SG_ERR_ARG;
    const unsigned grid = (unsigned)std::min((int64_t)LISTED_BLOCKS, n_episodes);
    noise_listed_kernel<<<grid, NM_THREADS, sizeof(double) * nsample * 11, (cudaStream_t)stream>>>(
        out, n_episodes, n_noise, nsample, W_dev, model_host[0], model_host[1], model_host[2], model_host[3],
        model_host[4], seed, key, list, count);
    return cudaGetLastError() == cudaSuccess ? SG_OK : SG_ERR_CUDA;
}

int sg_gen_scenarios_listed(int32_t *meal_minute, double *meal_grams, int64_t n_episodes, int32_t max_meals,
                            const int32_t *start_minute, int32_t horizon_minutes, uint64_t seed, const int64_t *key,
                            const int32_t *list, const int32_t *count, void *stream)
{
    if (!meal_minute || !meal_grams || !start_minute || !key || !list || !count || n_episodes <= 0 || max_meals <= 0 ||
        horizon_minutes <= 0)
        return SG_ERR_ARG;
    const unsigned grid = (unsigned)std::min((int64_t)LISTED_BLOCKS, (n_episodes + 3) / 4);
    scenario_listed_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(meal_minute, meal_grams, n_episodes, max_meals,
                                                                   start_minute, horizon_minutes, seed, key, list, count);
    return cudaGetLastError() == cudaSuccess ? SG_OK : SG_ERR_CUDA;
}

int sg_gen_episode_starts(int32_t *start_minute, double *x0, const double *table_x0_dev, const int32_t *patient_row,
                          int64_t n_episodes, int32_t n_draws, uint64_t seed, const int64_t *key, const int32_t *mask,
                          void *stream)
{
    if ((!start_minute && !x0) || !key || n_episodes <= 0 || n_draws < 0 || (x0 && (!table_x0_dev || !patient_row)))
        return SG_ERR_ARG;
    episode_start_kernel<<<blocks(n_episodes, 128), 128, 0, (cudaStream_t)stream>>>(
        start_minute, x0, table_x0_dev, patient_row, n_episodes, n_draws, seed, StreamSel{0, 1, key, mask});
    return cudaGetLastError() == cudaSuccess ? SG_OK : SG_ERR_CUDA;
}

int sg_gen_perturbed_params(double *params_soa, const double *table_dev, const int32_t *patient_row,
                            int64_t n_episodes, double amplitude, uint64_t seed, int64_t episode_offset,
                            void *stream)
{
    if (!params_soa || !table_dev || !patient_row || n_episodes <= 0) return SG_ERR_ARG;
    perturb_kernel<<<blocks(n_episodes, 128), 128, 0, (cudaStream_t)stream>>>(
        params_soa, table_dev, patient_row, n_episodes, amplitude, seed, episode_offset, 1);
    return cudaGetLastError() == cudaSuccess ? SG_OK : SG_ERR_CUDA;
}

int sg_gen_uniform(double *out, int64_t n, double lo, double hi, uint64_t seed, uint64_t offset, void *stream)
{
    if (!out || n <= 0) return SG_ERR_ARG;
    uniform_kernel<<<blocks(n, 256), 256, 0, (cudaStream_t)stream>>>(out, n, lo, hi, seed, offset);
    return cudaGetLastError() == cudaSuccess ? SG_OK : SG_ERR_CUDA;
}

int sg_vec_finish(const double *row, int64_t n_envs, int32_t done_row, int32_t obs_row, int32_t *episode_steps,
                  int32_t horizon_steps, const int64_t *episodes_started, int64_t env_offset, int64_t group_size,
                  int32_t key_shift, int32_t *mask, int64_t *key, double *terminal_obs, uint8_t *done, uint8_t *truncated,
                  int32_t *list, int32_t *count, void *stream)
{
    if (!row || !episode_steps || !episodes_started || !mask || !key || !terminal_obs || !done || !truncated ||
        n_envs <= 0 || n_envs > INT32_MAX || group_size < 1 || key_shift < 1 || key_shift > 40 || done_row < 0 ||
        obs_row < 0 || (list != nullptr) != (count != nullptr))
        return SG_ERR_ARG;
    if (count && cudaMemsetAsync(count, 0, sizeof(int32_t), (cudaStream_t)stream) != cudaSuccess) return SG_ERR_CUDA;
    vec_finish_kernel<<<blocks(n_envs, 256), 256, 0, (cudaStream_t)stream>>>(
        row, n_envs, done_row, obs_row, episode_steps, horizon_steps, episodes_started, env_offset, group_size, key_shift,
        mask, key, terminal_obs, done, truncated, list, count);
    return cudaGetLastError() == cudaSuccess ? SG_OK : SG_ERR_CUDA;
}

int sg_vec_restart(const int32_t *mask, int64_t n_envs, int64_t *episodes_started, int32_t *episode_steps,
                   const double *obs_new, const double *obs_step, double *obs_out, void *stream)
{
    if (!mask || !episodes_started || !episode_steps || !obs_new || !obs_step || !obs_out || n_envs <= 0) return SG_ERR_ARG;
    vec_restart_kernel<<<blocks(n_envs, 256), 256, 0, (cudaStream_t)stream>>>(mask, n_envs, episodes_started,
                                                                              episode_steps, obs_new, obs_step, obs_out);
    return cudaGetLastError() == cudaSuccess ? SG_OK : SG_ERR_CUDA;
}

}  // extern "C"

// ---- FP64 pipe probe: the measured DFMA peak bench.py uses as the roofline denominator ----------
namespace {
__global__ void __launch_bounds__(256) dfma_probe_kernel(double *out, int64_t iters)
{
    const double a = 1.0 + 1e-9 * threadIdx.x, b = 1e-12;
    double v0 = 1.0, v1 = 1.1, v2 = 1.2, v3 = 1.3, v4 = 1.4, v5 = 1.5, v6 = 1.6, v7 = 1.7;
    for (int64_t i = 0; i < iters; i++) {
        v0 = fma(v0, a, b); v1 = fma(v1, a, b); v2 = fma(v2, a, b); v3 = fma(v3, a, b);
        v4 = fma(v4, a, b); v5 = fma(v5, a, b); v6 = fma(v6, a, b); v7 = fma(v7, a, b);
    }
    out[(int64_t)blockIdx.x * blockDim.x + threadIdx.x] = ((v0 + v1) + (v2 + v3)) + ((v4 + v5) + (v6 + v7));
}
}  // namespace

extern "C" int sg_probe_dfma(double *out, int32_t n_blocks, int32_t n_threads, int64_t iters, void *stream)
{
    if (!out || n_blocks <= 0 || n_threads <= 0 || n_threads > 256 || iters <= 0) return SG_ERR_ARG;
    dfma_probe_kernel<<<n_blocks, n_threads, 0, (cudaStream_t)stream>>>(out, iters);
    return cudaGetLastError() == cudaSuccess ? SG_OK : SG_ERR_CUDA;
}
