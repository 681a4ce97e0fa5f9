// C-ABI of the batched simulator (include/simglucose_b200.h).  Host-side bookkeeping only: argument
// checks, the resident state allocation, and kernel launches on the caller's stream.
#include "../../include/simglucose_b200.h"
#include "sg_kernels.cuh"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

static_assert(SG_NTRACE == sgk::NTRACE, "trace columns");
static_assert(SG_NSTAT == sgk::NSTAT, "stat columns");
static_assert(SG_NCTRL == sgk::NCTRL, "controller constants");
static_assert(SG_NPOUT == sgk::NPOUT, "patient step rows");
static_assert(SG_NPARAM == 32 && SG_NX == 13, "table shapes");
static_assert(SG_ND == sgk::SD_COUNT && SG_NI == sgk::SI_COUNT, "resident state rows");
static_assert(SG_D_PID_PREV == sgk::SD_PID_PREV && SG_I_STATUS == sgk::SI_STATUS, "resident state order");

namespace {

thread_local char g_err[256] = "";

int fail(int code, const char *what, cudaError_t ce = cudaSuccess)
{
    if (ce != cudaSuccess) snprintf(g_err, sizeof g_err, "%s: %s", what, cudaGetErrorString(ce));
    else snprintf(g_err, sizeof g_err, "%s", what);
    return code;
}

}  // namespace

// For the entry points that live in other source files (sg_noalloc.cu): the same per-thread text sg_last_error returns.
namespace sgk {
void set_last_error(const char *what, cudaError_t ce) { fail(0, what, ce); }
}

struct sg_context {
    int device = 0;
    int64_t cap = 0, n = 0;
    uint32_t flags = 0;
    sgk::Hardware hw{};
    bool have_hw = false;
    double *sd = nullptr;
    int32_t *si = nullptr;
    double *stats = nullptr;
    double *table = nullptr, *table_x0 = nullptr;
    int32_t n_rows = 0;
    double *partials = nullptr;
    int32_t max_partials = 0;
    // borrowed from the last sg_reset
    const int32_t *patient_row = nullptr;
    const double *params_soa = nullptr;
    const int32_t *meal_minute = nullptr;
    const double *meal_grams = nullptr;
    int32_t max_meals = 0;
    const double *noise = nullptr;
    int64_t n_noise = 0;
    int64_t steps_done = 0;
    bool is_reset = false;
    int32_t t0 = 0;
    int64_t launches = 0;
    // heavy minutes put aside (sg_set_step_cap)
    int32_t step_cap = 0;
    double *defer_d = nullptr;
    int32_t *defer_i = nullptr, *defer_list = nullptr, *defer_count = nullptr;
    // schedule-coherent launch order (sg_set_coherence)
    int32_t coherence_steps = 0;       // env-steps between two sorts; 0 = episodes run in index order
    int32_t steps_since_sort = 0;
    bool perm_valid = false;
    int32_t *perm = nullptr, *sort_keys = nullptr, *sort_hist = nullptr;
    int32_t sort_n_keys = 0;           // keys of the current (segment, sample_time): what the sort kernels are told
    int32_t sort_hist_keys = 0;        // keys sort_hist was allocated for (>= sort_n_keys)
    int32_t sort_mode = 0;             // sg_set_sort_mode: 0 by next meal event, 1 adaptive (recent DOPRI5 steps)
    double *sort_prev_steps = nullptr;
    // episodes with positions of their own (sg_reset_masked): the host-side stream length check no longer applies
    bool per_episode_positions = false;
    int64_t steps_total = 0;           // env-steps launched since the last sg_reset (Philox counter of SG_CTRL_RANDOM)
};

namespace {

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev)
    {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

sgk::EpisodeArgs base_args(const sg_context *h)
{
    sgk::EpisodeArgs a{};
    a.n = h->n; a.cap = h->cap;
    a.sd = h->sd; a.si = h->si; a.stats = h->stats;
    a.params_soa = h->params_soa; a.table = h->table; a.table_x0 = h->table_x0;
    a.patient_row = h->patient_row;
    a.meal_minute = h->meal_minute; a.meal_grams = h->meal_grams; a.max_meals = h->max_meals;
    a.noise = h->noise; a.n_noise = h->n_noise;
    a.hw = h->hw;
    a.carry_h = (h->flags & SG_FLAG_CARRY_H) ? 1 : 0;
    a.step_cap = h->step_cap;
    a.perm = (h->coherence_steps > 0 && h->perm_valid) ? h->perm : nullptr;
    a.defer_d = h->defer_d; a.defer_i = h->defer_i; a.defer_list = h->defer_list; a.defer_count = h->defer_count;
    return a;
}

bool use_fma(const sg_context *h) { return (h->flags & SG_FLAG_FMA) != 0; }

// Buffers of the launch-order sort for a segment of `segment_steps` env-steps at a sensor period of `st` minutes.  The
// key space grows with segment_steps * st, so both sg_set_coherence and a later sg_set_hardware (another sample_time)
// come through here; h->sort_n_keys always matches what the next sort will be told.
int size_sort_buffers(sg_context *h, int32_t segment_steps, int32_t st, const char *who)
{
    static thread_local char msg[160];
    const int64_t window = (int64_t)segment_steps * st;
    if (window > 4096) {
        snprintf(msg, sizeof msg, "%s: at most 4096 minutes between two sorts (segment of %d env-steps x %d min)", who,
                 (int)segment_steps, (int)st);
        return fail(SG_ERR_ARG, msg);
    }
    const int32_t n_keys = sgc::key_count((int32_t)window);
    if (!h->perm) {
        cudaError_t ce = cudaMalloc(&h->perm, sizeof(int32_t) * h->cap);
        if (ce == cudaSuccess) ce = cudaMalloc(&h->sort_keys, sizeof(int32_t) * h->cap);
        if (ce != cudaSuccess) { snprintf(msg, sizeof msg, "%s: cudaMalloc", who); return fail(SG_ERR_NOMEM, msg, ce); }
    }
    if (n_keys > h->sort_hist_keys) {
        cudaFree(h->sort_hist);
        h->sort_hist = nullptr; h->sort_hist_keys = 0; h->sort_n_keys = 0;
        cudaError_t ce = cudaMalloc(&h->sort_hist, sizeof(int32_t) * n_keys);
        if (ce != cudaSuccess) { snprintf(msg, sizeof msg, "%s: cudaMalloc", who); return fail(SG_ERR_NOMEM, msg, ce); }
        h->sort_hist_keys = n_keys;
    }
    h->sort_n_keys = n_keys;
    return SG_OK;
}

}  // namespace

extern "C" {

int sg_version(void) { return SG_ABI_VERSION; }
const char *sg_last_error(void) { return g_err; }

int sg_create(int device, int64_t capacity, uint32_t flags, sg_handle *out)
{
    if (!out || capacity <= 0) return fail(SG_ERR_ARG, "sg_create: bad arguments");
    *out = nullptr;
    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev <= 0) return fail(SG_ERR_CUDA, "sg_create: no CUDA device", ce);
    if (device < 0 || device >= ndev) return fail(SG_ERR_ARG, "sg_create: device out of range");
    DeviceGuard g(device);
    if (!g.ok) return fail(SG_ERR_CUDA, "sg_create: cudaSetDevice failed");
    sg_context *h = new (std::nothrow) sg_context();
    if (!h) return fail(SG_ERR_NOMEM, "sg_create: host allocation failed");
    if (flags & ~(SG_FLAG_CARRY_H | SG_FLAG_FMA | SG_FLAG_FP32)) {
        delete h;
        return fail(SG_ERR_ARG, "sg_create: unknown flag");
    }
    h->device = device; h->cap = capacity; h->flags = flags;
    h->max_partials = sg_strict::reduce_blocks(capacity);
    ce = cudaMalloc(&h->sd, sizeof(double) * sgk::SD_COUNT * capacity);
    if (ce == cudaSuccess) ce = cudaMalloc(&h->si, sizeof(int32_t) * sgk::SI_COUNT * capacity);
    if (ce == cudaSuccess) ce = cudaMalloc(&h->stats, sizeof(double) * sgk::NSTAT * capacity);
    if (ce == cudaSuccess) ce = cudaMalloc(&h->partials, sizeof(double) * sgk::NSTAT * h->max_partials);
    if (ce != cudaSuccess) {
        sg_destroy(h);
        return fail(SG_ERR_NOMEM, "sg_create: cudaMalloc", ce);
    }
    *out = h;
    return SG_OK;
}

int sg_destroy(sg_handle h)
{
    if (!h) return SG_OK;
    DeviceGuard g(h->device);
    cudaFree(h->sd); cudaFree(h->si); cudaFree(h->stats); cudaFree(h->partials);
    cudaFree(h->table); cudaFree(h->table_x0);
    cudaFree(h->defer_d); cudaFree(h->defer_i); cudaFree(h->defer_list); cudaFree(h->defer_count);
    cudaFree(h->perm); cudaFree(h->sort_keys); cudaFree(h->sort_hist); cudaFree(h->sort_prev_steps);
    delete h;
    return SG_OK;
}

int sg_set_hardware(sg_handle h, const sg_hardware *hw)
{
    if (!h || !hw) return fail(SG_ERR_ARG, "sg_set_hardware: null argument");
    if (hw->sample_time < 1 || !(hw->pump[2] > 0.0) || !(hw->pump[5] > 0.0))
        return fail(SG_ERR_ARG, "sg_set_hardware: sample_time >= 1 and pump increments > 0 required");
    if (h->coherence_steps > 0 && hw->sample_time != h->hw.st) {
        // the sort keys count minutes: another sensor period is another key space (refused if it no longer fits)
        DeviceGuard g(h->device);
        const int rc = size_sort_buffers(h, h->coherence_steps, hw->sample_time, "sg_set_hardware (coherent launch order on)");
        if (rc != SG_OK) return rc;
        h->perm_valid = false;
    }
    h->hw.st = hw->sample_time;
    h->hw.smin = hw->sensor_min; h->hw.smax = hw->sensor_max;
    memcpy(h->hw.pump, hw->pump, sizeof h->hw.pump);
    h->have_hw = true;
    return SG_OK;
}

int sg_load_params(sg_handle h, const double *params_host, const double *x0_host, int32_t n_rows)
{
    if (!h || !params_host || !x0_host || n_rows <= 0) return fail(SG_ERR_ARG, "sg_load_params: bad arguments");
    DeviceGuard g(h->device);
    cudaFree(h->table); cudaFree(h->table_x0);
    h->table = h->table_x0 = nullptr; h->n_rows = 0;
    cudaError_t ce = cudaMalloc(&h->table, sizeof(double) * SG_NPARAM * n_rows);
    if (ce == cudaSuccess) ce = cudaMalloc(&h->table_x0, sizeof(double) * SG_NX * n_rows);
    if (ce != cudaSuccess) return fail(SG_ERR_NOMEM, "sg_load_params: cudaMalloc", ce);
    ce = cudaMemcpy(h->table, params_host, sizeof(double) * SG_NPARAM * n_rows, cudaMemcpyHostToDevice);
    if (ce == cudaSuccess)
        ce = cudaMemcpy(h->table_x0, x0_host, sizeof(double) * SG_NX * n_rows, cudaMemcpyHostToDevice);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_load_params: cudaMemcpy", ce);
    h->n_rows = n_rows;
    return SG_OK;
}

int sg_reset(sg_handle h, int64_t n_episodes, const int32_t *patient_row, const double *params_soa,
             const double *x0, const int32_t *meal_minute, const double *meal_grams, int32_t max_meals,
             const double *noise, int64_t n_noise, double *row0, void *stream)
{
    if (!h) return fail(SG_ERR_ARG, "sg_reset: null handle");
    if (!h->have_hw) return fail(SG_ERR_STATE, "sg_reset: call sg_set_hardware first");
    if (n_episodes <= 0 || n_episodes > h->cap) return fail(SG_ERR_ARG, "sg_reset: n_episodes outside 1..capacity");
    if (!params_soa && (!patient_row || !h->table)) return fail(SG_ERR_ARG, "sg_reset: need params_soa, or patient_row + sg_load_params");
    if (!x0 && (!patient_row || !h->table_x0)) return fail(SG_ERR_ARG, "sg_reset: need x0, or patient_row + sg_load_params");
    if (max_meals < 0 || (max_meals > 0 && (!meal_minute || !meal_grams))) return fail(SG_ERR_ARG, "sg_reset: meal table");
    if (!noise || n_noise < 2) return fail(SG_ERR_ARG, "sg_reset: the reset consumes two noise samples");
    DeviceGuard g(h->device);
    h->n = n_episodes;
    h->patient_row = patient_row; h->params_soa = params_soa;
    h->meal_minute = meal_minute; h->meal_grams = meal_grams; h->max_meals = max_meals;
    h->noise = noise; h->n_noise = n_noise;
    h->steps_done = 0;
    h->steps_total = 0;            // SG_CTRL_RANDOM counts env-steps from here: a reused handle draws what a fresh one does
    h->perm_valid = false;
    h->per_episode_positions = false;
    sgk::EpisodeArgs a = base_args(h);
    a.x0 = x0; a.row0 = row0; a.t0 = h->t0;
    cudaError_t ce = use_fma(h) ? sg_fma::launch_reset(a, (cudaStream_t)stream)
                                : sg_strict::launch_reset(a, (cudaStream_t)stream);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_reset: launch", ce);
    h->launches++;
    h->is_reset = true;
    return SG_OK;
}

// Counting sort of the episodes by their next meal event (sg_coherence.cu) -> h->perm.
static int sort_episodes(sg_handle h, cudaStream_t stream)
{
    sgc::SortArgs sa{};
    sa.n = h->n; sa.cap = h->cap; sa.sd = h->sd; sa.si = h->si;
    sa.meal_minute = h->meal_minute; sa.meal_grams = h->meal_grams; sa.max_meals = h->max_meals;
    sa.patient_row = h->patient_row;
    sa.mode = h->sort_mode == 1 ? 2 : 0;
    sa.stats = h->stats; sa.prev_steps = h->sort_prev_steps;
    sa.steps_between = h->coherence_steps; sa.st = h->hw.st;
    sa.window = h->coherence_steps * h->hw.st;
    sa.n_keys = h->sort_n_keys;
    if (!h->sort_hist || sgc::key_count(sa.window) != h->sort_n_keys || h->sort_n_keys > h->sort_hist_keys)
        return fail(SG_ERR_STATE, "launch-order sort: key space does not match the allocated histogram");
    sa.keys = h->sort_keys; sa.hist = h->sort_hist; sa.perm = h->perm;
    cudaError_t ce = sgc::launch_sort(sa, stream);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_set_coherence: sort launch", ce);
    h->launches += 3;
    h->perm_valid = true;
    h->steps_since_sort = 0;
    return SG_OK;
}

static int run_impl(sg_handle h, int32_t controller, const double *ctrl_host, const double *ctrl_rows,
                    const double *basal, const double *bolus, int64_t act_stride, int32_t n_steps,
                    double *trace, int32_t *counters, void *stream, const char *who,
                    const double *step_meals = nullptr)
{
    if (!h->is_reset) return fail(SG_ERR_STATE, who);
    // every episode started at the last sg_reset: the stream length is checked here; after sg_reset_masked the
    // episodes have positions of their own and the kernel's guard reports an exhausted stream per episode (-5)
    if (!h->per_episode_positions && h->steps_done + n_steps + 2 > h->n_noise)
        return fail(SG_ERR_ARG, "noise stream too short for the requested steps");
    if (n_steps == 0) return SG_OK;
    DeviceGuard g(h->device);
    sgk::EpisodeArgs a = base_args(h);
    a.controller = controller;
    a.step0 = h->steps_total;
    if (ctrl_host) memcpy(a.ctrl, ctrl_host, sizeof a.ctrl);
    a.ctrl_rows = ctrl_rows;
    a.act_basal = basal; a.act_bolus = bolus; a.act_stride = act_stride;
    a.n_steps = n_steps;
    a.trace = trace; a.counters = counters; a.step_meals = step_meals;
    // The run is cut into launches: one env-step each with a step cap (phase A / phase B pairs), the env-steps up to the
    // next sort with a coherent launch order, otherwise a single launch.
    int32_t k = 0;
    while (k < n_steps) {
        int32_t len = n_steps - k;
        if (h->step_cap > 0) len = 1;
        if (h->coherence_steps > 0) {
            if (!h->perm_valid || h->steps_since_sort >= h->coherence_steps) {
                int rc = sort_episodes(h, (cudaStream_t)stream);
                if (rc != SG_OK) return rc;
            }
            const int32_t left = h->coherence_steps - h->steps_since_sort;
            if (left < len) len = left;
        }
        sgk::EpisodeArgs b = a;
        b.perm = (h->coherence_steps > 0 && h->perm_valid) ? h->perm : nullptr;
        b.n_steps = len;
        b.step0 = h->steps_total;
        if (basal) { b.act_basal = basal + (int64_t)k * act_stride; b.act_bolus = bolus + (int64_t)k * act_stride; }
        if (trace) b.trace = trace + (int64_t)k * sgk::NTRACE * h->n;
        if (counters) b.counters = counters + (int64_t)k * h->hw.st * 2 * h->n;
        cudaError_t ce;
        if (h->step_cap > 0) {
            ce = use_fma(h) ? sg_fma::launch_run_capped(b, (cudaStream_t)stream)
                            : sg_strict::launch_run_capped(b, (cudaStream_t)stream);
            h->launches++;                                   // phase B
        } else if (h->flags & SG_FLAG_FP32) {
            ce = sg_fma::launch_run_fp32(b, (cudaStream_t)stream);
        } else {
            ce = use_fma(h) ? sg_fma::launch_run(b, (cudaStream_t)stream) : sg_strict::launch_run(b, (cudaStream_t)stream);
        }
        if (ce != cudaSuccess) return fail(SG_ERR_CUDA, who, ce);
        h->launches++;
        h->steps_done += len;
        h->steps_total += len;
        h->steps_since_sort += len;
        k += len;
    }
    return SG_OK;
}

int sg_run(sg_handle h, int32_t controller, const double *ctrl_host, const double *ctrl_rows,
           const double *actions, int32_t n_steps, double *trace, int32_t *counters, void *stream)
{
    if (!h) return fail(SG_ERR_ARG, "sg_run: null handle");
    if (n_steps < 0) return fail(SG_ERR_ARG, "sg_run: n_steps < 0");
    if (controller < SG_CTRL_ACTIONS || controller > SG_CTRL_RANDOM) return fail(SG_ERR_ARG, "sg_run: unknown controller");
    if (controller == SG_CTRL_ACTIONS && !actions) return fail(SG_ERR_ARG, "sg_run: actions required");
    if (controller != SG_CTRL_ACTIONS && !ctrl_host && !ctrl_rows) return fail(SG_ERR_ARG, "sg_run: controller constants required");
    if (controller == SG_CTRL_RANDOM) {
        // seed and first global index travel as doubles and are cast to uint64 in the kernel: whole numbers in [0, 2^53)
        if (!ctrl_host) return fail(SG_ERR_ARG, "sg_run: SG_CTRL_RANDOM takes its constants from ctrl_host");
        for (int i = 2; i < 4; i++)
            if (!(ctrl_host[i] >= 0.0 && ctrl_host[i] < 9007199254740992.0 && ctrl_host[i] == (double)(uint64_t)ctrl_host[i]))
                return fail(SG_ERR_ARG, "sg_run: SG_CTRL_RANDOM seed and first episode index must be whole numbers in [0, 2^53)");
    }
    return run_impl(h, controller, ctrl_host, ctrl_rows, actions, actions ? actions + h->n : nullptr, 2 * h->n,
                    n_steps, trace, counters, stream, "sg_run");
}

int sg_step(sg_handle h, const double *basal, const double *bolus, double *out, void *stream)
{
    if (!h) return fail(SG_ERR_ARG, "sg_step: null handle");
    if (!basal || !bolus) return fail(SG_ERR_ARG, "sg_step: basal and bolus required");
    return run_impl(h, SG_CTRL_ACTIONS, nullptr, nullptr, basal, bolus, 0, 1, out, nullptr, stream, "sg_step");
}

int sg_step_meals(sg_handle h, const double *basal, const double *bolus, const double *meals, double *out,
                  void *stream)
{
    if (!h) return fail(SG_ERR_ARG, "sg_step_meals: null handle");
    if (!basal || !bolus || !meals) return fail(SG_ERR_ARG, "sg_step_meals: basal, bolus and meals required");
    return run_impl(h, SG_CTRL_ACTIONS, nullptr, nullptr, basal, bolus, 0, 1, out, nullptr, stream,
                    "sg_step_meals", meals);
}

int sg_rebind_streams(sg_handle h, const int32_t *meal_minute, const double *meal_grams, int32_t max_meals,
                      const double *noise, int64_t n_noise)
{
    if (!h) return fail(SG_ERR_ARG, "sg_rebind_streams: null handle");
    if (!h->is_reset) return fail(SG_ERR_STATE, "sg_rebind_streams: call sg_reset first");
    if (max_meals < 0 || (max_meals > 0 && (!meal_minute || !meal_grams))) return fail(SG_ERR_ARG, "sg_rebind_streams: meal table");
    if (!noise || n_noise < h->steps_done + 2) return fail(SG_ERR_ARG, "sg_rebind_streams: noise stream shorter than what was consumed");
    h->meal_minute = meal_minute; h->meal_grams = meal_grams; h->max_meals = max_meals;
    h->noise = noise; h->n_noise = n_noise;
    return SG_OK;
}

int sg_set_step_cap(sg_handle h, int32_t cap)
{
    if (!h || cap < 0) return fail(SG_ERR_ARG, "sg_set_step_cap: bad arguments");
    if (cap > 0 && cap < 4) return fail(SG_ERR_ARG, "sg_set_step_cap: a cap below 4 steps would put ordinary minutes aside");
    if (cap > 0 && (h->flags & SG_FLAG_FP32))
        return fail(SG_ERR_ARG, "sg_set_step_cap: the heavy-minute kernels are FP64 only; not available with SG_FLAG_FP32");
    DeviceGuard g(h->device);
    if (cap > 0 && !h->defer_d) {
        cudaError_t ce = cudaMalloc(&h->defer_d, sizeof(double) * sgk::DEFER_ND * h->cap);
        if (ce == cudaSuccess) ce = cudaMalloc(&h->defer_i, sizeof(int32_t) * sgk::DEFER_NI * h->cap);
        if (ce == cudaSuccess) ce = cudaMalloc(&h->defer_list, sizeof(int32_t) * h->cap);
        if (ce == cudaSuccess) ce = cudaMalloc(&h->defer_count, sizeof(int32_t));
        if (ce != cudaSuccess) return fail(SG_ERR_NOMEM, "sg_set_step_cap: cudaMalloc", ce);
    }
    h->step_cap = cap;
    return SG_OK;
}

int sg_set_coherence(sg_handle h, int32_t segment_steps)
{
    if (!h || segment_steps < 0) return fail(SG_ERR_ARG, "sg_set_coherence: bad arguments");
    if (segment_steps > 0 && !h->have_hw) return fail(SG_ERR_STATE, "sg_set_coherence: call sg_set_hardware first");
    DeviceGuard g(h->device);
    if (segment_steps > 0) {
        const int rc = size_sort_buffers(h, segment_steps, h->hw.st, "sg_set_coherence");
        if (rc != SG_OK) return rc;
    }
    h->coherence_steps = segment_steps;
    h->perm_valid = false;
    return SG_OK;
}

int sg_set_sort_mode(sg_handle h, int32_t mode)
{
    if (!h || mode < 0 || mode > 1) return fail(SG_ERR_ARG, "sg_set_sort_mode: mode 0 (next meal event) or 1 (adaptive)");
    DeviceGuard g(h->device);
    if (mode == 1 && !h->sort_prev_steps) {
        cudaError_t ce = cudaMalloc(&h->sort_prev_steps, sizeof(double) * h->cap);
        if (ce == cudaSuccess) ce = cudaMemset(h->sort_prev_steps, 0, sizeof(double) * h->cap);
        if (ce != cudaSuccess) return fail(SG_ERR_NOMEM, "sg_set_sort_mode: cudaMalloc", ce);
    }
    h->sort_mode = mode;
    h->perm_valid = false;
    return SG_OK;
}

int sg_get_launch_order(sg_handle h, int32_t *perm_out, void *stream)
{
    if (!h || !perm_out) return fail(SG_ERR_ARG, "sg_get_launch_order: null argument");
    if (!h->is_reset || h->coherence_steps == 0 || !h->perm_valid)
        return fail(SG_ERR_STATE, "sg_get_launch_order: no sorted launch order (sg_set_coherence, then a step)");
    DeviceGuard g(h->device);
    cudaError_t ce = cudaMemcpyAsync(perm_out, h->perm, sizeof(int32_t) * h->n, cudaMemcpyDeviceToDevice, (cudaStream_t)stream);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_get_launch_order: copy", ce);
    return SG_OK;
}

int sg_reset_masked(sg_handle h, const int32_t *mask, const double *x0, double *row0, void *stream)
{
    if (!h || !mask) return fail(SG_ERR_ARG, "sg_reset_masked: null argument");
    if (!h->is_reset) return fail(SG_ERR_STATE, "sg_reset_masked: call sg_reset first");
    if (!x0 && (!h->patient_row || !h->table_x0)) return fail(SG_ERR_ARG, "sg_reset_masked: need x0, or patient_row + sg_load_params");
    DeviceGuard g(h->device);
    sgk::EpisodeArgs a = base_args(h);
    a.x0 = x0; a.row0 = row0; a.t0 = h->t0; a.reset_mask = mask;
    cudaError_t ce = use_fma(h) ? sg_fma::launch_reset(a, (cudaStream_t)stream)
                                : sg_strict::launch_reset(a, (cudaStream_t)stream);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_reset_masked: launch", ce);
    h->launches++;
    h->per_episode_positions = true;
    return SG_OK;
}

int sg_set_t0(sg_handle h, int32_t t0)
{
    if (!h || t0 < 0) return fail(SG_ERR_ARG, "sg_set_t0: bad arguments");
    h->t0 = t0;
    return SG_OK;
}

int sg_patient_step(sg_handle h, const double *cho, const double *insulin, double *out, void *stream)
{
    if (!h || !cho || !insulin) return fail(SG_ERR_ARG, "sg_patient_step: bad arguments");
    if (!h->is_reset) return fail(SG_ERR_STATE, "sg_patient_step: call sg_reset first");
    DeviceGuard g(h->device);
    sgk::PatientArgs pa{};
    pa.e = base_args(h);
    pa.cho = cho; pa.insulin = insulin; pa.out = out;
    cudaError_t ce = use_fma(h) ? sg_fma::launch_patient_step(pa, (cudaStream_t)stream)
                                : sg_strict::launch_patient_step(pa, (cudaStream_t)stream);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_patient_step: launch", ce);
    h->launches++;
    return SG_OK;
}

int sg_model(sg_handle h, int64_t n, const double *x, const double *cho, const double *insulin,
             const double *last_qsto, const double *last_food, const int32_t *patient_row,
             const double *params_soa, double *dxdt, void *stream)
{
    if (!h || n <= 0 || !x || !cho || !insulin || !last_qsto || !last_food || !dxdt)
        return fail(SG_ERR_ARG, "sg_model: bad arguments");
    if (!params_soa && (!patient_row || !h->table)) return fail(SG_ERR_ARG, "sg_model: need params_soa, or patient_row + sg_load_params");
    DeviceGuard g(h->device);
    sgk::ModelArgs m{};
    m.n = n; m.x = x; m.cho = cho; m.insulin = insulin; m.last_qsto = last_qsto; m.last_food = last_food;
    m.params_soa = params_soa; m.table = h->table; m.patient_row = patient_row; m.dxdt = dxdt;
    cudaError_t ce = use_fma(h) ? sg_fma::launch_model(m, (cudaStream_t)stream)
                                : sg_strict::launch_model(m, (cudaStream_t)stream);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_model: launch", ce);
    h->launches++;
    return SG_OK;
}

int sg_pump(const double *amount, double *out, int64_t n, double vmin, double vmax, double inc, void *stream)
{
    if (!amount || !out || n <= 0 || !(inc > 0.0)) return fail(SG_ERR_ARG, "sg_pump: bad arguments");
    sgk::ElementwiseArgs a{n, amount, out, {vmin, vmax, inc}, nullptr};
    cudaError_t ce = sg_strict::launch_pump(a, (cudaStream_t)stream);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_pump: launch", ce);
    return SG_OK;
}

int sg_risk_index(const double *bg, double *out, int64_t n, void *stream)
{
    if (!bg || !out || n <= 0) return fail(SG_ERR_ARG, "sg_risk_index: bad arguments");
    sgk::ElementwiseArgs a{n, bg, out, {0.0, 0.0, 0.0}, nullptr};
    cudaError_t ce = sg_strict::launch_risk(a, (cudaStream_t)stream);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_risk_index: launch", ce);
    return SG_OK;
}

int sg_risk_window(const double *bg, double *out, int64_t n, void *stream)
{
    if (!bg || !out || n <= 0) return fail(SG_ERR_ARG, "sg_risk_window: bad arguments");
    sgk::ElementwiseArgs a{n, bg, out, {0.0, 0.0, 0.0}, nullptr};
    cudaError_t ce = sg_strict::launch_risk_window(a, (cudaStream_t)stream);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_risk_window: launch", ce);
    return SG_OK;
}

int sg_cgm_measure(const double *bg, const double *noise, double *out, int64_t n, double sensor_min,
                   double sensor_max, void *stream)
{
    if (!bg || !noise || !out || n <= 0) return fail(SG_ERR_ARG, "sg_cgm_measure: bad arguments");
    sgk::ElementwiseArgs a{n, bg, out, {sensor_min, sensor_max, 0.0}, noise};
    cudaError_t ce = sg_strict::launch_cgm(a, (cudaStream_t)stream);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_cgm_measure: launch", ce);
    return SG_OK;
}

int sg_math_probe(int32_t kind, const double *a, const double *b, double *out, int64_t n, void *stream)
{
    if (kind < 0 || kind > 14 || !a || !b || !out || n <= 0) return fail(SG_ERR_ARG, "sg_math_probe: bad arguments");
    cudaError_t ce = sg_strict::launch_math_probe(kind, a, b, out, n, (cudaStream_t)stream);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_math_probe: launch", ce);
    return SG_OK;
}

static int copy_out(sg_handle h, void *dst, const void *src, size_t elem, int rows, void *stream, const char *who)
{
    if (!h || !dst) return fail(SG_ERR_ARG, who);
    if (!h->is_reset) return fail(SG_ERR_STATE, who);
    DeviceGuard g(h->device);
    cudaError_t ce = cudaMemcpy2DAsync(dst, elem * h->n, src, elem * h->cap, elem * h->n, rows,
                                       cudaMemcpyDeviceToDevice, (cudaStream_t)stream);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, who, ce);
    return SG_OK;
}

int sg_get_state(sg_handle h, double *x_out, void *stream)
{
    return copy_out(h, x_out, h ? h->sd : nullptr, sizeof(double), SG_NX, stream, "sg_get_state");
}

int sg_get_episode_state(sg_handle h, double *d_out, int32_t *i_out, void *stream)
{
    if (!h || (!d_out && !i_out)) return fail(SG_ERR_ARG, "sg_get_episode_state: null argument");
    if (d_out) {
        int rc = copy_out(h, d_out, h->sd, sizeof(double), SG_ND, stream, "sg_get_episode_state");
        if (rc != SG_OK) return rc;
    }
    if (i_out) return copy_out(h, i_out, h->si, sizeof(int32_t), SG_NI, stream, "sg_get_episode_state");
    return SG_OK;
}

int sg_set_episode_state(sg_handle h, const double *d_in, const int32_t *i_in, int64_t steps_done, void *stream)
{
    if (!h || (!d_in && !i_in) || steps_done < 0) return fail(SG_ERR_ARG, "sg_set_episode_state: bad arguments");
    if (!h->is_reset) return fail(SG_ERR_STATE, "sg_set_episode_state: call sg_reset first");
    DeviceGuard g(h->device);
    cudaError_t ce = cudaSuccess;
    if (d_in)
        ce = cudaMemcpy2DAsync(h->sd, sizeof(double) * h->cap, d_in, sizeof(double) * h->n, sizeof(double) * h->n,
                               SG_ND, cudaMemcpyDeviceToDevice, (cudaStream_t)stream);
    if (ce == cudaSuccess && i_in)
        ce = cudaMemcpy2DAsync(h->si, sizeof(int32_t) * h->cap, i_in, sizeof(int32_t) * h->n, sizeof(int32_t) * h->n,
                               SG_NI, cudaMemcpyDeviceToDevice, (cudaStream_t)stream);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_set_episode_state: copy", ce);
    h->steps_done = steps_done;
    return SG_OK;
}

int sg_get_stats(sg_handle h, double *stats_out, void *stream)
{
    return copy_out(h, stats_out, h ? h->stats : nullptr, sizeof(double), SG_NSTAT, stream, "sg_get_stats");
}

int sg_get_status(sg_handle h, int32_t *status_out, void *stream)
{
    return copy_out(h, status_out, h ? h->si + (int64_t)sgk::SI_STATUS * h->cap : nullptr, sizeof(int32_t), 1,
                    stream, "sg_get_status");
}

int sg_reduce_stats(sg_handle h, double *agg_out, void *stream)
{
    if (!h || !agg_out) return fail(SG_ERR_ARG, "sg_reduce_stats: null argument");
    if (!h->is_reset) return fail(SG_ERR_STATE, "sg_reduce_stats: call sg_reset first");
    DeviceGuard g(h->device);
    sgk::ReduceArgs r{};
    r.n = h->n; r.cap = h->cap; r.stats = h->stats; r.partials = h->partials; r.out = agg_out;
    r.n_blocks = sg_strict::reduce_blocks(h->n);
    cudaError_t ce = sg_strict::launch_reduce(r, (cudaStream_t)stream);
    if (ce != cudaSuccess) return fail(SG_ERR_CUDA, "sg_reduce_stats: launch", ce);
    h->launches += 2;
    return SG_OK;
}

int64_t sg_num_episodes(sg_handle h) { return h ? h->n : 0; }
int64_t sg_launch_count(sg_handle h) { return h ? h->launches : 0; }

}  // extern "C"
