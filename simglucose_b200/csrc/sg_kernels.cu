// Kernels of the batched closed-loop simulator.  Compiled twice (see build.py):
//   -fmad=false -DSG_NS=sg_strict   decision-faithful arithmetic (default mode)
//   -fmad=true  -DSG_NS=sg_fma      FMA contraction allowed (SG_FLAG_FMA)
//
// reset_kernel  T1DSimEnv.reset                      simulation/env.py:119-155
// run_kernel    SimObj.simulate / T1DSimEnv.step     simulation/sim_engine.py:29-39, env.py:48-117
//               with BBController / PIDController    controller/basal_bolus_ctrller.py:64-80,
//                                                    controller/pid_ctrller.py:17-36
// reduce_*      population statistics (warp shuffles -> block partials -> fixed-order pass)
#include "sg_kernels.cuh"
#include "sg_model.cuh"
#include "sg_philox.cuh"

#ifndef SG_NS
#error "compile with -DSG_NS=sg_strict or -DSG_NS=sg_fma"
#endif

namespace SG_NS {
using namespace sgk;

struct ParamLoad {
    const double *base;
    int64_t stride;
    __device__ __forceinline__ double operator()(int i) const { return __ldg(base + i * stride); }
};

__device__ __forceinline__ ParamLoad param_source(const EpisodeArgs &a, int64_t e)
{
    if (a.params_soa) return ParamLoad{a.params_soa + e, a.n};
    return ParamLoad{a.table + (int64_t)a.patient_row[e] * P_COUNT, 1};
}

// CGMSensor.measure at a sampling minute, cgm.py:28-33.
__device__ __forceinline__ double cgm_sample(const EpisodeArgs &a, int64_t e, double bg, int &noise_pos, int &status)
{
    // the C-ABI checks the stream length against the steps taken when every episode started together; episodes with
    // a position of their own (sg_reset_masked, sg_set_episode_state) are guarded here: a sample beyond the bound
    // stream reads as zero noise and marks the episode (sg_get_status: -5), it never reads out of bounds
    double nz = 0.0;
    if (noise_pos < a.n_noise) nz = a.noise[(int64_t)noise_pos * a.n + e];
    else if (status > 0) status = STATUS_STREAM_EXHAUSTED;
    double cgm = bg + nz;
    noise_pos++;
    cgm = (a.hw.smin > cgm) ? a.hw.smin : cgm;   // max(CGM, min)
    cgm = (a.hw.smax < cgm) ? a.hw.smax : cgm;   // min(CGM, max)
    return cgm;
}

__device__ __forceinline__ void stats_add_row(double (&s)[NSTAT], double bg, double lbgi, double hbgi,
                                              double ri, double reward, bool done)
{
    const double row = s[0];                      // index of the row being added
    s[0] = row + 1.0;
    s[1] += bg;
    s[2] = (bg < s[2]) ? bg : s[2];
    s[3] = (bg > s[3]) ? bg : s[3];
    s[4] += lbgi; s[5] += hbgi; s[6] += ri; s[7] += reward;
    if (bg < 50.0) s[8] += 1.0;
    if (bg < 70.0) s[9] += 1.0;
    if (bg >= 70.0 && bg <= 180.0) s[10] += 1.0;
    if (bg > 180.0) s[11] += 1.0;
    if (bg > 250.0) s[12] += 1.0;
    if (done && s[13] < 0.0) s[13] = row;
}

// The same on the resident statistics (global memory, one coalesced read-modify-write per env-step): the 16
// accumulators would otherwise hold 32 registers across the integrator, which needs them for scheduling.
__device__ __forceinline__ void stats_add_row_resident(double *st, int64_t cap, double bg, double lbgi, double hbgi,
                                                       double ri, double reward, bool done, int nstep, int nrej)
{
    double s[NSTAT];
#pragma unroll
    for (int i = 0; i < NSTAT; i++) s[i] = st[(int64_t)i * cap];
    stats_add_row(s, bg, lbgi, hbgi, ri, reward, done);
    s[14] += (double)nstep;
    s[15] += (double)nrej;
#pragma unroll
    for (int i = 0; i < NSTAT; i++) st[(int64_t)i * cap] = s[i];
}

__global__ void __launch_bounds__(BLOCK) reset_kernel(const EpisodeArgs a)
{
    const int64_t e = (int64_t)blockIdx.x * BLOCK + threadIdx.x;
    if (e >= a.n) return;
    if (a.reset_mask && a.reset_mask[e] == 0) return;       // sg_reset_masked: this episode keeps running
    const ParamLoad p = param_source(a, e);
    const double Vg = p(P_Vg);
    double x[NX];
    if (a.x0) {
#pragma unroll
        for (int i = 0; i < NX; i++) x[i] = a.x0[(int64_t)i * a.n + e];
    } else {
        const double *t = a.table_x0 + (int64_t)a.patient_row[e] * NX;
#pragma unroll
        for (int i = 0; i < NX; i++) x[i] = t[i];
    }
    // patient.reset(), t1dpatient.py:247-281
    double *sd = a.sd + e;
    int32_t *si = a.si + e;
#pragma unroll
    for (int i = 0; i < NX; i++) sd[(int64_t)(SD_X0 + i) * a.cap] = x[i];
    sd[(int64_t)SD_PLANNED * a.cap] = 0.0;
    sd[(int64_t)SD_LAST_CHO * a.cap] = 0.0;
    sd[(int64_t)SD_LAST_QSTO * a.cap] = x[0] + x[1];
    sd[(int64_t)SD_LAST_FOOD * a.cap] = 0.0;
    sd[(int64_t)SD_H_CARRY * a.cap] = 0.0;
    // env._reset() + env.reset(), env.py:119-155: two sensor samples at t = 0
    const double bg = x[12] / Vg;
    double lbgi, hbgi, ri;
    risk_index(bg, lbgi, hbgi, ri);
    // sensor.measure samples only when patient.t % sample_time == 0 (cgm.py:27); otherwise it returns the value held
    // since sensor.reset(), 0 (cgm.py:47-50), and no noise is consumed
    int noise_pos = 0, status = 1;
    double cgm0 = 0.0, obs = 0.0;
    if (a.t0 % a.hw.st == 0) {
        cgm0 = cgm_sample(a, e, bg, noise_pos, status);     // CGM_hist[0]
        obs = cgm_sample(a, e, bg, noise_pos, status);      // returned observation
    }
    sd[(int64_t)SD_LAST_CGM * a.cap] = obs;
    sd[(int64_t)SD_CGM_PREV * a.cap] = cgm0;
    sd[(int64_t)SD_OBS * a.cap] = obs;
    sd[(int64_t)SD_INFO_MEAL * a.cap] = 0.0;
    sd[(int64_t)SD_PID_INTEG * a.cap] = 0.0;                // controller.reset(), pid_ctrller.py:38-40
    sd[(int64_t)SD_PID_PREV * a.cap] = 0.0;
    si[(int64_t)SI_EATING * a.cap] = 0;
    si[(int64_t)SI_MINUTE * a.cap] = a.t0;
    si[(int64_t)SI_NOISE_POS * a.cap] = noise_pos;
    si[(int64_t)SI_MEAL_POS * a.cap] = 0;
    si[(int64_t)SI_STATUS * a.cap] = status;
    double s[NSTAT];
#pragma unroll
    for (int i = 0; i < NSTAT; i++) s[i] = 0.0;
    s[2] = INFINITY; s[3] = -INFINITY; s[13] = -1.0;
    stats_add_row(s, bg, lbgi, hbgi, ri, 0.0, false);
#pragma unroll
    for (int i = 0; i < NSTAT; i++) a.stats[(int64_t)i * a.cap + e] = s[i];
    if (a.row0) {
        double *r = a.row0 + e;
        const double nan = __longlong_as_double(0x7ff8000000000000LL);
        r[0 * a.n] = bg; r[1 * a.n] = cgm0; r[2 * a.n] = nan; r[3 * a.n] = nan;
        r[4 * a.n] = lbgi; r[5 * a.n] = hbgi; r[6 * a.n] = ri; r[7 * a.n] = obs;
        r[8 * a.n] = 0.0; r[9 * a.n] = 0.0; r[10 * a.n] = nan; r[11 * a.n] = nan;
    }
}

// MODE RUN_PLAIN: every minute is integrated to its end by the lane that owns the episode.
// MODE RUN_CAPPED (phase A of sg_set_step_cap, one env-step per launch): a minute that needs more than step_cap DOPRI5
//   steps is put aside -- integrator state to the defer scratch, episode appended to defer_list -- and the lane
//   idles for the rest of the env-step, so its warp is not held up.
// MODE RUN_RESUME (phase B): one thread per entry of defer_list finishes that minute and the rest of the env-step.
enum RunMode { RUN_PLAIN = 0, RUN_CAPPED = 1, RUN_RESUME = 2 };

template <int MODE>
__device__ __forceinline__ void episode_steps(const EpisodeArgs &a, const int64_t e, const unsigned lanes, double *smem)
{
    constexpr bool DEFER = MODE != RUN_PLAIN;
    const Column<BLOCK> c{smem + 2 * threadIdx.x};
    const ParamLoad p = param_source(a, e);
    stage_params(c, p);

    double *sd = a.sd + e;
    int32_t *si = a.si + e;
    double x[NX];
#pragma unroll
    for (int i = 0; i < NX; i++) x[i] = canonical_zero(sd[(int64_t)(SD_X0 + i) * a.cap]);
    Meal meal;
    meal.planned = sd[(int64_t)SD_PLANNED * a.cap];
    meal.last_cho = sd[(int64_t)SD_LAST_CHO * a.cap];
    meal.last_Qsto = sd[(int64_t)SD_LAST_QSTO * a.cap];
    meal.last_food = sd[(int64_t)SD_LAST_FOOD * a.cap];
    meal.eating = si[(int64_t)SI_EATING * a.cap];
    double h_carry = sd[(int64_t)SD_H_CARRY * a.cap];
    double last_cgm = sd[(int64_t)SD_LAST_CGM * a.cap];
    // cgm_prev, obs, info_meal and the PID state are only touched at env-step boundaries: they stay in the
    // resident state (global memory, L2 hits) instead of holding registers across the integrator
    int minute = si[(int64_t)SI_MINUTE * a.cap];
    int noise_pos = si[(int64_t)SI_NOISE_POS * a.cap];
    int meal_pos = si[(int64_t)SI_MEAL_POS * a.cap];
    int status = si[(int64_t)SI_STATUS * a.cap];
    int step_nstep = 0, step_nrej = 0;             // DOPRI5 steps / rejects of the current env-step

    const int st_i = a.hw.st;
    const double st = (double)st_i;
    // x / st (env.py:75-81) through the refined reciprocal: bit-identical to the IEEE quotient (sg_math.cuh div_r)
    const double r_st = sgm::rcp_nr(st);
    int next_mm = (meal_pos < a.max_meals) ? a.meal_minute[(int64_t)meal_pos * a.n + e] : -1;
    bool have_rprev = false;                           // H_RPREV holds risk_index(CGM_hist[-1]) from the step before

    // ---- minute put aside (phases A / B) ----
    Suspend sp{0.0, 0.0, 0.0, 0, 0, 0, 0};
    int j0 = 0;                                        // RUN_RESUME: the minute of the env-step to resume at
    double r_ins = 0.0, r_announced = 0.0, r_CHO = 0.0, r_insulin = 0.0, r_BG = 0.0, r_CGM = 0.0;
    if (MODE == RUN_RESUME) {
        const double *dd = a.defer_d + e;
        const int32_t *di = a.defer_i + e;
#pragma unroll
        for (int i = 0; i < NX; i++) c.set(H_K1 + i, dd[(int64_t)i * a.cap]);
        sp.h = dd[(int64_t)13 * a.cap]; sp.x = dd[(int64_t)14 * a.cap]; sp.lfacold = dd[(int64_t)15 * a.cap];
        r_ins = dd[(int64_t)16 * a.cap]; r_announced = dd[(int64_t)17 * a.cap];
        r_CHO = dd[(int64_t)18 * a.cap]; r_insulin = dd[(int64_t)19 * a.cap];
        r_BG = dd[(int64_t)20 * a.cap]; r_CGM = dd[(int64_t)21 * a.cap];
        sp.nstep = di[(int64_t)0 * a.cap]; sp.naccpt = di[(int64_t)1 * a.cap]; sp.nrej = di[(int64_t)2 * a.cap];
        sp.reject = di[(int64_t)3 * a.cap]; j0 = di[(int64_t)4 * a.cap];
        step_nstep = di[(int64_t)5 * a.cap]; step_nrej = di[(int64_t)6 * a.cap];
    }

    for (int k = 0; k < a.n_steps; k++) {
        // ---- controller.policy(observation, reward, done, **info) ----
        double a_basal = 0.0, a_bolus = 0.0;
        double ins, CHO = 0.0, insulin = 0.0, BG = 0.0, CGM = 0.0;
        if (MODE == RUN_RESUME) {
            ins = r_ins; CHO = r_CHO; insulin = r_insulin; BG = r_BG; CGM = r_CGM;
        } else {
        double ctrl[NCTRL];                            // re-read per env-step (L2 resident): not live across the integrator
#pragma unroll
        for (int i = 0; i < NCTRL; i++) ctrl[i] = a.ctrl_rows ? __ldg(a.ctrl_rows + (int64_t)i * a.n + e) : a.ctrl[i];
        const double obs = sd[(int64_t)SD_OBS * a.cap];
        if (a.controller == 1) {                       // BBController._bb_policy
            const double info_meal = sd[(int64_t)SD_INFO_MEAL * a.cap];
            // basal = u2ss * BW / 6000 of the CSV row the controller finds by patient NAME (basal_bolus_ctrller.py:52-64):
            // ctrl[3] when the caller supplies it (perturbed populations keep the nominal basal), else the episode's
            // own parameter row
            a_basal = (ctrl[3] > 0.0) ? ctrl[3] : c(H_BASAL);
            if (info_meal > 0.0) {
                const double corr = sgm::div_cold((obs > 150.0 ? 1.0 : 0.0) * (obs - ctrl[0]), ctrl[2]);
                a_bolus = sgm::div_cold(info_meal * st, ctrl[1]) + corr;
            } else {
                a_bolus = 0.0;
            }
            a_bolus = sgm::div_cold(a_bolus, st);
        } else if (a.controller == 2) {                // PIDController.policy
            const double pid_integ = sd[(int64_t)SD_PID_INTEG * a.cap];
            const double pid_prev = sd[(int64_t)SD_PID_PREV * a.cap];
            a_basal = ctrl[0] * (obs - ctrl[3]) + ctrl[1] * pid_integ + sgm::div_cold(ctrl[2] * (obs - pid_prev), st);
            sd[(int64_t)SD_PID_PREV * a.cap] = obs;
            sd[(int64_t)SD_PID_INTEG * a.cap] = pid_integ + (obs - ctrl[3]) * st;
            a_bolus = 0.0;
        } else if (a.controller == 3) {                // random basal rate ~ U(ctrl[0], ctrl[1]), bolus 0: the action
            // gym's action_space.sample() would hand the wrapper (simglucose_gym_env.py:41-46), drawn in the kernel from
            // Philox keyed by (seed = ctrl[2], global episode index = ctrl[3] + e, env-step index)
            sgp::Philox rng((uint64_t)ctrl[2], (uint64_t)ctrl[3] + (uint64_t)e, 0x6163746eu /* 'actn' */,
                            (uint32_t)(a.step0 + k));
            a_basal = ctrl[0] + (ctrl[1] - ctrl[0]) * rng.uniform();
            a_bolus = 0.0;
        } else {
            a_basal = a.act_basal[(int64_t)k * a.act_stride + e];
            a_bolus = a.act_bolus[(int64_t)k * a.act_stride + e];
        }
        // ---- env.step, env.py:66-117 ----
        ins = pump_quantise(a_basal, a.hw.pump[0], a.hw.pump[1], a.hw.pump[2]) +
              pump_quantise(a_bolus, a.hw.pump[3], a.hw.pump[4], a.hw.pump[5]);
        if (a.trace) {                                 // the action columns of the history row
            double *r = a.trace + (int64_t)k * NTRACE * a.n + e;
            r[10 * a.n] = a_basal; r[11 * a.n] = a_bolus;
        }
        }
        bool aside = false;                            // RUN_CAPPED: this lane's minute was put aside
        const double ins_kg = sgm::div_cold(ins * 6000.0, c(H_BW));  // t1dpatient.py:122, constant over the env-step
        const double ins_mean = sgm::div_r(ins, st, r_st);
        for (int j = 0; j < st_i; j++) {               // mini_step, env.py:48-64
            int mode = MINUTE_FRESH;
            if (MODE == RUN_CAPPED && aside) mode = MINUTE_IDLE;
            if (MODE == RUN_RESUME) mode = (j < j0) ? MINUTE_IDLE : (j == j0 ? MINUTE_RESUME : MINUTE_FRESH);
            double announced = 0.0;
            if (DEFER && mode == MINUTE_IDLE) {
                // nothing to prepare
            } else if (DEFER && mode == MINUTE_RESUME) {
                announced = r_announced;               // the meal machine already ran for this minute
            } else if (a.step_meals) {                 // scenario.get_action evaluated by the caller
                announced = a.step_meals[(int64_t)j * a.n + e];
            } else {
                while (next_mm >= 0 && next_mm < minute) {
                    meal_pos++;
                    next_mm = (meal_pos < a.max_meals) ? a.meal_minute[(int64_t)meal_pos * a.n + e] : -1;
                }
                if (next_mm == minute) announced = a.meal_grams[(int64_t)meal_pos * a.n + e];
            }
            Minute m{0.0, 0.0, false};
            if (!DEFER || mode != MINUTE_IDLE) {
                const double to_eat = (DEFER && mode == MINUTE_RESUME) ? meal.last_cho
                                                                        : meal_machine(meal, announced, x[0], x[1]);
                m = minute_setup_rate(c, to_eat, ins_kg, meal.last_Qsto, meal.last_food);
            }
            StepCount cnt{0, 0};
            const int idid = dopri5_minute<BLOCK, DEFER>(c, m, x, (double)minute, a.carry_h != 0, h_carry, cnt, lanes,
                                                         mode, MODE == RUN_CAPPED ? a.step_cap : 0, &sp);
            if (DEFER && mode == MINUTE_IDLE) continue;
            if (MODE == RUN_CAPPED && idid == IDID_SUSPENDED) {
                // ---- put the rest of the env-step aside for phase B ----
                aside = true;
                double *dd = a.defer_d + e;
                int32_t *di = a.defer_i + e;
#pragma unroll
                for (int i = 0; i < NX; i++) dd[(int64_t)i * a.cap] = c(H_K1 + i);
                dd[(int64_t)13 * a.cap] = sp.h; dd[(int64_t)14 * a.cap] = sp.x; dd[(int64_t)15 * a.cap] = sp.lfacold;
                dd[(int64_t)16 * a.cap] = ins; dd[(int64_t)17 * a.cap] = announced;
                dd[(int64_t)18 * a.cap] = CHO; dd[(int64_t)19 * a.cap] = insulin;
                dd[(int64_t)20 * a.cap] = BG; dd[(int64_t)21 * a.cap] = CGM;
                di[(int64_t)0 * a.cap] = sp.nstep; di[(int64_t)1 * a.cap] = sp.naccpt; di[(int64_t)2 * a.cap] = sp.nrej;
                di[(int64_t)3 * a.cap] = sp.reject; di[(int64_t)4 * a.cap] = j;
                di[(int64_t)5 * a.cap] = step_nstep; di[(int64_t)6 * a.cap] = step_nrej;
                a.defer_list[atomicAdd(a.defer_count, 1)] = (int32_t)e;
                continue;
            }
            if (idid < 0 && status > 0) status = idid;
            if (a.counters) {
                const int64_t row = (int64_t)(k * st_i + j) * 2;
                a.counters[row * a.n + e] = cnt.nstep;
                a.counters[(row + 1) * a.n + e] = cnt.nrej;
            }
            step_nstep += cnt.nstep;
            step_nrej += cnt.nrej;
            minute++;
            const double bg = sgm::div_r(x[12], c(H_Vg), c(H_rVg));      // t1dpatient.py:217-218
            if (minute % st_i == 0) last_cgm = cgm_sample(a, e, bg, noise_pos, status);   // cgm.py:27
            CHO += sgm::div_r(announced, st, r_st); insulin += ins_mean;
            BG += sgm::div_r(bg, st, r_st); CGM += sgm::div_r(last_cgm, st, r_st);
        }
        if (MODE == RUN_CAPPED && aside) continue;     // phase B closes this env-step
        double lbgi, hbgi, ri, r_prev, r_now, t0, t1;
        risk_index(BG, lbgi, hbgi, ri);                                  // env.py:85
        if (have_rprev) {
            r_prev = c(H_RPREV);
        } else {
            risk_index(sd[(int64_t)SD_CGM_PREV * a.cap], t0, t1, r_prev);    // risk_diff, env.py:27-33
        }
        risk_index(CGM, t0, t1, r_now);
        c.set(H_RPREV, r_now);
        have_rprev = true;
        const double reward = r_prev - r_now;
        const bool done = (BG < 70.0) || (BG > 350.0);                   // env.py:103
        sd[(int64_t)SD_CGM_PREV * a.cap] = CGM;
        sd[(int64_t)SD_OBS * a.cap] = CGM;
        sd[(int64_t)SD_INFO_MEAL * a.cap] = CHO;
        stats_add_row_resident(a.stats + e, a.cap, BG, lbgi, hbgi, ri, reward, done, step_nstep, step_nrej);
        step_nstep = 0; step_nrej = 0;
        if (a.trace) {
            double *r = a.trace + (int64_t)k * NTRACE * a.n + e;
            r[0 * a.n] = BG; r[1 * a.n] = CGM; r[2 * a.n] = CHO; r[3 * a.n] = insulin;
            r[4 * a.n] = lbgi; r[5 * a.n] = hbgi; r[6 * a.n] = ri; r[7 * a.n] = CGM;
            r[8 * a.n] = reward; r[9 * a.n] = done ? 1.0 : 0.0;
        }
    }

#pragma unroll
    for (int i = 0; i < NX; i++) sd[(int64_t)(SD_X0 + i) * a.cap] = x[i];
    sd[(int64_t)SD_PLANNED * a.cap] = meal.planned;
    sd[(int64_t)SD_LAST_CHO * a.cap] = meal.last_cho;
    sd[(int64_t)SD_LAST_QSTO * a.cap] = meal.last_Qsto;
    sd[(int64_t)SD_LAST_FOOD * a.cap] = meal.last_food;
    sd[(int64_t)SD_H_CARRY * a.cap] = h_carry;
    sd[(int64_t)SD_LAST_CGM * a.cap] = last_cgm;
    si[(int64_t)SI_EATING * a.cap] = meal.eating;
    si[(int64_t)SI_MINUTE * a.cap] = minute;
    si[(int64_t)SI_NOISE_POS * a.cap] = noise_pos;
    si[(int64_t)SI_MEAL_POS * a.cap] = meal_pos;
    si[(int64_t)SI_STATUS * a.cap] = status;
}

__global__ void __launch_bounds__(BLOCK) run_kernel(const EpisodeArgs a)
{
    extern __shared__ double smem[];   // H_COUNT rows x BLOCK columns
    const int64_t idx = (int64_t)blockIdx.x * BLOCK + threadIdx.x;
    const unsigned lanes = __ballot_sync(0xffffffffu, idx < a.n);   // the lanes that own an episode
    if (idx >= a.n) return;        // no block-level synchronisation below: columns are private
    const int e = a.perm ? a.perm[idx] : (int)idx;                 // schedule-coherent order (sg_set_coherence)
    episode_steps<RUN_PLAIN>(a, e, lanes, smem);
}

__global__ void __launch_bounds__(BLOCK) run_capped_kernel(const EpisodeArgs a)
{
    extern __shared__ double smem[];
    const int64_t idx = (int64_t)blockIdx.x * BLOCK + threadIdx.x;
    const unsigned lanes = __ballot_sync(0xffffffffu, idx < a.n);
    if (idx >= a.n) return;
    const int e = a.perm ? a.perm[idx] : (int)idx;
    episode_steps<RUN_CAPPED>(a, e, lanes, smem);
}

// Phase B: grid-stride over the episodes that phase A put aside (every warp is full of heavy minutes).
__global__ void __launch_bounds__(BLOCK) run_resume_kernel(const EpisodeArgs a)
{
    extern __shared__ double smem[];
    const int count = *a.defer_count;
    for (int64_t base = (int64_t)blockIdx.x * BLOCK; base < count; base += (int64_t)gridDim.x * BLOCK) {
        const int64_t idx = base + threadIdx.x;
        const unsigned lanes = __ballot_sync(0xffffffffu, idx < count);
        if (idx < count) episode_steps<RUN_RESUME>(a, (int64_t)a.defer_list[idx], lanes, smem);
    }
}

// T1DPatient.step (t1dpatient.py:82-116) for every episode: the patient-level seam of the reference
// (an env built from user-supplied sensor / pump / scenario objects drives the patient minute by minute).
__global__ void __launch_bounds__(BLOCK) patient_step_kernel(const PatientArgs pa)
{
    extern __shared__ double smem[];
    const EpisodeArgs &a = pa.e;
    const int64_t e = (int64_t)blockIdx.x * BLOCK + threadIdx.x;
    const unsigned lanes = __ballot_sync(0xffffffffu, e < a.n);
    if (e >= a.n) return;
    const Column<BLOCK> c{smem + 2 * threadIdx.x};
    const ParamLoad p = param_source(a, e);
    stage_params(c, p);
    double *sd = a.sd + e;
    int32_t *si = a.si + e;
    double x[NX];
#pragma unroll
    for (int i = 0; i < NX; i++) x[i] = canonical_zero(sd[(int64_t)(SD_X0 + i) * a.cap]);
    Meal meal;
    meal.planned = sd[(int64_t)SD_PLANNED * a.cap];
    meal.last_cho = sd[(int64_t)SD_LAST_CHO * a.cap];
    meal.last_Qsto = sd[(int64_t)SD_LAST_QSTO * a.cap];
    meal.last_food = sd[(int64_t)SD_LAST_FOOD * a.cap];
    meal.eating = si[(int64_t)SI_EATING * a.cap];
    double h_carry = sd[(int64_t)SD_H_CARRY * a.cap];
    int minute = si[(int64_t)SI_MINUTE * a.cap];
    int status = si[(int64_t)SI_STATUS * a.cap];

    const double to_eat = meal_machine(meal, pa.cho[e], x[0], x[1]);
    const Minute m = minute_setup(c, to_eat, pa.insulin[e], meal.last_Qsto, meal.last_food);
    StepCount cnt{0, 0};
    const int idid = dopri5_minute(c, m, x, (double)minute, a.carry_h != 0, h_carry, cnt, lanes);
    if (idid < 0 && status > 0) status = idid;
    minute++;

#pragma unroll
    for (int i = 0; i < NX; i++) sd[(int64_t)(SD_X0 + i) * a.cap] = x[i];
    sd[(int64_t)SD_PLANNED * a.cap] = meal.planned;
    sd[(int64_t)SD_LAST_CHO * a.cap] = meal.last_cho;
    sd[(int64_t)SD_LAST_QSTO * a.cap] = meal.last_Qsto;
    sd[(int64_t)SD_LAST_FOOD * a.cap] = meal.last_food;
    sd[(int64_t)SD_H_CARRY * a.cap] = h_carry;
    si[(int64_t)SI_EATING * a.cap] = meal.eating;
    si[(int64_t)SI_MINUTE * a.cap] = minute;
    si[(int64_t)SI_STATUS * a.cap] = status;
    if (pa.out) {
        double *o = pa.out + e;
#pragma unroll
        for (int i = 0; i < NX; i++) o[(int64_t)i * a.n] = x[i];
        o[(int64_t)13 * a.n] = x[12] / c(H_Vg);                 // observation.Gsub, t1dpatient.py:217-218
        o[(int64_t)14 * a.n] = to_eat;
        o[(int64_t)15 * a.n] = (double)idid;
        o[(int64_t)16 * a.n] = (double)cnt.nstep;
        o[(int64_t)17 * a.n] = meal.planned;
        o[(int64_t)18 * a.n] = meal.last_Qsto;
        o[(int64_t)19 * a.n] = meal.last_food;
        o[(int64_t)20 * a.n] = (double)meal.eating;
    }
}

// T1DPatient.model (t1dpatient.py:118-208) as a batched function evaluation.
__global__ void __launch_bounds__(BLOCK) model_kernel(const ModelArgs a)
{
    extern __shared__ double smem[];
    const int64_t e = (int64_t)blockIdx.x * BLOCK + threadIdx.x;
    if (e >= a.n) return;
    const Column<BLOCK> c{smem + 2 * threadIdx.x};
    const ParamLoad p = a.params_soa ? ParamLoad{a.params_soa + e, a.n}
                                     : ParamLoad{a.table + (int64_t)a.patient_row[e] * P_COUNT, 1};
    stage_params(c, p);
    double x[NX], dx[NX];
#pragma unroll
    for (int i = 0; i < NX; i++) x[i] = canonical_zero(a.x[(int64_t)i * a.n + e]);
    const Minute m = minute_setup(c, a.cho[e], a.insulin[e], a.last_qsto[e], a.last_food[e]);
    rhs(c, m, x, dx);
#pragma unroll
    for (int i = 0; i < NX; i++) a.dxdt[(int64_t)i * a.n + e] = dx[i];
}

// InsulinPump.basal / .bolus (pump.py:23-39) and risk_index([BG], 1) (risk.py:5-17), elementwise.
__global__ void __launch_bounds__(256) pump_kernel(const ElementwiseArgs a)
{
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i < a.n) a.out[i] = pump_quantise(a.in[i], a.p[0], a.p[1], a.p[2]);
}

__global__ void __launch_bounds__(256) risk_kernel(const ElementwiseArgs a)
{
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= a.n) return;
    double l, h, r;
    risk_index(a.in[i], l, h, r);
    a.out[i] = l; a.out[a.n + i] = h; a.out[2 * a.n + i] = r;
}

// The sampling branch of CGMSensor.measure (cgm.py:28-33), elementwise; p = sensor min, max.
__global__ void __launch_bounds__(256) cgm_kernel(const ElementwiseArgs a)
{
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= a.n) return;
    double cgm = a.in[i] + a.in2[i];
    cgm = (a.p[0] > cgm) ? a.p[0] : cgm;
    cgm = (a.p[1] < cgm) ? a.p[1] : cgm;
    a.out[i] = cgm;
}

// risk_index(BG, horizon) over a window (risk.py:5-17): means of 10 f^2 over the negative / positive f.
// One block; every thread accumulates a strided slice in index order, thread 0 adds the slices in order.
__global__ void __launch_bounds__(256) risk_window_kernel(const ElementwiseArgs a)
{
    __shared__ double part[256][4];
    double sl = 0.0, sh = 0.0, nl = 0.0, nh = 0.0;
    for (int64_t i = threadIdx.x; i < a.n; i += 256) {
        const double f = risk_f(a.in[i]);
        if (f < 0.0) { sl += 10.0 * (f * f); nl += 1.0; }
        if (f > 0.0) { sh += 10.0 * (f * f); nh += 1.0; }
    }
    part[threadIdx.x][0] = sl; part[threadIdx.x][1] = sh; part[threadIdx.x][2] = nl; part[threadIdx.x][3] = nh;
    __syncthreads();
    if (threadIdx.x == 0) {
        sl = sh = nl = nh = 0.0;
        for (int t = 0; t < 256; t++) { sl += part[t][0]; sh += part[t][1]; nl += part[t][2]; nh += part[t][3]; }
        double lbgi = nl > 0.0 ? sl / nl : 0.0;      // nan_to_num(mean of an empty selection) == 0
        double hbgi = nh > 0.0 ? sh / nh : 0.0;
        if (isinf(lbgi)) lbgi = DBL_MAX;
        if (isinf(hbgi)) hbgi = DBL_MAX;
        a.out[0] = lbgi; a.out[1] = hbgi; a.out[2] = lbgi + hbgi;
    }
}

// Accuracy probe of the helpers in sg_math.cuh against CUDA's libm (tests/test_gpu_math.py).
__global__ void __launch_bounds__(256) math_probe_kernel(int kind, const double *a, const double *b, double *out, int64_t n)
{
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const double u = a[i], v = b[i];
    double r;
    switch (kind) {
    case 0: r = sgm::div_nr(u, v); break;
    case 1: r = u / v; break;
    case 2: r = sgm::tanh_diff(u, v); break;
    case 3: r = tanh(u) - tanh(v); break;
    case 4: r = sgm::pow_pos(u, v); break;
    case 5: r = pow(u, v); break;
    case 6: r = sgm::exp_core(u); break;
    case 7: r = exp(u); break;
    case 8: r = sgm::log_core(u); break;
    case 9: r = log(u); break;
    case 10: r = 2.0 * sgm::tanh_sum_half(2.0 * u, 2.0 * v); break;      // tanh(u) - tanh(v) + 2 as the RHS forms it
    case 11: r = tanh(u) - tanh(v) + 2.0; break;
    case 12: r = sgm::rcp_ulp(u); break;
    case 13: r = (sgm::is_neg(u) ? 1.0 : 0.0) + (sgm::is_pos(u) ? 2.0 : 0.0); break;
    case 14: r = sgm::exp_tanh(u); break;
    default: r = log(u); break;
    }
    out[i] = r;
}

// ---- population statistics ------------------------------------------------------------------
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

constexpr int RBLOCK = 256;

__global__ void __launch_bounds__(RBLOCK) reduce_partial_kernel(const ReduceArgs a)
{
    __shared__ double warp_part[RBLOCK / 32][NSTAT];
    double v[NSTAT];
#pragma unroll
    for (int i = 0; i < NSTAT; i++) v[i] = identity(i);
    for (int64_t e = (int64_t)blockIdx.x * RBLOCK + threadIdx.x; e < a.n; e += (int64_t)gridDim.x * RBLOCK) {
#pragma unroll
        for (int i = 0; i < NSTAT; i++) {
            double u = a.stats[(int64_t)i * a.cap + e];
            if (i == 13) u = (u >= 0.0) ? 1.0 : 0.0;      // episodes that reported done
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
        a.partials[(int64_t)blockIdx.x * NSTAT + threadIdx.x] = acc;
    }
}

__global__ void reduce_final_kernel(const ReduceArgs a)
{
    if (threadIdx.x < NSTAT) {
        double acc = identity(threadIdx.x);
        for (int b = 0; b < a.n_blocks; b++) acc = combine(threadIdx.x, acc, a.partials[(int64_t)b * NSTAT + threadIdx.x]);
        a.out[threadIdx.x] = acc;
    }
}

// ---- launchers --------------------------------------------------------------------------------
static inline unsigned grid_for(int64_t n) { return (unsigned)((n + BLOCK - 1) / BLOCK); }

// cudaFuncAttributeMaxDynamicSharedMemorySize belongs to the (kernel, device) pair: a process that creates handles on
// several GPUs has to raise it on each of them.  `done` is the launcher's own per-device record.
constexpr int MAX_DEVICES = 64;
template <class Kernel>
static cudaError_t allow_dynamic_smem(Kernel kernel, int bytes, bool (&done)[MAX_DEVICES])
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    const bool tracked = dev >= 0 && dev < MAX_DEVICES;
    if (tracked && done[dev]) return cudaSuccess;
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess && tracked) done[dev] = true;
    return e;
}

cudaError_t launch_reset(const EpisodeArgs &a, cudaStream_t s)
{
    reset_kernel<<<grid_for(a.n), BLOCK, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_run(const EpisodeArgs &a, cudaStream_t s)
{
    constexpr int kSmem = H_COUNT * BLOCK * (int)sizeof(double);
    static bool done[MAX_DEVICES] = {};
    cudaError_t e = allow_dynamic_smem(run_kernel, kSmem, done);
    if (e != cudaSuccess) return e;
    run_kernel<<<grid_for(a.n), BLOCK, kSmem, s>>>(a);
    return cudaGetLastError();
}

// One env-step with heavy minutes put aside: counter reset, phase A over all episodes, phase B over the list.
cudaError_t launch_run_capped(const EpisodeArgs &a, cudaStream_t s)
{
    constexpr int kSmem = H_COUNT * BLOCK * (int)sizeof(double);
    static bool done_a[MAX_DEVICES] = {}, done_b[MAX_DEVICES] = {};
    cudaError_t e = allow_dynamic_smem(run_capped_kernel, kSmem, done_a);
    if (e == cudaSuccess) e = allow_dynamic_smem(run_resume_kernel, kSmem, done_b);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(a.defer_count, 0, sizeof(int32_t), s);
    if (e != cudaSuccess) return e;
    run_capped_kernel<<<grid_for(a.n), BLOCK, kSmem, s>>>(a);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    unsigned blocks = grid_for(a.n);
    if (blocks > 148u * 4u) blocks = 148u * 4u;            // one resident wave; the kernel strides over the list
    run_resume_kernel<<<blocks, BLOCK, kSmem, s>>>(a);
    return cudaGetLastError();
}

#ifdef SG_WITH_FP32
#include "sg_fp32.cuh"
#endif

cudaError_t launch_patient_step(const PatientArgs &a, cudaStream_t s)
{
    constexpr int kSmem = H_COUNT * BLOCK * (int)sizeof(double);
    static bool done[MAX_DEVICES] = {};
    cudaError_t e = allow_dynamic_smem(patient_step_kernel, kSmem, done);
    if (e != cudaSuccess) return e;
    patient_step_kernel<<<grid_for(a.e.n), BLOCK, kSmem, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_model(const ModelArgs &a, cudaStream_t s)
{
    constexpr int kSmem = H_COUNT * BLOCK * (int)sizeof(double);
    static bool done[MAX_DEVICES] = {};
    cudaError_t e = allow_dynamic_smem(model_kernel, kSmem, done);
    if (e != cudaSuccess) return e;
    model_kernel<<<grid_for(a.n), BLOCK, kSmem, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_pump(const ElementwiseArgs &a, cudaStream_t s)
{
    pump_kernel<<<(unsigned)((a.n + 255) / 256), 256, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_risk(const ElementwiseArgs &a, cudaStream_t s)
{
    risk_kernel<<<(unsigned)((a.n + 255) / 256), 256, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_cgm(const ElementwiseArgs &a, cudaStream_t s)
{
    cgm_kernel<<<(unsigned)((a.n + 255) / 256), 256, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_risk_window(const ElementwiseArgs &a, cudaStream_t s)
{
    risk_window_kernel<<<1, 256, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_math_probe(int kind, const double *a, const double *b, double *out, int64_t n, cudaStream_t s)
{
    math_probe_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(kind, a, b, out, n);
    return cudaGetLastError();
}

int reduce_blocks(int64_t n)
{
    int64_t b = (n + RBLOCK - 1) / RBLOCK;
    if (b > 148 * 4) b = 148 * 4;
    if (b < 1) b = 1;
    return (int)b;
}

cudaError_t launch_reduce(const ReduceArgs &a, cudaStream_t s)
{
    reduce_partial_kernel<<<a.n_blocks, RBLOCK, 0, s>>>(a);
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) return err;
    reduce_final_kernel<<<1, 32, 0, s>>>(a);
    return cudaGetLastError();
}

}  // namespace SG_NS
