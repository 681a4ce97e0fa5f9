/* simglucose_b200 -- C ABI of the B200-native batched patient simulator.
 *
 * The reference (Sawyerbatch/simglucose) is pure Python and has no FFI; its seams are Python
 * objects.  Each entry point below names the reference interface it
 * stands behind (paths relative to the reference root).  A reference-side binding is a ctypes
 * stub; the in-tree one is simglucose_b200/_capi.py.
 *
 * Conventions
 *   - plain C types only; every array argument is a DEVICE pointer unless its name ends in
 *     _host; arrays over episodes are structure-of-arrays: element (i, e) of a [N][E] array
 *     lives at ptr[i * n_episodes + e] so that one warp reads consecutive addresses;
 *   - the caller owns every buffer it passes; the handle owns the resident episode state it
 *     allocated in sg_create (nothing is allocated afterwards);
 *   - `stream` is a cudaStream_t passed as void*; calls are asynchronous and stream-ordered;
 *   - return value: SG_OK (0) or a negative sg_error; nothing throws across the ABI;
 *   - one host thread per handle at a time.
 */
#ifndef SIMGLUCOSE_B200_H
#define SIMGLUCOSE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SG_ABI_VERSION 1
#define SG_NX 13          /* patient state size, t1dpatient.py:252 (params.iloc[2:15]) */

/* The 32 columns of simglucose/params/vpatient_params.csv the hot path reads
 * (T1DPatient.model t1dpatient.py:118-208; u2ss/BW by BBController
 * controller/basal_bolus_ctrller.py:64; Vg by T1DPatient.observation t1dpatient.py:217-218). */
typedef enum {
    SG_P_BW = 0, SG_P_u2ss, SG_P_kmax, SG_P_kmin, SG_P_b, SG_P_d, SG_P_kabs, SG_P_f,
    SG_P_kp1, SG_P_kp2, SG_P_kp3, SG_P_Fsnc, SG_P_ke1, SG_P_ke2, SG_P_k1, SG_P_k2,
    SG_P_Vm0, SG_P_Vmx, SG_P_Km0, SG_P_m1, SG_P_m2, SG_P_m4, SG_P_m30, SG_P_ka1,
    SG_P_ka2, SG_P_Vi, SG_P_p2u, SG_P_Ib, SG_P_ki, SG_P_kd, SG_P_ksc, SG_P_Vg,
    SG_NPARAM
} sg_param;

typedef enum { SG_OK = 0, SG_ERR_ARG = -1, SG_ERR_CUDA = -2, SG_ERR_STATE = -3, SG_ERR_NOMEM = -4 } sg_error;

/* sg_create flags */
#define SG_FLAG_CARRY_H 1u   /* scipy 1.6.3 (Pipfile.lock) behaviour: the step size proposed at the end of one
                                minute starts the next (WORK(7)=H); default 0 = scipy >= 1.7: HINIT every minute */
#define SG_FLAG_FMA     2u   /* allow FMA contraction in the kernels (faster; not decision-identical to scipy) */
/* 4u was SG_FLAG_PAIR (two threads per episode; measured 0.80x and removed in round 2): rejected */

#define SG_FLAG_FP32    8u   /* optional single-precision mode: the patient ODE and its DOPRI5 integration run in FP32
                                (local time, uround 1.2e-7, rtol 1e-4, FMA); pump, sensor, meal machine, risk and
                                statistics stay FP64.  Looser bound: max |dBG| / max(|BG|, 1) <= 1e-3 over a day against
                                the FP64 path (tests/test_gpu_fp32.py; measured 1.2e-4).  Not the graded mode; sg_set_step_cap refuses it. */

/* Controller fused into sg_run. */
typedef enum {
    SG_CTRL_ACTIONS = 0,     /* per-step (basal, bolus) supplied by the caller: any Controller.policy
                                (controller/base.py:11-34) evaluated outside */
    SG_CTRL_BB = 1,          /* BBController._bb_policy, controller/basal_bolus_ctrller.py:64-80; ctrl = target, CR, CF,
                                basal.  basal (U/min) is u2ss * BW / 6000 of the CSV row the reference looks up by
                                patient NAME (basal_bolus_ctrller.py:52-64); 0 = take it from the episode's own
                                parameter row (identical for unperturbed cohort rows) */
    SG_CTRL_PID = 2,         /* PIDController.policy, controller/pid_ctrller.py:17-36; ctrl = P, I, D, target */
    SG_CTRL_RANDOM = 3       /* Action(basal ~ U(lo, hi), bolus 0): what gym's action_space.sample() feeds
                                T1DSimEnv._step (envs/simglucose_gym_env.py:41-46, 78-81), drawn inside the kernel from
                                Philox keyed by (seed, global episode index, env-step index counted from the last
                                sg_reset -- sg_reset_masked does not restart it, so environments that restart one
                                by one keep drawing fresh actions);
                                ctrl = lo, hi, seed, global index of episode 0 (whole numbers in [0, 2^53); anything
                                else is refused) */
} sg_controller;
#define SG_NCTRL 4

/* One history row per env-step: the columns of T1DSimEnv.show_history() (simulation/env.py:169-180)
 * followed by the Step tuple fields (env.py:106-117) and the action that was applied. */
typedef enum {
    SG_T_BG = 0, SG_T_CGM, SG_T_CHO, SG_T_INSULIN, SG_T_LBGI, SG_T_HBGI, SG_T_RISK,
    SG_T_OBS, SG_T_REWARD, SG_T_DONE, SG_T_BASAL, SG_T_BOLUS,
    SG_NTRACE
} sg_trace_col;

/* Per-episode statistics accumulated over history rows (row 0 = reset row included). */
typedef enum {
    SG_S_ROWS = 0, SG_S_SUM_BG, SG_S_MIN_BG, SG_S_MAX_BG, SG_S_SUM_LBGI, SG_S_SUM_HBGI,
    SG_S_SUM_RISK, SG_S_SUM_REWARD, SG_S_LT50, SG_S_LT70, SG_S_INRANGE, SG_S_GT180,
    SG_S_GT250, SG_S_FIRST_DONE, SG_S_DOPRI_STEPS, SG_S_DOPRI_REJECTS,
    SG_NSTAT
} sg_stat;

/* Sensor + pump rows (simglucose/params/sensor_params.csv, pump_params.csv) used by
 * CGMSensor.measure (sensor/cgm.py:26-36) and InsulinPump.basal/.bolus (actuator/pump.py:23-39). */
typedef struct {
    int32_t sample_time;     /* minutes between CGM samples == minutes per env-step (env.py:75) */
    double sensor_min, sensor_max;
    double pump[6];          /* min_basal, max_basal, inc_basal, min_bolus, max_bolus, inc_bolus */
} sg_hardware;

typedef struct sg_context *sg_handle;

/* Library / build identification; sg_version() == SG_ABI_VERSION. */
int sg_version(void);
const char *sg_last_error(void);

/* Allocates the resident state for up to `capacity` episodes on CUDA device `device`.
 * Stands where the reference builds one T1DSimEnv per patient (simulation/env.py:37-42,
 * simulation/user_interface.py:364-372). */
int sg_create(int device, int64_t capacity, uint32_t flags, sg_handle *out);
int sg_destroy(sg_handle h);

/* CGMSensor.withName / InsulinPump.withName (sensor/cgm.py:18-24, actuator/pump.py:16-21). */
int sg_set_hardware(sg_handle h, const sg_hardware *hw_host);

/* Parameter table: n_rows x SG_NPARAM (row-major, HOST) and the matching initial states
 * n_rows x SG_NX (vpatient_params.csv columns x0_1..x0_13).  T1DPatient.withID/withName
 * (patient/t1dpatient.py:44-68). */
int sg_load_params(sg_handle h, const double *params_host, const double *x0_host, int32_t n_rows);

/* T1DSimEnv.reset (simulation/env.py:119-155) for n_episodes episodes at once.
 *   patient_row [E]            row of the table for each episode
 *   params_soa  [SG_NPARAM][E] per-episode parameters (perturbed populations); NULL = use the table
 *   x0          [SG_NX][E]     post-reset patient state (T1DPatient.reset t1dpatient.py:247-281,
 *                              incl. random_init_bg); NULL = table x0
 *   meal_minute [M][E] int32   episode minute of each announced meal in firing order, -1 = none
 *   meal_grams  [M][E]         grams (scenario.get_action, simulation/scenario.py:33-42, scenario_gen.py:15-31)
 *   noise       [n_noise][E]   CGM noise stream (sensor/noise_gen.py:61-69); the reset consumes 2 samples,
 *                              every env-step one more
 *   row0        [SG_NTRACE][E] history row 0 (may be NULL)
 * The meal/noise/params buffers are BORROWED until the next sg_reset or sg_destroy. */
int sg_reset(sg_handle h, int64_t n_episodes, const int32_t *patient_row, const double *params_soa,
             const double *x0, const int32_t *meal_minute, const double *meal_grams, int32_t max_meals,
             const double *noise, int64_t n_noise, double *row0, void *stream);

/* T1DSimEnv.step (simulation/env.py:66-117), lockstep over all episodes: one env-step of
 * sample_time minutes with Action(basal[e], bolus[e]).  out [SG_NTRACE][E] = the new history row. */
int sg_step(sg_handle h, const double *basal, const double *bolus, double *out, void *stream);

/* Same as sg_step, with the meal announcements of this step supplied by the caller instead of the meal
 * table: meals [sample_time][E] = scenario.get_action(time).meal for each minute of the step (env.py:50).
 * This is the seam for arbitrary Scenario subclasses (simulation/scenario.py:10-21) evaluated on the host. */
int sg_step_meals(sg_handle h, const double *basal, const double *bolus, const double *meals, double *out,
                  void *stream);

/* Replaces the buffers borrowed by sg_reset without resetting the episodes (longer noise stream / meal
 * table as an open-ended episode proceeds).  Positions already consumed stay valid. */
int sg_rebind_streams(sg_handle h, const int32_t *meal_minute, const double *meal_grams, int32_t max_meals,
                      const double *noise, int64_t n_noise);

/* Heavy-minute compaction for populations whose episodes have meal schedules of their own (gym-style vector
 * environments, perturbed sweeps): a minute that needs more than `cap` DOPRI5 steps (a meal onset takes up to 36,
 * an ordinary minute 3) is put aside by the lane that owns it and finished, with the rest of that env-step, by a
 * second launch over the compacted list of such episodes, so that the other 31 lanes of its warp do not wait.
 * sg_step / sg_step_meals / sg_run then issue two launches per env-step.  Results are bit-identical to cap = 0
 * (the default, off).  cap >= 4.  The scratch (about 200 B per episode of capacity) is allocated here.
 * Pays for lockstep stepping of large populations (measured on 1,048,576 environments at meal times: 8.2 -> 5.2 ms
 * per env-step); the second launch has the latency of the longest minute, so small populations and fused
 * multi-step runs are faster with cap = 0. */
int sg_set_step_cap(sg_handle h, int32_t cap);

/* Schedule-coherent launch order for the same kind of population (the other answer to the same divergence, and the
 * one that also pays for fused multi-step runs): every `segment_steps` env-steps the episodes are counting-sorted by
 * the minute of their next meal event (episodes in the middle of a meal: by the minutes of eating left) and the
 * grams, and thread i of the following launches owns episode perm[i] -- the lanes of a warp then meet the meal onset,
 * the end of the meal and the bolus edges in the same minutes.  Only the thread -> episode mapping moves; every array
 * stays indexed by the episode and per-episode results are bit-identical to segment_steps = 0 (default, off).
 * sg_run is cut into one launch per segment (plus three small sort launches); sg_step re-sorts every segment_steps calls.
 * The scratch (8 B per episode of capacity + the histogram) is allocated here.  segment_steps * sample_time <= 4096. */
int sg_set_coherence(sg_handle h, int32_t segment_steps);

/* What the launch order of sg_set_coherence is sorted by.
 *   0 (default)  the next meal event, as described above: for orders that are kept for tens of env-steps;
 *   1            adaptive, for orders that are re-sorted every few env-steps (a lockstep vector environment whose
 *                environments restart one by one): an episode with a meal announced inside the segment is keyed by
 *                (minutes until it, grams); every other episode by the DOPRI5 steps per env-step it took since the
 *                previous sort (statistics row SG_STAT dopri_steps), restarted episodes last -- a fresh start costs
 *                8 - 9 steps in its first minute and 4 - 5 for the next twenty against 3 for a settled episode, and
 *                whatever else makes an episode expensive shows in the steps it just took.
 * Like the order itself this cannot reach a result.  8 B of scratch per episode of capacity for mode 1. */
int sg_set_sort_mode(sg_handle h, int32_t mode);
/* The current launch order perm_out [E] (a permutation of 0..E-1); SG_ERR_STATE before the first sorted launch. */
int sg_get_launch_order(sg_handle h, int32_t *perm_out, void *stream);

/* T1DSimEnv.reset for SOME episodes of a running population: the gym wrapper's reset on `done`
 * (envs/simglucose_gym_env.py:48-51 builds a new env and resets it), batched.  mask [E] int32: non-zero = reset this
 * episode (state, meal machine, controller state, statistics, stream positions back to 0, two noise samples
 * consumed); the others are untouched.  The caller has already overwritten the masked COLUMNS of the bound noise /
 * meal buffers (and x0 [SG_NX][E] if given; NULL = table x0) with the new episode's streams, e.g. with the keyed
 * generators below.  row0 [SG_NTRACE][E] (may be NULL) receives history row 0 of the masked episodes only.
 * From here on episodes have stream positions of their own: the host-side length check of sg_step / sg_run is
 * replaced by the kernel's own guard, which marks an episode that runs past its noise stream with status -5. */
int sg_reset_masked(sg_handle h, const int32_t *mask, const double *x0, double *row0, void *stream);

/* patient.t at reset: T1DPatient(..., t0=...) (patient/t1dpatient.py:27,277).  Default 0. */
int sg_set_t0(sg_handle h, int32_t t0);

/* T1DPatient.step (patient/t1dpatient.py:82-116) for every episode: meal state machine + one minute of
 * dopri5 with an explicit patient Action(CHO = grams announced this minute, insulin U/min).
 * out [SG_NPOUT][E] (may be NULL): x[13], Gsub (observation, t1dpatient.py:210-220), grams eaten this
 * minute, idid of this call, DOPRI5 steps taken, then the meal machine after the minute: planned_meal,
 * _last_Qsto, _last_foodtaken, is_eating (t1dpatient.py:88-104, 222-236). */
#define SG_NPOUT 21
int sg_patient_step(sg_handle h, const double *cho, const double *insulin, double *out, void *stream);

/* T1DPatient.model (patient/t1dpatient.py:118-208): dxdt [SG_NX][n] for n independent evaluations.
 * x [SG_NX][n]; cho, insulin, last_qsto, last_food [n]; parameters from params_soa [SG_NPARAM][n] or,
 * when NULL, from the loaded table by patient_row [n]. */
int sg_model(sg_handle h, int64_t n, const double *x, const double *cho, const double *insulin,
             const double *last_qsto, const double *last_food, const int32_t *patient_row,
             const double *params_soa, double *dxdt, void *stream);

/* InsulinPump.basal / .bolus (actuator/pump.py:23-39), elementwise on the current device. */
int sg_pump(const double *amount, double *out, int64_t n, double vmin, double vmax, double inc, void *stream);

/* risk_index([BG], 1) (analysis/risk.py:5-17), elementwise: out [3][n] = LBGI, HBGI, RI. */
int sg_risk_index(const double *bg, double *out, int64_t n, void *stream);

/* risk_index(BG, horizon) (analysis/risk.py:5-17) over a window: bg [n] holds BG[-horizon:]; out [3] = LBGI,
 * HBGI, RI with LBGI = mean of 10 f^2 over the samples with f < 0 (0 when there are none), HBGI likewise
 * for f > 0.  One block, fixed-order sums. */
int sg_risk_window(const double *bg, double *out, int64_t n, void *stream);

/* The sampling branch of CGMSensor.measure (sensor/cgm.py:28-33), elementwise:
 * out[i] = min(max(bg[i] + noise[i], sensor_min), sensor_max). */
int sg_cgm_measure(const double *bg, const double *noise, double *out, int64_t n, double sensor_min,
                   double sensor_max, void *stream);

/* SimObj.simulate (simulation/sim_engine.py:29-39) for every episode, n_steps env-steps fused in
 * one launch with the controller evaluated on device.
 *   ctrl_host [SG_NCTRL]      controller constants used for every episode when ctrl_rows is NULL
 *   ctrl_rows [SG_NCTRL][E]   per-episode constants (BB: target, CR, CF from Quest.csv)
 *   actions   [n_steps][2][E] (basal, bolus) when controller == SG_CTRL_ACTIONS
 *   trace     [n_steps][SG_NTRACE][E] history rows 1..n_steps (may be NULL)
 *   counters  [n_steps*sample_time][2][E] int32 (DOPRI5 steps, rejects) per minute (may be NULL) */
int sg_run(sg_handle h, int32_t controller, const double *ctrl_host, const double *ctrl_rows,
           const double *actions, int32_t n_steps, double *trace, int32_t *counters, void *stream);

/* Rows of the resident per-episode state as sg_get_episode_state returns it. */
typedef enum {
    SG_D_X = 0,              /* 13 rows: patient state                              t1dpatient.py:238-240 */
    SG_D_PLANNED = 13,       /* planned_meal                                        t1dpatient.py:225 */
    SG_D_LAST_CHO,           /* _last_action.CHO                                    t1dpatient.py:106 */
    SG_D_LAST_QSTO,          /* _last_Qsto                                          t1dpatient.py:90 */
    SG_D_LAST_FOOD,          /* _last_foodtaken                                     t1dpatient.py:91,99 */
    SG_D_H_CARRY,            /* dopri5 WORK(7) (SG_FLAG_CARRY_H only) */
    SG_D_LAST_CGM,           /* CGMSensor._last_CGM                                 cgm.py:33 */
    SG_D_CGM_PREV,           /* CGM_hist[-1]                                        env.py:96 */
    SG_D_OBS,                /* last observation.CGM                                env.py:105,142 */
    SG_D_INFO_MEAL,          /* info['meal'] of the last step                       env.py:112 */
    SG_D_PID_INTEG,          /* PIDController.integrated_state                      pid_ctrller.py:15 */
    SG_D_PID_PREV,           /* PIDController.prev_state                            pid_ctrller.py:16 */
    SG_ND
} sg_state_d;
typedef enum {
    SG_I_EATING = 0,         /* is_eating                                           t1dpatient.py:92,104 */
    SG_I_MINUTE,             /* patient.t as an integer minute                      t1dpatient.py:73-75 */
    SG_I_NOISE_POS,          /* noise samples consumed */
    SG_I_MEAL_POS,           /* next unconsumed meal-table entry */
    SG_I_STATUS,             /* scipy idid of the first failed minute, else 1 */
    SG_NI
} sg_state_i;

/* Resident state readers (T1DPatient.state t1dpatient.py:238-240; env histories). */
/* The whole resident state: d_out [SG_ND][E] and i_out [SG_NI][E] (either may be NULL). */
int sg_get_episode_state(sg_handle h, double *d_out, int32_t *i_out, void *stream);
/* The inverse: overwrite the resident state of the n_episodes of the last sg_reset (checkpoint / resume, and
 * handing episodes simulated in one batch back to per-patient handles).  Either pointer may be NULL.
 * steps_done = env-steps these episodes have already taken (what the noise-stream length check counts). */
int sg_set_episode_state(sg_handle h, const double *d_in, const int32_t *i_in, int64_t steps_done, void *stream);
int sg_get_state(sg_handle h, double *x_out /* [SG_NX][E] */, void *stream);
int sg_get_stats(sg_handle h, double *stats_out /* [SG_NSTAT][E] */, void *stream);
/* Per-episode solver status: 1 ok, -2 (more than 500 steps), -3 (step size too small): scipy's idid,
 * checked by the reference through _odesolver.successful() (t1dpatient.py:112-116); -5: the episode needed a CGM
 * noise sample beyond the bound stream (zero noise was used from there on). */
int sg_get_status(sg_handle h, int32_t *status_out /* [E] */, void *stream);

/* Population aggregate of the per-episode statistics, reduced with warp shuffles then a fixed-order
 * pass over per-block partials (deterministic): agg_out [SG_NSTAT] (sums; MIN_BG/MAX_BG extrema;
 * FIRST_DONE = number of episodes that ever reported done). */
int sg_reduce_stats(sg_handle h, double *agg_out, void *stream);

/* The same aggregate for callers that may not allocate device memory.  (No counterpart in the reference: it has no
 * device and no allocator; this exists for the hosts described next.)
 *
 * Who needs it: code that is handed device pointers by an embedding application but is not allowed to call
 * cudaMalloc itself -- a plugin under a fixed memory budget, a host whose pool allocator must account for every
 * byte, a real-time loop in which a driver allocation could block.  Such a caller cannot hold an sg_handle
 * (sg_create allocates the resident state and the reduction's partial-sum buffer), so this entry point takes the
 * per-episode statistics directly and keeps its partial sums in an array compiled into the library:
 * SG_REDUCE_STATIC_MAX_BLOCKS x SG_NSTAT doubles = 75,776 bytes, held by every GPU on which the library's code is
 * loaded until the process exits.  It calls no cudaMalloc / cudaFree and no allocating library routine.
 *
 *   stats     [SG_NSTAT][capacity] device, the layout sg_get_stats writes (capacity = row stride >= n_episodes)
 *   agg_out   [SG_NSTAT] device; bit-identical to what sg_reduce_stats gives for the same statistics
 *   limit     n_episodes <= SG_REDUCE_STATIC_MAX_EPISODES (the largest population BASELINE.json names, and the
 *             largest this entry point is specified for); more is refused with SG_ERR_ARG and a message in
 *             sg_last_error -- use sg_reduce_stats, or reduce in pieces and combine
 *   device    the CURRENT device of the calling thread (there is no handle to name one); stats / agg_out live there
 *   first call on each device: resolves the address of the built-in array on that device (cudaGetSymbolAddress).
 *             Make that first call during set-up, before any stream capture is begun.
 *   sharing   every call on a device uses the SAME array, on any stream and from any thread.  Each call overwrites
 *             the rows it later reads, so nothing has to be cleared between calls and stale rows of an earlier,
 *             larger call are never read; but two calls that can overlap on the GPU (different streams without an
 *             event between them, or two host threads) race on it and give garbage.  Keep them on one stream or
 *             order them yourself. */
#define SG_REDUCE_STATIC_MAX_BLOCKS 592
#define SG_REDUCE_STATIC_MAX_EPISODES 16777216
int sg_reduce_stats_static(const double *stats, int64_t n_episodes, int64_t capacity, double *agg_out, void *stream);

int64_t sg_num_episodes(sg_handle h);
/* Kernels launched through this handle since sg_create (bench.py's gpu_launches). */
int64_t sg_launch_count(sg_handle h);

/* ---- device-side generators of the per-episode input streams (distribution-compatible with the
 * reference's numpy generators; counter-based Philox4x32-10 keyed by (seed, (episode_offset + e) / group_size): the episodes of
 * one group -- e.g. the 30 patients simulated under one scenario/sensor seed -- share a stream) ---- */

/* CGMNoise (sensor/noise_gen.py:15-97): out [n_noise][E].  W_dev [nsample][11] is the constant matrix of
 * the cubic interp1d from eleven 15-minute knots to the sensor grid (nsample = floor(150/sample_time)+1,
 * noise_gen.py:40-46); model_host = PACF, gamma, lambda, delta, xi (sensor_params.csv). */
int sg_gen_noise(double *out, int64_t n_episodes, int64_t n_noise, int32_t nsample, const double *W_dev,
                 const double *model_host, uint64_t seed, int64_t episode_offset, int64_t group_size, void *stream);

/* RandomScenario.create_scenario (simulation/scenario_gen.py:33-60), one schedule per calendar day, as a
 * meal table meal_minute/meal_grams [max_meals][E] in firing order (-1 terminated). */
int sg_gen_scenarios(int32_t *meal_minute, double *meal_grams, int64_t n_episodes, int32_t max_meals,
                     int32_t start_minute, int32_t horizon_minutes, uint64_t seed, int64_t episode_offset,
                     int64_t group_size, void *stream);

/* Keyed variants for vector environments, whose episodes restart one by one (gym reset on `done`): key [E] int64
 * names the Philox stream of each episode (e.g. env index + num_envs * episodes started in that slot) instead of its
 * position, only the columns with mask[e] != 0 are written (mask NULL = all), and the scenario start is per episode
 * (start_minute [E], minute of day). */
int sg_gen_noise_keyed(double *out, int64_t n_episodes, int64_t n_noise, int32_t nsample, const double *W_dev,
                       const double *model_host, uint64_t seed, const int64_t *key, const int32_t *mask, void *stream);
int sg_gen_scenarios_keyed(int32_t *meal_minute, double *meal_grams, int64_t n_episodes, int32_t max_meals,
                           const int32_t *start_minute, int32_t horizon_minutes, uint64_t seed, const int64_t *key,
                           const int32_t *mask, void *stream);

/* Where and how an episode starts, as the gym wrapper draws it (envs/simglucose_gym_env.py:58-73):
 *   start_minute [E] (may be NULL)  60 * randint(0, 24)
 *   x0 [SG_NX][E]    (may be NULL)  T1DPatient.reset with random_init_bg=True (patient/t1dpatient.py:256-270): the table
 *                    state of patient_row[e] with x[3], x[4], x[12] ~ N(mean, 0.1 * mean), independent; n_draws
 *                    successive randomisations, each around the one before with the same deviates (the wrapper resets
 *                    the patient twice: n_draws = 2; 0 = the table state)
 * Distribution-compatible (Philox); the host replay for given gym seeds is streams.random_init_state / np_random. */
int sg_gen_episode_starts(int32_t *start_minute, double *x0, const double *table_x0_dev, const int32_t *patient_row,
                          int64_t n_episodes, int32_t n_draws, uint64_t seed, const int64_t *key, const int32_t *mask,
                          void *stream);

/* The keyed generators for a LIST of episodes (list [0 .. *count) on the device, as sg_vec_finish leaves it) instead of a
 * mask: the thread blocks stride over the list, so the launch takes (listed / resident blocks) episodes of time however the
 * listed episodes are spread over the population.  Same streams as the keyed (and masked) variants, bit for bit. */
int sg_gen_noise_listed(double *out, int64_t n_episodes, int64_t n_noise, int32_t nsample, const double *W_dev,
                        const double *model_host, uint64_t seed, const int64_t *key, const int32_t *list,
                        const int32_t *count, void *stream);
int sg_gen_scenarios_listed(int32_t *meal_minute, double *meal_grams, int64_t n_episodes, int32_t max_meals,
                            const int32_t *start_minute, int32_t horizon_minutes, uint64_t seed, const int64_t *key,
                            const int32_t *list, const int32_t *count, void *stream);

/* Perturbed population (BASELINE.json configs[3]): params_soa[i][e] = table[row[e]][i] * (1 + amplitude*u),
 * u ~ U(-1, 1) i.i.d. per parameter.  table_dev: [n_rows][SG_NPARAM] on the device. */
int sg_gen_perturbed_params(double *params_soa, const double *table_dev, const int32_t *patient_row,
                            int64_t n_episodes, double amplitude, uint64_t seed, int64_t episode_offset,
                            void *stream);

/* Book-keeping of a lockstep vector environment around per-environment resets -- the `if done: env.reset()` of an RL loop
 * over envs/simglucose_gym_env.py:48-51, for every environment at once (simglucose_b200/envs/vector_env.py).
 * sg_vec_finish, after an env-step: episode_steps[e] += 1; done[e] = row[done_row][e] != 0 (env.py:103); mask[e] = done or
 *   episode_steps[e] >= horizon_steps; truncated[e] = mask and not done; terminal_obs[e] = row[obs_row][e];
 *   key[e] = ((env_offset + e) / group_size << key_shift) | episodes_started[e], the Philox stream of the episode that
 *   would replace e (sg_gen_*_keyed).  row: the [SG_NTRACE][n_envs] history row of the step.  list / count (both or
 *   neither): the environments with mask != 0 gathered into list[0 .. *count) in no particular order (device memory,
 *   list [n_envs] int32, count [1] int32), for sg_gen_noise_listed / sg_gen_scenarios_listed.
 * sg_vec_restart, after sg_reset_masked(mask): episodes_started[e] += 1 and episode_steps[e] = 0 where mask[e] != 0;
 *   obs_out[e] = mask[e] ? obs_new[e] : obs_step[e]. */
int sg_vec_finish(const double *row, int64_t n_envs, int32_t done_row, int32_t obs_row, int32_t *episode_steps,
                  int32_t horizon_steps, const int64_t *episodes_started, int64_t env_offset, int64_t group_size,
                  int32_t key_shift, int32_t *mask, int64_t *key, double *terminal_obs, uint8_t *done, uint8_t *truncated,
                  int32_t *list, int32_t *count, void *stream);
int sg_vec_restart(const int32_t *mask, int64_t n_envs, int64_t *episodes_started, int32_t *episode_steps,
                   const double *obs_new, const double *obs_step, double *obs_out, void *stream);

/* out[i] ~ U(lo, hi): random basal actions (BASELINE.json configs[2]). */
int sg_gen_uniform(double *out, int64_t n, double lo, double hi, uint64_t seed, uint64_t offset, void *stream);

/* ---- post-hoc statistics of BG traces (simglucose/analysis/report.py), one thread per episode ----
 * bg: element (row r, episode e) at bg[r * row_stride + e] (e.g. the BG column of an sg_run trace: row_stride =
 * SG_NTRACE * E).  Any output may be NULL.
 *   percent_out [5][E]            percent_stats (report.py:74-92): 70<=BG<=180, BG>180, BG<70, BG>250, BG<50
 *   ri_out      [3][n_chunks][E]  risk_index_trace (report.py:95-133): LBGI, HBGI, RI of each chunk of chunk_rows
 *                                 rows (the reference uses 60 rows); n_chunks = ceil(n_rows / chunk_rows)
 *   ri_mean_out [3][E]            their means over the chunks (needs ri_out)
 *   cvga_out    [2][E]            CVGA_analysis (report.py:198-204): 2.5 and 97.5 percentiles (numpy's linear
 *                                 interpolation of exact order statistics), clipped to [50, 400] */
int sg_report_stats(const double *bg, int64_t row_stride, int64_t n_rows, int64_t n_episodes, int32_t chunk_rows,
                    double *percent_out, double *ri_out, double *ri_mean_out, double *cvga_out, void *stream);

/* CVGA zone counts (report.py:206-216) over the episodes: counts_out [5] = #A, #(A or B), #C, #D, #E. */
int sg_cvga_zones(const double *cvga, int64_t n_episodes, double *counts_out, void *stream);

/* Accuracy probe of the kernel's own FP64 helpers (csrc/sg_math.cuh) against CUDA's libm, elementwise:
 * kind 0 div_nr(a,b), 1 a/b, 2 tanh_diff(a,b), 3 tanh(a)-tanh(b), 4 pow_pos(a,b), 5 pow(a,b), 6 exp_core(a),
 * 7 exp(a), 8 log_core(a), 9 log(a), 10 2 * tanh_sum_half(2a, 2b) (tanh(a) - tanh(b) + 2 as the RHS forms it),
 * 11 tanh(a) - tanh(b) + 2, 12 rcp_ulp(a), 13 is_neg(a) + 2 * is_pos(a), 14 exp_tanh(a).  Test hook only. */
int sg_math_probe(int32_t kind, const double *a, const double *b, double *out, int64_t n, void *stream);

/* FP64 pipe probe: n_blocks x n_threads threads each run `iters` x 8 independent DFMA; out gets one double per
 * thread.  Flops = n_blocks * n_threads * iters * 16.  bench.py times it for the roofline denominator. */
int sg_probe_dfma(double *out, int32_t n_blocks, int32_t n_threads, int64_t iters, void *stream);

#ifdef __cplusplus
}
#endif
#endif
