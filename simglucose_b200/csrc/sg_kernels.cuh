// Kernel argument blocks and launch entry points shared by sg_kernels.cu (compiled twice: with
// -fmad=false as namespace sg_strict and with -fmad=true as namespace sg_fma) and sg_capi.cu.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace sgk {

// Resident per-episode state, structure-of-arrays: sd[i * cap + e], si[i * cap + e].
enum StateD {
    SD_X0 = 0,                 // 13 doubles: patient state (t1dpatient.py:238-240)
    SD_PLANNED = 13,           // planned_meal                     t1dpatient.py:225
    SD_LAST_CHO,               // _last_action.CHO                 t1dpatient.py:106
    SD_LAST_QSTO,              // _last_Qsto                       t1dpatient.py:90
    SD_LAST_FOOD,              // _last_foodtaken                  t1dpatient.py:91,99
    SD_H_CARRY,                // WORK(7) of dopri5 (only with SG_FLAG_CARRY_H)
    SD_LAST_CGM,               // CGMSensor._last_CGM              cgm.py:33
    SD_CGM_PREV,               // CGM_hist[-1]                     env.py:96
    SD_OBS,                    // last observation.CGM             env.py:105,142
    SD_INFO_MEAL,              // info['meal'] of the last step    env.py:112
    SD_PID_INTEG,              // PIDController.integrated_state   pid_ctrller.py:15
    SD_PID_PREV,               // PIDController.prev_state         pid_ctrller.py:16
    SD_COUNT
};
enum StateI {
    SI_EATING = 0,             // is_eating                        t1dpatient.py:92,104
    SI_MINUTE,                 // patient.t as an integer minute   t1dpatient.py:73-75
    SI_NOISE_POS,              // samples consumed from the noise stream
    SI_MEAL_POS,               // next unconsumed entry of the meal table
    SI_STATUS,                 // scipy idid of the first failed minute (-5: noise stream exhausted), else 1
    SI_COUNT
};

constexpr int NTRACE = 12;
constexpr int NSTAT = 16;
constexpr int NCTRL = 4;
constexpr int BLOCK = 64;
// scratch rows of an episode whose minute was put aside: k1[13], h, x, log(facold), ins, announced, CHO, insulin, BG, CGM
constexpr int DEFER_ND = 22;
// nstep, naccpt, nrej, reject, minute index inside the env-step, steps / rejects of the env-step so far
constexpr int DEFER_NI = 7;

struct Hardware {
    int st;
    double smin, smax;
    double pump[6];
};

constexpr int STATUS_STREAM_EXHAUSTED = -5;   // the episode asked for a noise sample beyond the bound stream

struct EpisodeArgs {
    int64_t n;                 // episodes
    int64_t cap;               // stride of the resident state arrays
    // Schedule-coherent launch order (sg_set_coherence): thread i owns episode perm[i]; null = identity.  Only the
    // thread -> episode mapping moves: every array stays indexed by the episode, so results do not depend on it.
    const int32_t *perm;
    double *sd;
    int32_t *si;
    double *stats;             // [NSTAT][cap]
    // parameters: per-episode SoA [32][n] if non-null, else table[row * 32 + i] by patient_row
    const double *params_soa;
    const double *table;
    const double *table_x0;    // [n_rows][13]
    const int32_t *patient_row;
    // borrowed inputs
    const int32_t *meal_minute;
    const double *meal_grams;
    int32_t max_meals;
    const double *noise;
    int64_t n_noise;
    Hardware hw;
    int32_t carry_h;
    // run
    int32_t controller;
    int64_t step0;             // env-steps this handle has launched since sg_reset (SG_CTRL_RANDOM: Philox counter)
    double ctrl[NCTRL];
    const double *ctrl_rows;   // [NCTRL][n] or null
    const double *act_basal;   // step k of episode e at act_basal[k * act_stride + e]
    const double *act_bolus;
    int64_t act_stride;
    int32_t n_steps;
    const double *step_meals;  // [st][n] grams announced in each minute of a single step (n_steps == 1), or null: table
    double *trace;             // [n_steps][NTRACE][n] or null
    int32_t *counters;         // [n_steps*st][2][n] or null
    // heavy minutes put aside (sg_set_step_cap): phase A fills, phase B consumes
    int32_t step_cap;          // DOPRI5 steps after which a minute is put aside; 0 = never
    double *defer_d;           // [DEFER_ND][cap]
    int32_t *defer_i;          // [DEFER_NI][cap]
    int32_t *defer_list;       // [cap] episodes put aside by the last phase A launch
    int32_t *defer_count;      // [1]
    // reset
    const double *x0;          // [13][n] or null
    double *row0;              // [NTRACE][n] or null
    int32_t t0;                // patient.t at reset (T1DPatient(t0=...), t1dpatient.py:27,277)
    const int32_t *reset_mask; // [n] or null: only episodes with a non-zero entry are reset (sg_reset_masked)
};

// T1DPatient.step for every episode: one minute with an explicit Action(CHO announced, insulin).
constexpr int NPOUT = 21;      // out rows: x[13], Gsub, to_eat, idid, dopri5 steps, planned, last_Qsto, last_food, eating
struct PatientArgs {
    EpisodeArgs e;
    const double *cho;         // [n] grams announced this minute
    const double *insulin;     // [n] U/min
    double *out;               // [NPOUT][n] or null
};

// T1DPatient.model evaluated for n independent (x, action, last_Qsto, last_foodtaken).
struct ModelArgs {
    int64_t n;
    const double *x;           // [13][n]
    const double *cho, *insulin, *last_qsto, *last_food;   // [n]
    const double *params_soa;  // [32][n] or null
    const double *table;
    const int32_t *patient_row;
    double *dxdt;              // [13][n]
};

struct ElementwiseArgs {
    int64_t n;
    const double *in;
    double *out;
    double p[3];
    const double *in2;         // second operand (cgm_kernel: noise)
};

struct ReduceArgs {
    int64_t n, cap;
    const double *stats;
    double *partials;          // [gridDim.x][NSTAT]
    double *out;               // [NSTAT]
    int32_t n_blocks;
};

#define SG_DECLARE_LAUNCHERS(NS)                                              \
    namespace NS {                                                            \
    cudaError_t launch_reset(const sgk::EpisodeArgs &a, cudaStream_t s);      \
    cudaError_t launch_run(const sgk::EpisodeArgs &a, cudaStream_t s);        \
    cudaError_t launch_run_capped(const sgk::EpisodeArgs &a, cudaStream_t s); \
    cudaError_t launch_reduce(const sgk::ReduceArgs &a, cudaStream_t s);      \
    cudaError_t launch_patient_step(const sgk::PatientArgs &a, cudaStream_t s); \
    cudaError_t launch_model(const sgk::ModelArgs &a, cudaStream_t s);        \
    cudaError_t launch_pump(const sgk::ElementwiseArgs &a, cudaStream_t s);   \
    cudaError_t launch_risk(const sgk::ElementwiseArgs &a, cudaStream_t s);   \
    cudaError_t launch_cgm(const sgk::ElementwiseArgs &a, cudaStream_t s);    \
    cudaError_t launch_risk_window(const sgk::ElementwiseArgs &a, cudaStream_t s); \
    cudaError_t launch_math_probe(int kind, const double *a, const double *b, double *out, int64_t n, cudaStream_t s); \
    int reduce_blocks(int64_t n);                                             \
    }

}  // namespace sgk

SG_DECLARE_LAUNCHERS(sg_strict)
SG_DECLARE_LAUNCHERS(sg_fma)
namespace sg_fma { cudaError_t launch_run_fp32(const sgk::EpisodeArgs &a, cudaStream_t s); }   // built with FMA contraction only

// Schedule-coherent launch order (sg_coherence.cu): a counting sort of the episodes by the minute of their next
// meal event, so that the lanes of a warp meet the expensive minutes (meal onset / offset, bolus) together.
namespace sgc {
constexpr int GRAM_BUCKETS = 128;
struct SortArgs {
    int64_t n, cap;
    const double *sd;
    const int32_t *si;
    const int32_t *meal_minute;    // [max_meals][n]
    const double *meal_grams;
    int32_t max_meals;
    const int32_t *patient_row;    // [n] or null
    int32_t mode;                  // 0: key (minute, grams); 1: key (minute, patient row, grams / 8); 2: adaptive
    // mode 2 (sg_set_sort_mode(h, 1)): episodes without a meal onset inside the window are keyed by the DOPRI5 steps
    // they took since the previous sort (statistics row 14; prev_steps keeps the last value seen per episode)
    const double *stats;           // [NSTAT][cap]
    double *prev_steps;            // [cap]
    int32_t steps_between;         // env-steps between two sorts (normalises the step count)
    int32_t st;                    // sample_time
    int32_t window;                // minutes ahead that are told apart (the segment until the next sort)
    int32_t n_keys;                // key_count(window)
    int32_t *keys;                 // [cap] scratch
    int32_t *hist;                 // [n_keys] scratch: counts, then running offsets
    int32_t *perm;                 // [cap] out: thread i of the next launches owns episode perm[i]
};
inline int32_t key_count(int32_t window) { return (window + 1) * GRAM_BUCKETS + 1; }
cudaError_t launch_sort(const SortArgs &a, cudaStream_t s);    // memset + 3 kernels
}  // namespace sgc
