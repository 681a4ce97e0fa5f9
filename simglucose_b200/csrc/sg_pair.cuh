// Warp-pair organisation of the episode kernel (included by sg_kernels.cu inside namespace SG_NS).
//
// 64-thread blocks own 32 patient-episodes: lane l of warp 0 (role G) and lane l of warp 1 (role I) work on
// the SAME episode.  The UVa/Padova model splits along its physiology (t1dpatient.py:118-208):
//   role G  x0..x4, x12   gut (with the tanh gastric emptying), glucose kinetics, s.c. glucose; reads x6, x8
//   role I  x5..x11       insulin kinetics / action and s.c. insulin; reads nothing of role G
// Both roles run the SAME instruction stream over seven state slots (role G pads one): only the RHS halves and
// the step-size control branch on the role.  That matters: the first version of this kernel gave each role a
// loop body of its own, 32 KB of hot code together, and lost 27 % to instruction fetch (profiles/r01_s30_*).
// Per RHS pass two doubles cross from I to G through a double-buffered mailbox in shared memory, one
// __syncthreads per pass.  The error / HINIT norms are sums over the 13 states in index order: role I reads the
// terms of role G from shared memory, adds everything in that order, runs the step-size controller and
// publishes the decision (a code word and h); G only follows.  Role I also owns the closed loop around the
// integrator (sensor, controller, pump, risk, statistics), role G the meal state machine.  Every state is
// integrated by exactly one thread with the operation order of the single-thread kernel, so in strict
// arithmetic the results are bit-identical to run_kernel (tests/test_gpu_pair.py).

namespace pair {

constexpr int PL = 32;                   // episodes per block
constexpr int PBLOCK = 64;               // threads per block: warp 0 = role G, warp 1 = role I
constexpr int NS = 7;                    // state slots per role
using Col = Column<PL, false>;       // plain layout: this kernel reads its rows one value at a time

// Rows of the episode's shared-memory column after the parameter rows (H_K1 is the first vector row of the
// single-thread layout).  Vector rows: [role][slot]; role G slots = x0, x1, x2, x3, x4, x12, pad; role I = x5..x11.
enum PairRow {
    R_K1 = H_K1, R_P7 = R_K1 + 2 * NS, R_PE = R_P7 + 2 * NS, R_Y = R_PE + 2 * NS,
    M_X6A = R_Y + 2 * NS, M_X8A,         // trial-state values role G reads, buffer 0
    M_X6B, M_X8B,                        //                                  buffer 1
    M_H, M_ACT,                          // decision of role I
    M_CHOA, M_CHOB,                      // running mean of the announced grams over the env-step (role G -> role I,
                                         // info['meal']), double-buffered by minute parity
    M_LAST_CGM,                          // CGMSensor._last_CGM (role I)
    PAIR_ROWS
};

// state index (t1dpatient.py order) -> row offset inside a vector
__device__ __forceinline__ int state_row(int i) { return i < 5 ? i : (i < 12 ? i + 2 : 5); }

enum ActBits { ACT_COMMIT = 1, ACT_BEGIN = 2, ACT_K1SMEM = 4, ACT_PROBE = 8, ACT_STOP = 16 };
enum Phase { PH_K1 = 0, PH_HINIT, PH_S2, PH_S3, PH_S4, PH_S5, PH_S6, PH_S7 };

// ---- the two halves of T1DPatient.model ---------------------------------------------------------------------
// slots: x0, x1, x2, x3, x4, x12, pad
__device__ __forceinline__ void rhs_g(const Col &c, double d, bool fed, const double (&x)[NS], double x6, double x8,
                                      double (&dx)[NS])
{
    const double kmax = c(H_kmax);
    const double qsto = x[0] + x[1];                                        // :126
    dx[0] = -kmax * x[0] + d;                                               // :133
    double kgut;
    if (fed) {                                                              // :135-142
        const double tdiff = sgm::tanh_diff(c(H_aa) * (qsto - c(H_bD)), c(H_cc) * (qsto - c(H_dD)));
        kgut = c(H_kmin) + c(H_hk) * (tdiff + 2.0);
    } else {
        kgut = kmax;
    }
    dx[1] = kmax * x[0] - x[1] * kgut;                                      // :145
    const double kabs = c(H_kabs);
    dx[2] = kgut * x[1] - kabs * x[2];                                      // :148
    const double Rat = sgm::div_r(c(H_fk) * x[2], c(H_BW), c(H_rBW));       // :151
    const double EGPt = c(H_kp1) - c(H_kp2) * x[3] - c(H_kp3) * x8;         // :153
    const double ke2 = c(H_ke2);
    const double Et = (x[3] > ke2) ? c(H_ke1) * (x[3] - ke2) : 0.0;         // :158-161
    const double egp = (0.0 > EGPt) ? 0.0 : EGPt;                           // :165
    const double k1 = c(H_k1), k2 = c(H_k2);
    dx[3] = mask_nonneg(x[3], egp + Rat - c(H_Fsnc) - Et - k1 * x[3] + k2 * x[4]);   // :165-167
    const double Vmt = c(H_Vm0) + c(H_Vmx) * x6;                            // :169
    const double Uidt = sgm::div_nr(Vmt * x[4], c(H_Km0) + x[4]);           // :171
    dx[4] = mask_nonneg(x[4], -Uidt + k1 * x[3] - k2 * x[4]);               // :172-173
    const double ksc = c(H_ksc);
    dx[5] = mask_nonneg(x[5], -ksc * x[5] + ksc * x[3]);                    // :201-202
    dx[6] = 0.0;
}

// slots: x5, x6, x7, x8, x9, x10, x11
__device__ __forceinline__ void rhs_i(const Col &c, double ins, const double (&x)[NS], double (&dx)[NS])
{
    const double ka2 = c(H_ka2), m2 = c(H_m2);
    dx[0] = mask_nonneg(x[0], -c(H_m24) * x[0] + c(H_m1) * x[4] + c(H_ka1) * x[5] + ka2 * x[6]);   // :176-179
    const double It = sgm::div_r(x[0], c(H_Vi), c(H_rVi));                  // :178
    const double p2u = c(H_p2u), ki = c(H_ki);
    dx[1] = -p2u * x[1] + p2u * (It - c(H_Ib));                             // :182
    dx[2] = -ki * (x[2] - It);                                              // :185
    dx[3] = -ki * (x[3] - x[2]);                                            // :187
    dx[4] = mask_nonneg(x[4], -c(H_m13) * x[4] + m2 * x[0]);                // :190-191
    const double kd = c(H_kd);
    dx[5] = mask_nonneg(x[5], ins - c(H_kad) * x[5]);                       // :194-195
    dx[6] = mask_nonneg(x[6], kd * x[5] - ka2 * x[6]);                      // :197-198
}

// ---- once-per-minute / once-per-step pieces of role I (not inlined: the loop body has to stay small) --------
struct HinitGuess2 { double h, dnf; };

// HINIT, first half, over all 13 states in index order (rows of both roles): h = 0.01 * |y| / |k1|.
static __device__ __noinline__ HinitGuess2 hinit_first_all(double *col, double hmax)
{
    const Col c{col};
    double dny = 0.0, dnf = 0.0;
#pragma unroll 1
    for (int i = 0; i < NX; i++) {
        const int r_ = state_row(i);
        const double yi = c(R_Y + r_);
        const double r = sgm::rcp_nr(ATOL + RTOL * fabs(yi));
        const double sk = ATOL + RTOL * fabs(yi);
        const double a = sgm::div_r(c(R_K1 + r_), sk, r), b = sgm::div_r(yi, sk, r);
        dnf = dnf + a * a;
        dny = dny + b * b;
    }
    double h = (dnf <= 1e-10 || dny <= 1e-10) ? 1.0e-6 : sqrt(dny / dnf) * 0.01;
    h = fmin(h, hmax);
    return HinitGuess2{h, dnf};
}

// HINIT, second half: second-derivative estimate from f(x + h, y + h*k1) (R_P7 rows) and k1 (R_K1 rows).
static __device__ __noinline__ double hinit_second_all(double *col, double h, double dnf, double hmax)
{
    const Col c{col};
    double der2 = 0.0;
#pragma unroll 1
    for (int i = 0; i < NX; i++) {
        const int r_ = state_row(i);
        const double sk = ATOL + RTOL * fabs(c(R_Y + r_));
        const double a = sgm::div_nr(c(R_P7 + r_) - c(R_K1 + r_), sk);
        der2 = der2 + a * a;
    }
    der2 = sqrt(der2) / h;
    const double der12 = fmax(fabs(der2), sqrt(dnf));
    const double h1 = (der12 <= 1e-15) ? fmax(1.0e-6, fabs(h) * 1.0e-3) : sgm::pow_pos(0.01 / der12, 1.0 / 5.0);
    return fmin(fmin(100.0 * fabs(h), h1), hmax);
}

// Integrator control state of role I.
struct Ctl {
    double x, xend, lfacold, dnf;
    int nstep, naccpt, nrej, idid;
    bool last, reject;
};
// hmax = |xend - x0| of one integrate(t + 1) call: exactly 1.0 for integer minutes below 2^52
constexpr double HMAX = 1.0;

// Start a step of size h at s.x (DOPCOR's loop head): returns the action bits, h may be shortened to hit xend.
__device__ __forceinline__ int begin_step(Ctl &s, double &h)
{
    if (s.nstep > NMAX) { s.idid = -2; return ACT_STOP; }
    if (0.1 * fabs(h) <= fabs(s.x) * UROUND) { s.idid = -3; return ACT_STOP; }
    if ((s.x + 1.01 * h - s.xend) > 0.0) { h = s.xend - s.x; s.last = true; }
    s.nstep++;
    return ACT_BEGIN;
}

// Error estimate of a finished step -> accept / reject, next step size (DOPCOR).  Returns the action bits.
__device__ __forceinline__ int control_step(Ctl &s, double errsum, double &h, double l4)
{
    const StepControl sc = step_control(errsum, s.lfacold, h);
    const double err = sc.err;
    double hnew = sc.hnew_accept;
    bool begin = false;
    int code;
    if (err <= 1.0) {
        s.lfacold = (err > 1.0e-4) ? sc.lerr : l4;           // facold = max(err, 1e-4)
        s.naccpt++;
        code = ACT_COMMIT;                                   // kt is k1 of the next step (FSAL)
        s.x = s.x + h;
        if (s.last) {
            s.idid = 1;
            code |= ACT_STOP;
        } else {
            if (fabs(hnew) > HMAX) hnew = HMAX;
            if (s.reject) hnew = sgm::min_sel(fabs(hnew), fabs(h));
            s.reject = false;
            begin = true;
        }
    } else {
        hnew = sc.hnew_reject;
        s.reject = true;
        if (s.naccpt >= 1) s.nrej++;
        s.last = false;
        code = ACT_K1SMEM;
        begin = true;
    }
    h = hnew;
    if (begin) code |= begin_step(s, h);
    return code;
}

// One call of ode('dopri5').integrate(t + 1) by the warp pair.  The accepted state stays in the R_Y rows of the
// episode's column between calls.  c: the episode's column; v: the same column shifted to this role's vector rows.
// u: CHO rate * 1000 (role G) / insulin rate per kg (role I).  par: mailbox buffer parity (kept across calls).
// hc: the episode's WORK(7) slot in the resident state (read / written by role I with SG_FLAG_CARRY_H only).
__device__ __forceinline__ int minute_pair(const Col &c, const Col &v, const int role, const bool valid, const double u,
                                           const bool fed, const double t0, const bool carry_h, double *hc,
                                           StepCount &cnt, int &par)
{
    constexpr unsigned FULL = 0xffffffffu;
    // The stage-4 sum lives only from stage 2 to stage 3: it waits in the R_PE rows, which are first written at
    // stage 3.  P5 is dead from stage 5 on: the squared error terms of stage 7 reuse its registers (sq).
    double yt[NS], kt[NS], P5[NS], P6[NS];
    double (&sq)[NS] = P5;
    double h = 0.0;
    int idid = valid ? 0 : 1, phase = PH_K1;
    Ctl s{t0, t0 + 1.0, c(H_L4), 0.0, 0, 0, 0, 0, false, false};

    if (valid) {
#pragma unroll
        for (int i = 0; i < NS; i++) yt[i] = v(R_Y + i);
        if (role == 1) { c.set(M_X6A + 2 * par, yt[1]); c.set(M_X8A + 2 * par, yt[3]); }
    }
    __syncthreads();

#pragma unroll 1
    while (__any_sync(FULL, idid == 0)) {
        const bool act = idid == 0;
        const bool ctl_phase = phase < PH_S2 || phase == PH_S7;
        const bool ctl_pass = __any_sync(FULL, act && ctl_phase);     // the same in both warps of the pair
        if (act) {
            if (role == 0) rhs_g(c, u, fed, yt, c(M_X6A + 2 * par), c(M_X8A + 2 * par), kt);
            else rhs_i(c, u, yt, kt);

            switch (phase) {
            case PH_K1:                                          // kt = k1 = f(x, y)
#pragma unroll
                for (int i = 0; i < NS; i++) v.set(R_K1 + i, kt[i]);
                break;
            case PH_HINIT:                                       // kt = f(x + h, y + h*k1)
#pragma unroll
                for (int i = 0; i < NS; i++) v.set(R_P7 + i, kt[i]);
                break;
            case PH_S2:                                          // kt = k2
#pragma unroll
                for (int i = 0; i < NS; i++) {
                    const double k1 = v(R_K1 + i);
                    yt[i] = v(R_Y + i) + h * (A31 * k1 + A32 * kt[i]);
                    v.set(R_PE + i, A41 * k1 + A42 * kt[i]);
                    P5[i] = A51 * k1 + A52 * kt[i];
                    P6[i] = A61 * k1 + A62 * kt[i];
                }
                phase = PH_S3;
                break;
            case PH_S3:                                          // kt = k3
#pragma unroll
                for (int i = 0; i < NS; i++) {
                    const double k1 = v(R_K1 + i);
                    yt[i] = v(R_Y + i) + h * (v(R_PE + i) + A43 * kt[i]);
                    P5[i] = P5[i] + A53 * kt[i];
                    P6[i] = P6[i] + A63 * kt[i];
                    v.set(R_P7 + i, A71 * k1 + A73 * kt[i]);
                    v.set(R_PE + i, E1 * k1 + E3 * kt[i]);
                }
                phase = PH_S4;
                break;
            case PH_S4:                                          // kt = k4
#pragma unroll
                for (int i = 0; i < NS; i++) {
                    yt[i] = v(R_Y + i) + h * (P5[i] + A54 * kt[i]);
                    P6[i] = P6[i] + A64 * kt[i];
                    v.set(R_P7 + i, v(R_P7 + i) + A74 * kt[i]);
                    v.set(R_PE + i, v(R_PE + i) + E4 * kt[i]);
                }
                phase = PH_S5;
                break;
            case PH_S5:                                          // kt = k5
#pragma unroll
                for (int i = 0; i < NS; i++) {
                    yt[i] = v(R_Y + i) + h * (P6[i] + A65 * kt[i]);
                    v.set(R_P7 + i, v(R_P7 + i) + A75 * kt[i]);
                    v.set(R_PE + i, v(R_PE + i) + E5 * kt[i]);
                }
                phase = PH_S6;
                break;
            case PH_S6:                                          // kt = k6; yt becomes the 5th-order solution
#pragma unroll
                for (int i = 0; i < NS; i++) {
                    yt[i] = v(R_Y + i) + h * (v(R_P7 + i) + A76 * kt[i]);
                    v.set(R_PE + i, v(R_PE + i) + E6 * kt[i]);
                }
                phase = PH_S7;
                break;
            default:                                             // kt = k7: squared scaled error terms
#pragma unroll
                for (int i = 0; i < NS; i++) {
                    const double e = (v(R_PE + i) + E7 * kt[i]) * h;
                    const double sk = ATOL + RTOL * sgm::max_sel(fabs(v(R_Y + i)), fabs(yt[i]));
                    const double a = sgm::div_nr(e, sk);
                    sq[i] = a * a;
                    v.set(R_PE + i, sq[i]);                      // role I reads the terms of role G from here
                }
                break;
            }
        }

        if (ctl_pass) {
            __syncthreads();                                     // the rows written by role G are visible
            int code = 0;
            if (role == 1 && act && ctl_phase) {
                if (phase == PH_K1) {
                    h = carry_h ? *hc : 0.0;
                    if (h == 0.0) {                              // HINIT, first half
                        const HinitGuess2 g = hinit_first_all(c.base, HMAX);
                        h = g.h;
                        s.dnf = g.dnf;
                        code = ACT_PROBE;
                    } else {
                        code = begin_step(s, h);
                    }
                } else if (phase == PH_HINIT) {                  // HINIT, second half
                    h = hinit_second_all(c.base, h, s.dnf, HMAX);
                    code = ACT_K1SMEM | begin_step(s, h);
                } else {                                         // error norm in state order: x0..x4, x5..x11, x12
                    double err = 0.0;
#pragma unroll
                    for (int i = 0; i < 5; i++) err = err + c(R_PE + i);
#pragma unroll
                    for (int i = 0; i < NS; i++) err = err + sq[i];
                    err = err + c(R_PE + 5);
                    code = control_step(s, err, h, c(H_L4));
                }
                c.set(M_H, h);
                c.set(M_ACT, (double)code);
            }
            __syncthreads();                                     // the decision is visible
            if (act && ctl_phase) {
                if (role == 0) { code = (int)c(M_ACT); h = c(M_H); }
                if (code & ACT_COMMIT) {
#pragma unroll
                    for (int i = 0; i < NS; i++) v.set(R_Y + i, yt[i]);
                }
                if (code & ACT_K1SMEM) {
#pragma unroll
                    for (int i = 0; i < NS; i++) kt[i] = v(R_K1 + i);
                }
                if (code & ACT_PROBE) {
#pragma unroll
                    for (int i = 0; i < NS; i++) yt[i] = v(R_Y + i) + h * kt[i];
                    phase = PH_HINIT;
                }
                if (code & ACT_BEGIN) {                          // kt = k1 of the step that starts here
                    const double ha = h * A21;
                    if (!(code & ACT_K1SMEM)) {
#pragma unroll
                        for (int i = 0; i < NS; i++) v.set(R_K1 + i, kt[i]);
                    }
#pragma unroll
                    for (int i = 0; i < NS; i++) yt[i] = v(R_Y + i) + ha * kt[i];
                    phase = PH_S2;
                }
                if (code & ACT_STOP) idid = (role == 1) ? s.idid : 1;
            }
        }

        par ^= 1;
        if (act && role == 1) { c.set(M_X6A + 2 * par, yt[1]); c.set(M_X8A + 2 * par, yt[3]); }
        __syncthreads();
    }
    if (role == 1) {
        if (carry_h && valid) *hc = h;
        cnt.nstep += s.nstep;
        cnt.nrej += s.nrej;
    }
    return idid;
}

// The episode loop of both roles (SimObj.simulate / T1DSimEnv.step, env.py:48-117).  role 0: meal state machine
// (t1dpatient.py:82-107) and the glucose side of the patient; role 1: sensor, controller, pump, risk, statistics
// and the insulin side.
__device__ __forceinline__ void episode_pair(const EpisodeArgs &a, const int64_t e, const bool valid, const int role,
                                             const Col &c)
{
    const Col v{c.base + role * NS * PL};
    double *sd = a.sd + e;
    int32_t *si = a.si + e;
    Meal meal{0.0, 0.0, 0.0, 0.0, 0};
    int minute = 0, meal_pos = 0, next_mm = -1, noise_pos = 0, status = 1;
    if (valid) {
        minute = si[(int64_t)SI_MINUTE * a.cap];
        if (role == 0) {
#pragma unroll
            for (int i = 0; i < 5; i++) v.set(R_Y + i, sd[(int64_t)(SD_X0 + i) * a.cap]);
            v.set(R_Y + 5, sd[(int64_t)(SD_X0 + 12) * a.cap]);
            v.set(R_Y + 6, 0.0);
            meal.planned = sd[(int64_t)SD_PLANNED * a.cap];
            meal.last_cho = sd[(int64_t)SD_LAST_CHO * a.cap];
            meal.last_Qsto = sd[(int64_t)SD_LAST_QSTO * a.cap];
            meal.last_food = sd[(int64_t)SD_LAST_FOOD * a.cap];
            meal.eating = si[(int64_t)SI_EATING * a.cap];
            meal_pos = si[(int64_t)SI_MEAL_POS * a.cap];
            next_mm = (meal_pos < a.max_meals) ? a.meal_minute[(int64_t)meal_pos * a.n + e] : -1;
        } else {
#pragma unroll
            for (int i = 0; i < NS; i++) v.set(R_Y + i, sd[(int64_t)(SD_X0 + 5 + i) * a.cap]);
            c.set(M_LAST_CGM, sd[(int64_t)SD_LAST_CGM * a.cap]);
            noise_pos = si[(int64_t)SI_NOISE_POS * a.cap];
            status = si[(int64_t)SI_STATUS * a.cap];
        }
    }
    const int st_i = a.hw.st;
    const double st = (double)st_i;
    int par = 0, cho_buf = 0, step_nstep = 0, step_nrej = 0;

    for (int k = 0; k < a.n_steps; k++) {
        double u = 0.0, CHO = 0.0, BG = 0.0, CGM = 0.0;
        bool fed = false;
        if (role == 1 && valid) {
            // ---- controller.policy(observation, reward, done, **info) ----
            double a_basal = 0.0, a_bolus = 0.0;
            double ctrl[NCTRL];
#pragma unroll
            for (int i = 0; i < NCTRL; i++) ctrl[i] = a.ctrl_rows ? __ldg(a.ctrl_rows + (int64_t)i * a.n + e) : a.ctrl[i];
            const double obs = sd[(int64_t)SD_OBS * a.cap];
            if (a.controller == 1) {                       // BBController._bb_policy
                const ParamLoad p = param_source(a, e);
                const double info_meal = sd[(int64_t)SD_INFO_MEAL * a.cap];
                a_basal = __ldg(&p.base[P_u2ss * p.stride]) * c(H_BW) / 6000.0;
                if (info_meal > 0.0) {
                    const double corr = (obs > 150.0 ? 1.0 : 0.0) * (obs - ctrl[0]) / ctrl[2];
                    a_bolus = (info_meal * st) / ctrl[1] + corr;
                } else {
                    a_bolus = 0.0;
                }
                a_bolus = a_bolus / st;
            } else if (a.controller == 2) {                // PIDController.policy
                const double pid_integ = sd[(int64_t)SD_PID_INTEG * a.cap];
                const double pid_prev = sd[(int64_t)SD_PID_PREV * a.cap];
                a_basal = ctrl[0] * (obs - ctrl[3]) + ctrl[1] * pid_integ + ctrl[2] * (obs - pid_prev) / st;
                sd[(int64_t)SD_PID_PREV * a.cap] = obs;
                sd[(int64_t)SD_PID_INTEG * a.cap] = pid_integ + (obs - ctrl[3]) * st;
                a_bolus = 0.0;
            } else {
                a_basal = a.act_basal[(int64_t)k * a.act_stride + e];
                a_bolus = a.act_bolus[(int64_t)k * a.act_stride + e];
            }
            // ---- env.step, env.py:66-117 ----
            const double ins = pump_quantise(a_basal, a.hw.pump[0], a.hw.pump[1], a.hw.pump[2]) +
                               pump_quantise(a_bolus, a.hw.pump[3], a.hw.pump[4], a.hw.pump[5]);
            u = ins * 6000.0 / c(H_BW);                     // t1dpatient.py:122
            // the pump command is constant over the env-step: its mean (env.py:78) does not wait for the minutes
            double insulin = 0.0;
            for (int j = 0; j < st_i; j++) insulin += ins / st;
            if (a.trace) {
                double *r = a.trace + (int64_t)k * NTRACE * a.n + e;
                r[3 * a.n] = insulin; r[10 * a.n] = a_basal; r[11 * a.n] = a_bolus;
            }
        }
        for (int j = 0; j < st_i; j++) {                   // mini_step, env.py:48-64
            if (role == 0 && valid) {
                double announced = 0.0;
                if (a.step_meals) {                        // scenario.get_action evaluated by the caller
                    announced = a.step_meals[(int64_t)j * a.n + e];
                } else {
                    while (next_mm >= 0 && next_mm < minute) {
                        meal_pos++;
                        next_mm = (meal_pos < a.max_meals) ? a.meal_minute[(int64_t)meal_pos * a.n + e] : -1;
                    }
                    if (next_mm == minute) announced = a.meal_grams[(int64_t)meal_pos * a.n + e];
                }
                CHO += announced / st;
                c.set(M_CHOA + cho_buf, CHO);
                const double to_eat = meal_machine(meal, announced, v(R_Y + 0), v(R_Y + 1));
                u = to_eat * 1000.0;                                         // t1dpatient.py:121
                const double Dbar = meal.last_Qsto + meal.last_food * 1000.0;
                fed = Dbar > 0.0;
                if (fed) {                                                   // t1dpatient.py:136-140
                    c.set(H_aa, 5.0 / 2.0 / (1.0 - c(H_b)) / Dbar);
                    c.set(H_cc, 5.0 / 2.0 / c(H_d) / Dbar);
                    c.set(H_bD, c(H_b) * Dbar);
                    c.set(H_dD, c(H_d) * Dbar);
                }
            }
            StepCount cnt{0, 0};
            const int idid = minute_pair(c, v, role, valid, u, fed, (double)minute, a.carry_h != 0,
                                         sd + (int64_t)SD_H_CARRY * a.cap, cnt, par);
            minute++;
            cho_buf ^= 1;                                      // role I may still be reading the other buffer
            if (role == 1 && valid) {
                if (idid < 0 && status > 0) status = idid;
                if (a.counters) {
                    const int64_t row = (int64_t)(k * st_i + j) * 2;
                    a.counters[row * a.n + e] = cnt.nstep;
                    a.counters[(row + 1) * a.n + e] = cnt.nrej;
                }
                step_nstep += cnt.nstep;
                step_nrej += cnt.nrej;
                const double bg = c(R_Y + 5) / c(H_Vg);                         // x12 / Vg, t1dpatient.py:217-218
                double last_cgm = c(M_LAST_CGM);
                if (minute % st_i == 0) {                                       // cgm.py:27
                    last_cgm = cgm_sample(a, e, bg, noise_pos);
                    c.set(M_LAST_CGM, last_cgm);
                }
                BG += bg / st; CGM += last_cgm / st;
            }
        }
        if (role == 1 && valid) {
            CHO = c(M_CHOA + (cho_buf ^ 1));                                 // published by role G in the last minute
            const double cgm_prev = sd[(int64_t)SD_CGM_PREV * a.cap];
            double lbgi, hbgi, ri, r_prev, r_now, t0, t1;
            risk_index(BG, lbgi, hbgi, ri);                                  // env.py:85
            risk_index(cgm_prev, t0, t1, r_prev);                            // risk_diff, env.py:27-33
            risk_index(CGM, t0, t1, r_now);
            const double reward = r_prev - r_now;
            const bool done = (BG < 70.0) || (BG > 350.0);                   // env.py:103
            sd[(int64_t)SD_CGM_PREV * a.cap] = CGM;
            sd[(int64_t)SD_OBS * a.cap] = CGM;
            sd[(int64_t)SD_INFO_MEAL * a.cap] = CHO;
            stats_add_row_resident(a.stats + e, a.cap, BG, lbgi, hbgi, ri, reward, done, step_nstep, step_nrej);
            step_nstep = 0; step_nrej = 0;
            if (a.trace) {
                double *r = a.trace + (int64_t)k * NTRACE * a.n + e;
                r[0 * a.n] = BG; r[1 * a.n] = CGM; r[2 * a.n] = CHO;
                r[4 * a.n] = lbgi; r[5 * a.n] = hbgi; r[6 * a.n] = ri; r[7 * a.n] = CGM;
                r[8 * a.n] = reward; r[9 * a.n] = done ? 1.0 : 0.0;
            }
        }
    }
    if (valid) {
        if (role == 0) {
#pragma unroll
            for (int i = 0; i < 5; i++) sd[(int64_t)(SD_X0 + i) * a.cap] = v(R_Y + i);
            sd[(int64_t)(SD_X0 + 12) * a.cap] = v(R_Y + 5);
            sd[(int64_t)SD_PLANNED * a.cap] = meal.planned;
            sd[(int64_t)SD_LAST_CHO * a.cap] = meal.last_cho;
            sd[(int64_t)SD_LAST_QSTO * a.cap] = meal.last_Qsto;
            sd[(int64_t)SD_LAST_FOOD * a.cap] = meal.last_food;
            si[(int64_t)SI_EATING * a.cap] = meal.eating;
            si[(int64_t)SI_MEAL_POS * a.cap] = meal_pos;
        } else {
#pragma unroll
            for (int i = 0; i < NS; i++) sd[(int64_t)(SD_X0 + 5 + i) * a.cap] = v(R_Y + i);
            sd[(int64_t)SD_LAST_CGM * a.cap] = c(M_LAST_CGM);
            si[(int64_t)SI_MINUTE * a.cap] = minute;
            si[(int64_t)SI_NOISE_POS * a.cap] = noise_pos;
            si[(int64_t)SI_STATUS * a.cap] = status;
        }
    }
}

}  // namespace pair

__global__ void __launch_bounds__(pair::PBLOCK, 8) run_pair_kernel(const EpisodeArgs a)
{
    extern __shared__ double smem[];   // PAIR_ROWS rows x 32 columns
    const int lane = threadIdx.x & 31, role = threadIdx.x >> 5;
    const int64_t e = (int64_t)blockIdx.x * pair::PL + lane;
    const bool valid = e < a.n;
    const pair::Col c{smem + lane};
    if (role == 0 && valid) stage_params(c, param_source(a, e));
    __syncthreads();
    pair::episode_pair(a, e, valid, role, c);
}

cudaError_t launch_run_pair(const EpisodeArgs &a, cudaStream_t s)
{
    constexpr int kSmem = pair::PAIR_ROWS * pair::PL * (int)sizeof(double);
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(run_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(run_pair_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    run_pair_kernel<<<(unsigned)((a.n + pair::PL - 1) / pair::PL), pair::PBLOCK, kSmem, s>>>(a);
    return cudaGetLastError();
}
