// Optional FP32 mode of the episode kernel (SG_FLAG_FP32; included by sg_kernels.cu inside namespace SG_NS).
//
// BASELINE.json asks for FP64 trajectories within 1e-6 of scipy and names an "optional FP32 mode" with a looser,
// stated bound: this is it.  The patient ODE (T1DPatient.model, t1dpatient.py:118-208) and its DOPRI5 integration
// run in single precision -- 128 FP32 lanes per SM against 64 FP64 lanes, FMA contraction, the seven stages held
// in registers -- while everything around the integrator that is integer / index work or feeds it (meal state
// machine, pump quantisation, sensor sampling, risk index, controllers, statistics) stays the FP64 code of
// run_kernel.  It is meant for reinforcement-learning populations (configs[2]) where a relative BG error of
// 1e-3 is far below sensor noise (Dexcom noise is several mg/dL); the bound is checked in tests/test_gpu_fp32.py
// (measured: 1.2e-4 over a day, 2.4e-4 with step-size carry; 2.64e9 patient-min/s on the bench workload, s37).
//
// What changes against the FP64 path, besides the type:
//   * time inside integrate(t, t + 1) is local (0..1): the model is autonomous (t is unused, t1dpatient.py:119) and
//     a float cannot resolve t + h for t in the thousands of minutes;
//   * uround = 1.2e-7 (float epsilon) in DOPCOR's "step size too small" test, and rtol = 1e-4 instead of scipy's
//     1e-6 (atol 1e-12 kept): at 1e-6 the embedded error estimate is float round-off -- measured 24.6 DOPRI5 steps
//     per minute against 3.0 in FP64, with minutes that exhaust nsteps = 500 (session s33);
//   * operation order is left to the compiler.

namespace fp32 {

constexpr int FBLOCK = 128;
constexpr float F_UROUND = 1.2e-7f;
#ifndef SG_FP32_RTOL
#define SG_FP32_RTOL 1e-4f
#endif
constexpr float F_RTOL = SG_FP32_RTOL;

enum FRow {
    F_kmax = 0, F_kmin, F_hk, F_b, F_d, F_kabs, F_fk, F_rBW, F_kp1, F_kp2, F_kp3, F_Fsnc, F_ke1, F_ke2, F_k1, F_k2,
    F_Vm0, F_Vmx, F_Km0, F_m24, F_m1, F_ka1, F_ka2, F_rVi, F_p2u, F_Ib, F_ki, F_m13, F_m2, F_kad, F_kd, F_ksc,
    F_COUNT
};

struct FCol {
    float *base;
    __device__ __forceinline__ float operator()(int i) const { return ((const volatile float *)base)[i * FBLOCK]; }
    __device__ __forceinline__ void set(int i, float v) const { base[i * FBLOCK] = v; }
};

struct FMinute {
    float d, ins, aa, cc, bD, dD;
    bool fed;
};

__device__ __forceinline__ float fmask(float x, float dx) { return x >= 0.0f ? dx : 0.0f; }

// T1DPatient.model in single precision.
__device__ __forceinline__ void rhs_f(const FCol &c, const FMinute &m, const float (&x)[NX], float (&dx)[NX])
{
    const float kmax = c(F_kmax);
    const float qsto = x[0] + x[1];
    dx[0] = -kmax * x[0] + m.d;
    float kgut = kmax;
    if (m.fed) {
        // tanh(u) - tanh(v) = 2 (e^2u - e^2v) / ((e^2u + 1)(e^2v + 1)) with hardware exp2 (2 ulp): ten instructions
        // instead of two branchy tanhf; arguments clamped where tanh has saturated in single precision
        const float u = fminf(fmaxf(2.0f * m.aa * (qsto - m.bD), -30.0f), 30.0f);
        const float v = fminf(fmaxf(2.0f * m.cc * (qsto - m.dD), -30.0f), 30.0f);
        const float eu = __expf(u), ev = __expf(v);
        const float t = __fdividef(2.0f * (eu - ev), (eu + 1.0f) * (ev + 1.0f));
        kgut = c(F_kmin) + c(F_hk) * (t + 2.0f);
    }
    dx[1] = kmax * x[0] - x[1] * kgut;
    dx[2] = kgut * x[1] - c(F_kabs) * x[2];
    const float Rat = c(F_fk) * x[2] * c(F_rBW);
    const float EGPt = c(F_kp1) - c(F_kp2) * x[3] - c(F_kp3) * x[8];
    const float ke2 = c(F_ke2);
    const float Et = (x[3] > ke2) ? c(F_ke1) * (x[3] - ke2) : 0.0f;
    const float k1 = c(F_k1), k2 = c(F_k2);
    dx[3] = fmask(x[3], fmaxf(EGPt, 0.0f) + Rat - c(F_Fsnc) - Et - k1 * x[3] + k2 * x[4]);
    const float Vmt = c(F_Vm0) + c(F_Vmx) * x[6];
    const float Uidt = __fdividef(Vmt * x[4], c(F_Km0) + x[4]);
    dx[4] = fmask(x[4], -Uidt + k1 * x[3] - k2 * x[4]);
    const float ka2 = c(F_ka2);
    dx[5] = fmask(x[5], -c(F_m24) * x[5] + c(F_m1) * x[9] + c(F_ka1) * x[10] + ka2 * x[11]);
    const float It = x[5] * c(F_rVi);
    const float p2u = c(F_p2u), ki = c(F_ki);
    dx[6] = -p2u * x[6] + p2u * (It - c(F_Ib));
    dx[7] = -ki * (x[7] - It);
    dx[8] = -ki * (x[8] - x[7]);
    dx[9] = fmask(x[9], -c(F_m13) * x[9] + c(F_m2) * x[5]);
    dx[10] = fmask(x[10], m.ins - c(F_kad) * x[10]);
    dx[11] = fmask(x[11], c(F_kd) * x[10] - ka2 * x[11]);
    const float ksc = c(F_ksc);
    dx[12] = fmask(x[12], -ksc * x[12] + ksc * x[3]);
}

template <class Load>
__device__ __forceinline__ void stage_params_f(const FCol &c, Load p)
{
    const double kmax = p(P_kmax), kmin = p(P_kmin), kabs = p(P_kabs);
    c.set(F_kmax, (float)kmax); c.set(F_kmin, (float)kmin); c.set(F_hk, (float)((kmax - kmin) / 2.0));
    c.set(F_b, (float)p(P_b)); c.set(F_d, (float)p(P_d)); c.set(F_kabs, (float)kabs);
    c.set(F_fk, (float)(p(P_f) * kabs)); c.set(F_rBW, (float)(1.0 / p(P_BW)));
    c.set(F_kp1, (float)p(P_kp1)); c.set(F_kp2, (float)p(P_kp2)); c.set(F_kp3, (float)p(P_kp3));
    c.set(F_Fsnc, (float)p(P_Fsnc)); c.set(F_ke1, (float)p(P_ke1)); c.set(F_ke2, (float)p(P_ke2));
    c.set(F_k1, (float)p(P_k1)); c.set(F_k2, (float)p(P_k2)); c.set(F_Vm0, (float)p(P_Vm0));
    c.set(F_Vmx, (float)p(P_Vmx)); c.set(F_Km0, (float)p(P_Km0)); c.set(F_m24, (float)(p(P_m2) + p(P_m4)));
    c.set(F_m1, (float)p(P_m1)); c.set(F_ka1, (float)p(P_ka1)); c.set(F_ka2, (float)p(P_ka2));
    c.set(F_rVi, (float)(1.0 / p(P_Vi))); c.set(F_p2u, (float)p(P_p2u)); c.set(F_Ib, (float)p(P_Ib));
    c.set(F_ki, (float)p(P_ki)); c.set(F_m13, (float)(p(P_m1) + p(P_m30))); c.set(F_m2, (float)p(P_m2));
    c.set(F_kad, (float)(p(P_ka1) + p(P_kd))); c.set(F_kd, (float)p(P_kd)); c.set(F_ksc, (float)p(P_ksc));
}

// One call of ode('dopri5').integrate(t + 1) in single precision: k1, HINIT (unless a step size is carried over),
// then the adaptive step loop of DOPCOR with all seven stages in registers.  Lanes that finish their minute wait
// for the warp (the loop condition is a vote), as in the FP64 kernel.  Returns scipy's idid.
__device__ __forceinline__ int dopri5_minute_f(const FCol &c, const FMinute &m, float (&y)[NX], bool carry_h,
                                               float &h_carry, StepCount &cnt, unsigned lanes)
{
    const float rtol = F_RTOL, atol = 1e-12f, hmax = 1.0f, xend = 1.0f;
    const float a21 = 0.2f, a31 = 3.0f / 40.0f, a32 = 9.0f / 40.0f, a41 = 44.0f / 45.0f, a42 = -56.0f / 15.0f,
                a43 = 32.0f / 9.0f, a51 = 19372.0f / 6561.0f, a52 = -25360.0f / 2187.0f, a53 = 64448.0f / 6561.0f,
                a54 = -212.0f / 729.0f, a61 = 9017.0f / 3168.0f, a62 = -355.0f / 33.0f, a63 = 46732.0f / 5247.0f,
                a64 = 49.0f / 176.0f, a65 = -5103.0f / 18656.0f, a71 = 35.0f / 384.0f, a73 = 500.0f / 1113.0f,
                a74 = 125.0f / 192.0f, a75 = -2187.0f / 6784.0f, a76 = 11.0f / 84.0f, e1 = 71.0f / 57600.0f,
                e3 = -71.0f / 16695.0f, e4 = 71.0f / 1920.0f, e5 = -17253.0f / 339200.0f, e6 = 22.0f / 525.0f,
                e7 = -1.0f / 40.0f;
    const float beta = 0.04f, safe = 0.9f, expo1 = 0.2f - 0.04f * 0.75f, facc1 = 5.0f, facc2 = 0.1f;
    float k1[NX], k2[NX], k3[NX], k4[NX], k5[NX], k6[NX], yt[NX];
    float x = 0.0f, facold = 1.0e-4f;
    bool last = false, reject = false;
    int nstep = 0, naccpt = 0, idid = 0;

    rhs_f(c, m, y, k1);
    float h = carry_h ? h_carry : 0.0f;
    if (h == 0.0f) {                                          // HINIT
        float dnf = 0.0f, dny = 0.0f;
#pragma unroll
        for (int i = 0; i < NX; i++) {
            const float rs = __frcp_rn(atol + rtol * fabsf(y[i]));
            const float a = k1[i] * rs, b = y[i] * rs;
            dnf += a * a; dny += b * b;
        }
        h = (dnf <= 1e-10f || dny <= 1e-10f) ? 1.0e-6f : sqrtf(dny / dnf) * 0.01f;
        h = fminf(h, hmax);
#pragma unroll
        for (int i = 0; i < NX; i++) yt[i] = y[i] + h * k1[i];
        rhs_f(c, m, yt, k2);
        float der2 = 0.0f;
#pragma unroll
        for (int i = 0; i < NX; i++) {
            const float a = (k2[i] - k1[i]) * __frcp_rn(atol + rtol * fabsf(y[i]));
            der2 += a * a;
        }
        der2 = sqrtf(der2) / h;
        const float der12 = fmaxf(fabsf(der2), sqrtf(dnf));
        const float h1 = (der12 <= 1e-15f) ? fmaxf(1.0e-6f, fabsf(h) * 1.0e-3f) : powf(0.01f / der12, 0.2f);
        h = fminf(fminf(100.0f * fabsf(h), h1), hmax);
    }

#pragma unroll 1
    while (__any_sync(lanes, idid == 0)) {
        if (idid != 0) continue;
        if (nstep > NMAX) { idid = -2; continue; }
        if (0.1f * fabsf(h) <= fabsf(x) * F_UROUND) { idid = -3; continue; }
        if ((x + 1.01f * h - xend) > 0.0f) { h = xend - x; last = true; }
        nstep++;
#pragma unroll
        for (int i = 0; i < NX; i++) yt[i] = y[i] + h * a21 * k1[i];
        rhs_f(c, m, yt, k2);
#pragma unroll
        for (int i = 0; i < NX; i++) yt[i] = y[i] + h * (a31 * k1[i] + a32 * k2[i]);
        rhs_f(c, m, yt, k3);
#pragma unroll
        for (int i = 0; i < NX; i++) yt[i] = y[i] + h * (a41 * k1[i] + a42 * k2[i] + a43 * k3[i]);
        rhs_f(c, m, yt, k4);
#pragma unroll
        for (int i = 0; i < NX; i++) yt[i] = y[i] + h * (a51 * k1[i] + a52 * k2[i] + a53 * k3[i] + a54 * k4[i]);
        rhs_f(c, m, yt, k5);
#pragma unroll
        for (int i = 0; i < NX; i++)
            yt[i] = y[i] + h * (a61 * k1[i] + a62 * k2[i] + a63 * k3[i] + a64 * k4[i] + a65 * k5[i]);
        rhs_f(c, m, yt, k6);
#pragma unroll
        for (int i = 0; i < NX; i++)
            yt[i] = y[i] + h * (a71 * k1[i] + a73 * k3[i] + a74 * k4[i] + a75 * k5[i] + a76 * k6[i]);
        rhs_f(c, m, yt, k2);                                  // k7 (FSAL)
        float err = 0.0f;
#pragma unroll
        for (int i = 0; i < NX; i++) {
            const float e = h * (e1 * k1[i] + e3 * k3[i] + e4 * k4[i] + e5 * k5[i] + e6 * k6[i] + e7 * k2[i]);
            const float a = __fdividef(e, atol + rtol * fmaxf(fabsf(y[i]), fabsf(yt[i])));   // MUFU.RCP: 1 ulp is plenty here
            err += a * a;
        }
        err = sqrtf(err * (1.0f / NX));
        const float fac11 = __powf(fmaxf(err, 1e-30f), expo1);      // step-size control needs a few digits only
        float fac = __fdividef(fac11, __powf(facold, beta));
        fac = fmaxf(facc2, fminf(facc1, fac / safe));
        float hnew = h / fac;
        if (err <= 1.0f) {
            facold = fmaxf(err, 1.0e-4f);
            naccpt++;
#pragma unroll
            for (int i = 0; i < NX; i++) { k1[i] = k2[i]; y[i] = yt[i]; }
            x += h;
            if (last) {
                idid = 1;
            } else {
                if (fabsf(hnew) > hmax) hnew = hmax;
                if (reject) hnew = fminf(fabsf(hnew), fabsf(h));
                reject = false;
            }
        } else {
            hnew = h / fminf(facc1, fac11 / safe);
            reject = true;
            if (naccpt >= 1) cnt.nrej++;
            last = false;
        }
        h = hnew;
    }
    if (carry_h) h_carry = h;
    cnt.nstep += nstep;
    return idid;
}

}  // namespace fp32

// SimObj.simulate / T1DSimEnv.step with the patient integrated in FP32; the closed loop around it is the FP64 code
// of run_kernel (same resident state, same trace / statistics layout, same status codes).
__global__ void __launch_bounds__(fp32::FBLOCK, 3) run_fp32_kernel(const EpisodeArgs a)
{
    using namespace fp32;
    extern __shared__ float fsmem[];   // F_COUNT rows x FBLOCK columns
    const int64_t idx = (int64_t)blockIdx.x * FBLOCK + threadIdx.x;
    const unsigned lanes = __ballot_sync(0xffffffffu, idx < a.n);
    if (idx >= a.n) return;
    const int64_t e = a.perm ? (int64_t)a.perm[idx] : idx;         // schedule-coherent order (sg_set_coherence)
    const FCol c{fsmem + threadIdx.x};
    const ParamLoad p = param_source(a, e);
    stage_params_f(c, p);
    const double BW = p(P_BW), Vg = p(P_Vg), bpar = p(P_b), dpar = p(P_d);

    double *sd = a.sd + e;
    int32_t *si = a.si + e;
    float x[NX];
#pragma unroll
    for (int i = 0; i < NX; i++) x[i] = (float)sd[(int64_t)(SD_X0 + i) * a.cap];
    Meal meal;
    meal.planned = sd[(int64_t)SD_PLANNED * a.cap];
    meal.last_cho = sd[(int64_t)SD_LAST_CHO * a.cap];
    meal.last_Qsto = sd[(int64_t)SD_LAST_QSTO * a.cap];
    meal.last_food = sd[(int64_t)SD_LAST_FOOD * a.cap];
    meal.eating = si[(int64_t)SI_EATING * a.cap];
    float h_carry = (float)sd[(int64_t)SD_H_CARRY * a.cap];
    double last_cgm = sd[(int64_t)SD_LAST_CGM * a.cap];
    int minute = si[(int64_t)SI_MINUTE * a.cap];
    int noise_pos = si[(int64_t)SI_NOISE_POS * a.cap];
    int meal_pos = si[(int64_t)SI_MEAL_POS * a.cap];
    int status = si[(int64_t)SI_STATUS * a.cap];
    const int st_i = a.hw.st;
    const double st = (double)st_i;
    const double r_st = sgm::rcp_nr(st), r_Vg = sgm::rcp_nr(Vg);       // exact quotients through div_r, as in run_kernel
    int next_mm = (meal_pos < a.max_meals) ? a.meal_minute[(int64_t)meal_pos * a.n + e] : -1;
    double Dbar_seen = 0.0, r_prev = 0.0;
    bool have_rprev = false;
    FMinute m;
    m.d = m.ins = m.aa = m.cc = m.bD = m.dD = 0.0f;
    m.fed = false;

    for (int k = 0; k < a.n_steps; k++) {
        // ---- controller.policy(observation, reward, done, **info) ----
        double a_basal = 0.0, a_bolus = 0.0;
        double ctrl[NCTRL];
#pragma unroll
        for (int i = 0; i < NCTRL; i++) ctrl[i] = a.ctrl_rows ? __ldg(a.ctrl_rows + (int64_t)i * a.n + e) : a.ctrl[i];
        const double obs = sd[(int64_t)SD_OBS * a.cap];
        if (a.controller == 1) {                       // BBController._bb_policy
            const double info_meal = sd[(int64_t)SD_INFO_MEAL * a.cap];
            a_basal = (ctrl[3] > 0.0) ? ctrl[3] : p(P_u2ss) * BW / 6000.0;   // see run_kernel
            if (info_meal > 0.0) {
                const double corr = (obs > 150.0 ? 1.0 : 0.0) * (obs - ctrl[0]) / ctrl[2];
                a_bolus = (info_meal * st) / ctrl[1] + corr;
            }
            a_bolus = a_bolus / st;
        } else if (a.controller == 2) {                // PIDController.policy
            const double pid_integ = sd[(int64_t)SD_PID_INTEG * a.cap];
            const double pid_prev = sd[(int64_t)SD_PID_PREV * a.cap];
            a_basal = ctrl[0] * (obs - ctrl[3]) + ctrl[1] * pid_integ + ctrl[2] * (obs - pid_prev) / st;
            sd[(int64_t)SD_PID_PREV * a.cap] = obs;
            sd[(int64_t)SD_PID_INTEG * a.cap] = pid_integ + (obs - ctrl[3]) * st;
        } else if (a.controller == 3) {                // random basal ~ U(ctrl[0], ctrl[1]), see run_kernel
            sgp::Philox rng((uint64_t)ctrl[2], (uint64_t)ctrl[3] + (uint64_t)e, 0x6163746eu, (uint32_t)(a.step0 + k));
            a_basal = ctrl[0] + (ctrl[1] - ctrl[0]) * rng.uniform();
        } else {
            a_basal = a.act_basal[(int64_t)k * a.act_stride + e];
            a_bolus = a.act_bolus[(int64_t)k * a.act_stride + e];
        }
        // ---- env.step, env.py:66-117 ----
        const double ins = pump_quantise(a_basal, a.hw.pump[0], a.hw.pump[1], a.hw.pump[2]) +
                           pump_quantise(a_bolus, a.hw.pump[3], a.hw.pump[4], a.hw.pump[5]);
        double CHO = 0.0, insulin = 0.0, BG = 0.0, CGM = 0.0;
        int step_nstep = 0, step_nrej = 0;
        m.ins = (float)(ins * 6000.0 / BW);                                  // t1dpatient.py:122
        const double ins_mean = sgm::div_r(ins, st, r_st);
        for (int j = 0; j < st_i; j++) {               // mini_step, env.py:48-64
            double announced = 0.0;
            if (a.step_meals) {
                announced = a.step_meals[(int64_t)j * a.n + e];
            } else {
                while (next_mm >= 0 && next_mm < minute) {
                    meal_pos++;
                    next_mm = (meal_pos < a.max_meals) ? a.meal_minute[(int64_t)meal_pos * a.n + e] : -1;
                }
                if (next_mm == minute) announced = a.meal_grams[(int64_t)meal_pos * a.n + e];
            }
            const double to_eat = meal_machine(meal, announced, (double)x[0], (double)x[1]);
            m.d = (float)(to_eat * 1000.0);                                  // t1dpatient.py:121
            const double Dbar = meal.last_Qsto + meal.last_food * 1000.0;    // t1dpatient.py:130
            m.fed = Dbar > 0.0;
            if (m.fed && Dbar != Dbar_seen) {                                // t1dpatient.py:136-140; Dbar only moves
                Dbar_seen = Dbar;                                            // while the patient eats
                m.aa = (float)(5.0 / 2.0 / (1.0 - bpar) / Dbar);
                m.cc = (float)(5.0 / 2.0 / dpar / Dbar);
                m.bD = (float)(bpar * Dbar);
                m.dD = (float)(dpar * Dbar);
            }
            StepCount cnt{0, 0};
            const int idid = dopri5_minute_f(c, m, x, a.carry_h != 0, h_carry, cnt, lanes);
            if (idid < 0 && status > 0) status = idid;
            if (a.counters) {
                const int64_t row = (int64_t)(k * st_i + j) * 2;
                a.counters[row * a.n + e] = cnt.nstep;
                a.counters[(row + 1) * a.n + e] = cnt.nrej;
            }
            step_nstep += cnt.nstep;
            step_nrej += cnt.nrej;
            minute++;
            const double bg = sgm::div_r((double)x[12], Vg, r_Vg);            // t1dpatient.py:217-218
            if (minute % st_i == 0) last_cgm = cgm_sample(a, e, bg, noise_pos, status);   // cgm.py:27
            CHO += sgm::div_r(announced, st, r_st); insulin += ins_mean;
            BG += sgm::div_r(bg, st, r_st); CGM += sgm::div_r(last_cgm, st, r_st);
        }
        double lbgi, hbgi, ri, r_now, t0, t1;
        risk_index(BG, lbgi, hbgi, ri);                                  // env.py:85
        if (!have_rprev) risk_index(sd[(int64_t)SD_CGM_PREV * a.cap], t0, t1, r_prev);   // risk_diff, env.py:27-33
        risk_index(CGM, t0, t1, r_now);
        const double reward = r_prev - r_now;
        r_prev = r_now;
        have_rprev = true;
        const bool done = (BG < 70.0) || (BG > 350.0);                   // env.py:103
        sd[(int64_t)SD_CGM_PREV * a.cap] = CGM;
        sd[(int64_t)SD_OBS * a.cap] = CGM;
        sd[(int64_t)SD_INFO_MEAL * a.cap] = CHO;
        stats_add_row_resident(a.stats + e, a.cap, BG, lbgi, hbgi, ri, reward, done, step_nstep, step_nrej);
        if (a.trace) {
            double *r = a.trace + (int64_t)k * NTRACE * a.n + e;
            r[0 * a.n] = BG; r[1 * a.n] = CGM; r[2 * a.n] = CHO; r[3 * a.n] = insulin;
            r[4 * a.n] = lbgi; r[5 * a.n] = hbgi; r[6 * a.n] = ri; r[7 * a.n] = CGM;
            r[8 * a.n] = reward; r[9 * a.n] = done ? 1.0 : 0.0; r[10 * a.n] = a_basal; r[11 * a.n] = a_bolus;
        }
    }

#pragma unroll
    for (int i = 0; i < NX; i++) sd[(int64_t)(SD_X0 + i) * a.cap] = (double)x[i];
    sd[(int64_t)SD_PLANNED * a.cap] = meal.planned;
    sd[(int64_t)SD_LAST_CHO * a.cap] = meal.last_cho;
    sd[(int64_t)SD_LAST_QSTO * a.cap] = meal.last_Qsto;
    sd[(int64_t)SD_LAST_FOOD * a.cap] = meal.last_food;
    sd[(int64_t)SD_H_CARRY * a.cap] = (double)h_carry;
    sd[(int64_t)SD_LAST_CGM * a.cap] = last_cgm;
    si[(int64_t)SI_EATING * a.cap] = meal.eating;
    si[(int64_t)SI_MINUTE * a.cap] = minute;
    si[(int64_t)SI_NOISE_POS * a.cap] = noise_pos;
    si[(int64_t)SI_MEAL_POS * a.cap] = meal_pos;
    si[(int64_t)SI_STATUS * a.cap] = status;
}

cudaError_t launch_run_fp32(const EpisodeArgs &a, cudaStream_t s)
{
    constexpr int kSmem = fp32::F_COUNT * fp32::FBLOCK * (int)sizeof(float);
    run_fp32_kernel<<<(unsigned)((a.n + fp32::FBLOCK - 1) / fp32::FBLOCK), fp32::FBLOCK, kSmem, s>>>(a);
    return cudaGetLastError();
}
