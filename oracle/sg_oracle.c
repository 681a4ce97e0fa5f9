/* TEST INFRASTRUCTURE ONLY -- see sg_oracle.h.  Plain C, IEEE double, no FMA contraction
 * (build with -ffp-contract=off), libm tanh/pow/log/sqrt.
 *
 * Every function cites the reference lines it restates (paths relative to the reference
 * root).  Operation order follows the Python source literally: Python evaluates
 * a*b/c as (a*b)/c and a+b+c as (a+b)+c.
 */
#include "sg_oracle.h"

#include <float.h>
#include <math.h>
#include <stddef.h>
#include <string.h>
#include <pthread.h>
#include <stdatomic.h>
#include <unistd.h>

#define NX 13

/* ------------------------------------------------------------------------------------------
 * T1DPatient.model -- simglucose/patient/t1dpatient.py:118-208
 * ---------------------------------------------------------------------------------------- */
void sgo_model(const double *x, double cho, double insulin_u, const double *p,
               double last_Qsto, double last_foodtaken, double *dxdt)
{
    double d = cho * 1000.0;                              /* :121 */
    double insulin = insulin_u * 6000.0 / p[SGO_BW];      /* :122 */
    double qsto = x[0] + x[1];                            /* :126 */
    double Dbar = last_Qsto + last_foodtaken * 1000.0;    /* :130 */
    double kgut;

    dxdt[0] = -p[SGO_kmax] * x[0] + d;                    /* :133 */

    if (Dbar > 0) {                                       /* :135-142 */
        double aa = 5.0 / 2.0 / (1.0 - p[SGO_b]) / Dbar;
        double cc = 5.0 / 2.0 / p[SGO_d] / Dbar;
        kgut = p[SGO_kmin] + (p[SGO_kmax] - p[SGO_kmin]) / 2.0 *
               (tanh(aa * (qsto - p[SGO_b] * Dbar)) - tanh(cc * (qsto - p[SGO_d] * Dbar)) + 2.0);
    } else {
        kgut = p[SGO_kmax];
    }

    dxdt[1] = p[SGO_kmax] * x[0] - x[1] * kgut;           /* :145 */
    dxdt[2] = kgut * x[1] - p[SGO_kabs] * x[2];           /* :148 */

    double Rat = p[SGO_f] * p[SGO_kabs] * x[2] / p[SGO_BW];            /* :151 */
    double EGPt = p[SGO_kp1] - p[SGO_kp2] * x[3] - p[SGO_kp3] * x[8];  /* :153 */
    double Uiit = p[SGO_Fsnc];                                         /* :155 */
    double Et = (x[3] > p[SGO_ke2]) ? p[SGO_ke1] * (x[3] - p[SGO_ke2]) : 0.0;   /* :158-161 */
    double egp_pos = (0.0 > EGPt) ? 0.0 : EGPt;           /* Python max(EGPt, 0) :165 */

    dxdt[3] = egp_pos + Rat - Uiit - Et - p[SGO_k1] * x[3] + p[SGO_k2] * x[4];  /* :165-166 */
    dxdt[3] = (x[3] >= 0 ? 1.0 : 0.0) * dxdt[3];                                /* :167 */

    double Vmt = p[SGO_Vm0] + p[SGO_Vmx] * x[6];          /* :169 */
    double Kmt = p[SGO_Km0];
    double Uidt = Vmt * x[4] / (Kmt + x[4]);              /* :171 */
    dxdt[4] = -Uidt + p[SGO_k1] * x[3] - p[SGO_k2] * x[4];
    dxdt[4] = (x[4] >= 0 ? 1.0 : 0.0) * dxdt[4];

    dxdt[5] = -(p[SGO_m2] + p[SGO_m4]) * x[5] + p[SGO_m1] * x[9] + p[SGO_ka1] * x[10] +
              p[SGO_ka2] * x[11];                         /* :176-177 */
    double It = x[5] / p[SGO_Vi];                         /* :178 */
    dxdt[5] = (x[5] >= 0 ? 1.0 : 0.0) * dxdt[5];

    dxdt[6] = -p[SGO_p2u] * x[6] + p[SGO_p2u] * (It - p[SGO_Ib]);      /* :182 */
    dxdt[7] = -p[SGO_ki] * (x[7] - It);                   /* :185 */
    dxdt[8] = -p[SGO_ki] * (x[8] - x[7]);                 /* :187 */

    dxdt[9] = -(p[SGO_m1] + p[SGO_m30]) * x[9] + p[SGO_m2] * x[5];     /* :190 */
    dxdt[9] = (x[9] >= 0 ? 1.0 : 0.0) * dxdt[9];

    dxdt[10] = insulin - (p[SGO_ka1] + p[SGO_kd]) * x[10];             /* :194 */
    dxdt[10] = (x[10] >= 0 ? 1.0 : 0.0) * dxdt[10];

    dxdt[11] = p[SGO_kd] * x[10] - p[SGO_ka2] * x[11];    /* :197 */
    dxdt[11] = (x[11] >= 0 ? 1.0 : 0.0) * dxdt[11];

    dxdt[12] = (-p[SGO_ksc] * x[12] + p[SGO_ksc] * x[3]); /* :201 */
    dxdt[12] = (x[12] >= 0 ? 1.0 : 0.0) * dxdt[12];
}

/* ------------------------------------------------------------------------------------------
 * InsulinPump.basal / .bolus -- simglucose/actuator/pump.py:23-39.  np.round is
 * round-half-to-even == rint() in the default rounding mode.
 * ---------------------------------------------------------------------------------------- */
double sgo_pump(double amount, double vmin, double vmax, double inc)
{
    double v = amount * 6000.0;
    v = rint(v / inc) * inc;
    v = v / 6000.0;
    /* Python min(v, vmax): vmax if vmax < v else v ; max(v, vmin): vmin if vmin > v else v */
    v = (vmax < v) ? vmax : v;
    v = (vmin > v) ? vmin : v;
    return v;
}

/* ------------------------------------------------------------------------------------------
 * risk_index([BG], 1) -- simglucose/analysis/risk.py:5-17.  A one-element array: the mean of
 * an empty selection is NaN -> nan_to_num -> 0; +inf -> DBL_MAX.
 * ---------------------------------------------------------------------------------------- */
static double nan_to_num(double v)
{
    if (isnan(v)) return 0.0;
    if (isinf(v)) return v > 0 ? DBL_MAX : -DBL_MAX;
    return v;
}

void sgo_risk(double bg, double *out)
{
    double f = 1.509 * (pow(log(bg), 1.084) - 5.381);
    double lbgi = NAN, hbgi = NAN;
    if (f < 0) lbgi = 10.0 * (f * f);
    if (f > 0) hbgi = 10.0 * (f * f);
    out[0] = nan_to_num(lbgi);
    out[1] = nan_to_num(hbgi);
    out[2] = out[0] + out[1];
}

/* ------------------------------------------------------------------------------------------
 * scipy.integrate.ode('dopri5') one-minute call -- call sites t1dpatient.py:276-277 (defaults:
 * rtol 1e-6, atol 1e-12, nsteps 500, first_step 0, max_step 0, safety .9, ifactor 10,
 * dfactor .2, beta 0 -> 0.04) and :110-113.  Body: Hairer, Norsett & Wanner DOPRI5
 * (DOPCOR + HINIT), third-party, restated from the published algorithm.
 * ---------------------------------------------------------------------------------------- */
static const double A21 = 0.2, A31 = 3.0 / 40.0, A32 = 9.0 / 40.0, A41 = 44.0 / 45.0,
    A42 = -56.0 / 15.0, A43 = 32.0 / 9.0, A51 = 19372.0 / 6561.0, A52 = -25360.0 / 2187.0,
    A53 = 64448.0 / 6561.0, A54 = -212.0 / 729.0, A61 = 9017.0 / 3168.0, A62 = -355.0 / 33.0,
    A63 = 46732.0 / 5247.0, A64 = 49.0 / 176.0, A65 = -5103.0 / 18656.0, A71 = 35.0 / 384.0,
    A73 = 500.0 / 1113.0, A74 = 125.0 / 192.0, A75 = -2187.0 / 6784.0, A76 = 11.0 / 84.0,
    E1 = 71.0 / 57600.0, E3 = -71.0 / 16695.0, E4 = 71.0 / 1920.0, E5 = -17253.0 / 339200.0,
    E6 = 22.0 / 525.0, E7 = -1.0 / 40.0;

#define RTOL 1e-6
#define ATOL 1e-12
#define NMAX 500
#define UROUND 2.3e-16

#define F(yy, out) sgo_model((yy), cho, insulin, p, last_Qsto, last_foodtaken, (out))

static double hinit(const double *y, const double *f0, double hmax, const double *p, double cho,
                    double insulin, double last_Qsto, double last_foodtaken)
{
    double dnf = 0.0, dny = 0.0, y1[NX], f1[NX];
    for (int i = 0; i < NX; i++) {
        double sk = ATOL + RTOL * fabs(y[i]);
        double a = f0[i] / sk, b = y[i] / sk;
        dnf = dnf + a * a;
        dny = dny + b * b;
    }
    double h;
    if (dnf <= 1e-10 || dny <= 1e-10) h = 1.0e-6;
    else h = sqrt(dny / dnf) * 0.01;
    h = fmin(h, hmax);
    for (int i = 0; i < NX; i++) y1[i] = y[i] + h * f0[i];
    F(y1, f1);
    double der2 = 0.0;
    for (int i = 0; i < NX; i++) {
        double sk = ATOL + RTOL * fabs(y[i]);
        double a = (f1[i] - f0[i]) / sk;
        der2 = der2 + a * a;
    }
    der2 = sqrt(der2) / h;
    double der12 = fmax(fabs(der2), sqrt(dnf));
    double h1;
    if (der12 <= 1e-15) h1 = fmax(1.0e-6, fabs(h) * 1.0e-3);
    else h1 = pow(0.01 / der12, 1.0 / 5.0);
    h = fmin(fmin(100.0 * fabs(h), h1), hmax);
    return h;
}

int sgo_dopri5_minute(double *y, double *t, const double *p, double cho, double insulin,
                      double last_Qsto, double last_foodtaken, int carry_h, double *h_carry,
                      int *counters)
{
    const double beta = 0.04, safe = 0.9, fac1 = 0.2, fac2 = 10.0;
    double x = *t, xend = *t + 1.0;
    double hmax = xend - x;
    double facold = 1.0e-4;
    double expo1 = 0.2 - beta * 0.75;
    double facc1 = 1.0 / fac1, facc2 = 1.0 / fac2;
    int last = 0, reject = 0, nstep = 0, naccpt = 0, nrejct = 0, nfcn = 0;
    double k1[NX], k2[NX], k3[NX], k4[NX], k5[NX], k6[NX], y1[NX], ysti[NX];
    double h = (carry_h ? *h_carry : 0.0);
    int idid = 0;

    F(y, k1);
    hmax = fabs(hmax);
    if (h == 0.0) h = hinit(y, k1, hmax, p, cho, insulin, last_Qsto, last_foodtaken);
    nfcn += 2;

    for (;;) {
        if (nstep > NMAX) { idid = -2; break; }
        if (0.1 * fabs(h) <= fabs(x) * UROUND) { idid = -3; break; }
        if ((x + 1.01 * h - xend) > 0.0) { h = xend - x; last = 1; }
        nstep++;
        for (int i = 0; i < NX; i++) y1[i] = y[i] + h * A21 * k1[i];
        F(y1, k2);
        for (int i = 0; i < NX; i++) y1[i] = y[i] + h * (A31 * k1[i] + A32 * k2[i]);
        F(y1, k3);
        for (int i = 0; i < NX; i++) y1[i] = y[i] + h * (A41 * k1[i] + A42 * k2[i] + A43 * k3[i]);
        F(y1, k4);
        for (int i = 0; i < NX; i++)
            y1[i] = y[i] + h * (A51 * k1[i] + A52 * k2[i] + A53 * k3[i] + A54 * k4[i]);
        F(y1, k5);
        for (int i = 0; i < NX; i++)
            ysti[i] = y[i] + h * (A61 * k1[i] + A62 * k2[i] + A63 * k3[i] + A64 * k4[i] + A65 * k5[i]);
        double xph = x + h;
        F(ysti, k6);
        for (int i = 0; i < NX; i++)
            y1[i] = y[i] + h * (A71 * k1[i] + A73 * k3[i] + A74 * k4[i] + A75 * k5[i] + A76 * k6[i]);
        F(y1, k2);
        for (int i = 0; i < NX; i++)
            k4[i] = (E1 * k1[i] + E3 * k3[i] + E4 * k4[i] + E5 * k5[i] + E6 * k6[i] + E7 * k2[i]) * h;
        nfcn += 6;
        double err = 0.0;
        for (int i = 0; i < NX; i++) {
            double sk = ATOL + RTOL * fmax(fabs(y[i]), fabs(y1[i]));
            double a = k4[i] / sk;
            err = err + a * a;
        }
        err = sqrt(err / (double)NX);
        double fac11 = pow(err, expo1);
        double fac = fac11 / pow(facold, beta);
        fac = fmax(facc2, fmin(facc1, fac / safe));
        double hnew = h / fac;
        if (err <= 1.0) {
            facold = fmax(err, 1.0e-4);
            naccpt++;
            /* stiffness detection (every NSTIFF=1000 accepted steps) can never fire:
             * nmax=500 steps per call */
            for (int i = 0; i < NX; i++) { k1[i] = k2[i]; y[i] = y1[i]; }
            x = xph;
            if (last) { h = hnew; idid = 1; break; }
            if (fabs(hnew) > hmax) hnew = hmax;
            if (reject) hnew = fmin(fabs(hnew), fabs(h));
            reject = 0;
        } else {
            hnew = h / fmin(facc1, fac11 / safe);
            reject = 1;
            if (naccpt >= 1) nrejct++;
            last = 0;
        }
        h = hnew;
    }
    *t = x;
    if (carry_h) *h_carry = h;
    if (counters) { counters[0] += nfcn; counters[1] += nstep; counters[2] += naccpt; counters[3] += nrejct; }
    return idid;
}

/* ------------------------------------------------------------------------------------------
 * Closed loop -- SimObj.simulate (simulation/sim_engine.py:29-39) over T1DSimEnv.reset/.step
 * (simulation/env.py:119-155, 48-117), T1DPatient.step (patient/t1dpatient.py:82-116,
 * 222-236), CGMSensor.measure (sensor/cgm.py:26-36), BBController._bb_policy
 * (controller/basal_bolus_ctrller.py:64-80), PIDController.policy (controller/pid_ctrller.py:
 * 17-36), risk_diff (env.py:27-33).
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    double x[NX];
    double t;
    double planned_meal, last_cho, last_Qsto, last_foodtaken;
    int is_eating;
    double h_carry;
    double last_cgm;
    int noise_pos;
} patient_state;

static double sensor_measure(const sgo_config *cfg, patient_state *s, const double *p,
                             const double *noise)
{
    /* cgm.py:27-36: `patient.t % sample_time == 0` on the solver's float time */
    if (fmod(s->t, (double)cfg->sample_time) == 0.0) {
        double bg = s->x[12] / p[SGO_Vg];
        double cgm = bg + noise[s->noise_pos++];
        cgm = (cfg->sensor_min > cgm) ? cfg->sensor_min : cgm;   /* max(CGM, min) */
        cgm = (cfg->sensor_max < cgm) ? cfg->sensor_max : cgm;   /* min(CGM, max) */
        s->last_cgm = cgm;
        return cgm;
    }
    return s->last_cgm;
}

static int patient_step(patient_state *s, const double *p, double meal, double insulin,
                        int carry_h, int *counters)
{
    /* _announce_meal, t1dpatient.py:222-236 */
    double to_eat;
    s->planned_meal += meal;
    if (s->planned_meal > 0) {
        to_eat = (s->planned_meal < 5.0) ? s->planned_meal : 5.0;   /* min(EAT_RATE, planned) */
        s->planned_meal -= to_eat;
        s->planned_meal = (0.0 > s->planned_meal) ? 0.0 : s->planned_meal;   /* max(0, planned) */
    } else {
        to_eat = 0.0;
    }
    /* t1dpatient.py:88-107 */
    if (to_eat > 0 && s->last_cho <= 0) {
        s->last_Qsto = s->x[0] + s->x[1];
        s->last_foodtaken = 0.0;
        s->is_eating = 1;
    }
    if (s->is_eating) s->last_foodtaken += to_eat;
    if (to_eat <= 0 && s->last_cho > 0) s->is_eating = 0;
    s->last_cho = to_eat;
    /* t1dpatient.py:110-113 */
    return sgo_dopri5_minute(s->x, &s->t, p, to_eat, insulin, s->last_Qsto, s->last_foodtaken,
                             carry_h, &s->h_carry, counters);
}

static double risk_total(double v)
{
    double r[3];
    sgo_risk(v, r);
    return r[2];
}

int sgo_run_episode(const sgo_config *cfg, const double *p, const double *x0, int n_steps,
                    const int *meal_minute, const double *meal_grams, int n_meals,
                    const double *noise, const double *actions,
                    double *trace, int *minute_counters, double *stats, double *final_state)
{
    patient_state s;
    const int st_i = cfg->sample_time;
    const double st = (double)st_i;
    int status = 1;
    double r[3];

    /* patient.reset(), t1dpatient.py:247-281 (x0 is the post-reset init_state) */
    memcpy(s.x, x0, sizeof s.x);
    s.t = (double)cfg->t0;     /* t1dpatient.py:277: self._odesolver.set_initial_value(self.init_state, self.t0) */
    s.planned_meal = 0.0; s.last_cho = 0.0;
    s.last_Qsto = x0[0] + x0[1]; s.last_foodtaken = 0.0; s.is_eating = 0;
    s.h_carry = 0.0; s.last_cgm = 0.0; s.noise_pos = 0;

    /* env._reset(), env.py:119-134 */
    double bg_hist = s.x[12] / p[SGO_Vg];
    sgo_risk(bg_hist, r);
    double cgm_hist_prev = sensor_measure(cfg, &s, p, noise);   /* CGM_hist[0] */
    /* env.reset(), env.py:141-143: second sample becomes the observation */
    double obs = sensor_measure(cfg, &s, p, noise);
    double reward = 0.0, done = 0.0, info_meal = 0.0;

    /* controller.reset(): PID state (pid_ctrller.py:38-40) */
    double pid_integ = 0.0, pid_prev = 0.0;

    double st_rows = 0, sum_bg = 0, min_bg = INFINITY, max_bg = -INFINITY, sum_l = 0, sum_h = 0,
           sum_r = 0, sum_rew = 0, lt50 = 0, lt70 = 0, inr = 0, gt180 = 0, gt250 = 0,
           first_done = -1, tot_steps = 0, tot_rej = 0;
    int mpos = 0;

    for (int k = 0;; k++) {
        /* row k of the history */
        double row_bg = bg_hist, row_cgm = cgm_hist_prev;
        st_rows += 1; sum_bg += row_bg;
        if (row_bg < min_bg) min_bg = row_bg;
        if (row_bg > max_bg) max_bg = row_bg;
        sum_l += r[0]; sum_h += r[1]; sum_r += r[2]; sum_rew += reward;
        if (row_bg < 50) lt50 += 1;
        if (row_bg < 70) lt70 += 1;
        if (row_bg >= 70 && row_bg <= 180) inr += 1;
        if (row_bg > 180) gt180 += 1;
        if (row_bg > 250) gt250 += 1;
        if (done != 0.0 && first_done < 0) first_done = k;
        if (trace) {
            double *tr = trace + (size_t)k * SGO_NTRACE;
            tr[SGO_T_BG] = row_bg; tr[SGO_T_CGM] = row_cgm;
            tr[SGO_T_CHO] = NAN; tr[SGO_T_INSULIN] = NAN;
            tr[SGO_T_LBGI] = r[0]; tr[SGO_T_HBGI] = r[1]; tr[SGO_T_RISK] = r[2];
            tr[SGO_T_OBS] = obs; tr[SGO_T_REWARD] = reward; tr[SGO_T_DONE] = done;
            tr[SGO_T_BASAL] = NAN; tr[SGO_T_BOLUS] = NAN;
        }
        if (k == n_steps) break;

        /* controller.policy(obs, reward, done, **info) */
        double a_basal, a_bolus;
        if (cfg->controller == SGO_CTRL_BB) {
            /* basal_bolus_ctrller.py:64-80; ctrl = target, CR, CF, basal.  The reference looks u2ss and BW up in the
             * CSV by patient NAME (basal_bolus_ctrller.py:52-58), not in the patient object: a patient built on a
             * perturbed parameter Series keeps the nominal basal.  ctrl[3] > 0 carries that nominal u2ss * BW / 6000;
             * 0 = the episode's own row (the same number for the 30 CSV rows). */
            a_basal = (cfg->ctrl[3] > 0.0) ? cfg->ctrl[3] : p[SGO_u2ss] * p[SGO_BW] / 6000.0;
            if (info_meal > 0) {
                double corr = (obs > 150 ? 1.0 : 0.0) * (obs - cfg->ctrl[0]) / cfg->ctrl[2];
                a_bolus = (info_meal * st) / cfg->ctrl[1] + corr;
            } else {
                a_bolus = 0.0;
            }
            a_bolus = a_bolus / st;
        } else if (cfg->controller == SGO_CTRL_PID) {
            /* pid_ctrller.py:22-35; ctrl = P, I, D, target */
            double bgc = obs;
            a_basal = cfg->ctrl[0] * (bgc - cfg->ctrl[3]) + cfg->ctrl[1] * pid_integ +
                      cfg->ctrl[2] * (bgc - pid_prev) / st;
            pid_prev = bgc;
            pid_integ += (bgc - cfg->ctrl[3]) * st;
            a_bolus = 0.0;
        } else {
            a_basal = actions[2 * k]; a_bolus = actions[2 * k + 1];
        }

        /* env.step, env.py:66-117 */
        double CHO = 0.0, insulin = 0.0, BG = 0.0, CGM = 0.0;
        for (int j = 0; j < st_i; j++) {
            /* mini_step, env.py:48-64; scenario.get_action(self.time) with time = start + patient.t */
            int minute = k * st_i + j;
            /* env.time = scenario.start_time + patient.t minutes (env.py:44-46): the table is in minutes after start_time */
            const int clock = cfg->t0 + minute;
            double meal = 0.0;
            while (mpos < n_meals && meal_minute[mpos] >= 0 && meal_minute[mpos] < clock) mpos++;
            if (mpos < n_meals && meal_minute[mpos] == clock) meal = meal_grams[mpos];
            double basal = sgo_pump(a_basal, cfg->pump[0], cfg->pump[1], cfg->pump[2]);
            double bolus = sgo_pump(a_bolus, cfg->pump[3], cfg->pump[4], cfg->pump[5]);
            double ins = basal + bolus;
            int c[4] = {0, 0, 0, 0};
            int idid = patient_step(&s, p, meal, ins, cfg->carry_h, c);
            if (minute_counters) memcpy(minute_counters + (size_t)minute * 4, c, sizeof c);
            tot_steps += c[1]; tot_rej += c[3];
            if (idid < 0 && status > 0) status = idid;
            double tmp_bg = s.x[12] / p[SGO_Vg];
            double tmp_cgm = sensor_measure(cfg, &s, p, noise);
            CHO += meal / st; insulin += ins / st; BG += tmp_bg / st; CGM += tmp_cgm / st;
        }
        if (trace) {
            double *tr = trace + (size_t)k * SGO_NTRACE;
            tr[SGO_T_CHO] = CHO; tr[SGO_T_INSULIN] = insulin;
            tr[SGO_T_BASAL] = a_basal; tr[SGO_T_BOLUS] = a_bolus;
        }
        sgo_risk(BG, r);
        /* reward_fun(CGM_hist[-window:]) with the default risk_diff, env.py:27-33,100-102 */
        reward = risk_total(cgm_hist_prev) - risk_total(CGM);
        done = (BG < 70 || BG > 350) ? 1.0 : 0.0;
        bg_hist = BG; cgm_hist_prev = CGM; obs = CGM; info_meal = CHO;
        if (status < 0) {   /* t1dpatient.py:112-116: the reference raises on the next step */
            if (trace)
                for (int kk = k + 1; kk <= n_steps; kk++)
                    for (int c2 = 0; c2 < SGO_NTRACE; c2++) trace[(size_t)kk * SGO_NTRACE + c2] = NAN;
            break;
        }
    }
    if (stats) {
        stats[SGO_S_ROWS] = st_rows; stats[SGO_S_SUM_BG] = sum_bg; stats[SGO_S_MIN_BG] = min_bg;
        stats[SGO_S_MAX_BG] = max_bg; stats[SGO_S_SUM_LBGI] = sum_l; stats[SGO_S_SUM_HBGI] = sum_h;
        stats[SGO_S_SUM_RISK] = sum_r; stats[SGO_S_SUM_REWARD] = sum_rew; stats[SGO_S_LT50] = lt50;
        stats[SGO_S_LT70] = lt70; stats[SGO_S_INRANGE] = inr; stats[SGO_S_GT180] = gt180;
        stats[SGO_S_GT250] = gt250; stats[SGO_S_FIRST_DONE] = first_done;
        stats[SGO_S_DOPRI_STEPS] = tot_steps; stats[SGO_S_DOPRI_REJECTS] = tot_rej;
    }
    if (final_state) memcpy(final_state, s.x, sizeof s.x);
    return status;
}

int sgo_max_threads(void)
{
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n < 1 ? 1 : (int)n;
}

typedef struct {
    const sgo_config *cfg;
    int n_episodes, n_steps, max_meals;
    const double *params, *x0, *meal_grams, *noise, *ctrl_rows;
    const int *meal_minute;
    double *stats;
    atomic_int next, failed;
} batch_job;

static void *batch_worker(void *arg)
{
    batch_job *j = (batch_job *)arg;
    for (;;) {
        int e = atomic_fetch_add(&j->next, 1);
        if (e >= j->n_episodes) break;
        sgo_config c = *j->cfg;
        if (j->ctrl_rows) memcpy(c.ctrl, j->ctrl_rows + (size_t)e * 6, sizeof c.ctrl);
        int nm = 0;
        const int *mm = j->meal_minute + (size_t)e * j->max_meals;
        while (nm < j->max_meals && mm[nm] >= 0) nm++;
        int rc = sgo_run_episode(&c, j->params + (size_t)e * SGO_NPARAM, j->x0 + (size_t)e * NX,
                                 j->n_steps, mm, j->meal_grams + (size_t)e * j->max_meals, nm,
                                 j->noise + (size_t)e * (j->n_steps + 2), NULL, NULL, NULL,
                                 j->stats + (size_t)e * SGO_NSTAT, NULL);
        if (rc < 0) atomic_fetch_add(&j->failed, 1);
    }
    return NULL;
}

/* The process-pool map of simulation/sim_engine.py:65-69, as a thread pool over episodes. */
int sgo_run_batch(const sgo_config *cfg, int n_episodes, const double *params, const double *x0,
                  int n_steps, const int *meal_minute, const double *meal_grams, int max_meals,
                  const double *noise, const double *ctrl_rows, double *stats, int n_threads)
{
    batch_job j = {cfg, n_episodes, n_steps, max_meals, params, x0, meal_grams, noise, ctrl_rows,
                   meal_minute, stats, 0, 0};
    if (n_threads < 1) n_threads = sgo_max_threads();
    if (n_threads > 1024) n_threads = 1024;
    pthread_t th[1024];
    int started = 0;
    for (int i = 1; i < n_threads; i++)
        if (pthread_create(&th[started], NULL, batch_worker, &j) == 0) started++;
    batch_worker(&j);
    for (int i = 0; i < started; i++) pthread_join(th[i], NULL);
    return atomic_load(&j.failed);
}
