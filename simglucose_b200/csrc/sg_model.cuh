// Device functions of the batched UVa/Padova-2008 patient simulator (sm_100a, FP64).
//
// One CUDA thread owns one patient-episode: the 13-state vector and the Dormand-Prince
// stages live in registers, the per-episode constants derived from the parameter row live in
// a shared-memory column private to the thread (sp[i * BLOCK], conflict-free: consecutive
// lanes hit consecutive 8-byte words).
//
// Reference behaviour reproduced here (paths relative to the reference root):
//   rhs()             T1DPatient.model              simglucose/patient/t1dpatient.py:118-208
//   dopri5_minute()   scipy ode('dopri5') as called at t1dpatient.py:276-277 (defaults) and
//                     :110-113 (integrate one minute) -- Hairer's DOPCOR + HINIT (third party)
//   meal_machine()    T1DPatient.step / _announce_meal          t1dpatient.py:82-107, 222-236
//   pump_quantise()   InsulinPump.basal / .bolus        simglucose/actuator/pump.py:23-39
//   risk_index()      risk_index([BG], 1)               simglucose/analysis/risk.py:5-17
//
// The operation ORDER follows the Python source (a*b/c == (a*b)/c, sums left to right) so that
// a build with -fmad=false makes the same accept/reject decisions as the scipy path; see
// DESIGN.md "Arithmetic contract".
#pragma once
#include <cfloat>
#include <cmath>
#include <cstdint>

#include "sg_math.cuh"

namespace sgk {

constexpr int NX = 13;

// Per-episode constants staged in shared memory (one column per thread).
enum Hoisted {
    // Rows 0 .. H_PAIR_END-1 are stored as PAIRS: the two values of rows (2k, 2k+1) of a thread are adjacent in shared
    // memory, so that the RHS fetches them with one LDS.128 (Column::pair).  Pair partners are neighbours in the
    // RHS's order of use.  aa..dD are per-minute values (rewritten when Dbar moves), rBW / rVi refined reciprocals
    // of constant divisors (sgm::rcp_nr): x / BW and x / Vi cost 3 instructions.
    H_aa = 0, H_bD, H_cc, H_dD, H_kmin, H_hk, H_kmax, H_kabs, H_fk, H_BW, H_rBW, H_kp1, H_kp2, H_kp3,
    H_ke2, H_ke1, H_k1, H_k2, H_Fsnc, H_Km0, H_Vm0, H_Vmx, H_ka2, H_m2, H_m24, H_m1, H_ka1, H_m13,
    H_Vi, H_rVi, H_p2u, H_ki, H_Ib, H_kd, H_kad, H_ksc,
    H_PAIR_END,
    H_b = H_PAIR_END, H_d, H_Vg,
    // per-minute / per-env-step bookkeeping kept out of registers: reciprocal of Vg (BG = x12 / Vg every minute), the
    // Dbar the H_aa..H_dD rows were computed for, the risk index of the previous env-step's CGM (risk_diff)
    H_rVg, H_Dbar, H_RPREV,
    H_L4,                      // log of 1e-4, the floor of DOPCOR's facold (step_control works on logarithms)
    H_BASAL,                   // u2ss * BW / 6000: BBController's basal from the episode's own row (basal_bolus_ctrller.py:64)
    // per-thread scratch vectors of the integrator, NV = 14 rows each (13 states + one pad row, so that states
    // (2k, 2k+1) form a pair): k1 of the current step (needed again only after a rejected step and by HINIT), the two
    // longest-lived stage accumulators (see dopri5_minute) and the accepted state y of the integrator
    H_K1, H_P7 = H_K1 + 14, H_PE = H_P7 + 14, H_Y = H_PE + 14,
    H_COUNT = H_Y + 14
};
constexpr int NV = 14;

// Order of the parameter row handed over the C-ABI (include/simglucose_b200.h, sg_param).
enum Param {
    P_BW = 0, P_u2ss, P_kmax, P_kmin, P_b, P_d, P_kabs, P_f, P_kp1, P_kp2, P_kp3, P_Fsnc,
    P_ke1, P_ke2, P_k1, P_k2, P_Vm0, P_Vmx, P_Km0, P_m1, P_m2, P_m4, P_m30, P_ka1, P_ka2,
    P_Vi, P_p2u, P_Ib, P_ki, P_kd, P_ksc, P_Vg,
    P_COUNT
};

static_assert(H_PAIR_END % 2 == 0 && H_K1 % 2 == 0 && H_COUNT % 2 == 0, "paired rows come in twos");

// The shared-memory column of one thread.
//   PAIRED (the single-thread kernels): rows (2k, 2k+1) of thread t are adjacent -- base = &smem[2 * t], row i at
//     base[(i & ~1) * BLOCK + (i & 1)] -- so that two parameters / two states move with one 16-byte access
//     (pair / set_pair: a warp's 32 x 16 bytes are consecutive).  An 8-byte access in this layout costs twice the
//     wavefronts, so the integrator loop only uses pairs.
//   plain (the warp-pair kernel): base = &smem[t], row i at base[i * BLOCK].
// volatile: every use re-reads shared memory.  Without it ptxas keeps the ~35 parameters of the eight inlined RHS
// copies in registers and spills Runge-Kutta stages to local memory instead.
template <int BLOCK, bool PAIRED = true>
struct Column {
    double *base;
    static __device__ __forceinline__ int at(int i) { return PAIRED ? (i & ~1) * BLOCK + (i & 1) : i * BLOCK; }
    __device__ __forceinline__ double operator()(int i) const { return ((const volatile double *)base)[at(i)]; }
    __device__ __forceinline__ void set(int i, double v) const { base[at(i)] = v; }
    // rows (i, i + 1), i even: one 16-byte access (PAIRED only)
    __device__ __forceinline__ double2 pair(int i) const
    {
        double2 r;
        asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];"
                     : "=d"(r.x), "=d"(r.y)
                     : "r"((unsigned)__cvta_generic_to_shared(base + i * BLOCK)));
        return r;
    }
    __device__ __forceinline__ void set_pair(int i, double2 v) const
    {
        asm volatile("st.shared.v2.f64 [%0], {%1, %2};"
                     :: "r"((unsigned)__cvta_generic_to_shared(base + i * BLOCK)), "d"(v.x), "d"(v.y) : "memory");
    }
};

// Constants that only depend on the parameter row (t1dpatient.py:138,151,176,190,194: the
// sub-expressions (kmax-kmin)/2, f*kabs, m2+m4, m1+m30, ka1+kd are evaluated first by Python's
// precedence, so hoisting them is bit-safe).
template <class ColT, class Load>
__device__ __forceinline__ void stage_params(const ColT &c, Load p)
{
    const double kmax = p(P_kmax), kmin = p(P_kmin), kabs = p(P_kabs);
    const double m1 = p(P_m1), m2 = p(P_m2), ka1 = p(P_ka1), kd = p(P_kd);
    c.set(H_kmax, kmax); c.set(H_kmin, kmin); c.set(H_hk, SG_TANH_SUM ? 2.0 * ((kmax - kmin) / 2.0) : (kmax - kmin) / 2.0);   // doubling is exact
    c.set(H_b, p(P_b)); c.set(H_d, p(P_d)); c.set(H_kabs, kabs); c.set(H_fk, p(P_f) * kabs);
    c.set(H_BW, p(P_BW)); c.set(H_kp1, p(P_kp1)); c.set(H_kp2, p(P_kp2)); c.set(H_kp3, p(P_kp3));
    c.set(H_Fsnc, p(P_Fsnc)); c.set(H_ke1, p(P_ke1)); c.set(H_ke2, p(P_ke2));
    c.set(H_k1, p(P_k1)); c.set(H_k2, p(P_k2)); c.set(H_Vm0, p(P_Vm0)); c.set(H_Vmx, p(P_Vmx));
    c.set(H_Km0, p(P_Km0)); c.set(H_m24, m2 + p(P_m4)); c.set(H_m1, m1); c.set(H_ka1, ka1);
    c.set(H_ka2, p(P_ka2)); c.set(H_Vi, p(P_Vi)); c.set(H_p2u, p(P_p2u)); c.set(H_Ib, p(P_Ib));
    c.set(H_ki, p(P_ki)); c.set(H_m13, m1 + p(P_m30)); c.set(H_m2, m2); c.set(H_kad, ka1 + kd);
    c.set(H_kd, kd); c.set(H_ksc, p(P_ksc)); c.set(H_Vg, p(P_Vg));
    c.set(H_aa, 0.0); c.set(H_cc, 0.0); c.set(H_bD, 0.0); c.set(H_dD, 0.0);
    c.set(H_rBW, sgm::rcp_nr(p(P_BW))); c.set(H_rVi, sgm::rcp_nr(p(P_Vi)));
    c.set(H_rVg, sgm::rcp_nr(p(P_Vg))); c.set(H_Dbar, 0.0); c.set(H_RPREV, 0.0);
    c.set(H_L4, sgm::log_pos(1.0e-4));
    c.set(H_BASAL, sgm::div_cold(p(P_u2ss) * p(P_BW), 6000.0));
}

// Inputs of the ODE that are constant over one simulated minute (t1dpatient.py:110:
// set_f_params(action, params, last_Qsto, last_foodtaken)).
struct Minute {
    double d;      // action.CHO * 1000                      t1dpatient.py:121
    double ins;    // action.insulin * 6000 / BW             t1dpatient.py:122
    bool fed;      // Dbar > 0                               t1dpatient.py:130,135
    double2 ab, cd;  // (2 aa, b*Dbar), (2 cc, d*Dbar) of the gastric-emptying law: the head of the RHS's longest chain, in
                     // registers; the slopes carry the factor 2 of tanh(u) = 1 - 2 / (e^2u + 1) (exact: a power of two)
};

// ins_per_kg: action.insulin * 6000 / BW (constant over an env-step: the caller divides once).
template <int BLOCK>
__device__ __forceinline__ Minute minute_setup_rate(const Column<BLOCK> &c, double cho, double ins_per_kg,
                                                    double last_Qsto, double last_food)
{
    Minute m;
    m.d = cho * 1000.0;
    m.ins = ins_per_kg;
    const double Dbar = last_Qsto + last_food * 1000.0;
    m.fed = SG_INT_SIGN ? sgm::is_pos(Dbar) : Dbar > 0.0;
    // Dbar only moves while the patient eats: the four quotients / products below are recomputed when it did
    // (same inputs, same results -- the cache cannot change a bit)
    if (m.fed && Dbar != c(H_Dbar)) {                       // t1dpatient.py:136-140
        c.set(H_Dbar, Dbar);
        c.set(H_aa, sgm::div_cold(sgm::div_cold(5.0 / 2.0, 1.0 - c(H_b)), Dbar));
        c.set(H_cc, sgm::div_cold(sgm::div_cold(5.0 / 2.0, c(H_d)), Dbar));
        c.set(H_bD, c(H_b) * Dbar);
        c.set(H_dD, c(H_d) * Dbar);
    }
    m.ab = c.pair(H_aa); m.cd = c.pair(H_cc);
    m.ab.x *= 2.0; m.cd.x *= 2.0;
    return m;
}

template <int BLOCK>
__device__ __forceinline__ Minute minute_setup(const Column<BLOCK> &c, double cho, double insulin,
                                               double last_Qsto, double last_food)
{
    Minute m;
    m.d = cho * 1000.0;
    m.ins = insulin * 6000.0 / c(H_BW);
    const double Dbar = last_Qsto + last_food * 1000.0;
    m.fed = SG_INT_SIGN ? sgm::is_pos(Dbar) : Dbar > 0.0;
    if (m.fed) {                                            // t1dpatient.py:136-140
        c.set(H_aa, 5.0 / 2.0 / (1.0 - c(H_b)) / Dbar);
        c.set(H_cc, 5.0 / 2.0 / c(H_d) / Dbar);
        c.set(H_bD, c(H_b) * Dbar);
        c.set(H_dD, c(H_d) * Dbar);
    }
    m.ab = c.pair(H_aa); m.cd = c.pair(H_cc);
    m.ab.x *= 2.0; m.cd.x *= 2.0;
    return m;
}

// States enter the kernels through canonical_zero: x + 0.0 turns a -0.0 (which only a caller can supply) into +0.0 and
// changes no other value.  From then on no state can BE -0.0 -- every trial and accepted state is a sum y + h * (..), and
// a sum is -0.0 only if both terms are -- so "x < 0" is the sign bit, one integer test of the high word per mask
// instead of a 64-bit compare.  ((x >= 0) * dxdt of the reference treats -0.0 like +0.0 as well.)
__device__ __forceinline__ double canonical_zero(double x) { return x + 0.0; }

__device__ __forceinline__ double mask_nonneg(double x, double dx)
{
    // (x >= 0) * dxdt, t1dpatient.py:167,173,179,191,195,198,202 (the sign of a zero result is
    // not preserved; it never reaches an output)
#if SG_INT_SIGN == 2
    return __double2hiint(x) < 0 ? 0.0 : dx;
#elif SG_INT_SIGN
    return sgm::is_neg(x) ? 0.0 : dx;
#else
    return x >= 0.0 ? dx : 0.0;
#endif
}

// T1DPatient.model, t1dpatient.py:118-208.  Parameters arrive two per shared-memory load (Column::pair).
template <int BLOCK>
__device__ __forceinline__ void rhs(const Column<BLOCK> &c, const Minute &m, const double (&x)[NX],
                                    double (&dx)[NX])
{
    const double2 kk = c.pair(H_kmax);                                      // kmax, kabs
    const double kmax = kk.x;
    const double qsto = x[0] + x[1];                                        // :126
    dx[0] = -kmax * x[0] + m.d;                                             // :133
    // :135-142, branch-free: tanh(aa*(qsto - b*Dbar)) - tanh(cc*(qsto - d*Dbar)) in one expression (sg_math.cuh),
    // evaluated for every lane and discarded by the select when Dbar == 0
    const double2 ab = m.ab, cd = m.cd;                                     // (2 aa, b*Dbar), (2 cc, d*Dbar)
    const double2 kh = c.pair(H_kmin);                                      // kmin, (kmax - kmin) / 2 (SG_TANH_SUM: twice that)
#if SG_TANH_SUM
    // kmin + hk * (tanh(..) - tanh(..) + 2) with the bracket as one transcendental quantity T (sg_math.cuh tanh_sum_half
    // returns T / 2; hk * T == (2 hk) * (T / 2) exactly, and the row H_hk holds 2 hk)
    const double kfed = kh.x + kh.y * sgm::tanh_sum_half(ab.x * (qsto - ab.y), cd.x * (qsto - cd.y));
#else
    const double tdiff = sgm::tanh_diff2(ab.x * (qsto - ab.y), cd.x * (qsto - cd.y));
    const double kfed = kh.x + kh.y * (tdiff + 2.0);
#endif
    const double kgut = m.fed ? kfed : kmax;
    dx[1] = kmax * x[0] - x[1] * kgut;                                      // :145
    dx[2] = kgut * x[1] - kk.y * x[2];                                      // :148
    const double2 fb = c.pair(H_fk), rk = c.pair(H_rBW);                    // (f*kabs, BW), (1/BW, kp1)
    const double Rat = sgm::div_r(fb.x * x[2], fb.y, rk.x);                 // :151
    const double2 kp = c.pair(H_kp2);                                       // kp2, kp3
    const double EGPt = rk.y - kp.x * x[3] - kp.y * x[8];                   // :153
    const double2 ke = c.pair(H_ke2);                                       // ke2, ke1
#if SG_INT_SIGN
    const double over = x[3] - ke.x;                                        // x3 > ke2 <=> x3 - ke2 > 0 (exact in IEEE)
#if SG_INT_SIGN == 2
    // x3 is never -0.0, so neither is the difference: over >= +0 exactly when the sign bit is clear, and ke1 * (+0) is the
    // 0 of the else branch
    const double Et = __double2hiint(over) < 0 ? 0.0 : ke.y * over;        // :158-161
#else
    const double Et = sgm::is_pos(over) ? ke.y * over : 0.0;                // :158-161
#endif
    // (EGPt = kp1 - .. - .. is -0.0 only if kp1 is: the sign bit decides here too; max(-0.0, 0) and 0.0 are equal)
    const double egp = (SG_INT_SIGN >= 2 ? __double2hiint(EGPt) < 0 : sgm::is_neg(EGPt)) ? 0.0 : EGPt;   // max(EGPt, 0) :165
#else
    const double Et = (x[3] > ke.x) ? ke.y * (x[3] - ke.x) : 0.0;           // :158-161
    const double egp = (0.0 > EGPt) ? 0.0 : EGPt;                           // max(EGPt, 0) :165
#endif
    const double2 k12 = c.pair(H_k1), fk = c.pair(H_Fsnc);                  // (k1, k2), (Fsnc, Km0)
    dx[3] = egp + Rat - fk.x - Et - k12.x * x[3] + k12.y * x[4];            // :165-166
    const double2 vm = c.pair(H_Vm0);                                       // Vm0, Vmx
    const double Vmt = vm.x + vm.y * x[6];                                  // :169
    const double Uidt = sgm::div_nr(Vmt * x[4], fk.y + x[4]);               // :171
    dx[4] = -Uidt + k12.x * x[3] - k12.y * x[4];                            // :172
    const double2 am = c.pair(H_ka2), mm = c.pair(H_m24), km = c.pair(H_ka1);   // (ka2, m2), (m2+m4, m1), (ka1, m1+m30)
    dx[5] = -mm.x * x[5] + mm.y * x[9] + km.x * x[10] + am.x * x[11];       // :176
    const double2 vi = c.pair(H_Vi);                                        // Vi, 1/Vi
    const double It = sgm::div_r(x[5], vi.x, vi.y);                         // :178
    const double2 pk = c.pair(H_p2u), ik = c.pair(H_Ib);                    // (p2u, ki), (Ib, kd)
    dx[6] = -pk.x * x[6] + pk.x * (It - ik.x);                              // :182
    dx[7] = -pk.y * (x[7] - It);                                            // :185
    dx[8] = -pk.y * (x[8] - x[7]);                                          // :187
    dx[9] = -km.y * x[9] + am.y * x[5];                                     // :190
    const double2 ak = c.pair(H_kad);                                       // ka1 + kd, ksc
    dx[10] = m.ins - ak.x * x[10];                                          // :194
    dx[11] = ik.y * x[10] - am.x * x[11];                                   // :197
    dx[12] = -ak.y * x[12] + ak.y * x[3];                                   // :201
    // (x >= 0) * dxdt on seven states (:167,173,179,191,195,198,202).  Or-ing the sign bits and running these selects
    // only in a warp that holds a negative state was measured in s41: 15 instructions fewer per pass and 3 % SLOWER
    // (the vote and branch end the basic block in which the RHS tail overlaps the stage arithmetic).
    dx[3] = mask_nonneg(x[3], dx[3]);
    dx[4] = mask_nonneg(x[4], dx[4]);
    dx[5] = mask_nonneg(x[5], dx[5]);
    dx[9] = mask_nonneg(x[9], dx[9]);
    dx[10] = mask_nonneg(x[10], dx[10]);
    dx[11] = mask_nonneg(x[11], dx[11]);
    dx[12] = mask_nonneg(x[12], dx[12]);
}

// InsulinPump.basal / .bolus, pump.py:23-39.  np.round == round-half-even == rint().
__device__ __forceinline__ double pump_quantise(double amount, double vmin, double vmax, double inc)
{
    double v = amount * 6000.0;
    v = rint(sgm::div_cold(v, inc)) * inc;
    v = sgm::div_cold(v, 6000.0);
    v = (vmax < v) ? vmax : v;     // Python min(v, vmax)
    v = (vmin > v) ? vmin : v;     // Python max(v, vmin)
    return v;
}

// f(BG) = 1.509 * (ln(BG)^1.084 - 5.381), risk.py:9-10.  log and pow come from sg_math.cuh (the code the step-size
// controller already keeps in the instruction cache; CUDA's pow is ~180 instructions per inlined call and the episode
// kernels evaluate the risk two or three times per env-step) and the function is not inlined.  NaN, negative BG and
// 0 < BG < 1 (log < 0 -> pow NaN) return NaN; BG == 0 gives (-inf)^1.084 = +inf; BG == 1 gives 0^1.084 = 0; +inf stays
// +inf (tests/test_gpu_math.py checks these against the oracle's libm evaluation).
static __device__ __noinline__ double risk_f(double bg)
{
    if (!(bg >= 1.0))                                  // log(+-0) = -inf and pow(-inf, 1.084) = +inf, as in libm / numpy
        return __longlong_as_double(bg == 0.0 ? 0x7ff0000000000000ll : 0x7ff8000000000000ll);
    if (bg > DBL_MAX) return bg;
    const double lg = sgm::log_core(bg);
    const double pw = (lg > 0.0) ? sgm::pow_pos(lg, 1.084) : 0.0;          // the subroutine HINIT already uses
    return 1.509 * (pw - 5.381);
}

// risk_index([BG], 1), risk.py:5-17 (NaN -> 0 via nan_to_num; +inf -> DBL_MAX).
__device__ __forceinline__ void risk_index(double bg, double &lbgi, double &hbgi, double &ri)
{
    const double f = risk_f(bg);
    lbgi = (f < 0.0) ? 10.0 * (f * f) : 0.0;
    hbgi = (f > 0.0) ? 10.0 * (f * f) : 0.0;
    if (isinf(lbgi)) lbgi = DBL_MAX;
    if (isinf(hbgi)) hbgi = DBL_MAX;
    ri = lbgi + hbgi;
}

// Meal state machine of T1DPatient.step (t1dpatient.py:82-107) with _announce_meal (:222-236).
struct Meal {
    double planned, last_cho, last_Qsto, last_food;
    int eating;
};

__device__ __forceinline__ double meal_machine(Meal &s, double announced, double x0, double x1)
{
    double to_eat;
    s.planned += announced;
    if (s.planned > 0.0) {
        to_eat = (s.planned < 5.0) ? s.planned : 5.0;       // min(EAT_RATE, planned)
        s.planned -= to_eat;
        s.planned = (0.0 > s.planned) ? 0.0 : s.planned;
    } else {
        to_eat = 0.0;
    }
    if (to_eat > 0.0 && s.last_cho <= 0.0) {                // onset :88-92
        s.last_Qsto = x0 + x1;
        s.last_food = 0.0;
        s.eating = 1;
    }
    if (s.eating) s.last_food += to_eat;                    // :98-99
    if (to_eat <= 0.0 && s.last_cho > 0.0) s.eating = 0;    // :102-104
    s.last_cho = to_eat;
    return to_eat;
}

// ---- DOPRI5 (Hairer, Norsett & Wanner; scipy _dop.dopri5), scipy defaults ----------------
// The coefficients live in constant memory so that DMUL/DFMA/DADD read them as c[bank][offset] operands;
// as 64-bit immediates each use costs two extra UMOV instructions (14 % of the instructions executed by
// the first version of this kernel, profiles/r01_*).
enum Coef {
    K_A21, K_A31, K_A32, K_A41, K_A42, K_A43, K_A51, K_A52, K_A53, K_A54, K_A61, K_A62, K_A63, K_A64, K_A65,
    K_A71, K_A73, K_A74, K_A75, K_A76, K_E1, K_E3, K_E4, K_E5, K_E6, K_E7, K_RTOL, K_ATOL, K_COUNT
};
static __constant__ double KC[K_COUNT] = {
    0.2, 3.0 / 40.0, 9.0 / 40.0, 44.0 / 45.0, -56.0 / 15.0, 32.0 / 9.0, 19372.0 / 6561.0, -25360.0 / 2187.0,
    64448.0 / 6561.0, -212.0 / 729.0, 9017.0 / 3168.0, -355.0 / 33.0, 46732.0 / 5247.0, 49.0 / 176.0,
    -5103.0 / 18656.0, 35.0 / 384.0, 500.0 / 1113.0, 125.0 / 192.0, -2187.0 / 6784.0, 11.0 / 84.0,
    71.0 / 57600.0, -71.0 / 16695.0, 71.0 / 1920.0, -17253.0 / 339200.0, 22.0 / 525.0, -1.0 / 40.0,
    1e-6, 1e-12};
#define A21 KC[K_A21]
#define A31 KC[K_A31]
#define A32 KC[K_A32]
#define A41 KC[K_A41]
#define A42 KC[K_A42]
#define A43 KC[K_A43]
#define A51 KC[K_A51]
#define A52 KC[K_A52]
#define A53 KC[K_A53]
#define A54 KC[K_A54]
#define A61 KC[K_A61]
#define A62 KC[K_A62]
#define A63 KC[K_A63]
#define A64 KC[K_A64]
#define A65 KC[K_A65]
#define A71 KC[K_A71]
#define A73 KC[K_A73]
#define A74 KC[K_A74]
#define A75 KC[K_A75]
#define A76 KC[K_A76]
#define E1 KC[K_E1]
#define E3 KC[K_E3]
#define E4 KC[K_E4]
#define E5 KC[K_E5]
#define E6 KC[K_E6]
#define E7 KC[K_E7]
#define RTOL KC[K_RTOL]
#define ATOL KC[K_ATOL]
constexpr double UROUND = 2.3e-16;
constexpr int NMAX = 500;

struct StepCount { int nstep, nrej; };

// Integrator state of a minute that was put aside at the beginning of a DOPRI5 step (heavy minutes -- a meal onset
// costs up to 36 steps against 3 -- are finished by a second kernel over the compacted list of such episodes, so
// that the other 31 lanes of the warp do not wait for them).  k1 of the step travels in the H_K1 slots.
struct Suspend {
    double h, x, lfacold;
    int nstep, naccpt, nrej, reject;
};
enum MinuteMode { MINUTE_IDLE = 0, MINUTE_FRESH = 1, MINUTE_RESUME = 2 };
constexpr int IDID_SUSPENDED = 2;

// ---- pieces of the integrator that run once per minute / once per step and are kept OUT of the inlined loop
// body (not inlined: the loop has to fit the instruction cache, see sg_math.cuh pow_pos) --------------------

struct HinitGuess { double h, dnf; };

// Loops over the 13 states two at a time (the pair layout of Column): i = 0, 2, .., 12; state i + 1 == 13 is the pad.
#define SG_FOR_PAIRS(i) _Pragma("unroll") for (int i = 0; i < NV; i += 2)

template <int BLOCK>
__device__ __forceinline__ void store_vec(const Column<BLOCK> &c, int row, const double (&v)[NX])
{
    SG_FOR_PAIRS(i) c.set_pair(row + i, make_double2(v[i], (i + 1 < NX) ? v[i + 1 < NX ? i + 1 : i] : 0.0));
}

template <int BLOCK>
__device__ __forceinline__ void load_vec(const Column<BLOCK> &c, int row, double (&v)[NX])
{
    SG_FOR_PAIRS(i) {
        const double2 t = c.pair(row + i);
        v[i] = t.x;
        if (i + 1 < NX) v[i + 1 < NX ? i + 1 : i] = t.y;
    }
}

// HINIT, first half: h = 0.01 * |y| / |k1| in the scaled norm.  Reads k1 (H_K1) and y (H_Y) from the column and
// leaves the reciprocal scales in the H_PE rows for the second half.
// Both halves run once per minute, so their loops over the seven state pairs are NOT unrolled: 0.7 KB of code each
// instead of 4.9 / 2.6 KB next to a 27 KB integrator loop in a 32 KB instruction cache.  The pad state of the last
// pair (y = k = 0) goes through the same arithmetic and adds +0.0 to the sums, which changes no bit.
template <int BLOCK>
__device__ __noinline__ HinitGuess hinit_first(double *col, double hmax)
{
    const Column<BLOCK> c{col};
    double dny = 0.0, dnf = 0.0;
#pragma unroll 1
    for (int i = 0; i < NV; i += 2) {
        const double2 y2 = c.pair(H_Y + i), k2 = c.pair(H_K1 + i);
        double2 r2;
#if SG_FAST_HINIT
        // the quotients by the scale through a reciprocal refined to an ulp: HINIT is a heuristic guess of the first step
        // size from two norms, a real-valued function like the error norm (sg_math.cuh rcp_ulp)
        {
            const double yi = y2.x;
            const double r = sgm::rcp_ulp(ATOL + RTOL * fabs(yi));
            const double a = k2.x * r, b = yi * r;
            dnf = dnf + a * a;
            dny = dny + b * b;
            r2.x = r;
        }
        {
            const double yi = y2.y;
            const double r = sgm::rcp_ulp(ATOL + RTOL * fabs(yi));
            const double a = k2.y * r, b = yi * r;
            dnf = dnf + a * a;
            dny = dny + b * b;
            r2.y = r;
        }
#else
        {
            const double yi = y2.x;
            const double r = sgm::rcp_nr(ATOL + RTOL * fabs(yi));
            const double sk = ATOL + RTOL * fabs(yi);
            const double a = sgm::div_r(k2.x, sk, r), b = sgm::div_r(yi, sk, r);
            dnf = dnf + a * a;
            dny = dny + b * b;
            r2.x = r;
        }
        {
            const double yi = y2.y;
            const double r = sgm::rcp_nr(ATOL + RTOL * fabs(yi));
            const double sk = ATOL + RTOL * fabs(yi);
            const double a = sgm::div_r(k2.y, sk, r), b = sgm::div_r(yi, sk, r);
            dnf = dnf + a * a;
            dny = dny + b * b;
            r2.y = r;
        }
#endif
        c.set_pair(H_PE + i, r2);                            // the second half divides by the same scales
    }
    // dny / dnf through div_nr: bit-identical for these operands (both above 1e-10)
    double h = (dnf <= 1e-10 || dny <= 1e-10) ? 1.0e-6 : sqrt(sgm::div_nr(dny, dnf)) * 0.01;
    h = fmin(h, hmax);
    return HinitGuess{h, dnf};
}

// HINIT, second half: second-derivative estimate from f(x + h, y + h*k1) (H_P7) and k1 (H_K1).
template <int BLOCK>
__device__ __noinline__ double hinit_second(double *col, double h, double dnf, double hmax)
{
    const Column<BLOCK> c{col};
    double der2 = 0.0;
#pragma unroll 1
    for (int i = 0; i < NV; i += 2) {
#if SG_FAST_HINIT
        const double2 k2 = c.pair(H_K1 + i), f2 = c.pair(H_P7 + i), r2 = c.pair(H_PE + i);
#else
        const double2 y2 = c.pair(H_Y + i), k2 = c.pair(H_K1 + i), f2 = c.pair(H_P7 + i), r2 = c.pair(H_PE + i);
#endif
#if SG_FAST_HINIT
        {
            const double a = (f2.x - k2.x) * r2.x;                // reciprocal left by the first half
            der2 = der2 + a * a;
        }
        {
            const double a = (f2.y - k2.y) * r2.y;
            der2 = der2 + a * a;
        }
#else
        {
            const double sk = ATOL + RTOL * fabs(y2.x);
            const double a = sgm::div_r(f2.x - k2.x, sk, r2.x);   // reciprocal left by the first half
            der2 = der2 + a * a;
        }
        {
            const double sk = ATOL + RTOL * fabs(y2.y);
            const double a = sgm::div_r(f2.y - k2.y, sk, r2.y);
            der2 = der2 + a * a;
        }
#endif
    }
    der2 = sgm::div_nr(sqrt(der2), h);                       // h >= 1e-6
    const double der12 = fmax(fabs(der2), sqrt(dnf));
    const double h1 = (der12 <= 1e-15) ? fmax(1.0e-6, fabs(h) * 1.0e-3)
                                       : sgm::pow_pos(sgm::div_nr(0.01, sgm::min_sel(der12, 1e300)), 1.0 / 5.0);
    return fmin(fmin(100.0 * fabs(h), h1), hmax);
}

#if SG_FAST_ERRNORM
#define SG_ERR_QUOT(e, sk) ((e) * sgm::rcp_ulp(sk))
#else
#define SG_ERR_QUOT(e, sk) sgm::div_nr(e, sk)
#endif

struct StepControl { bool accept, above_floor; double hnew_accept, hnew_reject, lerr; };   // hnew_reject: SG_MERGED_POW == 0 only

// Step-size controller of DOPCOR (Lund stabilisation, beta = 0.04) from the summed squared scaled errors.
// lfacold: log of facold (log_pos).  Returns the log of err as well: the caller keeps it as the next lfacold.
// accept = (err <= 1), above_floor = (err > 1e-4) with err = sqrt(errsum / 13).
static __device__ __forceinline__ StepControl step_control(double errsum, double lfacold, double h)
{
    const double beta = 0.04, safe = 0.9;
    const double expo1 = 0.2 - beta * 0.75;
    const double facc1 = 1.0 / 0.2, facc2 = 1.0 / 10.0;
    StepControl r;
#if SG_FAST_CONTROL
    // err = sqrt(errsum / 13) is never formed: log(err) = (log(errsum) - log 13) / 2; the two comparisons of err are
    // comparisons of errsum with the largest doubles x for which sqrt(fl(x / 13)) <= 1 resp. <= 1e-4 (found by
    // bisection over the doubles; IEEE division and square root are monotone), so the decisions are those of DOPCOR
    // bit for bit.  hnew = h / max(facc2, min(facc1, fac / safe)) = h * min(10, max(0.2, safe / fac)) with
    // 1 / fac = facold^beta / err^expo1 = exp(beta ln facold - expo1 ln err): one exp, no division -- a
    // transcendental result defined to an ulp like the powers it replaces.
    const double lerr = 0.5 * (sgm::log_pos(errsum) - 0x1.485042b318c51p+1);
    r.lerr = lerr;
    r.accept = errsum <= 0x1.a000000000002p+3;
    r.above_floor = errsum > 0x1.172c417c771f0p-23;
    const double g = sgm::exp_core(beta * lfacold - expo1 * lerr);
    r.hnew_accept = h * sgm::min_sel(10.0, sgm::max_sel(0.2, safe * g));
    r.hnew_reject = 0.0;
#else
    const double err = sqrt(sgm::div_nr(errsum, (double)NX));
    const double lerr = sgm::log_pos(err);
    r.accept = err <= 1.0;
    r.above_floor = err > 1.0e-4;
    r.lerr = lerr;
#if SG_MERGED_POW
    // fac11 / facold^beta = err^expo1 / facold^beta with ONE exp: exp(expo1 ln err - beta ln facold).  The two powers are
    // transcendental results defined to an ulp each, and so is their quotient; everything around it stays exact
    // arithmetic.  The rejected-step formula needs err^expo1 alone: step_reject, on the (rare) rejected steps only.
    double fac = sgm::exp_core(expo1 * lerr - beta * lfacold);
    fac = sgm::max_sel(facc2, sgm::min_sel(facc1, sgm::div_nr(fac, safe)));
    r.hnew_accept = sgm::div_nr(h, fac);
    r.hnew_reject = 0.0;
#else
    const double fac11 = sgm::pow_of_log(lerr, expo1);
    double fac = sgm::div_nr(fac11, sgm::pow_of_log(lfacold, beta));
    fac = sgm::max_sel(facc2, sgm::min_sel(facc1, sgm::div_nr(fac, safe)));
    r.hnew_accept = sgm::div_nr(h, fac);
    r.hnew_reject = sgm::div_nr(h, sgm::min_sel(facc1, sgm::div_nr(fac11, safe)));
#endif
#endif
    return r;
}

// hnew after a rejected step: h / min(facc1, err^expo1 / safe) (DOPCOR).
static __device__ __noinline__ double step_reject(double lerr, double h)
{
    const double expo1 = 0.2 - 0.04 * 0.75, safe = 0.9, facc1 = 1.0 / 0.2;
#if SG_FAST_CONTROL
    return h * sgm::max_sel(0.2, safe * sgm::exp_core(-expo1 * lerr));       // h / min(5, err^expo1 / safe)
#else
    const double fac11 = sgm::pow_of_log(lerr, expo1);
    return sgm::div_nr(h, sgm::min_sel(facc1, sgm::div_nr(fac11, safe)));
#endif
}

// One call of ode('dopri5').integrate(t + 1): k1 = f(y), HINIT (initial step size guess) unless a step
// size is carried over, then the adaptive step loop.  Returns idid (1 ok, -2 nmax, -3 step too small).
// h_carry: WORK(7) of the Fortran (kept across calls by scipy 1.6.3, zero in scipy >= 1.7).
//
// Shape of the code (sized by profiles/r01_ubench_fp64_icache.log: DFMA latency 8 cycles, one FP64 warp
// instruction per 2 cycles per sub-partition, instruction fetch collapses once the loop outgrows 32 KB):
//   * a phase machine around ONE inlined copy of the RHS -- every pass evaluates kt = f(yt), then the arm of
//     the current phase consumes kt.  All lanes of a warp walk the phases together (uniform switch).
//   * the stages are never stored.  Hairer's stage sums are evaluated left to right, so each later stage's
//     sum can be advanced as soon as a k arrives:  P_s += a_sj * k_j  for every s > j.  That is the same
//     sequence of roundings as  a_s1*k1 + a_s2*k2 + ...  and leaves no k -> register copies.  P3..P6 live in
//     registers; the two longest-lived sums (stage 7 and the error estimator) live in the thread's
//     shared-memory column together with k1, which is needed again only by HINIT and after a rejection.
//
// DEFER (compile time): mode selects what this lane does in the call -- MINUTE_FRESH a whole minute, MINUTE_IDLE
// nothing (it only takes part in the warp votes), MINUTE_RESUME the rest of a minute put aside earlier (sp, k1 in
// H_K1); with cap > 0 a minute that is about to start its (cap+1)-th step is put aside: sp and H_K1 are filled
// and IDID_SUSPENDED is returned.
template <int BLOCK, bool DEFER = false>
__device__ __forceinline__ int dopri5_minute(const Column<BLOCK> &c, const Minute &m, double (&y)[NX],
                                             double t0, bool carry_h, double &h_carry, StepCount &cnt,
                                             unsigned lanes, int mode = MINUTE_FRESH, int cap = 0,
                                             Suspend *sp = nullptr)
{
    enum { PH_K1 = 1, PH_HINIT = 2, PH_RESUME = 4, PH_S2 = 8, PH_S3 = 16, PH_S4 = 32, PH_S5 = 64, PH_S6 = 128, PH_S7 = 256 };
    double x = t0;
    const double xend = t0 + 1.0;
    const double hmax = fabs(xend - t0);
    double lfacold = c(H_L4);                          // facold = 1e-4
    bool last = false, reject = false;
    int nstep = 0, naccpt = 0, idid = 0, phase = PH_K1;
    double yt[NX], kt[NX], P4[NX], P5[NX], P6[NX];
    double h = 0.0, dnf = 0.0;

    if (DEFER && mode == MINUTE_IDLE) idid = 1;
    if (!DEFER || mode != MINUTE_IDLE) {
#pragma unroll
        for (int i = 0; i < NX; i++) yt[i] = y[i];
        store_vec(c, H_Y, y);
    }
    if (DEFER && mode == MINUTE_RESUME) {
        h = sp->h; x = sp->x; lfacold = sp->lfacold;
        nstep = sp->nstep; naccpt = sp->naccpt; cnt.nrej = sp->nrej; reject = sp->reject != 0;
        phase = PH_RESUME;
    }

    // The loop condition is a warp vote: lanes whose minute is complete wait here instead of running ahead
    // into the next minute on their own (without it ptxas leaves the warp split for the rest of the episode:
    // 12.8 of 32 lanes active per issued instruction, profiles/r01_i1_*).
    // Every arm walks the states two at a time (SG_FOR_PAIRS): one 16-byte shared-memory access per vector and pair;
    // "lo" / "hi" below are the two states of a pair, and the arithmetic per state is unchanged.
#pragma unroll 1
    while (__any_sync(lanes, idid == 0)) {
        if (idid != 0) continue;
        rhs(c, m, yt, kt);
        double2 yy[NV / 2];                                  // y for the arm, fetched ahead of the dispatch
        SG_FOR_PAIRS(i) yy[i / 2] = c.pair(H_Y + i);

        bool begin_step = false;
        // Phase dispatch as a tree of bit tests (the phases are one-hot): a dense switch becomes a jump table, whose
        // constant-bank load and indirect branch cost more than three predictable compares (profiles/r01_s38_*).
        if (phase & (PH_S2 | PH_S3 | PH_S4)) {
            if (phase & PH_S2) {                                          // kt = k2; the sums start here from k1 (shared memory)
            SG_FOR_PAIRS(i) {
                const double2 k2 = c.pair(H_K1 + i), y2 = yy[i / 2];
                {
                    const double k1 = k2.x;
                    yt[i] = y2.x + h * (A31 * k1 + A32 * kt[i]);
                    P4[i] = A41 * k1 + A42 * kt[i];
                    P5[i] = A51 * k1 + A52 * kt[i];
                    P6[i] = A61 * k1 + A62 * kt[i];
                }
                if (i + 1 < NX) {
                    constexpr int j = 0;
                    const int q = (i + 1 < NX ? i + 1 : i) + j;
                    const double k1 = k2.y;
                    yt[q] = y2.y + h * (A31 * k1 + A32 * kt[q]);
                    P4[q] = A41 * k1 + A42 * kt[q];
                    P5[q] = A51 * k1 + A52 * kt[q];
                    P6[q] = A61 * k1 + A62 * kt[q];
                }
            }
            phase = PH_S3;
            } else if (phase & PH_S3) {                                          // kt = k3
            // Register hand-over: P4 is consumed by this stage, so its registers take the stage-7 sum from here on
            // (P4 = A71 k1 + A73 k3 + ..); P5 is consumed by the next stage and takes the error sum there.  Only the
            // error sum's first two terms travel through shared memory (H_PE), for one stage.  Same operations in the
            // same order as with both sums in shared memory: bit-identical, 84 of 98 LDS / STS.128 per step fewer.
            SG_FOR_PAIRS(i) {
                const double2 k2 = c.pair(H_K1 + i), y2 = yy[i / 2];
                double2 pe = make_double2(0.0, 0.0);
                {
                    const double k1 = k2.x;
                    yt[i] = y2.x + h * (P4[i] + A43 * kt[i]);
                    P5[i] = P5[i] + A53 * kt[i];
                    P6[i] = P6[i] + A63 * kt[i];
                    P4[i] = A71 * k1 + A73 * kt[i];
                    pe.x = E1 * k1 + E3 * kt[i];
                }
                if (i + 1 < NX) {
                    const int q = i + 1 < NX ? i + 1 : i;
                    const double k1 = k2.y;
                    yt[q] = y2.y + h * (P4[q] + A43 * kt[q]);
                    P5[q] = P5[q] + A53 * kt[q];
                    P6[q] = P6[q] + A63 * kt[q];
                    P4[q] = A71 * k1 + A73 * kt[q];
                    pe.y = E1 * k1 + E3 * kt[q];
                }
                c.set_pair(H_PE + i, pe);
            }
            phase = PH_S4;
            } else {                                          // kt = k4
            SG_FOR_PAIRS(i) {
                const double2 y2 = yy[i / 2], pe = c.pair(H_PE + i);
                {
                    yt[i] = y2.x + h * (P5[i] + A54 * kt[i]);
                    P6[i] = P6[i] + A64 * kt[i];
                    P4[i] = P4[i] + A74 * kt[i];
                    P5[i] = pe.x + E4 * kt[i];
                }
                if (i + 1 < NX) {
                    const int q = i + 1 < NX ? i + 1 : i;
                    yt[q] = y2.y + h * (P5[q] + A54 * kt[q]);
                    P6[q] = P6[q] + A64 * kt[q];
                    P4[q] = P4[q] + A74 * kt[q];
                    P5[q] = pe.y + E4 * kt[q];
                }
            }
            phase = PH_S5;
            }
        } else if (phase & (PH_S5 | PH_S6 | PH_S7)) {
            if (phase & PH_S5) {                                          // kt = k5
            SG_FOR_PAIRS(i) {
                const double2 y2 = yy[i / 2];
                {
                    yt[i] = y2.x + h * (P6[i] + A65 * kt[i]);
                    P4[i] = P4[i] + A75 * kt[i];
                    P5[i] = P5[i] + E5 * kt[i];
                }
                if (i + 1 < NX) {
                    const int q = i + 1 < NX ? i + 1 : i;
                    yt[q] = y2.y + h * (P6[q] + A65 * kt[q]);
                    P4[q] = P4[q] + A75 * kt[q];
                    P5[q] = P5[q] + E5 * kt[q];
                }
            }
            phase = PH_S6;
            } else if (phase & PH_S6) {                                          // kt = k6; yt becomes the 5th-order solution
            SG_FOR_PAIRS(i) {
                const double2 y2 = yy[i / 2];
                {
                    yt[i] = y2.x + h * (P4[i] + A76 * kt[i]);
                    P5[i] = P5[i] + E6 * kt[i];
                }
                if (i + 1 < NX) {
                    const int q = i + 1 < NX ? i + 1 : i;
                    yt[q] = y2.y + h * (P4[q] + A76 * kt[q]);
                    P5[q] = P5[q] + E6 * kt[q];
                }
            }
            phase = PH_S7;
            } else {                                           // kt = k7 = f(x + h, y1): error estimate + control
            double err = 0.0;
#if SG_FAST_ERRNORM == 2
            // the 13 reciprocals advance together, one refinement level at a time: 13 independent chains for the
            // scheduler instead of one quotient after the other (the sums P4..P6 are dead here, registers are free)
            double sk[NX], rr[NX], ee[NX];
            SG_FOR_PAIRS(i) {
                const double2 y2 = yy[i / 2];
                // max(|y|, |ynew|): select the operand, take the magnitude inside the product (an operand modifier)
                sk[i] = ATOL + RTOL * fabs(fabs(y2.x) > fabs(yt[i]) ? y2.x : yt[i]);
                if (i + 1 < NX) {
                    const int q = i + 1 < NX ? i + 1 : i;
                    sk[q] = ATOL + RTOL * fabs(fabs(y2.y) > fabs(yt[q]) ? y2.y : yt[q]);
                }
            }
#pragma unroll
            for (int i = 0; i < NX; i++) rr[i] = sgm::rcp_seed(sk[i]);
#pragma unroll
            for (int i = 0; i < NX; i++) ee[i] = fma(rr[i], -sk[i], 1.0);
#pragma unroll
            for (int i = 0; i < NX; i++) ee[i] = fma(ee[i], ee[i], ee[i]);
#pragma unroll
            for (int i = 0; i < NX; i++) rr[i] = fma(rr[i], ee[i], rr[i]);
#pragma unroll
            for (int i = 0; i < NX; i++) {
                const double a = ((P5[i] + E7 * kt[i]) * h) * rr[i];
                err = err + a * a;
            }
#else
            SG_FOR_PAIRS(i) {
                const double2 y2 = yy[i / 2];
                {
                    const double e = (P5[i] + E7 * kt[i]) * h;
                    const double sk = ATOL + RTOL * sgm::max_sel(fabs(y2.x), fabs(yt[i]));
                    const double a = SG_ERR_QUOT(e, sk);
                    err = err + a * a;
                }
                if (i + 1 < NX) {
                    const int q = i + 1 < NX ? i + 1 : i;
                    const double e = (P5[q] + E7 * kt[q]) * h;
                    const double sk = ATOL + RTOL * sgm::max_sel(fabs(y2.y), fabs(yt[q]));
                    const double a = SG_ERR_QUOT(e, sk);
                    err = err + a * a;
                }
            }
#endif
            const StepControl sc = step_control(err, lfacold, h);
            double hnew = sc.hnew_accept;
            if (sc.accept) {
                lfacold = sc.above_floor ? sc.lerr : c(H_L4);        // facold = max(err, 1e-4)
                naccpt++;
                store_vec(c, H_Y, yt);                       // kt is k1 of the next step (FSAL)
                x = x + h;
                if (last) {
                    idid = 1;
                } else {
                    if (fabs(hnew) > hmax) hnew = hmax;
                    if (reject) hnew = sgm::min_sel(fabs(hnew), fabs(h));
                    reject = false;
                    begin_step = true;
                }
            } else {
#if SG_MERGED_POW
                hnew = step_reject(sc.lerr, h);
#else
                hnew = sc.hnew_reject;
#endif
                reject = true;
                if (naccpt >= 1) cnt.nrej++;
                last = false;
                load_vec(c, H_K1, kt);
                begin_step = true;
            }
            h = hnew;
            }
        } else {
            if (phase & PH_K1) {                                        // kt = k1 = f(x, y)
            h = carry_h ? h_carry : 0.0;
            if (h == 0.0) {                                  // HINIT, first half
                store_vec(c, H_K1, kt);
                const HinitGuess g = hinit_first<BLOCK>(c.base, hmax);
                h = g.h;
                dnf = g.dnf;
                SG_FOR_PAIRS(i) {
                    const double2 y2 = yy[i / 2];
                    yt[i] = y2.x + h * kt[i];
                    if (i + 1 < NX) yt[i + 1 < NX ? i + 1 : i] = y2.y + h * kt[i + 1 < NX ? i + 1 : i];
                }
                phase = PH_HINIT;
            } else {
                begin_step = true;
            }
            } else if (phase & PH_HINIT) {                                     // kt = f(x + h, y + h*k1)
            store_vec(c, H_P7, kt);
            h = hinit_second<BLOCK>(c.base, h, dnf, hmax);
            load_vec(c, H_K1, kt);
            begin_step = true;
            } else {                                      // (the RHS of this pass is not used) k1 comes back from H_K1
            load_vec(c, H_K1, kt);
            begin_step = true;
            }
        }
        if (begin_step) {                                    // kt = k1 of the step that starts here
            if (DEFER && cap > 0 && nstep >= cap) {          // put the rest of this minute aside
                store_vec(c, H_K1, kt);
                sp->h = h; sp->x = x; sp->lfacold = lfacold;
                sp->nstep = nstep; sp->naccpt = naccpt; sp->nrej = cnt.nrej; sp->reject = reject ? 1 : 0;
                idid = IDID_SUSPENDED;
            }
            else if (nstep > NMAX) idid = -2;
            else if (0.1 * fabs(h) <= fabs(x) * UROUND) idid = -3;
            else {
                if ((x + 1.01 * h - xend) > 0.0) { h = xend - x; last = true; }
                nstep++;
                const double ha = h * A21;
                store_vec(c, H_K1, kt);
                SG_FOR_PAIRS(i) {
                    const double2 y2 = c.pair(H_Y + i);
                    yt[i] = y2.x + ha * kt[i];
                    if (i + 1 < NX) yt[i + 1 < NX ? i + 1 : i] = y2.y + ha * kt[i + 1 < NX ? i + 1 : i];
                }
                phase = PH_S2;
            }
        }
    }
    if (DEFER && mode == MINUTE_IDLE) return 1;
    load_vec(c, H_Y, y);
    if (DEFER && idid == IDID_SUSPENDED) return idid;
    if (carry_h) h_carry = h;
    cnt.nstep += nstep;
    return idid;
}

}  // namespace sgk
