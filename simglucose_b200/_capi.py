"""ctypes binding of include/simglucose_b200.h (the C-ABI of the CUDA library).

There is no CPU fallback: if the shared library is missing or does not load, importing a symbol from
it raises.  `load()` never builds anything; run `python -m simglucose_b200.build` (or
`__graft_entry__.build()`) first.
"""
import ctypes as C
import os

from .build import LIB_PATH

SG_ABI_VERSION = 1
NX = 13
NPARAM = 32
NTRACE = 12
NSTAT = 16
NCTRL = 4
NPOUT = 21
ND, NI = 24, 5

PARAMS = ['BW', 'u2ss', 'kmax', 'kmin', 'b', 'd', 'kabs', 'f', 'kp1', 'kp2', 'kp3', 'Fsnc', 'ke1', 'ke2',
          'k1', 'k2', 'Vm0', 'Vmx', 'Km0', 'm1', 'm2', 'm4', 'm30', 'ka1', 'ka2', 'Vi', 'p2u', 'Ib', 'ki',
          'kd', 'ksc', 'Vg']
TRACE_COLS = ['BG', 'CGM', 'CHO', 'insulin', 'LBGI', 'HBGI', 'Risk', 'obs', 'reward', 'done', 'basal', 'bolus']
STAT_COLS = ['rows', 'sum_bg', 'min_bg', 'max_bg', 'sum_lbgi', 'sum_hbgi', 'sum_risk', 'sum_reward', 'lt50',
             'lt70', 'inrange', 'gt180', 'gt250', 'first_done', 'dopri_steps', 'dopri_rejects']
CTRL_ACTIONS, CTRL_BB, CTRL_PID, CTRL_RANDOM = 0, 1, 2, 3
FLAG_CARRY_H, FLAG_FMA, FLAG_FP32 = 1, 2, 8
STATUS_STREAM_EXHAUSTED = -5
REDUCE_STATIC_MAX_EPISODES = 16777216     # SG_REDUCE_STATIC_MAX_EPISODES

# every symbol include/simglucose_b200.h declares
SYMBOLS = ['sg_version', 'sg_last_error', 'sg_create', 'sg_destroy', 'sg_set_hardware', 'sg_load_params',
           'sg_reset', 'sg_step', 'sg_run', 'sg_get_state', 'sg_get_stats', 'sg_get_status',
           'sg_reduce_stats', 'sg_num_episodes', 'sg_launch_count',
           'sg_step_meals', 'sg_rebind_streams', 'sg_set_t0', 'sg_set_step_cap', 'sg_set_coherence', 'sg_set_sort_mode', 'sg_get_launch_order',
           'sg_reset_masked', 'sg_patient_step', 'sg_model', 'sg_pump',
           'sg_risk_index', 'sg_risk_window', 'sg_cgm_measure', 'sg_get_episode_state', 'sg_set_episode_state', 'sg_math_probe',
           'sg_report_stats', 'sg_cvga_zones', 'sg_reduce_stats_static',
           'sg_gen_noise', 'sg_gen_scenarios', 'sg_gen_perturbed_params', 'sg_gen_uniform', 'sg_probe_dfma',
           'sg_gen_noise_keyed', 'sg_gen_scenarios_keyed', 'sg_gen_episode_starts', 'sg_vec_finish', 'sg_vec_restart',
           'sg_gen_noise_listed', 'sg_gen_scenarios_listed']


class Hardware(C.Structure):
    _fields_ = [('sample_time', C.c_int32), ('sensor_min', C.c_double), ('sensor_max', C.c_double),
                ('pump', C.c_double * 6)]


class SgError(RuntimeError):
    pass


_lib = None


def load():
    """dlopen the in-tree CUDA library and declare the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SgError('CUDA library %s is missing: run `python -m simglucose_b200.build` '
                      '(there is no CPU fallback)' % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, dp, ip = C.c_void_p, C.c_void_p, C.c_void_p      # device pointers travel as integers
    L.sg_version.restype = C.c_int
    L.sg_last_error.restype = C.c_char_p
    L.sg_create.argtypes = [C.c_int, C.c_int64, C.c_uint32, C.POINTER(C.c_void_p)]
    L.sg_destroy.argtypes = [vp]
    L.sg_set_hardware.argtypes = [vp, C.POINTER(Hardware)]
    L.sg_load_params.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int32]
    L.sg_reset.argtypes = [vp, C.c_int64, ip, dp, dp, ip, dp, C.c_int32, dp, C.c_int64, dp, vp]
    L.sg_step.argtypes = [vp, dp, dp, dp, vp]
    L.sg_run.argtypes = [vp, C.c_int32, C.POINTER(C.c_double), dp, dp, C.c_int32, dp, ip, vp]
    L.sg_step_meals.argtypes = [vp, dp, dp, dp, dp, vp]
    L.sg_rebind_streams.argtypes = [vp, ip, dp, C.c_int32, dp, C.c_int64]
    L.sg_set_t0.argtypes = [vp, C.c_int32]
    L.sg_set_step_cap.argtypes = [vp, C.c_int32]
    L.sg_set_coherence.argtypes = [vp, C.c_int32]
    L.sg_set_sort_mode.argtypes = [vp, C.c_int32]
    L.sg_get_launch_order.argtypes = [vp, ip, vp]
    L.sg_reset_masked.argtypes = [vp, ip, dp, dp, vp]
    L.sg_patient_step.argtypes = [vp, dp, dp, dp, vp]
    L.sg_model.argtypes = [vp, C.c_int64, dp, dp, dp, dp, dp, ip, dp, dp, vp]
    L.sg_pump.argtypes = [dp, dp, C.c_int64, C.c_double, C.c_double, C.c_double, vp]
    L.sg_risk_index.argtypes = [dp, dp, C.c_int64, vp]
    L.sg_risk_window.argtypes = [dp, dp, C.c_int64, vp]
    L.sg_cgm_measure.argtypes = [dp, dp, dp, C.c_int64, C.c_double, C.c_double, vp]
    L.sg_get_episode_state.argtypes = [vp, dp, ip, vp]
    L.sg_set_episode_state.argtypes = [vp, dp, ip, C.c_int64, vp]
    L.sg_math_probe.argtypes = [C.c_int32, dp, dp, dp, C.c_int64, vp]
    L.sg_report_stats.argtypes = [dp, C.c_int64, C.c_int64, C.c_int64, C.c_int32, dp, dp, dp, dp, vp]
    L.sg_cvga_zones.argtypes = [dp, C.c_int64, dp, vp]
    L.sg_get_state.argtypes = [vp, dp, vp]
    L.sg_get_stats.argtypes = [vp, dp, vp]
    L.sg_get_status.argtypes = [vp, ip, vp]
    L.sg_reduce_stats.argtypes = [vp, dp, vp]
    L.sg_reduce_stats_static.argtypes = [dp, C.c_int64, C.c_int64, dp, vp]
    L.sg_num_episodes.argtypes = [vp]
    L.sg_num_episodes.restype = C.c_int64
    L.sg_launch_count.argtypes = [vp]
    L.sg_launch_count.restype = C.c_int64
    if True:
        L.sg_gen_noise.argtypes = [dp, C.c_int64, C.c_int64, C.c_int32, dp, C.POINTER(C.c_double), C.c_uint64, C.c_int64, C.c_int64, vp]
        L.sg_gen_scenarios.argtypes = [ip, dp, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_uint64, C.c_int64, C.c_int64, vp]
        L.sg_gen_perturbed_params.argtypes = [dp, dp, ip, C.c_int64, C.c_double, C.c_uint64, C.c_int64, vp]
        L.sg_probe_dfma.argtypes = [dp, C.c_int32, C.c_int32, C.c_int64, vp]
        L.sg_gen_uniform.argtypes = [dp, C.c_int64, C.c_double, C.c_double, C.c_uint64, C.c_uint64, vp]
        L.sg_gen_noise_keyed.argtypes = [dp, C.c_int64, C.c_int64, C.c_int32, dp, C.POINTER(C.c_double), C.c_uint64, ip, ip, vp]
        L.sg_gen_scenarios_keyed.argtypes = [ip, dp, C.c_int64, C.c_int32, ip, C.c_int32, C.c_uint64, ip, ip, vp]
        L.sg_gen_episode_starts.argtypes = [ip, dp, dp, ip, C.c_int64, C.c_int32, C.c_uint64, ip, ip, vp]
        L.sg_vec_finish.argtypes = [dp, C.c_int64, C.c_int32, C.c_int32, ip, C.c_int32, ip, C.c_int64, C.c_int64, C.c_int32,
                                    ip, ip, dp, ip, ip, ip, ip, vp]
        L.sg_gen_noise_listed.argtypes = [dp, C.c_int64, C.c_int64, C.c_int32, dp, C.POINTER(C.c_double), C.c_uint64, ip, ip, ip, vp]
        L.sg_gen_scenarios_listed.argtypes = [ip, dp, C.c_int64, C.c_int32, ip, C.c_int32, C.c_uint64, ip, ip, ip, vp]
        L.sg_vec_restart.argtypes = [ip, C.c_int64, ip, ip, dp, dp, dp, vp]
    for name in SYMBOLS:
        f = getattr(L, name, None)
        if f is not None and f.restype is C.c_int and name != 'sg_version':
            f.restype = C.c_int
    if L.sg_version() != SG_ABI_VERSION:
        raise SgError('ABI version mismatch: library %d, binding %d' % (L.sg_version(), SG_ABI_VERSION))
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise SgError('simglucose_b200 C-ABI error %d: %s' % (rc, load().sg_last_error().decode()))
