"""TEST INFRASTRUCTURE ONLY -- ctypes front-end of the CPU oracle (oracle/sg_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module; the product package never does.
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, '_ref', 'libsg_oracle.so')
_TABLES = os.path.join(_HERE, '..', 'simglucose_b200', 'params', 'tables.json')

HOT_PARAMS = ['BW', 'u2ss', 'kmax', 'kmin', 'b', 'd', 'kabs', 'f', 'kp1', 'kp2', 'kp3', 'Fsnc',
              'ke1', 'ke2', 'k1', 'k2', 'Vm0', 'Vmx', 'Km0', 'm1', 'm2', 'm4', 'm30', 'ka1',
              'ka2', 'Vi', 'p2u', 'Ib', 'ki', 'kd', 'ksc', 'Vg']
TRACE_COLS = ['BG', 'CGM', 'CHO', 'insulin', 'LBGI', 'HBGI', 'Risk', 'obs', 'reward', 'done',
              'basal', 'bolus']
STAT_COLS = ['rows', 'sum_bg', 'min_bg', 'max_bg', 'sum_lbgi', 'sum_hbgi', 'sum_risk',
             'sum_reward', 'lt50', 'lt70', 'inrange', 'gt180', 'gt250', 'first_done',
             'dopri_steps', 'dopri_rejects']
CTRL = {'actions': 0, 'bb': 1, 'pid': 2}


class Config(C.Structure):
    _fields_ = [('sample_time', C.c_int), ('sensor_min', C.c_double), ('sensor_max', C.c_double),
                ('pump', C.c_double * 6), ('controller', C.c_int), ('ctrl', C.c_double * 6),
                ('carry_h', C.c_int), ('t0', C.c_int)]


def build():
    """Compile the oracle (gcc, -ffp-contract=off) into oracle/_ref/."""
    subprocess.run(['make', '-s', '-C', _HERE], check=True)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
        L.sgo_model.argtypes = [dp, C.c_double, C.c_double, dp, C.c_double, C.c_double, dp]
        L.sgo_model.restype = None
        L.sgo_pump.argtypes = [C.c_double] * 4
        L.sgo_pump.restype = C.c_double
        L.sgo_risk.argtypes = [C.c_double, dp]
        L.sgo_risk.restype = None
        L.sgo_run_episode.argtypes = [C.POINTER(Config), dp, dp, C.c_int, ip, dp, C.c_int, dp, dp,
                                      dp, ip, dp, dp]
        L.sgo_run_episode.restype = C.c_int
        L.sgo_run_batch.argtypes = [C.POINTER(Config), C.c_int, dp, dp, C.c_int, ip, dp, C.c_int,
                                    dp, dp, dp, C.c_int]
        L.sgo_run_batch.restype = C.c_int
        L.sgo_max_threads.restype = C.c_int
        L.sgo_dopri5_minute.argtypes = [dp, dp, dp, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int, dp, ip]
        L.sgo_dopri5_minute.restype = C.c_int
        _lib = L
    return _lib


_tables = None


def tables():
    global _tables
    if _tables is None:
        with open(_TABLES) as f:
            _tables = json.load(f)
    return _tables


def _table(name):
    t = tables()[name]
    return t['columns'], t['rows']


def patient_names():
    cols, rows = _table('vpatient_params')
    return [r[0] for r in rows]


def patient_params(row):
    """(hot params[32], x0[13]) of CSV row `row` (0-based)."""
    cols, rows = _table('vpatient_params')
    r = rows[row]
    p = np.array([r[cols.index(c)] for c in HOT_PARAMS], dtype=np.float64)
    x0 = np.array(r[2:15], dtype=np.float64)
    return p, x0


def quest(row):
    cols, rows = _table('Quest')
    r = rows[row]
    return float(r[cols.index('CR')]), float(r[cols.index('CF')])


def sensor(name='Dexcom'):
    cols, rows = _table('sensor_params')
    r = [x for x in rows if x[0] == name][0]
    return dict(zip(cols, r))


def pump(name='Insulet'):
    cols, rows = _table('pump_params')
    r = [x for x in rows if x[0] == name][0]
    return dict(zip(cols, r))


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int))


def make_config(controller='bb', ctrl=(140.0, 1.0, 1.0), sensor_name='Dexcom', pump_name='Insulet',
                carry_h=0, t0=0):
    s, pm = sensor(sensor_name), pump(pump_name)
    cfg = Config()
    cfg.sample_time = int(s['sample_time'])
    cfg.sensor_min, cfg.sensor_max = s['min'], s['max']
    cfg.pump[:] = [pm['min_basal'], pm['max_basal'], pm['inc_basal'], pm['min_bolus'],
                   pm['max_bolus'], pm['inc_bolus']]
    cfg.controller = CTRL[controller]
    c = list(ctrl) + [0.0] * (6 - len(ctrl))
    cfg.ctrl[:] = c
    cfg.carry_h = int(carry_h)
    cfg.t0 = int(t0)
    return cfg


def model(x, cho, insulin, params, last_Qsto, last_foodtaken):
    x = np.ascontiguousarray(x, dtype=np.float64)
    p = np.ascontiguousarray(params, dtype=np.float64)
    out = np.zeros(13)
    lib().sgo_model(_dp(x), cho, insulin, _dp(p), last_Qsto, last_foodtaken, _dp(out))
    return out


def pump_quantise(amount, vmin, vmax, inc):
    return lib().sgo_pump(amount, vmin, vmax, inc)


def risk(bg):
    out = np.zeros(3)
    lib().sgo_risk(float(bg), _dp(out))
    return tuple(out)


def run_episode(params, x0, n_steps, meal_minute, meal_grams, noise, controller='bb',
                ctrl=(140.0, 1.0, 1.0), actions=None, sensor_name='Dexcom', pump_name='Insulet',
                carry_h=0, t0=0):
    """One closed-loop episode; returns dict(trace columns..., counters, stats, final_state, status).
    t0: T1DPatient(t0=...); meal_minute counts from scenario.start_time."""
    cfg = make_config(controller, ctrl, sensor_name, pump_name, carry_h, t0)
    st = cfg.sample_time
    p = np.ascontiguousarray(params, dtype=np.float64)
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    mm = np.ascontiguousarray(meal_minute, dtype=np.int32)
    mg = np.ascontiguousarray(meal_grams, dtype=np.float64)
    nz = np.ascontiguousarray(noise, dtype=np.float64)
    assert nz.size >= n_steps + 2
    act = None
    if controller == 'actions':
        act = np.ascontiguousarray(actions, dtype=np.float64).reshape(n_steps, 2)
    trace = np.zeros((n_steps + 1, len(TRACE_COLS)))
    counters = np.zeros((n_steps * st, 4), dtype=np.int32)
    stats = np.zeros(len(STAT_COLS))
    final = np.zeros(13)
    rc = lib().sgo_run_episode(C.byref(cfg), _dp(p), _dp(x0), n_steps, _ip(mm), _dp(mg), mm.size,
                               _dp(nz), _dp(act) if act is not None else None, _dp(trace),
                               _ip(counters), _dp(stats), _dp(final))
    out = {c: trace[:, i].copy() for i, c in enumerate(TRACE_COLS)}
    out.update(counters=counters, stats=dict(zip(STAT_COLS, stats)), final_state=final, status=rc)
    return out


def run_batch(params, x0, n_steps, meal_minute, meal_grams, noise, controller='bb',
              ctrl=(140.0, 1.0, 1.0), ctrl_rows=None, sensor_name='Dexcom', pump_name='Insulet',
              carry_h=0, n_threads=0):
    """Thread-pool batch of episodes (episode-major arrays); returns (stats[E,16], n_failed)."""
    cfg = make_config(controller, ctrl, sensor_name, pump_name, carry_h)
    p = np.ascontiguousarray(params, dtype=np.float64)
    E = p.shape[0]
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    mm = np.ascontiguousarray(meal_minute, dtype=np.int32).reshape(E, -1)
    mg = np.ascontiguousarray(meal_grams, dtype=np.float64).reshape(E, -1)
    nz = np.ascontiguousarray(noise, dtype=np.float64).reshape(E, n_steps + 2)
    cr = None if ctrl_rows is None else np.ascontiguousarray(ctrl_rows, dtype=np.float64).reshape(E, 6)
    stats = np.zeros((E, len(STAT_COLS)))
    failed = lib().sgo_run_batch(C.byref(cfg), E, _dp(p), _dp(x0), n_steps, _ip(mm), _dp(mg),
                                 mm.shape[1], _dp(nz), _dp(cr) if cr is not None else None,
                                 _dp(stats), int(n_threads))
    return stats, failed


def max_threads():
    return lib().sgo_max_threads()
