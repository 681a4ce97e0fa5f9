"""Per-episode input streams: CGM noise, meal schedules, initial states, gym seeding.

Two families:
  * host generators that replay the reference's numpy MT19937 streams for a given seed (used by the
    drop-in T1DSimEnv / gym facade and by the parity tests), and
  * wrappers of the CUDA generator kernels (csrc/sg_generators.cu: counter-based Philox, same
    distributions) for populations too large to generate on the host.

Reference: simglucose/sensor/noise_gen.py:15-97, simglucose/simulation/scenario_gen.py:10-75,
simglucose/simulation/scenario.py:24-59, simglucose/patient/t1dpatient.py:247-270,
simglucose/envs/simglucose_gym_env.py:58-73 (+ gym==0.9.4 utils/seeding.py, third party).
"""
import ctypes as C
import hashlib
import math
import struct
from datetime import datetime, timedelta

import numpy as np

from . import params as P

MDL_SAMPLE_TIME = 15      # noise_gen.py:17
PRECOMPUTE = 10           # noise_gen.py:16


# ---------------------------------------------------------------------------------------------
# CGM noise
# ---------------------------------------------------------------------------------------------
def _not_a_knot_cubic(xk, yk, xq):
    """Cubic spline through (xk, yk) with not-a-knot ends, evaluated at xq.  Same interpolant as
    scipy.interpolate.interp1d(kind='cubic') (noise_gen.py:45); numpy only."""
    xk = np.asarray(xk, float)
    yk = np.asarray(yk, float)
    n = len(xk)
    h = np.diff(xk)
    A = np.zeros((n, n))
    rhs = np.zeros(n)
    # unknowns: second derivatives M_i
    for i in range(1, n - 1):
        A[i, i - 1] = h[i - 1]
        A[i, i] = 2.0 * (h[i - 1] + h[i])
        A[i, i + 1] = h[i]
        rhs[i] = 6.0 * ((yk[i + 1] - yk[i]) / h[i] - (yk[i] - yk[i - 1]) / h[i - 1])
    # not-a-knot: third derivative continuous at xk[1] and xk[-2]
    A[0, 0], A[0, 1], A[0, 2] = h[1], -(h[0] + h[1]), h[0]
    A[-1, -3], A[-1, -2], A[-1, -1] = h[-1], -(h[-2] + h[-1]), h[-2]
    M = np.linalg.solve(A, rhs)
    xq = np.asarray(xq, float)
    j = np.clip(np.searchsorted(xk, xq, side='right') - 1, 0, n - 2)
    hj = h[j]
    a = (xk[j + 1] - xq) / hj
    b = (xq - xk[j]) / hj
    return a * yk[j] + b * yk[j + 1] + ((a ** 3 - a) * M[j] + (b ** 3 - b) * M[j + 1]) * hj * hj / 6.0


def noise_window_samples(sample_time):
    """Samples produced per 150-minute window incl. the carried first one (noise_gen.py:40-42)."""
    return int(math.floor(PRECOMPUTE * MDL_SAMPLE_TIME / sample_time)) + 1


def interp_matrix(sample_time):
    """W [nsample, 11]: the cubic interp1d of noise_gen.py:38-46 over fixed uniform knots is a constant
    linear map from the eleven 15-minute knots to the sensor grid."""
    nsample = noise_window_samples(sample_time)
    t15 = np.arange(PRECOMPUTE + 1) * float(MDL_SAMPLE_TIME)
    t = np.arange(nsample) * float(sample_time)
    W = np.zeros((nsample, PRECOMPUTE + 1))
    for j in range(PRECOMPUTE + 1):
        e = np.zeros(PRECOMPUTE + 1)
        e[j] = 1.0
        W[:, j] = _not_a_knot_cubic(t15, e, t)
    return W


def cgm_noise_stream(sensor, seed, n):
    """First n samples of CGMNoise(params, seed) (noise_gen.py:15-69), replaying numpy's MT19937.

    Uses scipy's interp1d per window exactly like the reference when scipy is importable (bit-identical
    stream); otherwise the equivalent constant matrix (differs by ~1e-14)."""
    s = P.sensor_params(sensor) if isinstance(sensor, str) else sensor
    st = s['sample_time']
    nsample = noise_window_samples(st)
    per_window = nsample - 1
    n_windows = max(1, -(-int(n) // per_window))
    rg = np.random.RandomState(seed)
    z = rg.randn(1 + PRECOMPUTE * n_windows)                 # noise15_iter draws, in order
    e = np.empty_like(z)
    e[0] = z[0]                                              # noise_gen.py:84-85
    pacf = s['PACF']
    for k in range(1, len(z)):
        e[k] = pacf * (e[k - 1] + z[k])                      # noise_gen.py:86-88
    eps = s['xi'] + s['lambda'] * np.sinh((e - s['gamma']) / s['delta'])   # noise_gen.py:11-12
    try:
        from scipy.interpolate import interp1d
    except Exception:                                        # pragma: no cover
        interp1d = None
    t15 = np.array(range(0, PRECOMPUTE + 1)) * MDL_SAMPLE_TIME
    t = np.array(range(0, nsample)) * st
    W = None if interp1d is not None else interp_matrix(st)
    # all windows at once: knots [11, n_windows]; scipy's spline solve treats the columns independently, and the result
    # equals the reference's one interp1d per window bit for bit (checked by tests/test_host_logic.py against streams
    # recorded from the reference)
    idx = np.arange(PRECOMPUTE + 1)[:, None] + PRECOMPUTE * np.arange(n_windows)[None, :]
    knots = eps[idx]
    vals = interp1d(t15, knots, kind='cubic', axis=0)(t) if interp1d is not None else W @ knots     # [nsample, n_windows]
    return vals[1:].T.reshape(-1)[:int(n)]


# ---------------------------------------------------------------------------------------------
# Scenarios -> meal tables (episode minute, grams)
# ---------------------------------------------------------------------------------------------
_PROB = [0.95, 0.3, 0.95, 0.3, 0.95, 0.3]                      # scenario_gen.py:38
_T_LB = np.array([5, 9, 10, 14, 16, 20]) * 60
_T_UB = np.array([9, 10, 14, 16, 20, 23]) * 60
_T_MU = np.array([7, 9.5, 12, 15, 18, 21.5]) * 60
_T_SD = np.array([60, 30, 60, 30, 60, 30])
_A_MU = [45, 10, 70, 10, 80, 10]
_A_SD = [10, 5, 10, 5, 10, 5]


def _create_day(rg):
    """RandomScenario.create_scenario (scenario_gen.py:33-60): -> (times [min of day], grams)."""
    from scipy.stats import truncnorm
    times, amounts = [], []
    for p, tlb, tub, tbar, tsd, mbar, msd in zip(_PROB, _T_LB, _T_UB, _T_MU, _T_SD, _A_MU, _A_SD):
        if rg.rand() < p:
            tmeal = np.round(truncnorm.rvs(a=(tlb - tbar) / tsd, b=(tub - tbar) / tsd, loc=tbar, scale=tsd,
                                           random_state=rg))
            times.append(float(tmeal))
            amounts.append(max(round(rg.normal(mbar, msd)), 0))
    return times, amounts


def random_scenario_table(seed, start_minute_of_day, n_minutes):
    """Meal table of RandomScenario(start_time, seed) as the env queries it once per episode minute
    (env.py:50): a schedule is drawn at reset() and again whenever the queried time is midnight
    (scenario_gen.py:20-22, 62-64).  Returns (minute int32[], grams float64[]) in firing order."""
    rg = np.random.RandomState(seed)
    times, amounts = _create_day(rg)
    mm, gg = [], []
    for m in range(int(n_minutes)):
        tod = (start_minute_of_day + m) % 1440
        if tod == 0:
            times, amounts = _create_day(rg)
        if float(tod) in times:
            g = amounts[times.index(float(tod))]
            if g != 0:
                mm.append(m)
                gg.append(float(g))
    return np.array(mm, dtype=np.int32), np.array(gg, dtype=np.float64)


def custom_scenario_table(scenario, start_time, n_minutes=None):
    """CustomScenario (scenario.py:24-59): list of (time, grams); a number is HOURS AFTER start_time,
    a timedelta an offset from start_time, a datetime absolute.  First match wins on duplicates."""
    mm, gg, seen = [], [], set()
    for t, g in scenario:
        if isinstance(t, (int, float)):
            when = start_time + timedelta(minutes=round(t * 60.0))
        elif isinstance(t, timedelta):
            t_sec = t.total_seconds()
            when = start_time + timedelta(minutes=round(t_sec / 60.0))
        elif isinstance(t, datetime):
            when = t
        else:
            raise ValueError('Expect time to be int, float, timedelta, datetime')
        delta = (when - start_time).total_seconds() / 60.0
        if delta != int(delta) or delta < 0 or int(delta) in seen:
            continue
        if n_minutes is not None and delta >= n_minutes:
            continue
        seen.add(int(delta))
        if g != 0:
            mm.append(int(delta))
            gg.append(float(g))
    order = np.argsort(np.array(mm, dtype=np.int64), kind='stable') if mm else []
    return (np.array(mm, dtype=np.int32)[order] if mm else np.zeros(0, np.int32),
            np.array(gg, dtype=np.float64)[order] if mm else np.zeros(0, np.float64))


def pad_tables(tables, min_width=1):
    """[(minute[], grams[]), ...] -> SoA (meal_minute [M, E] int32 with -1 padding, meal_grams [M, E])."""
    M = max(min_width, max((len(t[0]) for t in tables), default=0))
    mm = np.full((M, len(tables)), -1, dtype=np.int32)
    mg = np.zeros((M, len(tables)), dtype=np.float64)
    for e, (m, g) in enumerate(tables):
        mm[:len(m), e] = m
        mg[:len(g), e] = g
    return mm, mg


# ---------------------------------------------------------------------------------------------
# Initial state and gym seeding
# ---------------------------------------------------------------------------------------------
def random_init_state(x0, seed):
    """T1DPatient.reset with random_init_bg=True (t1dpatient.py:256-270): x[3], x[4], x[12] drawn from
    multivariate_normal(mean, diag(0.1*mean)) with RandomState(seed)."""
    x = np.array(x0, dtype=np.float64).copy()
    rs = np.random.RandomState(seed)
    mean = [1.0 * x[3], 1.0 * x[4], 1.0 * x[12]]
    cov = np.diag([0.1 * x[3], 0.1 * x[4], 0.1 * x[12]])
    bg = rs.multivariate_normal(mean, cov)
    x[3], x[4], x[12] = 1.0 * bg[0], 1.0 * bg[1], 1.0 * bg[2]
    return x


def _bigint_from_bytes(b):
    sizeof_int = 4
    padding = sizeof_int - len(b) % sizeof_int
    b += b'\0' * padding
    int_count = int(len(b) / sizeof_int)
    unpacked = struct.unpack('{}I'.format(int_count), b)
    accum = 0
    for i, val in enumerate(unpacked):
        accum += 2 ** (sizeof_int * 8 * i) * val
    return accum


def _int_list_from_bigint(bigint):
    if bigint < 0:
        raise ValueError('Seed must be non-negative, not {}'.format(bigint))
    elif bigint == 0:
        return [0]
    ints = []
    while bigint > 0:
        bigint, mod = divmod(bigint, 2 ** 32)
        ints.append(mod)
    return ints


def hash_seed(seed=None, max_bytes=8):
    """gym 0.9.4 utils.seeding.hash_seed (third party; semantics pinned by the reference's
    tests/test_seed.py:19,23)."""
    if seed is None:
        seed = create_seed(max_bytes=max_bytes)
    h = hashlib.sha512(str(seed).encode('utf8')).digest()
    return _bigint_from_bytes(h[:max_bytes])


def create_seed(a=None, max_bytes=8):
    import os
    if a is None:
        a = _bigint_from_bytes(os.urandom(max_bytes))
    elif isinstance(a, str):
        a = a.encode('utf8')
        a += hashlib.sha512(a).digest()
        a = _bigint_from_bytes(a[:max_bytes])
    elif isinstance(a, int):
        a = a % 2 ** (8 * max_bytes)
    else:
        raise ValueError('Invalid type for seed: {} ({})'.format(type(a), a))
    return a


def np_random(seed=None):
    """gym 0.9.4 utils.seeding.np_random: (RandomState, seed)."""
    if seed is not None and not (isinstance(seed, int) and 0 <= seed):
        raise ValueError('Seed must be a non-negative integer or omitted, not {}'.format(seed))
    seed = create_seed(seed)
    rng = np.random.RandomState()
    rng.seed(_int_list_from_bigint(hash_seed(seed)))
    return rng, seed


# ---------------------------------------------------------------------------------------------
# Device generators (CUDA, Philox): same distributions, any population size
# ---------------------------------------------------------------------------------------------
def _stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def device_noise(n_episodes, n_noise, sensor='Dexcom', seed=0, episode_offset=0, group_size=1, device=None, out=None):
    """[n_noise, E] float64 device tensor of CGM noise (sg_gen_noise)."""
    import torch
    from . import _capi
    lib = _capi.load()
    s = P.sensor_params(sensor) if isinstance(sensor, str) else sensor
    dev = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
    W = torch.from_numpy(interp_matrix(s['sample_time'])).to(dev)
    if out is None:
        out = torch.empty(n_noise, n_episodes, dtype=torch.float64, device=dev)
    model = (C.c_double * 5)(s['PACF'], s['gamma'], s['lambda'], s['delta'], s['xi'])
    with torch.cuda.device(dev):
        _capi.check(lib.sg_gen_noise(C.c_void_p(out.data_ptr()), n_episodes, n_noise, W.shape[0],
                                     C.c_void_p(W.data_ptr()), model, seed, episode_offset, group_size, _stream()))
        torch.cuda.current_stream().synchronize()      # W may be freed after return
    return out


def device_scenarios(n_episodes, horizon_minutes, start_minute=0, seed=0, episode_offset=0, group_size=1, device=None):
    """(meal_minute [M, E] int32, meal_grams [M, E]) device tensors of RandomScenario-distributed meals."""
    import torch
    from . import _capi
    lib = _capi.load()
    dev = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
    n_days = (start_minute + horizon_minutes + 1439) // 1440
    M = 6 * n_days
    mm = torch.empty(M, n_episodes, dtype=torch.int32, device=dev)
    mg = torch.empty(M, n_episodes, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _capi.check(lib.sg_gen_scenarios(C.c_void_p(mm.data_ptr()), C.c_void_p(mg.data_ptr()), n_episodes, M,
                                         int(start_minute), int(horizon_minutes), seed, episode_offset, group_size, _stream()))
    return mm, mg


def _noise_model(s):
    return (C.c_double * 5)(s['PACF'], s['gamma'], s['lambda'], s['delta'], s['xi'])


def device_noise_keyed(out, key, sensor='Dexcom', seed=0, mask=None, W=None, listed=None):
    """Rewrite columns of the noise buffer out [n_noise, E] in place: episode e draws the Philox stream named by
    key[e] (int64 device tensor); only columns with mask[e] != 0 (int32 device tensor, None = all) are written.
    W: the device copy of interp_matrix(sample_time), for callers that generate every step (no upload, no sync).
    listed = (list [E] int32, count [1] int32) device tensors: write the columns list[0 .. count) instead (what
    sg_vec_finish gathers; sg_gen_noise_listed) -- same streams."""
    import torch
    from . import _capi
    lib = _capi.load()
    s = P.sensor_params(sensor) if isinstance(sensor, str) else sensor
    dev = out.device
    own_W = W is None
    if own_W:
        W = torch.from_numpy(interp_matrix(s['sample_time'])).to(dev)
    assert out.is_contiguous() and key.dtype == torch.int64 and (mask is None or mask.dtype == torch.int32)
    if listed is not None:
        assert not own_W and listed[0].dtype == torch.int32 and listed[1].dtype == torch.int32
        with torch.cuda.device(dev):
            _capi.check(lib.sg_gen_noise_listed(C.c_void_p(out.data_ptr()), out.shape[1], out.shape[0], W.shape[0],
                                                C.c_void_p(W.data_ptr()), _noise_model(s), seed, C.c_void_p(key.data_ptr()),
                                                C.c_void_p(listed[0].data_ptr()), C.c_void_p(listed[1].data_ptr()), _stream()))
        return out
    with torch.cuda.device(dev):
        _capi.check(lib.sg_gen_noise_keyed(C.c_void_p(out.data_ptr()), out.shape[1], out.shape[0], W.shape[0],
                                           C.c_void_p(W.data_ptr()), _noise_model(s), seed, C.c_void_p(key.data_ptr()),
                                           C.c_void_p(mask.data_ptr()) if mask is not None else None, _stream()))
        if own_W:
            torch.cuda.current_stream().synchronize()      # W may be freed after return
    return out


def device_scenarios_keyed(meal_minute, meal_grams, start_minute, horizon_minutes, key, seed=0, mask=None, listed=None):
    """Rewrite columns of the meal table (meal_minute / meal_grams [M, E]) in place: RandomScenario-distributed meals
    of an episode that starts at minute-of-day start_minute[e] (int32 device tensor); stream key[e], columns mask[e]
    -- or the columns list[0 .. count) with listed = (list, count) (sg_gen_scenarios_listed)."""
    import torch
    from . import _capi
    lib = _capi.load()
    assert meal_minute.is_contiguous() and meal_grams.is_contiguous() and start_minute.dtype == torch.int32
    if listed is not None:
        with torch.cuda.device(meal_minute.device):
            _capi.check(lib.sg_gen_scenarios_listed(C.c_void_p(meal_minute.data_ptr()), C.c_void_p(meal_grams.data_ptr()),
                                                    meal_minute.shape[1], meal_minute.shape[0],
                                                    C.c_void_p(start_minute.data_ptr()), int(horizon_minutes), seed,
                                                    C.c_void_p(key.data_ptr()), C.c_void_p(listed[0].data_ptr()),
                                                    C.c_void_p(listed[1].data_ptr()), _stream()))
        return meal_minute, meal_grams
    with torch.cuda.device(meal_minute.device):
        _capi.check(lib.sg_gen_scenarios_keyed(C.c_void_p(meal_minute.data_ptr()), C.c_void_p(meal_grams.data_ptr()),
                                               meal_minute.shape[1], meal_minute.shape[0],
                                               C.c_void_p(start_minute.data_ptr()), int(horizon_minutes), seed,
                                               C.c_void_p(key.data_ptr()),
                                               C.c_void_p(mask.data_ptr()) if mask is not None else None, _stream()))
    return meal_minute, meal_grams


def device_episode_starts(key, patient_row=None, table_x0=None, n_draws=2, seed=0, mask=None, start_minute=None, x0=None):
    """Start hour and randomised initial state of every episode with mask[e] != 0, as the gym wrapper draws them
    (simglucose_gym_env.py:58-73; sg_gen_episode_starts).  start_minute [E] int32 / x0 [13, E] are written in place
    (allocated when None and wanted: x0 is produced when patient_row is given).  Returns (start_minute, x0)."""
    import torch
    from . import _capi
    lib = _capi.load()
    dev = key.device
    n = key.numel()
    if start_minute is None:
        start_minute = torch.zeros(n, dtype=torch.int32, device=dev)
    if x0 is None and patient_row is not None:
        x0 = torch.zeros(13, n, dtype=torch.float64, device=dev)
    tx = None
    if x0 is not None:
        tx = table_x0 if table_x0 is not None else torch.from_numpy(P.cohort_table()[1]).to(dev)
        assert tx.is_contiguous() and patient_row.dtype == torch.int32
    with torch.cuda.device(dev):
        _capi.check(lib.sg_gen_episode_starts(C.c_void_p(start_minute.data_ptr()),
                                              C.c_void_p(x0.data_ptr()) if x0 is not None else None,
                                              C.c_void_p(tx.data_ptr()) if tx is not None else None,
                                              C.c_void_p(patient_row.data_ptr()) if patient_row is not None else None,
                                              n, int(n_draws), seed, C.c_void_p(key.data_ptr()),
                                              C.c_void_p(mask.data_ptr()) if mask is not None else None, _stream()))
        if table_x0 is None and tx is not None:
            torch.cuda.current_stream().synchronize()
    return start_minute, x0


def device_perturbed_params(patient_row, amplitude=0.1, seed=1234, episode_offset=0, table=None):
    """[32, E] device tensor: table[row] * (1 + amplitude * U(-1, 1)) per parameter (configs[3])."""
    import torch
    from . import _capi
    lib = _capi.load()
    dev = patient_row.device
    if table is None:
        table = P.cohort_table()[0]
    tbl = torch.from_numpy(np.ascontiguousarray(table, dtype=np.float64)).to(dev)
    n = patient_row.numel()
    out = torch.empty(32, n, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _capi.check(lib.sg_gen_perturbed_params(C.c_void_p(out.data_ptr()), C.c_void_p(tbl.data_ptr()),
                                                C.c_void_p(patient_row.data_ptr()), n, float(amplitude), seed,
                                                episode_offset, _stream()))
        torch.cuda.current_stream().synchronize()
    return out


def device_uniform(n, lo, hi, seed=0, offset=0, device=None, out=None):
    import torch
    from . import _capi
    lib = _capi.load()
    dev = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
    if out is None:
        out = torch.empty(n, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _capi.check(lib.sg_gen_uniform(C.c_void_p(out.data_ptr()), out.numel(), float(lo), float(hi), seed, offset,
                                       _stream()))
    return out
