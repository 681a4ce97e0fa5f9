"""TEST INFRASTRUCTURE ONLY -- never imported by the product package.

Runs the UNMODIFIED reference (Sawyerbatch/simglucose, mounted read-only at
/root/reference) in the build container through the shims in oracle/shims and
records, for one closed-loop or open-loop episode:

  * the INPUTS the batched engine consumes (parameter row index, x0, meal
    table, CGM noise stream, controller spec or per-step actions), and
  * the OUTPUTS the reference produced (history frame columns, per-step
    observation / reward / done, per-minute DOPRI5 step counters, final state).

The reference cannot travel to the GPU box, so `make_golden.py` commits these
records as fixtures under tests/golden/.

Shims: stub `gym` (simglucose/__init__.py:1 imports
it), stub `matplotlib` (sensor/noise_gen.py:6, simulation/rendering.py:1-2),
and `init_state=params.iloc[2:15].to_numpy(float)` because pandas 3 raises at
patient/t1dpatient.py:272 (positional Series[int]).
"""
import os
import sys
import warnings
from datetime import datetime, timedelta

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_ROOT = os.environ.get('SIMGLUCOSE_REFERENCE', '/root/reference')


def _activate():
    shims = os.path.join(_HERE, 'shims')
    for p in (REFERENCE_ROOT, shims):
        if p not in sys.path:
            sys.path.insert(0, p)
    warnings.filterwarnings('ignore')


def reference_available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, 'simglucose'))


def make_patient(name, params_override=None, **kw):
    """T1DPatient.withName (patient/t1dpatient.py:57-68) with the ndarray
    init_state shim.  params_override: {column: value} written into the parameter Series before the
    patient is built -- T1DPatient.__init__ (t1dpatient.py:22-42) takes any Series; the Name (what
    BBController looks u2ss / BW / CR / CF up by, basal_bolus_ctrller.py:52-64) and the x0 columns stay
    those of the CSV row."""
    _activate()
    import pandas as pd
    from simglucose.patient.t1dpatient import T1DPatient, PATIENT_PARA_FILE
    csv = pd.read_csv(PATIENT_PARA_FILE)
    params = csv.loc[csv.Name == name].squeeze().copy()
    x0 = params.iloc[2:15].to_numpy(float).copy()
    if params_override:
        for col, val in params_override.items():
            assert col in params.index and not col.startswith('x0_'), col
            params[col] = float(val)
    return T1DPatient(params, init_state=x0, **kw)


def patient_names():
    _activate()
    import pandas as pd
    from simglucose.patient.t1dpatient import PATIENT_PARA_FILE
    return list(pd.read_csv(PATIENT_PARA_FILE).Name)


def noise_stream(sensor_name, seed, n):
    """First n samples of the reference CGMNoise stream
    (sensor/noise_gen.py:15-69) for a sensor/seed."""
    _activate()
    import pandas as pd
    from simglucose.sensor.cgm import SENSOR_PARA_FILE
    from simglucose.sensor.noise_gen import CGMNoise
    sp = pd.read_csv(SENSOR_PARA_FILE)
    params = sp.loc[sp.Name == sensor_name].squeeze()
    gen = CGMNoise(params, seed=seed)
    return np.array([next(gen) for _ in range(n)], dtype=np.float64)


def run_episode(patient_name, n_steps, controller=('bb', 140.0), scenario=('random', 1),
                start_hour=0, sensor_name='Dexcom', sensor_seed=1, pump_name='Insulet',
                actions=None, random_init_bg=False, patient_seed=None, t0=0, params_override=None):
    """Run the reference closed loop for n_steps env-steps exactly as
    SimObj.simulate does (simulation/sim_engine.py:29-39), recording inputs and
    outputs.

    controller: ('bb', target) | ('pid', P, I, D, target) | ('actions',) with
                `actions` an (n_steps, 2) array of (basal, bolus) U/min.
    scenario:   ('random', seed) | ('custom', [(hours_after_start, grams), ...])
    """
    _activate()
    from simglucose.simulation.env import T1DSimEnv
    from simglucose.controller.base import Action
    from simglucose.controller.basal_bolus_ctrller import BBController
    from simglucose.controller.pid_ctrller import PIDController
    from simglucose.sensor.cgm import CGMSensor
    from simglucose.actuator.pump import InsulinPump
    from simglucose.simulation.scenario_gen import RandomScenario
    from simglucose.simulation.scenario import CustomScenario

    start_time = datetime(2018, 1, 1, start_hour, 0, 0)
    patient = make_patient(patient_name, params_override=params_override, random_init_bg=random_init_bg,
                           seed=patient_seed, t0=t0)
    sensor = CGMSensor.withName(sensor_name, seed=sensor_seed)
    pump = InsulinPump.withName(pump_name)
    if scenario[0] == 'random':
        scen = RandomScenario(start_time=start_time, seed=scenario[1])
    else:
        scen = CustomScenario(start_time=start_time, scenario=list(scenario[1]))

    # record every non-zero meal the env sees, keyed by episode minute
    meals = []
    _orig_get_action = scen.get_action

    def _rec_get_action(t):
        a = _orig_get_action(t)
        if a.meal != 0:
            minute = int(round((t - start_time).total_seconds() / 60.0))
            meals.append((minute, float(a.meal)))
        return a
    scen.get_action = _rec_get_action

    # per-minute DOPRI5 counters (scipy iwork[16:20] = nfcn, nstep, naccpt, nrejct)
    counters = []
    _orig_step = patient.step.__func__

    def _rec_step(action):
        _orig_step(patient, action)
        iw = patient._odesolver._integrator.iwork
        counters.append((int(iw[16]), int(iw[17]), int(iw[18]), int(iw[19])))
    patient.step = _rec_step

    env = T1DSimEnv(patient, sensor, pump, scen)

    if controller[0] == 'bb':
        ctrl = BBController(target=controller[1])
    elif controller[0] == 'pid':
        ctrl = PIDController(P=controller[1], I=controller[2], D=controller[3], target=controller[4])
    else:
        ctrl = None
        actions = np.asarray(actions, dtype=np.float64)

    if ctrl is not None:
        ctrl.reset()
    meals.clear()
    counters.clear()
    obs, reward, done, info = env.reset()
    # env.reset() re-creates the solver; x0 is the post-reset state
    x0 = np.array(patient.state, dtype=np.float64).copy()
    sample_time = int(env.sample_time)
    obs_l, rew_l, done_l, act_l = [float(obs.CGM)], [float(reward)], [bool(done)], []
    for k in range(n_steps):
        if ctrl is not None:
            a = ctrl.policy(obs, reward, done, **info)
        else:
            a = Action(basal=float(actions[k, 0]), bolus=float(actions[k, 1]))
        act_l.append((float(a.basal), float(a.bolus)))
        obs, reward, done, info = env.step(a)
        obs_l.append(float(obs.CGM))
        rew_l.append(float(reward))
        done_l.append(bool(done))
    df = env.show_history()
    csv_row = patient_names().index(patient_name)
    n_noise = n_steps + 2
    rec = dict(
        patient_row=np.int32(csv_row),
        sample_time=np.int32(sample_time),
        start_minute=np.int32(start_hour * 60),
        t0=np.int32(t0),
        n_steps=np.int32(n_steps),
        x0=x0,
        meal_minute=np.array([m for m, _ in meals], dtype=np.int32),
        meal_grams=np.array([g for _, g in meals], dtype=np.float64),
        noise=noise_stream(sensor_name, sensor_seed, n_noise),
        actions=np.array(act_l, dtype=np.float64),
        BG=df['BG'].to_numpy(float), CGM=df['CGM'].to_numpy(float),
        CHO=df['CHO'].to_numpy(float), insulin=df['insulin'].to_numpy(float),
        LBGI=df['LBGI'].to_numpy(float), HBGI=df['HBGI'].to_numpy(float),
        Risk=df['Risk'].to_numpy(float),
        obs=np.array(obs_l), reward=np.array(rew_l), done=np.array(done_l),
        dop_counters=np.array(counters, dtype=np.int32),
        final_state=np.array(patient.state, dtype=np.float64),
    )
    return rec


def run_gym(seed, n_steps, basal_actions, patient_name='adolescent#001'):
    """Drive the reference gym wrapper (envs/simglucose_gym_env.py:18-85) exactly as
    tests/test_seed.py:14-25 does: env.seed(seed); env.reset(); then n_steps of
    env.step(basal).  Records the seeding chain, start hour, post-reset x0, meal table,
    noise stream and the per-step outputs."""
    _activate()
    import pandas as pd
    from simglucose.patient import t1dpatient as _tp
    from simglucose.envs.simglucose_gym_env import T1DSimEnv as GymEnv
    from simglucose.sensor.noise_gen import CGMNoise

    def _with_name(cls, name, **kw):
        csv = pd.read_csv(_tp.PATIENT_PARA_FILE)
        params = csv.loc[csv.Name == name].squeeze()
        return cls(params, init_state=params.iloc[2:15].to_numpy(float).copy(), **kw)
    _tp.T1DPatient.withName = classmethod(_with_name)

    env = GymEnv(patient_name=patient_name)
    seeds = env.seed(seed)
    # wrap the scenario of the env that reset() is about to build: patch after reset instead
    obs0 = env.reset()
    inner = env.env
    start_time = inner.scenario.start_time
    meals = []
    _orig = inner.scenario.get_action

    def _rec(t):
        a = _orig(t)
        if a.meal != 0:
            meals.append((int(round((t - start_time).total_seconds() / 60.0)), float(a.meal)))
        return a
    inner.scenario.get_action = _rec
    x0 = np.array(inner.patient.state, dtype=np.float64).copy()
    sensor_seed = inner.sensor.seed
    obs_l, rew_l, done_l, bg_l, st_l = [float(obs0.CGM)], [0.0], [False], [float(inner.BG_hist[0])], []
    for k in range(n_steps):
        obs, reward, done, info = env.step(float(basal_actions[k]))
        obs_l.append(float(obs.CGM)); rew_l.append(float(reward)); done_l.append(bool(done))
        bg_l.append(float(info['bg']))
        st_l.append(np.array(info['patient_state'], dtype=np.float64).copy())
    gen = CGMNoise(inner.sensor._params, seed=sensor_seed)
    noise = np.array([next(gen) for _ in range(n_steps + 2)])
    csv_row = patient_names().index(patient_name)
    return dict(seed=np.int64(seed), seeds=np.array(seeds, dtype=np.int64), start_hour=np.int32(start_time.hour),
                patient_row=np.int32(csv_row), x0=x0, sensor_seed=np.int64(sensor_seed),
                scenario_seed=np.int64(inner.scenario.seed), patient_seed=np.int64(inner.patient.seed),
                meal_minute=np.array([m for m, _ in meals], dtype=np.int32),
                meal_grams=np.array([g for _, g in meals], dtype=np.float64), noise=noise,
                actions=np.asarray(basal_actions, dtype=np.float64)[:n_steps],
                obs=np.array(obs_l), reward=np.array(rew_l), done=np.array(done_l), bg=np.array(bg_l),
                states=np.array(st_l), info_keys=np.array(sorted(info.keys())))
