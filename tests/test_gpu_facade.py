"""GPU: the drop-in Python surface (T1DPatient / T1DSimEnv / gym env / batch_sim / controllers / vector env)
against the golden fixtures recorded from the unmodified reference, written the way the reference's own
tests drive it (tests/test_sim_engine.py, test_seed.py, test_reset.py, test_pid_controller.py,
test_reward_fun.py of the reference).

Tolerance as in test_gpu_parity.py: max |d| / max(|ref|, 1) <= 1e-6 for BG / CGM, exact for announced CHO,
pump-quantised insulin (<= 1 ulp) and done flags.
"""
import os
from datetime import datetime, timedelta

import numpy as np
import pytest

import helpers as H
from helpers import O

pytestmark = pytest.mark.gpu

TOL = H.REL_TOL
CUSTOM = [(7, 45), (12, 70), (16, 15), (18, 80), (23, 10)]      # reference tests/test_sim_engine.py:55


def _names():
    from simglucose_b200 import params as P
    return P.patient_names()


def _build_env(name, start_hour, sensor_seed, scenario, sensor_name='Dexcom', pump_name='Insulet', **patient_kw):
    from simglucose_b200.actuator.pump import InsulinPump
    from simglucose_b200.patient.t1dpatient import T1DPatient
    from simglucose_b200.sensor.cgm import CGMSensor
    from simglucose_b200.simulation.env import T1DSimEnv
    from simglucose_b200.simulation.scenario import CustomScenario
    from simglucose_b200.simulation.scenario_gen import RandomScenario
    start = datetime(2018, 1, 1, start_hour, 0, 0)
    patient = T1DPatient.withName(name, **patient_kw)
    sensor = CGMSensor.withName(sensor_name, seed=sensor_seed)
    pump = InsulinPump.withName(pump_name)
    scen = RandomScenario(start_time=start, seed=scenario[1]) if scenario[0] == 'random' else \
        CustomScenario(start_time=start, scenario=list(scenario[1]))
    return T1DSimEnv(patient, sensor, pump, scen)


def _check_frame(df, d, i, tol=TOL):
    assert list(df.columns) == ['BG', 'CGM', 'CHO', 'insulin', 'LBGI', 'HBGI', 'Risk'] and df.index.name == 'Time'
    assert H.rel_err(df['BG'].to_numpy(), d['BG'][i]) <= tol
    assert H.rel_err(df['CGM'].to_numpy(), d['CGM'][i]) <= tol
    assert np.array_equal(df['CHO'].to_numpy()[:-1], d['CHO'][i][:-1])
    assert np.isnan(df['CHO'].to_numpy()[-1]) and np.isnan(df['insulin'].to_numpy()[-1])
    assert np.max(np.abs(df['insulin'].to_numpy()[:-1] - d['insulin'][i][:-1])) <= 1e-15
    for c in ('LBGI', 'HBGI', 'Risk'):
        assert H.rel_err(df[c].to_numpy(), d[c][i]) <= 2e-5


def test_env_loop_with_host_controller_matches_reference():
    """SimObj.simulate's loop written out by hand (sim_engine.py:29-39): controller.policy on the host, one
    step-kernel launch per env.step; custom_scenario fixture episode 3 (adult#007, start 06:00, seed 2)."""
    from simglucose_b200.controller.basal_bolus_ctrller import BBController
    d = H.load('custom_scenario')
    i = 3
    env = _build_env(_names()[int(d['patient_row'][i])], 6, 2, ('custom', CUSTOM))
    ctrl = BBController(target=140)
    ctrl.reset()
    obs, reward, done, info = env.reset()
    assert sorted(info.keys()) == ['bg', 'hbgi', 'lbgi', 'meal', 'patient_name', 'patient_state', 'risk',
                                   'sample_time', 'time']
    assert reward == 0 and done is False and info['meal'] == 0
    obs_l, rew_l, done_l = [obs.CGM], [reward], [done]
    for _ in range(int(d['n_steps'][i])):
        action = ctrl.policy(obs, reward, done, **info)
        obs, reward, done, info = env.step(action)
        obs_l.append(obs.CGM); rew_l.append(reward); done_l.append(done)
    assert env.time == datetime(2018, 1, 1, 6) + timedelta(minutes=3 * int(d['n_steps'][i]))
    _check_frame(env.show_history(), d, i)
    assert H.rel_err(obs_l, d['obs'][i]) <= TOL
    assert H.rel_err(rew_l, d['reward'][i]) <= 1e-5
    assert np.array_equal(np.array(done_l, dtype=bool), d['done'][i])
    assert H.rel_err(info['patient_state'], d['final_state'][i]) <= TOL
    assert H.rel_err(env.patient.state, d['final_state'][i]) <= TOL
    assert env.patient.t == 3.0 * int(d['n_steps'][i])


def test_late_start_t0_matches_reference():
    """T1DPatient(t0=...) (t1dpatient.py:27, 277): the patient's clock starts late, and with it env.time, the meal lookup
    and the sensor's sampling minute.  Fixtures recorded from the reference with t0 = 1, 2, 3, 7: for t0 % sample_time != 0
    the reset returns the sensor's held 0 and consumes no noise (cgm.py:27, 47-50).  Both seams: the loop with the host
    controller, and SimObj.simulate's fused launch, whose step count is ceil((minutes - t0) / sample_time)."""
    from simglucose_b200.controller.basal_bolus_ctrller import BBController
    from simglucose_b200.simulation.sim_engine import SimObj, _n_steps
    d = H.load('late_start')
    for i in range(len(d['t0'])):
        t0, n = int(d['t0'][i]), int(d['n_steps'][i])
        name = _names()[int(d['patient_row'][i])]
        env = _build_env(name, 6, 4, ('custom', CUSTOM), t0=t0)
        ctrl = BBController(target=140)
        ctrl.reset()
        obs, reward, done, info = env.reset()
        obs_l = [obs.CGM]
        for _ in range(n):
            obs, reward, done, info = env.step(ctrl.policy(obs, reward, done, **info))
            obs_l.append(obs.CGM)
        assert env.time == datetime(2018, 1, 1, 6) + timedelta(minutes=t0 + 3 * n) and env.patient.t == t0 + 3.0 * n
        _check_frame(env.show_history(), d, i)
        assert obs_l[0] == d['obs'][i][0] and H.rel_err(obs_l, d['obs'][i]) <= TOL
        assert H.rel_err(env.patient.state, d['final_state'][i]) <= TOL
        # the fused episode: sim_time chosen so that the reference's loop takes exactly n steps
        env2 = _build_env(name, 6, 4, ('custom', CUSTOM), t0=t0)
        sim_time = timedelta(minutes=t0 + 3 * n - 1)
        assert _n_steps(env2, sim_time) == n
        s = SimObj(env2, BBController(target=140), sim_time, animate=False, path=None)
        s.simulate()
        _check_frame(s.results(), d, i)


def test_batch_sim_fused_matches_reference_and_writes_csv(tmp_path):
    """batch_sim over SimObj's (sim_engine.py:65-76): BB controllers fuse into one episode launch."""
    import pandas as pd
    from simglucose_b200.controller.basal_bolus_ctrller import BBController
    from simglucose_b200.simulation.sim_engine import SimObj, batch_sim
    d = H.load('custom_scenario')
    names = _names()
    sims = []
    for i in range(3):                         # start 00:00, sensor seed 1
        env = _build_env(names[int(d['patient_row'][i])], 0, 1, ('custom', CUSTOM))
        sims.append(SimObj(env, BBController(), timedelta(days=1), animate=False, path=str(tmp_path)))
    launches0 = [s.env.patient._engine.launches for s in sims]
    results = batch_sim(sims, parallel=True)
    assert len(results) == 3
    for i, (s, df) in enumerate(zip(sims, results)):
        _check_frame(df, d, i)
        assert len(df) == 481 and df.index[0] == datetime(2018, 1, 1) and df.index[-1] == datetime(2018, 1, 2)
        f = os.path.join(str(tmp_path), s.env.patient.name + '.csv')
        back = pd.read_csv(f)
        assert list(back.columns) == ['Time', 'BG', 'CGM', 'CHO', 'insulin', 'LBGI', 'HBGI', 'Risk']
        np.testing.assert_allclose(back['BG'].to_numpy(), df['BG'].to_numpy(), rtol=1e-12)
        # the per-patient handles did no stepping of their own: the episodes ran as one population
        assert s.env.patient._engine.launches - launches0[i] <= 2
        assert H.rel_err(s.env.patient.state, d['final_state'][i]) <= TOL
    # the env can be stepped on from where the batch left it, like after the reference's loop
    e = sims[0].env
    from simglucose_b200.controller.base import Action
    step = e.step(Action(basal=0.01, bolus=0))
    assert e.time == datetime(2018, 1, 2, 0, 3) and np.isfinite(step.observation.CGM)


def test_same_result_single_and_batch():
    """The reference's own property (tests/test_sim_engine.py:25-85): sim() one by one == batch_sim()."""
    from simglucose_b200.controller.basal_bolus_ctrller import BBController
    from simglucose_b200.simulation.sim_engine import SimObj, batch_sim, sim
    names = ['adolescent#001', 'adult#001', 'child#001']

    def make(path):
        return [SimObj(_build_env(n, 6, 1, ('random', 1)), BBController(), timedelta(hours=6), animate=False,
                       path=path) for n in names]
    import tempfile
    with tempfile.TemporaryDirectory() as tmp:
        single = [sim(s) for s in make(tmp)]
        batch = batch_sim(make(tmp), parallel=True)
        lock = make(tmp)
        for s in lock:
            s._simulate_lockstep()
    for a, b, c in zip(single, batch, lock):
        assert a.equals(b)
        np.testing.assert_allclose(a['BG'].to_numpy(), c.results()['BG'].to_numpy(), rtol=1e-12)
        assert np.array_equal(a['CHO'].to_numpy()[:-1], c.results()['CHO'].to_numpy()[:-1])


def test_pid_simobj_matches_reference():
    """examples/run_pid_controller.py / tests/test_pid_controller.py of the reference: PID fused on device."""
    from simglucose_b200.controller.pid_ctrller import PIDController
    from simglucose_b200.simulation.sim_engine import SimObj
    d = H.load('cohort_pid_1day')
    for i in (3, 33, 95):                       # (patient, seed) = (3, 0), (3, 1), (5, 3); incl. child rows
        seed, row = divmod(i, 30)
        env = _build_env(_names()[row], 0, seed, ('random', seed))
        assert int(d['patient_row'][i]) == row
        ctrl = PIDController(P=0.001, I=0.00001, D=0.001, target=140)
        s = SimObj(env, ctrl, timedelta(days=1), animate=False, path=None)
        s.simulate()
        _check_frame(s.results(), d, i)
        assert ctrl.prev_state == s.results()['CGM'].to_numpy()[-2]      # last observation the controller saw


def test_gym_env_seeding_and_steps_match_reference():
    """tests/test_seed.py:14-25 known answers (start hours 23 and 14) plus stepped outputs."""
    from simglucose_b200.envs import T1DSimEnv
    d = H.load('gym_seeded')
    names = _names()
    for i in range(d['seed'].shape[0]):
        env = T1DSimEnv(patient_name=names[int(d['patient_row'][i])])
        seeds = env.seed(int(d['seed'][i]))
        assert [int(s) for s in seeds] == [int(s) for s in d['seeds'][i]]
        obs = env.reset()
        assert env.env.scenario.start_time == datetime(2018, 1, 1, int(d['start_hour'][i]), 0, 0)
        assert env.env.sensor.seed == int(d['sensor_seed'][i]) and env.env.scenario.seed == int(d['scenario_seed'][i])
        assert np.max(np.abs(env.env.patient.state - d['x0'][i])) <= 1e-9 * np.max(np.abs(d['x0'][i]))
        obs_l, rew_l, done_l, bg_l = [obs.CGM], [0.0], [False], [env.env.BG_hist[0]]
        for k in range(d['actions'].shape[1]):
            o, r, dn, info = env.step(float(d['actions'][i, k]))
            obs_l.append(o.CGM); rew_l.append(r); done_l.append(dn); bg_l.append(info['bg'])
            if k in (0, 17, 59):
                assert H.rel_err(info['patient_state'], d['states'][i, k]) <= TOL
        assert sorted(info.keys()) == sorted(str(k) for k in d['info_keys'][i])
        assert H.rel_err(obs_l, d['obs'][i]) <= TOL and H.rel_err(bg_l, d['bg'][i]) <= TOL
        assert H.rel_err(rew_l, d['reward'][i]) <= 1e-5
        assert np.array_equal(np.array(done_l, dtype=bool), d['done'][i])
    assert env.action_space.shape == (1,) and float(env.action_space.high[0]) == 30.0
    assert env.observation_space.shape == (1,)
    if int(d['seed'][0]) == 0 and int(d['seed'][1]) == 1000:
        assert int(d['start_hour'][0]) == 23 and int(d['start_hour'][1]) == 14       # tests/test_seed.py:19,23


def test_gym_reset_semantics():
    """tests/test_reset.py of the reference: resets differ under a fixed seed, deterministically."""
    from simglucose_b200.envs import T1DSimEnv
    env = T1DSimEnv(patient_name='adult#002')
    env.seed(0)
    o0 = env.reset(); t0 = env.env.scenario.start_time; s0 = env.env.scenario.scenario
    o1 = env.reset(); t1 = env.env.scenario.start_time; s1 = env.env.scenario.scenario
    assert o0 != o1 and t0 != t1 and s0 != s1
    env.seed(0)
    o2 = env.reset(); t2 = env.env.scenario.start_time; s2 = env.env.scenario.scenario
    o3 = env.reset(); t3 = env.env.scenario.start_time; s3 = env.env.scenario.scenario
    assert (o0, t0, s0) == (o2, t2, s2) and (o1, t1, s1) == (o3, t3, s3)
    env.seed(1)
    o4 = env.reset()
    assert o4 != o0 and env.env.scenario.start_time != t0


def test_custom_reward_function():
    """tests/test_reward_fun.py / examples/custom_reward_function.py: reward_fun sees the last hour of CGM."""
    from simglucose_b200.envs import T1DSimEnv
    seen = []

    def custom_reward(bg_last_hour):
        seen.append(list(bg_last_hour))
        return 1 if bg_last_hour[-1] > 180 else (-1 if bg_last_hour[-1] < 70 else 0)
    env = T1DSimEnv(patient_name='adolescent#002', reward_fun=custom_reward)
    env.seed(3)
    env.reset()
    for k in range(25):
        o, r, dn, info = env.step(0.01)
        assert r == custom_reward(env.env.CGM_hist[-20:])
    assert len(seen[0]) == 2 and max(len(s) for s in seen) == 20       # int(60 / 3) samples, env.py:100-102


def test_model_masks_at_signed_zeros_and_negative_states():
    """(x >= 0) * dxdt (t1dpatient.py:167-202): -0.0 counts as non-negative in the reference, negative states freeze
    their derivative.  The kernels test the sign BIT of the state after turning a caller's -0.0 into +0.0 at entry
    (csrc/sg_model.cuh canonical_zero): T1DPatient.model against the oracle for zero, -0.0, tiny and negative entries in
    the seven masked states, and max(EGPt, 0) / the renal threshold on either side."""
    from simglucose_b200.patient.t1dpatient import Action, T1DPatient
    p = T1DPatient.withID(7)
    prm, x0 = O.patient_params(6)
    masked = [3, 4, 5, 9, 10, 11, 12]
    cases = []
    for v in (0.0, -0.0, 5e-324, -5e-324, -1e-300, -3.5e-12, -1.0, 2.2250738585072014e-308):
        x = x0.copy()
        x[masked] = v
        cases.append(x)
    for k in masked:
        x = x0.copy()
        x[k] = -0.0
        cases.append(x)
        x = x0.copy()
        x[k] = -x[k] - 1.0
        cases.append(x)
    hi = x0.copy(); hi[3] = 5000.0; hi[8] = 900.0          # EGPt < 0 and x3 above the renal threshold ke2 = 339
    at = x0.copy(); at[3] = float(p._params['ke2'])        # exactly at the threshold: Et = 0
    cases += [hi, at]
    for x in cases:
        for cho, lq, lf in ((0.0, 0.0, 0.0), (5.0, 120.0, 35.0)):
            dx = T1DPatient.model(0.0, x, Action(CHO=cho, insulin=0.02), p._params, lq, lf)
            ref = O.model(x, cho, 0.02, prm, lq, lf)
            assert np.array_equal(np.asarray(dx)[masked] == 0.0, np.asarray(ref)[masked] == 0.0), (x, dx, ref)
            assert H.rel_err(dx, ref) <= 1e-12, (x, dx, ref)


def test_patient_step_and_model_against_oracle():
    """T1DPatient.step / .model (t1dpatient.py:82-208) minute by minute, with a meal and a bolus."""
    from simglucose_b200.patient.t1dpatient import Action, T1DPatient
    for pid in (1, 15, 23):
        p = T1DPatient.withID(pid)
        prm, x0 = O.patient_params(pid - 1)
        assert p.name == O.patient_names()[pid - 1] and np.array_equal(p.state, x0) and p.t == 0
        basal = float(p._params['u2ss']) * float(p._params['BW']) / 6000
        y, t, hc = x0.copy(), O.C.c_double(0.0), O.C.c_double(0.0)
        planned = 0.0; last_cho = 0.0; lq = x0[0] + x0[1]; lf = 0.0; eating = False
        for m in range(60):
            cho = 60.0 if m == 5 else 0.0
            ins = basal + (2.0 if m == 8 else 0.0)
            dx = T1DPatient.model(p.t, p.state, Action(CHO=min(5.0, planned + cho), insulin=ins), p._params,
                                  p._last_Qsto, p._last_foodtaken)
            ref = O.model(p.state, min(5.0, planned + cho), ins, prm, p._last_Qsto, p._last_foodtaken)
            assert H.rel_err(dx, ref) <= 1e-12
            p.step(Action(CHO=cho, insulin=ins))
            # oracle minute: same meal machine (t1dpatient.py:82-107) then sgo_dopri5_minute
            planned += cho
            to_eat = min(5.0, planned) if planned > 0 else 0.0
            planned = max(0.0, planned - to_eat)
            if to_eat > 0 and last_cho <= 0:
                lq = y[0] + y[1]; lf = 0.0; eating = True
            if eating:
                lf += to_eat
            if to_eat <= 0 and last_cho > 0:
                eating = False
            last_cho = to_eat
            cnt = np.zeros(4, dtype=np.int32)
            rc = O.lib().sgo_dopri5_minute(y.ctypes.data_as(O.C.POINTER(O.C.c_double)), O.C.byref(t),
                                           prm.ctypes.data_as(O.C.POINTER(O.C.c_double)), to_eat, ins, lq, lf, 0,
                                           O.C.byref(hc), cnt.ctypes.data_as(O.C.POINTER(O.C.c_int)))
            assert rc == 1
            assert H.rel_err(p.state, y) <= 1e-9
            assert p.t == m + 1 and p.is_eating == eating and p.planned_meal == planned
            assert p._last_foodtaken == lf and p._last_Qsto == lq
        assert abs(p.observation.Gsub - p.state[12] / prm[31]) == 0.0


def test_foreign_sensor_takes_the_unfused_path_with_the_same_result():
    """A user subclass of CGMSensor must be honoured (duck-typed seam): minute-by-minute choreography."""
    from simglucose_b200.controller.basal_bolus_ctrller import BBController
    from simglucose_b200.sensor.cgm import CGMSensor

    class MySensor(CGMSensor):
        calls = 0

        def measure(self, patient):
            MySensor.calls += 1
            return CGMSensor.measure(self, patient)
    fused = _build_env('adult#003', 6, 5, ('random', 5))
    plain = _build_env('adult#003', 6, 5, ('random', 5))
    plain.sensor = MySensor.withName('Dexcom', seed=5)
    assert fused.fused and not plain.fused
    out = []
    for env in (fused, plain):
        ctrl = BBController()
        obs, reward, done, info = env.reset()
        for _ in range(40):
            obs, reward, done, info = env.step(ctrl.policy(obs, reward, done, **info))
        out.append(env.show_history())
    assert MySensor.calls >= 120
    np.testing.assert_allclose(out[0]['BG'].to_numpy(), out[1]['BG'].to_numpy(), rtol=1e-13)
    np.testing.assert_allclose(out[0]['CGM'].to_numpy(), out[1]['CGM'].to_numpy(), rtol=1e-13)
    np.testing.assert_allclose(out[0]['Risk'].to_numpy(), out[1]['Risk'].to_numpy(), rtol=1e-9, atol=1e-12)
    assert np.array_equal(out[0]['insulin'].to_numpy()[:-1], out[1]['insulin'].to_numpy()[:-1])


def test_pump_sensor_risk_objects():
    from simglucose_b200.actuator.pump import InsulinPump
    from simglucose_b200.analysis.risk import risk_index
    pump = InsulinPump.withName('Insulet')
    assert pump.basal(0.0139375) == O.pump_quantise(0.0139375, 0, 30, 0.05)
    assert pump.bolus(45.0) == 30.0 and pump.basal(-1.0) == 0.0
    bg = np.array([40.0, 90.0, 112.5, 180.0, 400.0, 0.5])
    lb, hb, ri = risk_index(bg, 6)
    f = 1.509 * (np.log(bg[:5]) ** 1.084 - 5.381)
    assert abs(lb - np.mean(10 * f[f < 0] ** 2)) <= 1e-9 and abs(hb - np.mean(10 * f[f > 0] ** 2)) <= 1e-9
    assert risk_index([0.5], 1) == (0.0, 0.0, 0.0)
    l1, h1, r1 = risk_index(bg, 1)                      # only the last sample
    assert (l1, h1, r1) == (0.0, 0.0, 0.0)


def test_vector_env_reference_streams_match_single_envs():
    """T1DSimVecEnv (lockstep, one launch per step) == independent single T1DSimEnv's on the same seeds."""
    import torch
    from simglucose_b200.controller.base import Action
    from simglucose_b200.envs import T1DSimVecEnv
    names = ['adolescent#003', 'adult#004', 'child#005', 'adult#009']
    venv = T1DSimVecEnv(4, patient_names=names, seed=11, start_hour=6, horizon_steps=60, stream_mode='reference')
    obs = venv.reset()
    rng = np.random.RandomState(0)
    acts = rng.uniform(0.0, 0.05, size=(60, 4))
    singles = [_build_env(n, 6, 11 + e, ('random', 11 + e)) for e, n in enumerate(names)]
    first = [s.reset() for s in singles]
    assert np.allclose(obs.cpu().numpy(), [f.observation.CGM for f in first], rtol=0, atol=0)
    for k in range(60):
        o, r, dn, info = venv.step(torch.from_numpy(acts[k]))
        o, r, dn, bg = o.cpu().numpy(), r.cpu().numpy(), dn.cpu().numpy(), info['bg'].cpu().numpy()
        for e, s in enumerate(singles):
            st = s.step(Action(basal=acts[k, e], bolus=0))
            assert st.observation.CGM == o[e] and st.reward == r[e] and st.done == bool(dn[e]) and st.info['bg'] == bg[e]
    with pytest.raises(RuntimeError):
        venv.step(torch.zeros(4, dtype=torch.float64))
    agg = venv.statistics().cpu().numpy()
    assert agg[0] == 4 * 61
    venv.close()


def test_vector_env_device_streams():
    """Device (Philox) streams: deterministic per seed, group sharing, fused rollout == stepping."""
    import torch
    from simglucose_b200.envs import T1DSimVecEnv
    a = T1DSimVecEnv(600, seed=5, horizon_steps=40, group_size=30)
    b = T1DSimVecEnv(600, seed=5, horizon_steps=40, group_size=30)
    oa, ob = a.reset(), b.reset()
    assert torch.equal(oa, ob)
    # the 30 patients of one group share the noise stream: CGM - BG identical within a group at reset
    nz = (a.info['cgm'] - a.info['bg']).cpu().numpy().reshape(20, 30)
    assert np.all(np.abs(nz - nz[:, :1]) < 1e-9) and np.unique(np.round(nz[:, 0], 9)).size == 20
    basal = torch.full((600,), 0.012, dtype=torch.float64, device=a.device)
    rows = []
    for _ in range(40):
        o, r, dn, info = a.step(basal)
        rows.append(torch.stack([o, r, info['bg']]).clone())
    acts = torch.stack([basal, torch.zeros_like(basal)]).expand(40, 2, 600).contiguous()
    out = b.rollout(40, controller='actions', actions=acts, trace=True)['trace']
    from simglucose_b200 import _capi
    T = _capi.TRACE_COLS.index
    assert torch.equal(torch.stack(rows)[:, 0], out[:, T('obs')]) and torch.equal(torch.stack(rows)[:, 2], out[:, T('BG')])
    assert torch.equal(a.state, b.state)
    assert torch.isfinite(out[:, T('BG')]).all()
    a.close(); b.close()


def test_vector_env_custom_reward_window():
    """reward_fun hook of the vector env (env.py:66, 100-102): the callable sees CGM_hist[-20:] as a [n, E] device
    tensor, oldest first, reset sample included.  A tensor restatement of risk_diff (env.py:27-33) through the
    device risk kernel must reproduce the fused reward; a window statistic checks length and order."""
    import torch
    from simglucose_b200 import ops
    from simglucose_b200.envs import T1DSimVecEnv
    seen = []

    def risk_diff_t(w):
        seen.append(w.clone())
        if w.shape[0] < 2:
            return torch.zeros(w.shape[1], dtype=torch.float64, device=w.device)
        return ops.risk_index_elementwise(w[-2])[2] - ops.risk_index_elementwise(w[-1])[2]

    a = T1DSimVecEnv(90, seed=3, horizon_steps=30, group_size=30)
    b = T1DSimVecEnv(90, seed=3, horizon_steps=30, group_size=30, reward_fun=risk_diff_t)
    a.reset(); b.reset()
    assert b.window_size == 20
    basal = torch.full((90,), 0.012, dtype=torch.float64, device=a.device)
    hist = [b.info['cgm'].clone()]
    for k in range(30):
        _, ra, _, _ = a.step(basal)
        _, rb, _, info = b.step(basal)
        hist.append(info['cgm'].clone())
        assert torch.allclose(ra, rb, rtol=0, atol=1e-12)
        w = seen[-1]
        want = torch.stack(hist[-20:])
        assert w.shape[0] == min(k + 2, 20) and torch.equal(w, want)
    a.close(); b.close()


def test_simulate_keyword_interface(tmp_path):
    """user_interface.simulate(...) with every argument given (tests/test_ui.py of the reference drives the same
    call through mocked prompts): per-patient CSVs + the three report CSVs; numbers == the cohort fixture."""
    import pandas as pd
    from simglucose_b200.controller.basal_bolus_ctrller import BBController
    from simglucose_b200.simulation.scenario_gen import RandomScenario
    from simglucose_b200.simulation.user_interface import simulate
    d = H.load('cohort_bb_3day')
    names = [_names()[i] for i in (0, 12, 25)]
    start = datetime(2018, 1, 1, 0, 0, 0)
    res = simulate(sim_time=timedelta(days=3), scenario=RandomScenario(start_time=start, seed=1), controller=BBController(),
                   patient_names=names, cgm_name='Dexcom', cgm_seed=1, insulin_pump_name='Insulet', start_time=start,
                   save_path=str(tmp_path), animate=False, parallel=True)
    assert list(res.index) == names and res.shape == (3, 8)
    for j, i in enumerate((0, 12, 25)):
        back = pd.read_csv(os.path.join(str(tmp_path), names[j] + '.csv'))
        assert H.rel_err(back['BG'].to_numpy(), d['BG'][i]) <= TOL
        bg = d['BG'][i]
        assert abs(res.iloc[j]['BG>180'] - (bg > 180).sum() / len(bg) * 100) <= 100.0 / len(bg) + 1e-9
    for f in ('performance_stats.csv', 'risk_trace.csv', 'CVGA_stats.csv'):
        assert os.path.exists(os.path.join(str(tmp_path), f))
    with pytest.raises(ValueError):
        simulate(sim_time=timedelta(days=1))
