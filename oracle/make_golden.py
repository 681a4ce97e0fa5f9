"""TEST INFRASTRUCTURE ONLY -- generate tests/golden/*.npz by running the UNMODIFIED
reference (through oracle/ref_harness.py) in the build container.

    python oracle/make_golden.py [group ...]       # groups: see GROUPS below

Every fixture records the inputs the batched engine consumes (patient row, x0, meal table,
noise stream, controller spec / actions) and what the reference produced (history columns,
observation / reward / done per step, per-minute DOPRI5 counters, final state).  The
reference cannot travel to the GPU box; these files can.
"""
import os
import sys
from multiprocessing import Pool

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, _HERE)
import ref_harness as H  # noqa: E402

OUT = os.path.join(_HERE, '..', 'tests', 'golden')
CUSTOM = [(7, 45), (12, 70), (16, 15), (18, 80), (23, 10)]   # tests/test_sim_engine.py:55


def _pack(recs):
    """list of per-episode dicts -> dict of stacked / object-free arrays."""
    out = {}
    keys = recs[0].keys()
    n = len(recs)
    for k in keys:
        vals = [np.asarray(r[k]) for r in recs]
        if k in ('meal_minute', 'meal_grams'):
            m = max(1, max(v.size for v in vals))
            fill = -1 if k == 'meal_minute' else 0.0
            a = np.full((n, m), fill, dtype=vals[0].dtype)
            for i, v in enumerate(vals):
                a[i, :v.size] = v
            out[k] = a
        elif k == 'dop_counters':
            out[k] = np.stack(vals).astype(np.int16)
        else:
            out[k] = np.stack(vals)
    return out


def _episode(job):
    name, n_steps, kw = job
    rec = H.run_episode(name, n_steps, **kw)
    rec['controller_kind'] = np.array(kw.get('controller', ('bb', 140.0))[0])
    c = list(kw.get('controller', ('bb', 140.0))[1:])
    rec['controller_params'] = np.array(c + [0.0] * (4 - len(c)), dtype=np.float64)
    return rec


def _gym(job):
    seed, n, name = job
    rng = np.random.RandomState(seed + 7)
    return H.run_gym(seed, n, rng.uniform(0.0, 0.05, size=n), patient_name=name)


def _save(name, recs, **extra):
    os.makedirs(OUT, exist_ok=True)
    d = _pack(recs)
    d.update(extra)
    path = os.path.join(OUT, name + '.npz')
    np.savez_compressed(path, **d)
    print('wrote', os.path.normpath(path), '%.1f KB' % (os.path.getsize(path) / 1024.0), flush=True)


def g_golden_config(pool):
    """tests/test_sim_engine.py:87-113 (the config behind tests/sim_results.csv) + the file itself."""
    import pandas as pd
    recs = pool.map(_episode, [('adolescent#001', 960, dict(controller=('bb', 140.0), scenario=('random', 1),
                                                            start_hour=0, sensor_seed=1))])
    df = pd.read_csv(os.path.join(H.REFERENCE_ROOT, 'tests', 'sim_results.csv'))
    cols = ['BG', 'CGM', 'CHO', 'insulin', 'LBGI', 'HBGI', 'Risk']
    _save('golden_config', recs, sim_results_csv=df[cols].to_numpy(float), sim_results_cols=np.array(cols))


def g_custom(pool):
    """tests/test_sim_engine.py:55 CustomScenario, 1 day, BB (BASELINE.json configs[0])."""
    jobs = [(n, 480, dict(controller=('bb', 140.0), scenario=('custom', CUSTOM), start_hour=0, sensor_seed=1))
            for n in ('adolescent#001', 'adult#007', 'child#002')]
    jobs.append(('adult#007', 480, dict(controller=('bb', 140.0), scenario=('custom', CUSTOM), start_hour=6,
                                        sensor_seed=2)))
    _save('custom_scenario', pool.map(_episode, jobs))


def g_late_start(pool):
    """T1DPatient(t0=...) (patient/t1dpatient.py:27, 277): the patient's clock -- and with it env.time, the meal lookup
    and the sensor's sampling minute (cgm.py:27) -- starts late.  t0 = 3 keeps the sampling phase, t0 = 1 and 2 shift it:
    the reset then returns the sensor's held 0 (cgm.py:47-50) and the noise stream is consumed from its first sample."""
    jobs = [(n, 160, dict(controller=('bb', 140.0), scenario=('custom', CUSTOM), start_hour=6, sensor_seed=4, t0=t0))
            for n, t0 in (('adult#002', 1), ('child#005', 2), ('adolescent#003', 3), ('adult#009', 7))]
    _save('late_start', pool.map(_episode, jobs))


PERTURB_SEED, PERTURB_AMPLITUDE = 1234, 0.1        # bench.py config3() / BASELINE.json configs[3]


def _perturbed_jobs(indices, n_steps):
    """Episodes of the configs[3] population by GLOBAL index g: patient row g % 30, the parameter row the device
    generator draws for g (oracle/philox_host.py restates sg_gen_perturbed_params), RandomScenario / CGMSensor seeded
    with g (independent streams per episode), BBController (nominal basal, CR, CF by patient name), CSV x0."""
    import oracle as O
    import philox_host as PH
    names = H.patient_names()
    jobs, rows = [], []
    for g in indices:
        nominal, _ = O.patient_params(g % 30)
        row = PH.perturbed_row(nominal, PERTURB_AMPLITUDE, PERTURB_SEED, g)
        rows.append(row)
        jobs.append((names[g % 30], n_steps, dict(controller=('bb', 140.0), scenario=('random', g), start_hour=0,
                                                  sensor_seed=g, params_override=dict(zip(O.HOT_PARAMS, row)))))
    return jobs, rows


def _perturbed(pool, name, indices, n_steps):
    jobs, rows = _perturbed_jobs(indices, n_steps)
    recs = pool.map(_episode, jobs)
    for r, g, row in zip(recs, indices, rows):
        r['global_index'] = np.int64(g)
        r['params'] = row
    _save(name, recs, perturb_seed=np.int64(PERTURB_SEED), perturb_amplitude=np.float64(PERTURB_AMPLITUDE))


def g_perturbed(pool):
    """BASELINE.json configs[3] subset: the UNMODIFIED reference run on +-10 % perturbed parameter rows -- global
    episodes 0..59 x 1 day (every patient twice; a dozen of them dip below 50 mg/dL on the first day)."""
    _perturbed(pool, 'perturbed_1day', list(range(60)), 480)


# 7 days: two rows that stay physiological (7, 33) and three that the reference drives to BG ~ 0, where x3, x4, x12
# ride the (x >= 0) * dxdt masks of t1dpatient.py:167, 173, 202 (found by scanning episodes 0..399 with the C oracle)
PERTURBED_WEEK = [7, 22, 33, 57, 297]


def g_perturbed_week(pool):
    _perturbed(pool, 'perturbed_7day', PERTURBED_WEEK, 7 * 480)


def g_cohort_bb(pool):
    """BASELINE.json configs[1]: all 30 patients x 3 days, RandomScenario seed 1, BBController."""
    names = H.patient_names()
    jobs = [(n, 1440, dict(controller=('bb', 140.0), scenario=('random', 1), start_hour=0, sensor_seed=1))
            for n in names]
    _save('cohort_bb_3day', pool.map(_episode, jobs))


def g_cohort_pid(pool):
    """BASELINE.json configs[4] subset: 30 patients x seeds 0..3, PID 1e-3/1e-5/1e-3 (examples/run_pid_controller.py:5), 1 day."""
    names = H.patient_names()
    jobs = [(n, 480, dict(controller=('pid', 1e-3, 1e-5, 1e-3, 140.0), scenario=('random', s), start_hour=0,
                          sensor_seed=s)) for s in range(4) for n in names]
    _save('cohort_pid_1day', pool.map(_episode, jobs))


def g_open_loop(pool):
    """Open-loop random basal actions (BASELINE.json configs[2] style), random_init_bg, odd start hours."""
    names = H.patient_names()
    jobs = []
    for i, idx in enumerate([0, 4, 9, 10, 15, 19, 20, 24, 29, 7, 13, 27]):
        rng = np.random.RandomState(100 + i)
        acts = np.stack([rng.uniform(0.0, 0.05, 240), np.where(rng.rand(240) < 0.03, rng.uniform(0, 2.0, 240), 0.0)], 1)
        jobs.append((names[idx], 240, dict(controller=('actions',), actions=acts, scenario=('random', 10 + i),
                                           start_hour=(5 * i) % 24, sensor_seed=50 + i, random_init_bg=True,
                                           patient_seed=200 + i)))
    _save('open_loop_actions', pool.map(_episode, jobs))


def g_hardware(pool):
    """Other sensors / pump: GuardianRT (5 min), Navigator (1 min), Cozmo."""
    jobs = [('adult#003', 144, dict(controller=('bb', 140.0), scenario=('random', 3), start_hour=6, sensor_seed=3,
                                    sensor_name='GuardianRT', pump_name='Cozmo')),
            ('child#006', 360, dict(controller=('bb', 140.0), scenario=('random', 4), start_hour=6, sensor_seed=4,
                                    sensor_name='Navigator', pump_name='Insulet')),
            ('adolescent#005', 144, dict(controller=('pid', 1e-3, 1e-5, 1e-3, 140.0), scenario=('random', 5),
                                         start_hour=3, sensor_seed=5, sensor_name='GuardianRT', pump_name='Insulet'))]
    recs = pool.map(_episode, jobs)
    for r, j in zip(recs, jobs):
        r['sensor_name'] = np.array(j[2]['sensor_name'])
        r['pump_name'] = np.array(j[2]['pump_name'])
    for i, r in enumerate(recs):
        _save('hardware_%d' % i, [r])


def g_gym(pool):
    """gym wrapper: seeding chain known answers (tests/test_seed.py:19,23) and stepped outputs."""
    jobs = [(0, 60, 'adolescent#001'), (1000, 60, 'adolescent#001'), (7, 60, 'adult#002'), (123, 60, 'child#009')]
    _save('gym_seeded', pool.map(_gym, jobs))


def g_noise(pool):
    """CGMNoise streams (sensor/noise_gen.py) and RandomScenario schedules (scenario_gen.py) for generator parity."""
    H._activate()
    from datetime import datetime, timedelta
    from simglucose.simulation.scenario_gen import RandomScenario
    out = {}
    for sname in ('Dexcom', 'GuardianRT', 'Navigator'):
        out['noise_' + sname] = np.stack([H.noise_stream(sname, s, 400) for s in range(6)])
    sched_min, sched_g = [], []
    for s in range(8):
        for hour in (0, 6):
            start = datetime(2018, 1, 1, hour)
            sc = RandomScenario(start_time=start, seed=s)
            mm, gg = [], []
            for m in range(3 * 1440):
                a = sc.get_action(start + timedelta(minutes=m))
                if a.meal != 0:
                    mm.append(m); gg.append(float(a.meal))
            mm = (mm + [-1] * 24)[:24]; gg = (gg + [0.0] * 24)[:24]
            sched_min.append(mm); sched_g.append(gg)
    out['sched_seed_hour'] = np.array([(s, h) for s in range(8) for h in (0, 6)], dtype=np.int32)
    out['sched_minute'] = np.array(sched_min, dtype=np.int32)
    out['sched_grams'] = np.array(sched_g, dtype=np.float64)
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, 'generators.npz')
    np.savez_compressed(path, **out)
    print('wrote', os.path.normpath(path), flush=True)


def g_report(pool):
    """analysis/report.py (percent_stats :74-92, risk_index_trace :95-133, CVGA_analysis :198-217) run by the
    reference's own code on the BG traces of the cohort fixtures (plotting calls made inert)."""
    H._activate()
    import pandas as pd
    pd.DataFrame.plot = lambda self, *a, **k: None          # no matplotlib in the build container
    from simglucose.analysis import report as R

    class _NumpyCompat(object):
        """np.mean(DataFrame) is the per-column mean under the reference's pinned pandas 1.2.4 (it calls
        DataFrame.mean()); pandas >= 2 reduces over both axes.  Everything else is numpy itself."""
        def __getattr__(self, name):
            return getattr(np, name)

        @staticmethod
        def mean(a, *args, **kw):
            return a.mean() if isinstance(a, pd.DataFrame) else np.mean(a, *args, **kw)
    R.np = _NumpyCompat()
    out = {}
    for tag, fixture in (('bb', 'cohort_bb_3day'), ('pid', 'cohort_pid_1day')):
        d = np.load(os.path.join(OUT, fixture + '.npz'))
        n = 30
        BG = pd.DataFrame(d['BG'][:n].T, columns=H.patient_names()[:n])        # [time x patient], as df.unstack(level=0).BG
        pstats, _, _ = R.percent_stats(BG)
        ri_per_hour, ri_mean, _, _ = R.risk_index_trace(BG, visualize=False)
        BG_min, BG_max, A, B, C, D_, E = R.CVGA_analysis(BG)
        out[tag + '_BG'] = BG.to_numpy(float)
        out[tag + '_percent'] = pstats.to_numpy(float)                          # [patient x 5]
        out[tag + '_percent_cols'] = np.array(list(pstats.columns))
        out[tag + '_ri_per_hour'] = ri_per_hour.to_numpy(float)                 # [(metric, patient) x chunk]
        out[tag + '_ri_mean'] = ri_mean.to_numpy(float)                         # [patient x 3]
        out[tag + '_ri_mean_cols'] = np.array(list(ri_mean.columns))
        out[tag + '_cvga_min'] = np.asarray(BG_min, float)
        out[tag + '_cvga_max'] = np.asarray(BG_max, float)
        out[tag + '_zones'] = np.array([A, B, C, D_, E], dtype=float)
    path = os.path.join(OUT, 'report_stats.npz')
    np.savez_compressed(path, **out)
    print('wrote', os.path.normpath(path), '%.1f KB' % (os.path.getsize(path) / 1024.0), flush=True)


def _days(job):
    """n day schedules drawn by the reference's RandomScenario.create_scenario (scenario_gen.py:33-60) from one
    RandomState(seed): per slot (identified by its time window) the meal minute and grams, -1 where no meal."""
    seed, n = job
    H._activate()
    from datetime import datetime
    from simglucose.simulation.scenario_gen import RandomScenario
    sc = RandomScenario(start_time=datetime(2018, 1, 1, 0, 0, 0), seed=seed)
    ub = np.array([9, 10, 14, 16, 20, 23]) * 60
    t_out = np.full((n, 6), -1, dtype=np.int16)
    g_out = np.full((n, 6), -1, dtype=np.int16)
    for d in range(n):
        day = sc.create_scenario()
        for t, g in zip(day['meal']['time'], day['meal']['amount']):
            k = int(np.searchsorted(ub, t, side='left'))
            if t_out[d, k] < 0:                          # a second meal in the same window can only come from a shared
                t_out[d, k] = int(t); g_out[d, k] = int(g)   # boundary minute; keep the first
    return t_out, g_out


def g_generator_samples(pool):
    """Large samples of the reference's own stochastic generators, for the distribution tests of the device (Philox)
    generators (tests/test_gpu_generators.py): 24,000 day schedules of RandomScenario."""
    jobs = [(900 + k, 1500) for k in range(16)]
    res = pool.map(_days, jobs)
    out = dict(day_minute=np.concatenate([r[0] for r in res]), day_grams=np.concatenate([r[1] for r in res]),
               day_seeds=np.array([j[0] for j in jobs], dtype=np.int32))
    path = os.path.join(OUT, 'generator_samples.npz')
    np.savez_compressed(path, **out)
    print('wrote', os.path.normpath(path), '%.1f KB' % (os.path.getsize(path) / 1024.0), flush=True)


GROUPS = dict(golden_config=g_golden_config, custom=g_custom, cohort_bb=g_cohort_bb, cohort_pid=g_cohort_pid,
              open_loop=g_open_loop, hardware=g_hardware, gym=g_gym, noise=g_noise, report=g_report,
              generator_samples=g_generator_samples, late_start=g_late_start, perturbed=g_perturbed,
              perturbed_week=g_perturbed_week)

if __name__ == '__main__':
    groups = sys.argv[1:] or list(GROUPS)
    with Pool(int(os.environ.get('GOLDEN_PROCS', os.cpu_count() or 1))) as pool:
        for g in groups:
            GROUPS[g](pool)
