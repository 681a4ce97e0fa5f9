"""TEST / MEASUREMENT INFRASTRUCTURE ONLY -- never imported by the product package.

Times the UNMODIFIED Python reference (Sawyerbatch/simglucose) on this machine's host cores: the path
`batch_sim(sim_instances, parallel=True)` of simglucose/simulation/sim_engine.py:65-69 -- one SimObj per patient,
mapped over a process pool, each process running the stock `SimObj.simulate` loop (sim_engine.py:29-39) on the stock
T1DSimEnv / T1DPatient / scipy dopri5 code.

    python oracle/ref_python_bench.py [--patients 30] [--hours 24] [--procs N]      -> one JSON line

The reference is imported from baseline/_ref/ (installed there by __graft_entry__.build() with
`pip install --no-deps --target baseline/_ref <copy of /root/reference>`; git-ignored, it travels to the GPU box with the
snapshot) or from $SIMGLUCOSE_REFERENCE.  Three shims are needed in this image: stub `gym` and
`matplotlib` packages (oracle/shims; both are import-time dependencies only) and an ndarray `init_state` because pandas 3
rejects `Series[0]` at patient/t1dpatient.py:272.  pathos is not installed; `multiprocess.Pool` is the pool pathos wraps
(fallback: the standard library's multiprocessing with fork).  sim() itself is not used because it writes CSV files;
the worker calls the same SimObj.simulate() and returns results().
"""
import argparse
import json
import os
import sys
import time

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)


def _reference_root():
    env = os.environ.get('SIMGLUCOSE_REFERENCE')
    for cand in (env, os.path.join(ROOT, 'baseline', '_ref'), '/root/reference'):
        if cand and os.path.isdir(os.path.join(cand, 'simglucose')):
            return cand
    return None


def _build_sim(job):
    """One SimObj exactly as examples/ and simulation/user_interface.py:364-372 of the reference build it."""
    name, hours, seed = job
    from datetime import datetime, timedelta
    import ref_harness as H
    from simglucose.actuator.pump import InsulinPump
    from simglucose.controller.basal_bolus_ctrller import BBController
    from simglucose.sensor.cgm import CGMSensor
    from simglucose.simulation.env import T1DSimEnv
    from simglucose.simulation.scenario_gen import RandomScenario
    from simglucose.simulation.sim_engine import SimObj
    start = datetime(2018, 1, 1, 0, 0, 0)
    patient = H.make_patient(name)
    env = T1DSimEnv(patient, CGMSensor.withName('Dexcom', seed=seed), InsulinPump.withName('Insulet'),
                    RandomScenario(start_time=start, seed=seed))
    return SimObj(env, BBController(), timedelta(hours=hours), animate=False, path=None)


def _worker(job):
    sim = _build_sim(job)
    t0 = time.perf_counter()
    sim.simulate()                                   # the stock loop, sim_engine.py:29-39
    df = sim.results()
    return job[0], len(df), float(df['BG'].iloc[-1]), time.perf_counter() - t0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--patients', type=int, default=30)
    ap.add_argument('--hours', type=float, default=24.0)
    ap.add_argument('--procs', type=int, default=0)
    ap.add_argument('--seed', type=int, default=1)
    args = ap.parse_args()
    ref = _reference_root()
    if ref is None:
        print(json.dumps({'unavailable': 'no reference tree (baseline/_ref missing: run __graft_entry__.build() where '
                                         '/root/reference exists)'}))
        return
    os.environ['SIMGLUCOSE_REFERENCE'] = ref
    sys.path.insert(0, _HERE)
    import ref_harness as H
    H.REFERENCE_ROOT = ref
    H._activate()
    import logging
    logging.disable(logging.CRITICAL)
    import numpy, pandas, scipy
    names = H.patient_names()[:args.patients]
    procs = args.procs or os.cpu_count() or 1
    procs = max(1, min(procs, len(names)))
    jobs = [(n, args.hours, args.seed) for n in names]
    pool_kind = 'multiprocess.Pool'
    try:
        import multiprocess as mp
    except Exception:
        import multiprocessing as mp
        pool_kind = 'multiprocessing.Pool'
    t0 = time.perf_counter()
    if procs == 1:
        res = [_worker(j) for j in jobs]
        pool_kind = 'serial'
    else:
        with mp.Pool(procs) as pool:
            res = pool.map(_worker, jobs, chunksize=1)
    wall = time.perf_counter() - t0
    minutes = sum((rows - 1) * 3 for _, rows, _, _ in res)
    print(json.dumps({
        'value': minutes / wall, 'unit': 'patient-sim-minutes/sec', 'kind': 'reference', 'cores': procs,
        'host_cpus': os.cpu_count(), 'pool': pool_kind, 'wall_s': wall,
        'per_process_rate': minutes / sum(t for _, _, _, t in res),
        'sample': '%d virtual patients x %.0f h, BBController, RandomScenario seed %d, Dexcom seed %d, Insulet: the '
                  'unmodified Python reference (SimObj.simulate mapped over %s(%d), sim_engine.py:29-39, 65-69)'
                  % (len(names), args.hours, args.seed, args.seed, pool_kind, procs),
        'patient_minutes': minutes, 'final_bg_first_patient': res[0][2],
        'versions': {'python': sys.version.split()[0], 'numpy': numpy.__version__, 'scipy': scipy.__version__,
                     'pandas': pandas.__version__},
        'reference_root': os.path.relpath(ref, ROOT) if ref.startswith(ROOT) else ref}))


if __name__ == '__main__':
    main()
