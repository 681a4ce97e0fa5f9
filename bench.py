#!/usr/bin/env python
"""bench.py -- patient-sim-minutes/s of the batched closed loop on N B200s (one process per GPU).

Headline workload (BASELINE.json configs[1], filled out to occupy a GPU as SURVEY.md 8d "Config 2" says): the 30
virtual patients x R scenario/sensor seed groups x 3 simulated days, Dexcom + Insulet, BBController with the
Quest.csv CR/CF, RandomScenario-distributed meals and CGMNoise-distributed sensor noise (Philox streams generated
on device and pinned to the reference's distributions by tests/test_gpu_generators.py; the 30 patients of one seed
group share their streams exactly as RandomScenario(seed=s) / CGMSensor(seed=s) do in the reference).  One "step" =
one pass over all episodes: reset + 1440 fused env-steps + population statistics + the statistics all-reduce.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config 1|3]

--config 3 runs BASELINE.json configs[3] at its stated size instead: 16,777,216 episodes with +-10 % perturbed
parameters x 7 days, sharded over the ranks (distributed.shard_range), the NCCL statistics reduce inside the timed
region.

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for the definition of every key.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# Algorithmic FP64 operations: +,-,*,/,compare,max,abs and transcendental = 1 each
F_RHS = 84
F_STEP = 6 * F_RHS + 65 * 13 + 20          # 1369 per attempted DOPRI5 step
F_INIT = 2 * F_RHS + 244                   # 412 per minute: k1 + HINIT (scipy >= 1.7 mode)
F_INIT_CARRY = F_RHS                       # carry_h mode
F_MISC = 40                                # pump, meal machine, CGM, risk/3, controller/3 per minute

METRIC = 'patient-sim-minutes/sec'
TRAFFIC_FILE = os.path.join(ROOT, 'profiles', 'run_kernel_traffic.json')
# the sources run_kernel is compiled from: an ncu DRAM capture is only quoted for the kernel it was taken on
KERNEL_SOURCES = ['sg_model.cuh', 'sg_math.cuh', 'sg_kernels.cu', 'sg_kernels.cuh', 'sg_philox.cuh']

_OUT = None


def claim_stdout():
    """stdout carries exactly one JSON line: keep a private handle on it and point fd 1 at stderr, so that banners
    printed by libraries (NCCL prints its version with printf at communicator creation) land on stderr."""
    global _OUT
    if _OUT is None:
        sys.stdout.flush()
        _OUT = os.fdopen(os.dup(1), 'w')
        os.dup2(2, 1)


def emit(line):
    out = _OUT or sys.stdout
    out.write(json.dumps(line) + '\n')
    out.flush()


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--config', type=int, default=1, choices=[1, 3],
                    help='1: the headline cohort workload; 3: the 16.7M-episode perturbed population sweep at full size')
    ap.add_argument('--groups', type=int, default=8840,
                    help='seed groups of 30 patients per GPU (8840 x 30 episodes = 7 waves of 148 SMs x 256 resident threads)')
    ap.add_argument('--days', type=int, default=3)
    ap.add_argument('--fma', action='store_true', help='SG_FLAG_FMA kernels (not the default arithmetic)')
    ap.add_argument('--carry-h', action='store_true')
    ap.add_argument('--fp32', action='store_true', help='optional FP32 mode (SG_FLAG_FP32): NOT the graded arithmetic')
    ap.add_argument('--coherence', type=int, default=-1,
                    help='env-steps between two sorts of the launch order (sg_set_coherence); -1 = the default of the '
                         'workload (40 for resident populations; 0 for the one-wave shards of the e2e leg, where every '
                         'launch boundary costs a tail)')
    ap.add_argument('--adaptive-sort', action='store_true',
                    help='sort the launch order by the DOPRI5 steps every episode just took (sg_set_sort_mode 1) instead of '
                         'by its next meal event')
    ap.add_argument('--cpu-seconds', type=float, default=15.0, help='target duration of the cpu_baseline sample')
    ap.add_argument('--no-cpu', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-traces', action='store_true', help='skip the e2e leg that copies the trajectories out')
    ap.add_argument('--no-configs', action='store_true', help='skip the reduced runs of the other BASELINE.json configs')
    ap.add_argument('--no-ref-python', action='store_true', help='skip timing the Python reference on the host cores')
    ap.add_argument('--ref-python-hours', type=float, default=24.0)
    ap.add_argument('--shards', type=int, default=0,
                    help='host-resident shards of the e2e leg (engine.run_shards); 0 = one resident wave each '
                         '(148 SMs x 256 episodes = 1262 seed groups)')
    ap.add_argument('--episodes', type=int, default=16777216, help='--config 3: total episodes over all ranks')
    ap.add_argument('--sweep-days', type=int, default=7, help='--config 3: simulated days')
    ap.add_argument('--wave', type=int, default=2097152, help='--config 3: episodes resident per GPU at a time')
    ap.add_argument('--c3-episodes', type=int, default=16777216,
                    help='episodes of the configs[3] entry of the default line (its stated size by default)')
    ap.add_argument('--c3-days', type=int, default=7, help='simulated days of the configs[3] entry of the default line')
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
         'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '200'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for ln in self.lines:
            f = [x.strip() for x in ln.split(',')]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'power_w_max': max(pw) if pw else None, 'samples': len(sm), 'reasons': sorted(reasons)}


# ---------------------------------------------------------------------------------------------
# workload helpers
# ---------------------------------------------------------------------------------------------
def controller_rows(rows_np):
    """BBController constants per episode: target, CR, CF (Quest.csv) and the basal of the CSV row the reference finds
    by patient name (basal_bolus_ctrller.py:52-64)."""
    from simglucose_b200 import params as P
    q = P.quest_table()
    tbl = P.cohort_table()[0]
    basal = tbl[:, 1] * tbl[:, 0] / 6000.0                 # u2ss * BW / 6000
    return np.stack([np.full(len(rows_np), 140.0), q[rows_np, 0], q[rows_np, 1], basal[rows_np]])


def controller_rows_device(rows):
    import torch
    t = torch.from_numpy(controller_rows(np.arange(30))).to(rows.device)       # [4, 30]
    return t[:, rows.long()].contiguous()


def flops_per_pass(n_minutes_total, dopri_steps_total, carry_h):
    return n_minutes_total * ((F_INIT_CARRY if carry_h else F_INIT) + F_MISC) + F_STEP * dopri_steps_total


def kernel_sha16():
    h = hashlib.sha256()
    for f in KERNEL_SOURCES:
        with open(os.path.join(ROOT, 'simglucose_b200', 'csrc', f), 'rb') as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def measured_traffic(groups, days, coherence):
    """DRAM bytes of the run_kernel launches of ONE pass (with a sorted launch order a pass is cut into launches of
    `coherence` env-steps; their bytes are summed) from the committed ncu capture -- quoted only when the capture was
    taken on the kernel sources of this tree (hash) and on this workload."""
    try:
        t = json.load(open(TRAFFIC_FILE))
    except Exception:
        return None, 'no capture committed (profiles/run_kernel_traffic.json)'
    if t.get('kernel_sha16') != kernel_sha16():
        return None, 'capture %s is of an older kernel (sha %s, tree %s): not quoted' % (
            t.get('source'), t.get('kernel_sha16'), kernel_sha16())
    if t.get('groups') != groups or t.get('days') != days or t.get('coherence_steps', 0) != coherence:
        return None, 'capture is of another workload (%s groups x %s days, coherence %s)' % (
            t.get('groups'), t.get('days'), t.get('coherence_steps', 0))
    return int(t['dram_bytes_per_pass']), ('ncu dram__bytes_read.sum + dram__bytes_write.sum over the %d run_kernel '
                                           'launches of one pass, %s' % (t.get('launches_per_pass', 1), t.get('source')))


def cpu_sample(args, noise_h, mm_h, mg_h, rows_h, n_steps, threads, target_s):
    """Time the CPU oracle (oracle/sg_oracle.c, a C port of the reference path; pthread pool standing in
    for the pathos process pool) on the first episodes of the same workload."""
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import oracle as O
    from simglucose_b200 import params as P
    tbl, x0 = P.cohort_table()
    E_all = rows_h.shape[0]

    def run(ne):
        r = rows_h[:ne]
        t0 = time.perf_counter()
        st, failed = O.run_batch(tbl[r], x0[r], n_steps, mm_h[:, :ne].T.copy(), mg_h[:, :ne].T.copy(),
                                 noise_h[:, :ne].T.copy(), controller='bb', ctrl_rows=controller_rows(r).T.copy()[:, [0, 1, 2, 3, 3, 3]],
                                 carry_h=int(args.carry_h), n_threads=threads)
        return time.perf_counter() - t0, st, failed
    ne = min(E_all, max(threads, 30))
    dt, st, _ = run(ne)                                      # calibration pass
    rate = ne * n_steps * 3 / dt
    ne2 = int(min(E_all, max(ne, rate * target_s / (n_steps * 3))))
    ne2 = max(threads, (ne2 // 30) * 30 or ne2)
    dt, st, failed = run(ne2)
    return {'value': ne2 * n_steps * 3 / dt, 'unit': METRIC, 'cores': threads, 'kind': 'port',
            'sample': '%d episodes (first %d seed groups of the bench workload) x %d min in %.1f s, '
                      'oracle/sg_oracle.c over %d pthreads' % (ne2, ne2 // 30, n_steps * 3, dt, threads),
            'failed': int(failed), 'mean_dopri_steps_per_min': float(st[:, 14].sum() / (ne2 * n_steps * 3))}


def reference_python(hours):
    """The UNMODIFIED Python reference on this box's host cores (oracle/ref_python_bench.py: SimObj.simulate mapped
    over a process pool, 30 patients, BBController) -- the number north_star asks to see beside the GPU's."""
    cmd = [sys.executable, os.path.join(ROOT, 'oracle', 'ref_python_bench.py'), '--hours', str(hours)]
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
        line = [ln for ln in r.stdout.strip().split('\n') if ln.startswith('{')][-1]
        return json.loads(line)
    except Exception as e:                                   # noqa: BLE001
        return {'unavailable': 'oracle/ref_python_bench.py failed: %r' % (e,)}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path on the box's host cores.  The arm is the C
    oracle port (oracle/sg_oracle.c, cpu_baseline.kind 'port') on all host threads: the Python reference itself does
    ~1e2 patient-min/s per core, so a bounded sample of it cannot cover a 3-day cohort in minutes; it is timed as
    cpu_baseline.reference_python (30 patients x 1 day) in the default run and in this one.  Each step is a bounded
    sample of the bench workload (same patients, controller, scenario / noise distributions, generated on the host with
    the reference-compatible MT19937 generators), sized by a calibration pass to a few seconds -- the metric is a
    rate per patient-minute, so the sample size does not enter it."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import oracle as O
    from simglucose_b200 import streams as S, params as P
    threads = O.max_threads()
    n_steps = args.days * 480
    n_min = n_steps * 3
    tbl, x0 = P.cohort_table()

    def make(groups):
        E = groups * 30
        rows = (np.arange(E) % 30).astype(np.int32)
        noise = np.repeat(np.stack([S.cgm_noise_stream('Dexcom', g, n_steps + 2) for g in range(groups)], 0), 30, axis=0)
        tabs = [S.random_scenario_table(g, 0, n_min) for g in range(groups)]
        mm, mg = S.pad_tables([tabs[e // 30] for e in range(E)])
        ctrl = controller_rows(rows).T.copy()[:, [0, 1, 2, 3, 3, 3]]
        return dict(E=E, P=tbl[rows], x0=x0[rows], mm=mm.T.copy(), mg=mg.T.copy(), noise=noise, ctrl=ctrl)

    def run(w):
        t0 = time.perf_counter()
        st, failed = O.run_batch(w['P'], w['x0'], n_steps, w['mm'], w['mg'], w['noise'], controller='bb',
                                 ctrl_rows=w['ctrl'], carry_h=int(args.carry_h), n_threads=threads)
        return time.perf_counter() - t0, st, failed

    cal = make(max(1, (4 * threads + 29) // 30))
    run(cal)                                                          # library load, thread start-up
    dt, _, _ = run(cal)                                               # calibration
    rate = cal['E'] * n_min / dt
    groups = int(max(1, min(4096, round(rate * 4.0 / n_min / 30))))   # ~4 s of CPU work per step
    w = make(groups)
    times, st, failed = [], None, 0
    for it in range(args.warmup + args.steps):
        dt, st, failed = run(w)
        if it >= args.warmup:
            times.append(dt)
    sec = float(np.mean(times))
    v = w['E'] * n_min / sec
    sample = ('%d episodes (30 patients x %d seed groups) x %d min per step, %.2f s per step, oracle/sg_oracle.c '
              'over %d pthreads' % (w['E'], groups, n_min, sec, threads))
    cpu = {'value': v, 'unit': METRIC, 'cores': threads, 'kind': 'port', 'sample': sample, 'failed': int(failed),
           'mean_dopri_steps_per_min': float(st[:, 14].sum() / (w['E'] * n_min))}
    if not args.no_ref_python:
        cpu['reference_python'] = reference_python(args.ref_python_hours)
    line = {'impl': 'reference', 'metric': METRIC, 'value': v, 'unit': METRIC, 'n_gpus': args.gpus,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * sec,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': {'workload': 'configs[1] cohort: 30 virtual patients x seed groups x %d-day RandomScenario, '
                                   'BBController, Dexcom, Insulet (bounded CPU sample: %d seed groups)' % (args.days, groups),
                       'episodes': w['E'], 'minutes_per_episode': n_min, 'carry_h': bool(args.carry_h)},
            'cpu_baseline': cpu,
            'e2e': {'value': v, 'unit': METRIC, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}
    emit(line)


# ---------------------------------------------------------------------------------------------
# the other BASELINE.json configurations (configs[2] and [4] at their stated sizes in the default run; --config 3 runs configs[3] in full)
# ---------------------------------------------------------------------------------------------
def _timed(fn, dev, warm=1, reps=1):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    return e0.elapsed_time(e1) * 1e-3 / reps


def config0(local):
    """configs[0]: the single-environment path.  (a) latency of T1DSimEnv.step through the drop-in facade, host
    BBController in the loop, one kernel launch + one device->host read per env-step; (b) the drop-in call itself:
    batch_sim of 30 SimObj x 3 days, from Python objects to 30 DataFrames (host stream replay, one fused launch, trace
    D2H, DataFrame assembly -- the reference's actual deliverable, sim_engine.py:44-49, 56-62)."""
    from datetime import datetime, timedelta
    import torch
    from simglucose_b200 import params as P
    from simglucose_b200.actuator.pump import InsulinPump
    from simglucose_b200.controller.basal_bolus_ctrller import BBController
    from simglucose_b200.patient.t1dpatient import T1DPatient
    from simglucose_b200.sensor.cgm import CGMSensor
    from simglucose_b200.simulation.env import T1DSimEnv
    from simglucose_b200.simulation.scenario_gen import RandomScenario
    from simglucose_b200.simulation.sim_engine import SimObj, batch_sim
    start = datetime(2018, 1, 1, 0, 0, 0)

    def env_of(name):
        return T1DSimEnv(T1DPatient.withName(name, device=local), CGMSensor.withName('Dexcom', seed=1),
                         InsulinPump.withName('Insulet'), RandomScenario(start_time=start, seed=1))
    env, ctrl = env_of('adolescent#001'), BBController()
    obs, reward, done, info = env.reset()
    lat = []
    for k in range(220):
        t0 = time.perf_counter()
        action = ctrl.policy(obs, reward, done, **info)
        obs, reward, done, info = env.step(action)
        lat.append(time.perf_counter() - t0)
    lat = np.array(lat[20:]) * 1e3
    sims = [SimObj(env_of(n), BBController(), timedelta(days=3), animate=False, path=None) for n in P.patient_names()]
    import contextlib
    import io
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        frames = batch_sim(sims, parallel=True)
    torch.cuda.synchronize()
    sec = time.perf_counter() - t0
    rows = sum(len(f) for f in frames)
    return {'config': 'configs[0] single-environment path through the drop-in facade',
            'env_step_latency_ms': {'median': float(np.median(lat)), 'p90': float(np.percentile(lat, 90)),
                                    'what': 'controller.policy + T1DSimEnv.step (3 simulated minutes), adolescent#001, '
                                            'one launch and one D2H read per step; the Python reference takes ~30 ms'},
            'batch_sim_30x3days': {'seconds': sec, 'value': 30 * 4320 / sec, 'unit': METRIC, 'frames': len(frames),
                                   'rows': rows, 'columns': list(frames[0].columns),
                                   'what': 'batch_sim(30 SimObj, 3 days): host MT19937 stream replay, one fused launch, '
                                           'trace D2H, 30 DataFrames assembled (e2e incl. trajectories)',
                                   'mean_bg_first': float(frames[0]['BG'].mean())}}


def config2(local, envs=1048576, n_steps=480, fp32=False, coherence=None, auto_reset=True, gym_seeding=True):
    """configs[2]: gym 'simglucose-v0' as a lockstep vector env -- every environment is the gym wrapper (own seeds,
    start hour, randomised initial state), random basal actions drawn in the step kernel, risk_diff reward, every
    environment restarted on its own when it reports done or reaches the horizon."""
    import torch
    from simglucose_b200.distributed import summarize
    from simglucose_b200.envs import T1DSimVecEnv
    E = int(envs) // 64 * 64
    env = T1DSimVecEnv(E, seed=0, horizon_steps=480, stream_mode='device', gym_seeding=gym_seeding, auto_reset=auto_reset,
                       fp32=fp32, device=local, coherence=coherence)
    ret = torch.zeros(E, dtype=torch.float64, device=env.device)
    n_done = torch.zeros((), dtype=torch.float64, device=env.device)

    def episode():
        env.reset()
        for k in range(n_steps):
            obs, reward, done, info = env.step_random(0.0, 0.05)
            ret.add_(reward)
            n_done.add_(done.sum())
    l0 = env.engine.launches
    sec = _timed(episode, env.device, warm=0, reps=1)
    launches = env.engine.launches - l0
    agg = summarize(env.statistics())
    started = int(env.episodes_started.sum().item())
    failed = int((env.engine.status() < 0).sum().item())
    coh = env.coherence
    env.close()
    return {'config': 'configs[2] lockstep vector env (gym simglucose-v0 batched): %d environments x %d env-steps, basal ~ '
                      'U(0, 0.05) U/min drawn in the step kernel, risk_diff reward, gym seeding per environment (start '
                      'hour, random_init_bg), per-environment reset on done / horizon%s'
                      % (E, n_steps, ', OPTIONAL FP32 INTEGRATOR (not the graded arithmetic)' if fp32 else ''),
            'value': E * n_steps * 3 / sec, 'unit': METRIC, 'seconds': sec, 'env_steps_per_s': E * n_steps / sec,
            'ms_per_env_step': 1e3 * sec / n_steps, 'library_launches': int(launches), 'coherence_steps': coh,
            'episodes_started': started, 'done_events': float(n_done.item()), 'solver_failures': failed,
            'population_current_episodes': agg}


def config3(local, rank, world, total, days, wave, coherence, seed=1234, carry_h=False, fma=False, adaptive=False):
    """configs[3]: `total` episodes with +-10 % perturbed parameters x `days`, BBController (nominal basal by name),
    independent scenario / sensor streams per episode, sharded over the ranks in contiguous ranges and simulated in
    resident waves; parameters and streams are generated on device per wave, keyed by GLOBAL episode index, so the
    population does not depend on how it is sharded.  Returns (seconds of this rank, aggregate tensor [NSTAT] after the
    all-reduce, failures, launches); the caller times across ranks."""
    import torch
    import torch.distributed as dist
    from simglucose_b200 import _capi, streams as S
    from simglucose_b200.distributed import allreduce_stats, shard_range
    from simglucose_b200.engine import BatchEngine
    dev = torch.device('cuda', local)
    lo, hi = shard_range(total, rank, world)
    n_steps = days * 480
    cap = min(wave, hi - lo)
    eng = BatchEngine(cap, device=local, coherence=coherence, carry_h=carry_h, fma=fma, adaptive_sort=adaptive)
    noise_buf = torch.empty((n_steps + 2) * cap, dtype=torch.float64, device=dev)
    agg = torch.zeros(_capi.NSTAT, dtype=torch.float64, device=dev)
    agg[2], agg[3] = float('inf'), float('-inf')
    failed = torch.zeros(1, dtype=torch.float64, device=dev)
    wave_agg = torch.empty(_capi.NSTAT, dtype=torch.float64, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    gen_ms, l0 = 0.0, eng.launches
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize(dev)
    e0.record()
    for w_lo in range(lo, hi, wave):
        n = min(wave, hi - w_lo)
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        rows = ((torch.arange(n, device=dev, dtype=torch.int64) + w_lo) % 30).to(torch.int32)
        params = S.device_perturbed_params(rows, amplitude=0.1, seed=seed, episode_offset=w_lo)
        noise = S.device_noise(n, n_steps + 2, 'Dexcom', seed=seed, episode_offset=w_lo, group_size=1, device=dev,
                               out=noise_buf[:(n_steps + 2) * n].view(n_steps + 2, n))
        mm, mg = S.device_scenarios(n, n_steps * 3, start_minute=0, seed=seed, episode_offset=w_lo, group_size=1, device=dev)
        ctrl_rows = controller_rows_device(rows)
        g1.record()
        eng.reset(rows, noise, mm, mg, params=params, row0=False)
        eng.run(n_steps, controller='bb', ctrl_rows=ctrl_rows, trace=False)
        eng.reduce_stats(wave_agg)
        agg[[0, 1, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15]] += wave_agg[[0, 1, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15]]
        agg[2] = torch.minimum(agg[2], wave_agg[2])
        agg[3] = torch.maximum(agg[3], wave_agg[3])
        failed += (eng.status() < 0).sum()
        torch.cuda.synchronize(dev)
        gen_ms += g0.elapsed_time(g1)
    # the one collective of the path, inside the timed region: the product's own function
    allreduce_stats(agg)
    if world > 1:
        dist.all_reduce(failed, op=dist.ReduceOp.SUM)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1)
    launches = eng.launches - l0
    eng.close()
    del noise_buf
    torch.cuda.empty_cache()
    return ms * 1e-3, agg, int(failed.item()), int(launches), gen_ms * 1e-3, (lo, hi)


def config4(local, seeds=10000):
    """configs[4]: closed-loop PIDController (examples/run_pid_controller.py gains) fused on device, 30 patients x seeds."""
    import torch
    from simglucose_b200 import streams as S
    from simglucose_b200.distributed import summarize
    from simglucose_b200.engine import BatchEngine
    n_steps, E = 480, 30 * int(seeds)
    dev = torch.device('cuda', local)
    rows = (torch.arange(E, device=dev, dtype=torch.int64) % 30).to(torch.int32)
    noise = S.device_noise(E, n_steps + 2, 'Dexcom', seed=0, group_size=30, device=dev)
    mm, mg = S.device_scenarios(E, n_steps * 3, start_minute=0, seed=0, group_size=30, device=dev)
    eng = BatchEngine(E, device=local)          # (the sorted launch order does not pay here: 1.32e9 against 1.45e9, session X)

    def run():
        eng.reset(rows, noise, mm, mg, row0=False)
        eng.run(n_steps, controller='pid', ctrl=(1e-3, 1e-5, 1e-3, 140.0), trace=False)
        return eng.reduce_stats()
    sec = _timed(run, dev, warm=1, reps=2)
    agg = summarize(eng.reduce_stats())
    eng.close()
    return {'config': 'configs[4] PIDController (P 1e-3, I 1e-5, D 1e-3, target 140) fused on device, 30 patients x %d '
                      'seeds x 1 day' % seeds,
            'value': E * n_steps * 3 / sec, 'unit': METRIC, 'seconds': sec, 'population': agg}


def vector_hex(agg):
    """The aggregate vector, bit for bit (float.hex), for comparing runs at different numbers of GPUs."""
    return [float(v).hex() for v in agg.detach().cpu().tolist()]


def run_config3_line(args, rank, world, local):
    """--config 3: BASELINE.json configs[3] at its stated size as the measured workload."""
    import torch
    import torch.distributed as dist
    from simglucose_b200 import _capi
    from simglucose_b200.distributed import summarize
    dev = torch.device('cuda', local)
    coherence = 40 if args.coherence < 0 else args.coherence
    n_min = args.sweep_days * 1440
    # warm-up: the same kernels on a short horizon (a full sweep takes tens of seconds; the kernels are warm after one wave)
    for _ in range(max(1, min(args.warmup, 2))):
        config3(local, rank, world, 65536 * world, 1, 65536, coherence, carry_h=args.carry_h, fma=args.fma)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    secs, aggs = [], []
    for _ in range(max(1, args.steps)):
        sec, agg, failed, launches, gen_s, (lo, hi) = config3(local, rank, world, args.episodes, args.sweep_days, args.wave,
                                                              coherence, carry_h=args.carry_h, fma=args.fma)
        t = torch.tensor([sec], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        secs.append(float(t.item()))
        aggs.append(agg)
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        return
    sec = float(np.mean(secs))
    a = aggs[-1].cpu().numpy()
    fl = flops_per_pass(args.episodes * n_min, a[14], args.carry_h)
    line = {'metric': METRIC, 'value': args.episodes * n_min / sec, 'unit': METRIC, 'n_gpus': world, 'steps': len(secs),
            'warmup': max(1, min(args.warmup, 2)), 'ms_per_step': 1e3 * sec, 'higher_is_better': True, 'scaling': 'strong',
            'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': {'workload': 'configs[3] population sweep: %d episodes, vpatient_params perturbed +-10 %% (Philox seed '
                                   '1234), %d days, BBController, independent RandomScenario / CGMNoise streams per '
                                   'episode, sharded over %d GPU(s) in contiguous ranges, waves of %d resident episodes'
                                   % (args.episodes, args.sweep_days, world, args.wave),
                       'episodes': args.episodes, 'minutes_per_episode': n_min, 'coherence_steps': coherence,
                       'arithmetic': 'fma' if args.fma else 'strict (no FMA contraction)', 'carry_h': bool(args.carry_h),
                       'parallelism': 'episodes sharded by distributed.shard_range, dp%d; one NCCL statistics all-reduce '
                                      '(distributed.allreduce_stats) inside the timed region' % world,
                       'timed_region': 'per wave: parameter / stream generation on device, reset, fused episode kernel, '
                                       'statistics reduction; then the all-reduce',
                       'warmup_note': 'warm-up passes run the same kernels on 65,536 episodes x 1 day per rank',
                       'l2': 'inputs (%.1f GB noise per wave) larger than L2' % ((args.sweep_days * 480 + 2) * min(args.wave, args.episodes // world) * 8 / 1e9)},
            'roofline': {'bound': 'fp64', 'achieved': fl / sec / 1e12 / world, 'unit': 'TFLOP/s per GPU incl. generation and '
                         'reduction time', 'peak': None, 'frac': None, 'traffic': None,
                         'dopri_steps_per_minute': float(a[14] / (args.episodes * n_min)),
                         'note': 'the roofline fraction of run_kernel is reported by the default (configs[1]) run'},
            'cpu_baseline': None,
            'e2e': None, 'gpu_launches': int(launches), 'clocks': clocks,
            'stream_generation_s_rank0': gen_s, 'solver_failures': failed,
            'population': summarize(aggs[-1]), 'population_vector_hex': vector_hex(aggs[-1]),
            'seconds_per_sweep': secs}
    emit(line)


def main():
    args = parse()
    claim_stdout()
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if args.impl == 'reference':
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    from simglucose_b200 import _capi, streams as S, params as P
    from simglucose_b200.distributed import allreduce_stats, summarize
    from simglucose_b200.engine import BatchEngine
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device: there is no CPU fallback')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        # NCCL prints its version banner on stdout at NCCL_DEBUG=VERSION; stdout carries the one JSON line
        if os.environ.get('NCCL_DEBUG', '').upper() in ('', 'VERSION'):
            os.environ['NCCL_DEBUG'] = 'WARN'
        dist.init_process_group('nccl', device_id=dev)
    lib = _capi.load()
    if args.config == 3:
        run_config3_line(args, rank, world, local)
        if world > 1:
            dist.destroy_process_group()
        return

    E = args.groups * 30
    n_steps = args.days * 480
    n_min = n_steps * 3
    ep0 = rank * E                                    # weak scaling: every rank simulates its own E episodes
    coherence = 40 if args.coherence < 0 else args.coherence
    coherence_e2e = 0 if args.coherence < 0 else args.coherence
    rows = (torch.arange(E, device=dev, dtype=torch.int64) % 30).to(torch.int32)
    rows_h = rows.cpu().numpy()
    ctrl_rows = controller_rows_device(rows)
    # inputs resident in HBM before the timed region
    # (generated on device, once, outside the timed region; SURVEY.md 8d asks for its cost to be reported: the first
    # call also pays the library's module load, the second is the generation itself)
    S.device_noise(30, 8, 'Dexcom', seed=1, device=dev)
    torch.cuda.synchronize(dev)
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    noise = S.device_noise(E, n_steps + 2, 'Dexcom', seed=1, episode_offset=ep0, group_size=30, device=dev)
    mm, mg = S.device_scenarios(E, n_min, start_minute=0, seed=1, episode_offset=ep0, group_size=30, device=dev)
    g1.record()
    torch.cuda.synchronize(dev)
    stream_generation_s = g0.elapsed_time(g1) * 1e-3
    eng = BatchEngine(E, device=local, carry_h=args.carry_h, fma=args.fma, fp32=args.fp32, coherence=coherence,
                      adaptive_sort=args.adaptive_sort)
    agg = torch.zeros(_capi.NSTAT, dtype=torch.float64, device=dev)
    agg_all = torch.zeros(_capi.NSTAT, dtype=torch.float64, device=dev)

    def one_pass():
        eng.reset(rows, noise, mm, mg, row0=False)
        eng.run(n_steps, controller='bb', ctrl_rows=ctrl_rows, trace=False)
        eng.reduce_stats(agg)
        agg_all.copy_(agg)
        allreduce_stats(agg_all)                      # the one collective of the path (NCCL), every pass

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        one_pass()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = eng.launches
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * args.steps + 2)]
    ev[0].record()
    for k in range(args.steps):
        eng.reset(rows, noise, mm, mg, row0=False)
        ev[2 * k + 1].record()
        eng.run(n_steps, controller='bb', ctrl_rows=ctrl_rows, trace=False)
        ev[2 * k + 2].record()
        eng.reduce_stats(agg)
        agg_all.copy_(agg)
        allreduce_stats(agg_all)
    ev[-1].record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = eng.launches - launches0
    total_ms = ev[0].elapsed_time(ev[-1])
    kern_ms = float(np.mean([ev[2 * k + 1].elapsed_time(ev[2 * k + 2]) for k in range(args.steps)]))
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    a = agg.cpu().numpy()
    value = world * E * n_min * args.steps / (total_ms * 1e-3)

    # ---- roofline of the dominant kernel (run_kernel): FP64 pipe ----
    fl = flops_per_pass(E * n_min, a[14], args.carry_h)
    probe_out = torch.empty(148 * 8 * 256, dtype=torch.float64, device=dev)
    iters = 200000
    for _ in range(2):
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record()
        _capi.check(lib.sg_probe_dfma(probe_out.data_ptr(), 148 * 8, 256, iters, torch.cuda.current_stream().cuda_stream))
        p1.record()
        torch.cuda.synchronize()
        probe_ms = p0.elapsed_time(p1)
    peak_tf = 148 * 8 * 256 * iters * 16 / (probe_ms * 1e-3) / 1e12
    ach_tf = fl / (kern_ms * 1e-3) / 1e12
    # algorithmic HBM bytes of one pass: noise, meals, statistics once; state in / out, controller rows, patient row and
    # launch order once per launch (the parameter table itself is 8 KB and stays in L2)
    n_launch_b = (n_steps + coherence - 1) // coherence if coherence > 0 else 1
    hbm_bytes = E * ((n_steps + 2) * 8 + mm.shape[0] * 12 + 2 * 16 * 8 +
                     n_launch_b * (2 * (24 * 8 + 5 * 4) + 4 * 8 + 4 + (4 if coherence > 0 else 0)))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    traffic, traffic_note = measured_traffic(args.groups, args.days, coherence)
    n_launch = (n_steps + coherence - 1) // coherence if coherence > 0 else 1
    roofline = {'bound': 'fp64', 'achieved': ach_tf, 'peak': peak_tf, 'unit': 'TFLOP/s', 'frac': ach_tf / peak_tf,
                'traffic': traffic, 'traffic_unit': 'bytes per pass' if traffic else None, 'traffic_source': traffic_note,
                'kernel': 'run_kernel', 'kernel_ms': kern_ms, 'kernel_sha16': kernel_sha16(),
                'launches_per_pass': n_launch,
                'per': 'pass = %d run_kernel launch(es) of %s env-steps over all episodes; achieved = algorithmic flops of '
                       'the pass / summed launch time (CUDA events around eng.run, which also holds the %d sort launches)'
                       % (n_launch, coherence if coherence > 0 else n_steps, 3 * n_launch if coherence > 0 else 0),
                'peak_source': 'DFMA chain probe timed in this run (sg_probe_dfma); MEASURED_PEAKS.json has no FP64 entry',
                'algorithmic_flops_per_launch': fl, 'dopri_steps_per_minute': float(a[14] / (E * n_min)),
                'hbm': {'achieved_gbs': hbm_bytes / (kern_ms * 1e-3) / 1e9, 'peak_gbs': peaks.get('hbm_gbs', 6650.0),
                        'peak_source': 'MEASURED_PEAKS.json' if peaks else 'fallback',
                        'algorithmic_bytes_per_pass': hbm_bytes}}

    if args.fp32:                                     # the optional mode runs on the FP32 pipe: its own (nominal) ceiling
        sm_ghz = 1.965
        roofline.update({'bound': 'fp32', 'peak': 148 * 128 * 2 * sm_ghz / 1e3, 'frac': ach_tf / (148 * 128 * 2 * sm_ghz / 1e3),
                         'kernel': 'run_fp32_kernel',
                         'peak_source': 'nominal 148 SMs x 128 FP32 lanes x 2 flop x 1.965 GHz (optional mode, not graded)'})

    # ---- end to end: host buffers in, statistics out, through the public API ----
    e2e = None
    if not args.no_e2e:
        # the population lives in pinned host memory as shards of whole seed groups (public API: engine.ShardRunner,
        # which uploads shard k+1 on one stream while shard k computes on another)
        from simglucose_b200.engine import ShardRunner
        n_sh = max(1, min(args.shards or int(round(args.groups / 1262.0)), args.groups))
        cuts = [30 * (args.groups * i // n_sh) for i in range(n_sh + 1)]
        shards, h2d = [], 0
        for lo, hi in zip(cuts[:-1], cuts[1:]):
            sh = {'patient_row': rows[lo:hi], 'noise': noise[:, lo:hi], 'meal_minute': mm[:, lo:hi],
                  'meal_grams': mg[:, lo:hi], 'ctrl_rows': ctrl_rows[:, lo:hi]}
            sh = {k: torch.empty(v.shape, dtype=v.dtype).pin_memory().copy_(v) for k, v in sh.items()}
            h2d += sum(v.numel() * v.element_size() for v in sh.values())
            shards.append(sh)
        stats_h = [torch.empty(_capi.NSTAT, hi - lo, dtype=torch.float64).pin_memory()
                   for lo, hi in zip(cuts[:-1], cuts[1:])]
        d2h = sum(t.numel() * 8 for t in stats_h) + n_sh * _capi.NSTAT * 8
        e2e_agg = [None]
        runner = ShardRunner(max(hi - lo for lo, hi in zip(cuts[:-1], cuts[1:])), device=local, slots=2,
                             carry_h=args.carry_h, fma=args.fma, fp32=args.fp32, coherence=coherence_e2e)

        def e2e_pass():
            _, e2e_agg[0] = runner.run(shards, n_steps, controller='bb', stats_out=stats_h)
            torch.cuda.synchronize()
        e2e_pass()
        barrier()
        t0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        n_e2e = max(1, min(args.steps, 2))
        for _ in range(n_e2e):
            e2e_pass()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        e2e = {'value': world * E * n_min * n_e2e / (float(ms.item()) * 1e-3), 'unit': METRIC,
               'h2d_bytes_per_step': int(h2d), 'd2h_bytes_per_step': int(d2h), 'steps': n_e2e,
               'wall_s': time.perf_counter() - t0, 'api': 'simglucose_b200.engine.ShardRunner.run',
               'shards': n_sh, 'coherence_steps': coherence_e2e, 'rows_check': float(e2e_agg[0][0])}
        # ---- the same with the trajectories copied out: BG, CGM, CHO, insulin of every env-step (the columns of the
        # reference's result DataFrame, sim_engine.py:44-49) into pinned host memory, for the first shards ----
        if not args.no_traces and world == 1:
            try:
                n_tr = min(n_sh, 3)
                tr_h = [torch.empty(n_steps, 4, hi - lo, dtype=torch.float64).pin_memory()
                        for lo, hi in zip(cuts[:n_tr], cuts[1:n_tr + 1])]
                ep_tr = cuts[n_tr] - cuts[0]

                def tr_pass():
                    runner.run(shards[:n_tr], n_steps, controller='bb', stats_out=stats_h[:n_tr], traces_out=tr_h)
                    torch.cuda.synchronize()
                tr_pass()
                barrier()
                t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0.record()
                tr_pass()
                t1.record()
                barrier()
                ms = torch.tensor([t0.elapsed_time(t1)], dtype=torch.float64, device=dev)
                if world > 1:
                    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
                e2e['traces'] = {'value': world * ep_tr * n_min / (float(ms.item()) * 1e-3), 'unit': METRIC,
                                 'episodes_per_gpu': ep_tr, 'shards': n_tr, 'columns': ['BG', 'CGM', 'CHO', 'insulin'],
                                 'h2d_bytes': int(sum(v.numel() * v.element_size() for sh in shards[:n_tr] for v in sh.values())),
                                 'd2h_bytes': int(sum(t.numel() * 8 for t in tr_h) + sum(t.numel() * 8 for t in stats_h[:n_tr])),
                                 'what': 'ShardRunner.run(traces_out=...): host inputs in, per-episode statistics AND the '
                                         'trajectories (one row per env-step) out, device -> host copy overlapped with the '
                                         'next shard; bounded to the first %d shards' % n_tr,
                                 'bg_check': float(tr_h[0][-1, 0].mean())}
                del tr_h
            except Exception as ex:                          # noqa: BLE001  (pinned host memory is the usual suspect)
                e2e['traces'] = {'error': repr(ex)}
        runner.close()
        del shards, stats_h

    population = summarize(agg_all)
    headline_vec = vector_hex(agg_all)
    eng.close()
    del noise, mm, mg, eng
    torch.cuda.empty_cache()

    # ---- the other BASELINE.json configurations: [2], [3] and [4] at their stated sizes, [0] through the
    # single-environment facade ----
    configs = None
    if not args.no_configs:
        configs = {}
        # configs[3] is the multi-GPU configuration: a reduced population (262,144 episodes x 1 day) sharded over ALL
        # ranks with the statistics all-reduce inside its timed region, at every N
        # at its STATED size by default (16,777,216 episodes x 7 days: 97 / 48 / 24 / 12 s on 1 / 2 / 4 / 8 GPUs), after a
        # small warm-up; --c3-episodes / --c3-days shrink it for quick runs
        coh3 = 40 if args.coherence < 0 else args.coherence
        c3_total, c3_days = int(args.c3_episodes), int(args.c3_days)
        config3(local, rank, world, 262144, 1, 2097152, coh3)                      # warm-up
        sec3, agg3, failed3, launches3, gen3, _ = config3(local, rank, world, c3_total, c3_days, args.wave, coh3)
        t3 = torch.tensor([sec3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t3, op=dist.ReduceOp.MAX)
        configs['configs[3]'] = {
            'config': 'configs[3]%s: %d episodes with +-10 %% perturbed parameters (Philox seed 1234, keyed by global '
                      'episode index) x %d day(s), BBController (nominal basal by name), independent streams, sharded '
                      'over %d GPU(s) (shard_range) in resident waves of %d, distributed.allreduce_stats inside the timed '
                      'region; the integer entries and extrema of population_vector_hex do not depend on the number of '
                      'GPUs (tools/check_scale_vectors.py)'
                      % (' at its stated size' if (c3_total, c3_days) == (16777216, 7) else ' reduced', c3_total, c3_days,
                         world, args.wave),
            'episodes': c3_total, 'days': c3_days,
            'value': c3_total * 1440.0 * c3_days / float(t3.item()), 'unit': METRIC, 'seconds': float(t3.item()), 'n_gpus': world,
            'coherence_steps': coh3, 'solver_failures': failed3, 'library_launches_rank0': launches3,
            'includes': 'parameter / stream generation on device (%.3f s on rank 0), reset, episode kernel, reduction, '
                        'all-reduce' % gen3,
            'population': summarize(agg3), 'population_vector_hex': vector_hex(agg3)}
        if world == 1:
            try:
                configs['configs[0]'] = config0(local)
            except Exception as e:                              # noqa: BLE001
                configs['configs[0]'] = {'error': repr(e)}
            torch.cuda.empty_cache()
            configs['configs[2]'] = config2(local, envs=1048576, n_steps=480)              # at its stated size: 1.3 s
            torch.cuda.empty_cache()
            configs['configs[4]'] = config4(local, seeds=10000)
            torch.cuda.empty_cache()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        sys.path.insert(0, os.path.join(ROOT, 'oracle'))
        import oracle as O
        ns = min(E, 30 * 512)
        nz = S.device_noise(ns, n_steps + 2, 'Dexcom', seed=1, episode_offset=ep0, group_size=30, device=dev)
        m2, g2 = S.device_scenarios(ns, n_min, start_minute=0, seed=1, episode_offset=ep0, group_size=30, device=dev)
        cpu = cpu_sample(args, nz.cpu().numpy(), m2.cpu().numpy(), g2.cpu().numpy(), rows_h[:ns], n_steps,
                         O.max_threads(), args.cpu_seconds)
        if not args.no_ref_python:
            cpu['reference_python'] = reference_python(args.ref_python_hours)

    if rank == 0:
        line = {'metric': METRIC, 'value': value, 'unit': METRIC, 'n_gpus': world, 'steps': args.steps,
                'warmup': args.warmup, 'ms_per_step': total_ms / args.steps, 'higher_is_better': True,
                'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32' if args.fp32 else 'f64', 'data': 'synthetic',
                'config': {'workload': 'configs[1] cohort: 30 virtual patients x %d seed groups per GPU x %d-day '
                                       'RandomScenario, BBController, Dexcom, Insulet' % (args.groups, args.days),
                           'episodes_per_gpu': E, 'minutes_per_episode': n_min, 'arithmetic': ('fp32 integrator (optional mode, looser bound)' if args.fp32 else
                                          'fma' if args.fma else
                                          'strict part: patient model RHS and Runge-Kutta stage sums, every operation rounded '
                                          'separately in the reference\'s order (no FMA contraction); <= 2 ulp part: scaled norms '
                                          'of HINIT and of the error estimate (reciprocal products), step-size controller (exp of '
                                          'log), tanh bracket of the gastric-emptying law, risk index'),
                           'carry_h': bool(args.carry_h), 'coherence_steps': coherence,
                           'stream_generation_s_rank0': stream_generation_s,
                           'stream_generation': 'CGM noise and meal schedules of the whole population, generated on device '
                                                'once before the timed region (not part of value / e2e)',
                           'parallelism': 'episodes sharded, dp%d; distributed.allreduce_stats (NCCL) once per step inside '
                                          'the timed region' % world,
                           'l2': 'inputs (%.1f GB noise per GPU) larger than L2' % ((n_steps + 2) * E * 8 / 1e9)},
                'roofline': roofline, 'cpu_baseline': cpu, 'e2e': e2e, 'gpu_launches': int(launches),
                'clocks': clocks, 'population': population, 'population_vector_hex': headline_vec, 'configs': configs}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
