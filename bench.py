#!/usr/bin/env python
"""bench.py -- patient-sim-minutes/s of the batched closed loop on N B200s (one process per GPU).

Workload (BASELINE.json configs[1], filled out to occupy a GPU as SURVEY.md 8d "Config 2" says): the 30
virtual patients x R scenario/sensor seed groups x 3 simulated days, Dexcom + Insulet, BBController with
the Quest.csv CR/CF, RandomScenario-distributed meals and CGMNoise-distributed sensor noise (Philox
streams generated on device; the 30 patients of one seed group share their streams exactly as
RandomScenario(seed=s) / CGMSensor(seed=s) do in the reference).  One "step" = one pass over all
episodes: reset + 1440 fused env-steps + population statistics.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for the definition of every key.
"""
import argparse
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
    ap.add_argument('--groups', type=int, default=8840,
                    help='seed groups of 30 patients per GPU (8840 x 30 episodes = 7 waves of 148 SMs x 256 resident threads)')
    ap.add_argument('--days', type=int, default=3)
    ap.add_argument('--fma', action='store_true', help='SG_FLAG_FMA kernels (not the default arithmetic)')
    ap.add_argument('--carry-h', action='store_true')
    ap.add_argument('--fp32', action='store_true', help='optional FP32 mode (SG_FLAG_FP32): NOT the graded arithmetic')
    ap.add_argument('--cpu-seconds', type=float, default=15.0, help='target duration of the cpu_baseline sample')
    ap.add_argument('--no-cpu', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--shards', type=int, default=0,
                    help='host-resident shards of the e2e leg (engine.run_shards); 0 = one resident wave each '
                         '(148 SMs x 256 episodes = 1262 seed groups)')
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
# workload
# ---------------------------------------------------------------------------------------------
def controller_rows(rows_np):
    from simglucose_b200 import params as P
    q = P.quest_table()
    return np.stack([np.full(len(rows_np), 140.0), q[rows_np, 0], q[rows_np, 1], np.zeros(len(rows_np))])


def flops_per_pass(n_minutes_total, dopri_steps_total, carry_h):
    return n_minutes_total * ((F_INIT_CARRY if carry_h else F_INIT) + F_MISC) + F_STEP * dopri_steps_total


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


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path on the box's host cores.  The
    reference is pure Python and cannot travel to the GPU box, so this is the C oracle port
    (oracle/sg_oracle.c, cpu_baseline.kind 'port') on all host threads.  Each step is a bounded sample of
    the bench workload (same patients, controller, scenario / noise distributions, generated on the host
    with the reference-compatible MT19937 generators), sized by a calibration pass to a few seconds."""
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
    line = {'impl': 'reference', 'metric': METRIC, 'value': v, 'unit': METRIC, 'n_gpus': args.gpus,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * sec,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': {'workload': 'configs[1] cohort: 30 virtual patients x seed groups x %d-day RandomScenario, '
                                   'BBController, Dexcom, Insulet (bounded CPU sample: %d seed groups)' % (args.days, groups),
                       'episodes': w['E'], 'minutes_per_episode': n_min, 'carry_h': bool(args.carry_h)},
            'cpu_baseline': cpu,
            'e2e': {'value': v, 'unit': METRIC, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}
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

    E = args.groups * 30
    n_steps = args.days * 480
    n_min = n_steps * 3
    ep0 = rank * E                                    # weak scaling: every rank simulates its own E episodes
    rows = (torch.arange(E, device=dev, dtype=torch.int64) % 30).to(torch.int32)
    rows_h = rows.cpu().numpy()
    ctrl_rows = torch.from_numpy(controller_rows(rows_h)).to(dev)
    # inputs resident in HBM before the timed region
    noise = S.device_noise(E, n_steps + 2, 'Dexcom', seed=1, episode_offset=ep0, group_size=30, device=dev)
    mm, mg = S.device_scenarios(E, n_min, start_minute=0, seed=1, episode_offset=ep0, group_size=30, device=dev)
    eng = BatchEngine(E, device=local, carry_h=args.carry_h, fma=args.fma, fp32=args.fp32)
    agg = torch.zeros(_capi.NSTAT, dtype=torch.float64, device=dev)

    def one_pass():
        eng.reset(rows, noise, mm, mg, row0=False)
        eng.run(n_steps, controller='bb', ctrl_rows=ctrl_rows, trace=False)
        eng.reduce_stats(agg)

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
    ev[-1].record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = eng.launches - launches0
    total_ms = ev[0].elapsed_time(ev[-1])
    kern_ms = float(np.mean([ev[2 * k + 1].elapsed_time(ev[2 * k + 2]) for k in range(args.steps)]))
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    agg_all = agg.clone()
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        # the one collective of the path: population statistics (sums; BG extrema separately)
        mn, mx = agg[2:3].clone(), agg[3:4].clone()
        dist.all_reduce(agg_all, op=dist.ReduceOp.SUM)
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        agg_all[2], agg_all[3] = mn[0], mx[0]
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
    # algorithmic HBM bytes of one pass: state in/out, noise, meals, params, stats
    hbm_bytes = E * (2 * (24 * 8 + 5 * 4) + 2 * 16 * 8 + (n_steps + 2) * 8 + mm.shape[0] * 12 + 32 * 8 + 4 * 8 + 4)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    # DRAM bytes of one run_kernel launch from the committed ncu capture (profiles/r01_s60_run_kernel_dram_traffic.csv:
    # dram__bytes_read.sum + dram__bytes_write.sum = 3,264,205,824 + 96,096,000 on the default workload -- 0.996 of the
    # algorithmic bytes below; the traffic does not depend on the arithmetic mode).  Other workloads: not captured.
    default_workload = (args.groups == 8840 and args.days == 3)
    traffic = 3264205824 + 96096000 if default_workload else None
    roofline = {'bound': 'fp64', 'achieved': ach_tf, 'peak': peak_tf, 'unit': 'TFLOP/s', 'frac': ach_tf / peak_tf,
                'traffic': traffic, 'traffic_unit': 'bytes per launch (ncu, session s60)' if traffic else None,
                'kernel': 'run_kernel', 'kernel_ms': kern_ms,
                'peak_source': 'DFMA chain probe timed in this run (sg_probe_dfma); MEASURED_PEAKS.json has no FP64 entry',
                'algorithmic_flops_per_launch': fl, 'dopri_steps_per_minute': float(a[14] / (E * n_min)),
                'hbm': {'achieved_gbs': hbm_bytes / (kern_ms * 1e-3) / 1e9, 'peak_gbs': peaks.get('hbm_gbs', 6650.0),
                        'peak_source': 'MEASURED_PEAKS.json' if peaks else 'fallback',
                        'algorithmic_bytes_per_launch': hbm_bytes}}

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
                             carry_h=args.carry_h, fma=args.fma, fp32=args.fp32)

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
               'shards': n_sh, 'rows_check': float(e2e_agg[0][0])}
        runner.close()

    cpu = None
    if rank == 0 and not args.no_cpu:
        sys.path.insert(0, os.path.join(ROOT, 'oracle'))
        import oracle as O
        ns = min(E, 30 * 512)
        cpu = cpu_sample(args, noise[:, :ns].cpu().numpy(), mm[:, :ns].cpu().numpy(), mg[:, :ns].cpu().numpy(),
                         rows_h[:ns], n_steps, O.max_threads(), args.cpu_seconds)

    if rank == 0:
        aa = agg_all.cpu().numpy()
        line = {'metric': METRIC, 'value': value, 'unit': METRIC, 'n_gpus': world, 'steps': args.steps,
                'warmup': args.warmup, 'ms_per_step': total_ms / args.steps, 'higher_is_better': True,
                'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32' if args.fp32 else 'f64', 'data': 'synthetic',
                'config': {'workload': 'configs[1] cohort: 30 virtual patients x %d seed groups per GPU x %d-day '
                                       'RandomScenario, BBController, Dexcom, Insulet' % (args.groups, args.days),
                           'episodes_per_gpu': E, 'minutes_per_episode': n_min, 'arithmetic': ('fp32 integrator (optional mode, looser bound)' if args.fp32 else
                                          'fma' if args.fma else 'strict (no FMA contraction)'),
                           'carry_h': bool(args.carry_h), 'parallelism': 'episodes sharded, dp%d' % world,
                           'l2': 'inputs (%.1f GB noise per GPU) larger than L2' % (noise.numel() * 8 / 1e9)},
                'roofline': roofline, 'cpu_baseline': cpu, 'e2e': e2e, 'gpu_launches': int(launches),
                'clocks': clocks,
                'population': {'mean_bg': float(aa[1] / aa[0]), 'min_bg': float(aa[2]), 'max_bg': float(aa[3]),
                               'time_in_range': float(aa[10] / aa[0]), 'episodes_done': float(aa[13]),
                               'rows': float(aa[0])}}
        emit(line)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
