#!/usr/bin/env python
"""Throughput of the other BASELINE.json configurations on one GPU (bench.py measures configs[1]):

  configs[2]  gym-style lockstep vector env: 1,048,576 episodes x 480 env-steps, random basal actions
              U(0, 0.05) U/min drawn on device every step, risk_diff reward (one launch per env-step)
  configs[3]  population sweep: parameters perturbed +-10 %, 7 days, BBController -- one shard chunk of
              262,144 episodes (the full 16.7 M episodes are 64 such chunks spread over the GPUs)
  configs[4]  closed-loop PID fused on device: 30 patients x 10,000 seeds x 1 day

    python tools/bench_configs.py [--scale 1.0] > profiles/rNN_configs.jsonl

One JSON line per configuration: patient-sim-minutes/s (CUDA events), kernel launches, population statistics.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from simglucose_b200 import _capi, params as P, streams as S  # noqa: E402
from simglucose_b200.distributed import summarize  # noqa: E402
from simglucose_b200.engine import BatchEngine  # noqa: E402
from simglucose_b200.envs import T1DSimVecEnv  # noqa: E402


def timed(fn, warm=1, reps=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / reps


def config2(scale, fp32=False):
    E, n_steps = int(1048576 * scale) // 64 * 64, 480
    env = T1DSimVecEnv(E, seed=0, start_hour=0, horizon_steps=n_steps, stream_mode='device', group_size=1, fp32=fp32)
    basal = torch.empty(E, dtype=torch.float64, device=env.device)
    state = {'k': 0}

    def episode():
        env.reset()
        ret = torch.zeros(E, dtype=torch.float64, device=env.device)
        for k in range(n_steps):
            S.device_uniform(E, 0.0, 0.05, seed=0, offset=state['k'] * E, out=basal)
            state['k'] += 1
            obs, reward, done, info = env.step(basal)
            ret += reward
        return ret
    l0 = env.engine.launches
    sec = timed(episode, warm=1, reps=1)
    launches = (env.engine.launches - l0) // 2
    agg = summarize(env.statistics())
    env.close()
    return {'config': 'configs[2] lockstep vector env, %d episodes x %d env-steps, basal ~ U(0, 0.05) U/min on device, '
                      'risk_diff reward, no auto-reset, step_cap %d%s' % (E, n_steps, env.step_cap,
                                                                         ', OPTIONAL FP32 INTEGRATOR (not the graded arithmetic)' if fp32 else ''),
            'value': E * n_steps * 3 / sec, 'unit': 'patient-sim-minutes/sec', 'seconds': sec, 'env_steps_per_s': E * n_steps / sec,
            'gpu_launches': int(launches), 'ms_per_env_step': 1e3 * sec / n_steps, 'population': agg}


def config3(scale, cap=0):
    E, days = int(262144 * scale) // 64 * 64, 7
    n_steps = days * 480
    dev = torch.device('cuda', 0)
    rows = (torch.arange(E, device=dev, dtype=torch.int64) % 30).to(torch.int32)
    params = S.device_perturbed_params(rows, amplitude=0.1, seed=1234)
    q = torch.from_numpy(P.quest_table()).to(dev)[rows.long()]
    ctrl_rows = torch.stack([torch.full_like(q[:, 0], 140.0), q[:, 0], q[:, 1], torch.zeros_like(q[:, 0])]).contiguous()
    noise = S.device_noise(E, n_steps + 2, 'Dexcom', seed=1234, group_size=1, device=dev)
    mm, mg = S.device_scenarios(E, n_steps * 3, start_minute=0, seed=1234, group_size=1, device=dev)
    eng = BatchEngine(E)
    eng.set_step_cap(cap)

    def run():
        eng.reset(rows, noise, mm, mg, params=params, row0=False)
        eng.run(n_steps, controller='bb', ctrl_rows=ctrl_rows, trace=False)
        return eng.reduce_stats()
    sec = timed(run, warm=1, reps=1)
    agg = summarize(eng.reduce_stats())
    failed = int((eng.status() < 0).sum().item())
    eng.close()
    return {'config': 'configs[3] perturbed population (+-10 %% per parameter, Philox seed 1234), %d episodes x %d days, '
                      'BBController, independent scenario / sensor streams per episode (one shard chunk), step_cap %d' % (E, days, cap),
            'value': E * n_steps * 3 / sec, 'unit': 'patient-sim-minutes/sec', 'seconds': sec,
            'gpu_launches': 4 if cap == 0 else 3 + 2 * n_steps, 'solver_failures': failed, 'population': agg}


def config4(scale):
    seeds, n_steps = int(10000 * scale), 480
    E = 30 * seeds
    dev = torch.device('cuda', 0)
    rows = (torch.arange(E, device=dev, dtype=torch.int64) % 30).to(torch.int32)
    noise = S.device_noise(E, n_steps + 2, 'Dexcom', seed=0, group_size=30, device=dev)
    mm, mg = S.device_scenarios(E, n_steps * 3, start_minute=0, seed=0, group_size=30, device=dev)
    eng = BatchEngine(E)

    def run():
        eng.reset(rows, noise, mm, mg, row0=False)
        eng.run(n_steps, controller='pid', ctrl=(1e-3, 1e-5, 1e-3, 140.0), trace=False)
        return eng.reduce_stats()
    sec = timed(run, warm=1, reps=2)
    agg = summarize(eng.reduce_stats())
    eng.close()
    return {'config': 'configs[4] PIDController (P 1e-3, I 1e-5, D 1e-3, target 140) fused on device, 30 patients x %d seeds x '
                      '1 day' % seeds,
            'value': E * n_steps * 3 / sec, 'unit': 'patient-sim-minutes/sec', 'seconds': sec, 'gpu_launches': 4, 'population': agg}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--scale', type=float, default=1.0)
    ap.add_argument('--only', default='')
    args = ap.parse_args()
    assert torch.cuda.is_available(), 'needs a CUDA device (there is no CPU fallback)'
    _capi.load()
    for name, fn in (('2', config2), ('2f', lambda sc: config2(sc, fp32=True)), ('3', config3), ('4', config4)):
        if args.only and name not in args.only.split(','):
            continue
        print(json.dumps(fn(args.scale)), flush=True)
        torch.cuda.empty_cache()


if __name__ == '__main__':
    main()
