"""Where a lockstep env-step of the vector environment spends its time (configs[2]): CUDA-event timing of
  (a) the bare step kernel loop (engine.run(1, 'random')), (b) T1DSimVecEnv.step_random without resets,
  (c) with gym seeding + per-environment resets; run under `ncu --metrics gpu__time_duration.sum` for the launch list.
    python tools/lockstep_probe.py [envs] [steps] [mode]"""
import sys
import time

import torch

sys.path.insert(0, '/root/repo')
from simglucose_b200.envs import T1DSimVecEnv  # noqa: E402

E = int(sys.argv[1]) if len(sys.argv) > 1 else 1048576
n_steps = int(sys.argv[2]) if len(sys.argv) > 2 else 60
modes = sys.argv[3].split(',') if len(sys.argv) > 3 else ['kernel', 'plain', 'gym']


def timed(fn):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record(); fn(); e1.record(); t1 = time.perf_counter()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1), (t1 - t0) * 1e3


for mode in modes:
    # 'gym@4a': the gym environment with the launch order re-sorted every 4 env-steps in adaptive mode ('gym@40': by meals)
    mode, _, sort = mode.partition('@')
    gym = mode == 'gym'
    kw = {}
    if sort:
        kw = dict(coherence=int(sort.rstrip('a')), adaptive_sort=sort.endswith('a'))
    env = T1DSimVecEnv(E, seed=0, horizon_steps=480, gym_seeding=gym, auto_reset=gym, **kw)
    mode = mode + ('@' + sort if sort else '')
    env.reset()
    for _ in range(125):                                  # past the first sorts and into the steady state of the resets
        env.step_random(0.0, 0.05)
    if mode == 'kernel':
        def loop():
            for _ in range(n_steps):
                env.engine.run(1, controller='random', ctrl=(0.0, 0.05, 0.0, 0.0), trace=env._row)
    else:
        def loop():
            for _ in range(n_steps):
                env.step_random(0.0, 0.05)
    ms, host = timed(loop)
    print('%-6s %d envs: %.3f ms per env-step on the device, %.3f ms of host time per step, %.3g patient-min/s'
          % (mode, E, ms / n_steps, host / n_steps, E * n_steps * 3 / ms * 1e3), flush=True)
    env.close()
