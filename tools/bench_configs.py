#!/usr/bin/env python
"""Full-size runs of the BASELINE.json configurations other than the headline, on one GPU (bench.py carries reduced
versions of the same functions in its `configs` block; configs[3] at its stated size over several GPUs is
`bench.py --config 3` under torchrun):

  configs[0]  single-environment facade: T1DSimEnv.step latency, batch_sim of 30 SimObj x 3 days incl. DataFrames
  configs[2]  gym 'simglucose-v0' as a lockstep vector env: 1,048,576 environments x 480 env-steps, gym seeding per
              environment, random basal actions drawn in the step kernel, per-environment reset on done
  configs[3]  one resident wave of the perturbed population sweep: 2,097,152 episodes x 7 days (1/8 of the full sweep)
  configs[4]  closed-loop PID fused on device: 30 patients x 10,000 seeds x 1 day

    python tools/bench_configs.py [--only 0,2,2f,2plain,3,3plain,4] [--scale 1.0] > profiles/rNN_configs.jsonl

One JSON line per configuration.  "plain" variants switch the schedule-coherent launch order off (the round-1 behaviour)
so that its effect is on record.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
from simglucose_b200 import _capi  # noqa: E402
from simglucose_b200.distributed import summarize  # noqa: E402


def c3(scale, coherence, adaptive=False):
    n = int(2097152 * scale) // 64 * 64
    bench.config3(0, 0, 1, max(65536, n // 32), 1, n, coherence, adaptive=adaptive)                       # warm-up
    sec, agg, failed, launches, gen_s, _ = bench.config3(0, 0, 1, n, 7, n, coherence, adaptive=adaptive)
    return {'config': 'configs[3] one resident wave: %d episodes with +-10 %% perturbed parameters x 7 days, BBController, '
                      'independent streams, coherence_steps %d%s' % (n, coherence, ' (adaptive sort)' if adaptive else ''),
            'value': n * 10080 / sec, 'unit': bench.METRIC, 'seconds': sec, 'stream_generation_s': gen_s,
            'library_launches': launches, 'solver_failures': failed, 'population': summarize(agg),
            'population_vector_hex': bench.vector_hex(agg)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--scale', type=float, default=1.0)
    ap.add_argument('--only', default='0,2,3,4')
    args = ap.parse_args()
    assert torch.cuda.is_available(), 'needs a CUDA device (there is no CPU fallback)'
    _capi.load()
    sc = args.scale
    runs = {
        '0': lambda: bench.config0(0),
        '2': lambda: bench.config2(0, envs=int(1048576 * sc), n_steps=480),
        '2plain': lambda: bench.config2(0, envs=int(1048576 * sc), n_steps=480, coherence=0),
        '2noreset': lambda: bench.config2(0, envs=int(1048576 * sc), n_steps=480, auto_reset=False, gym_seeding=False),
        '2f': lambda: bench.config2(0, envs=int(1048576 * sc), n_steps=480, fp32=True),
        '3': lambda: c3(sc, 40),
        '3plain': lambda: c3(sc, 0),
        '3a10': lambda: c3(sc, 10, True), '3a20': lambda: c3(sc, 20, True), '3a40': lambda: c3(sc, 40, True),
        '4': lambda: bench.config4(0, seeds=int(10000 * sc)),
    }
    for name in args.only.split(','):
        print(json.dumps(runs[name]()), flush=True)
        torch.cuda.empty_cache()


if __name__ == '__main__':
    main()
