#!/usr/bin/env python
"""cProfile of the drop-in call batch_sim(30 SimObj x 3 days) (BASELINE.json configs[0] / bench.config0): where the
0.8 s of host time go.  python tools/profile_batch_sim.py [top]"""
import contextlib
import cProfile
import io
import os
import pstats
import sys
import time
from datetime import datetime, timedelta

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from simglucose_b200 import params as P  # noqa: E402
from simglucose_b200.actuator.pump import InsulinPump  # noqa: E402
from simglucose_b200.controller.basal_bolus_ctrller import BBController  # noqa: E402
from simglucose_b200.patient.t1dpatient import T1DPatient  # noqa: E402
from simglucose_b200.sensor.cgm import CGMSensor  # noqa: E402
from simglucose_b200.simulation.env import T1DSimEnv  # noqa: E402
from simglucose_b200.simulation.scenario_gen import RandomScenario  # noqa: E402
from simglucose_b200.simulation.sim_engine import SimObj, batch_sim  # noqa: E402

start = datetime(2018, 1, 1, 0, 0, 0)


def build():
    return [SimObj(T1DSimEnv(T1DPatient.withName(n), CGMSensor.withName('Dexcom', seed=1), InsulinPump.withName('Insulet'),
                             RandomScenario(start_time=start, seed=1)), BBController(), timedelta(days=3), animate=False,
                   path=None) for n in P.patient_names()]


def run(sims):
    with contextlib.redirect_stdout(io.StringIO()):
        return batch_sim(sims, parallel=True)


run(build())                                               # warm-up: library load, CUDA context
t0 = time.perf_counter(); sims = build(); t1 = time.perf_counter()
pr = cProfile.Profile()
pr.enable()
frames = run(sims)
pr.disable()
t2 = time.perf_counter()
print('build 30 SimObj: %.3f s; batch_sim: %.3f s (under cProfile); rows %d' % (t1 - t0, t2 - t1, sum(len(f) for f in frames)))
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats('cumulative').print_stats(int(sys.argv[1]) if len(sys.argv) > 1 else 35)
print(s.getvalue()[:6000])
