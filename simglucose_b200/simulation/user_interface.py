"""simulate(...) -- the programmatic (keyword) form of simglucose/simulation/user_interface.py:303-385.

The reference's interactive input() prompts are a CLI and are not reproduced: every argument the reference
would have prompted for must be given (a missing one raises ValueError naming it).  The work itself --
one env per patient, deep copies of the scenario and the controller, batch_sim, report -- follows the
reference; batch_sim runs all patients as one population on the GPU.
"""
import copy
from datetime import datetime

import pandas as pd

from ..actuator.pump import InsulinPump
from ..analysis.report import report
from ..patient.t1dpatient import T1DPatient
from ..sensor.cgm import CGMSensor
from .env import T1DSimEnv
from .sim_engine import SimObj, batch_sim


def simulate(sim_time=None, scenario=None, controller=None, patient_names=[], cgm_name=None, cgm_seed=None,
             insulin_pump_name=None, start_time=None, save_path=None, animate=None, parallel=None):
    """sim_time: timedelta; scenario: a Scenario; controller: a Controller; patient_names: list of names;
    cgm_name / cgm_seed / insulin_pump_name: hardware; save_path: directory for <patient>.csv and the report
    CSVs.  Returns the performance_stats DataFrame (percent in range + mean risk indices per patient)."""
    required = dict(sim_time=sim_time, scenario=scenario, controller=controller, cgm_name=cgm_name,
                    cgm_seed=cgm_seed, insulin_pump_name=insulin_pump_name, save_path=save_path)
    missing = [k for k, v in required.items() if v is None] + ([] if patient_names else ['patient_names'])
    if missing:
        raise ValueError('simulate(): no interactive prompts in this build; missing arguments: ' + ', '.join(missing))
    animate = False if animate is None else animate           # headless: render() is a no-op either way
    parallel = True if parallel is None else parallel
    if start_time is not None and not isinstance(start_time, datetime):
        raise ValueError('start_time must be a datetime')

    def build_env(pname):
        scen = copy.deepcopy(scenario)
        return T1DSimEnv(T1DPatient.withName(pname), CGMSensor.withName(cgm_name, seed=cgm_seed),
                         InsulinPump.withName(insulin_pump_name), scen)
    envs = [build_env(p) for p in patient_names]
    ctrllers = [copy.deepcopy(controller) for _ in envs]
    sim_instances = [SimObj(e, c, sim_time, animate=animate, path=save_path) for e, c in zip(envs, ctrllers)]
    results = batch_sim(sim_instances, parallel=parallel)
    df = pd.concat(results, keys=[s.env.patient.name for s in sim_instances])
    results, ri_per_hour, zone_stats, figs, axes = report(df, save_path)
    return results
