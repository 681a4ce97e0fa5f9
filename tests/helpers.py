"""Shared helpers of the test-suite: golden fixture access and oracle access (tests only)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, 'tests', 'golden')
sys.path.insert(0, os.path.join(ROOT, 'oracle'))

import oracle as O  # noqa: E402  (test infrastructure)

REL_TOL = 1e-6      # BASELINE.json: max |dBG| / max(|BG|, 1 mg/dL) <= 1e-6 in FP64


def load(name):
    return np.load(os.path.join(GOLDEN, name + '.npz'))


def rel_err(a, b):
    """max |a-b| / max(|b|, 1) ignoring positions where both are NaN (1 mg/dL floor)."""
    a, b = np.asarray(a, float), np.asarray(b, float)
    both_nan = np.isnan(a) & np.isnan(b)
    assert not np.any(np.isnan(a) ^ np.isnan(b)), 'NaN pattern differs'
    d = np.where(both_nan, 0.0, np.abs(a - b) / np.maximum(np.abs(np.where(both_nan, 0.0, b)), 1.0))
    return float(d.max()) if d.size else 0.0


def controller_of(d, i):
    """(kind, ctrl tuple) of episode i of a fixture, in engine/oracle convention."""
    kind = str(d['controller_kind'][i])
    cp = d['controller_params'][i]
    row = int(d['patient_row'][i])
    if kind == 'bb':
        cr, cf = O.quest(row)
        if 'params' in d.files:
            # perturbed parameter rows: the reference's BBController still finds u2ss and BW by patient NAME
            # (basal_bolus_ctrller.py:52-64), so the nominal basal travels with the controller constants
            nominal, _ = O.patient_params(row)
            return 'bb', (float(cp[0]), cr, cf, float(nominal[1] * nominal[0] / 6000.0))
        return 'bb', (float(cp[0]), cr, cf)
    if kind.startswith('pi'):
        return 'pid', tuple(float(v) for v in cp[:4])
    return 'actions', (0.0,)


def hardware_of(d, i):
    sn = str(d['sensor_name'][i]) if 'sensor_name' in d.files else 'Dexcom'
    pn = str(d['pump_name'][i]) if 'pump_name' in d.files else 'Insulet'
    return sn, pn


def oracle_episode(d, i, carry_h=0):
    row = int(d['patient_row'][i])
    p, _ = O.patient_params(row)
    if 'params' in d.files:
        p = d['params'][i]                      # the parameter Series the reference patient was built on
    kind, ctrl = controller_of(d, i)
    sn, pn = hardware_of(d, i)
    mm, mg = d['meal_minute'][i], d['meal_grams'][i]
    k = mm >= 0
    return O.run_episode(p, d['x0'][i], int(d['n_steps'][i]), mm[k], mg[k], d['noise'][i], controller=kind,
                         ctrl=ctrl, actions=d['actions'][i], sensor_name=sn, pump_name=pn, carry_h=carry_h,
                         t0=int(d['t0'][i]) if 't0' in d else 0)
