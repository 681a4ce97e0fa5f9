"""Parameter tables of the UVa/Padova virtual cohort (data imported by tools/import_params.py from
the reference's simglucose/params/*.csv: vpatient_params 30x62, Quest 30x5, pump_params, sensor_params)."""
import json
import os

import numpy as np

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'tables.json')
_doc = None

# the 32 columns the hot path reads, in C-ABI order (include/simglucose_b200.h: sg_param)
HOT_PARAMS = ['BW', 'u2ss', 'kmax', 'kmin', 'b', 'd', 'kabs', 'f', 'kp1', 'kp2', 'kp3', 'Fsnc', 'ke1', 'ke2',
              'k1', 'k2', 'Vm0', 'Vmx', 'Km0', 'm1', 'm2', 'm4', 'm30', 'ka1', 'ka2', 'Vi', 'p2u', 'Ib', 'ki',
              'kd', 'ksc', 'Vg']


def tables():
    global _doc
    if _doc is None:
        with open(_PATH) as f:
            _doc = json.load(f)
    return _doc


def _rows(name):
    t = tables()[name]
    return t['columns'], t['rows']


def patient_names():
    return [r[0] for r in _rows('vpatient_params')[1]]


def patient_row(name_or_id):
    """0-based table row from a name ('adult#003') or a 1-based id (T1DPatient.withID semantics)."""
    if isinstance(name_or_id, str):
        return patient_names().index(name_or_id)
    return int(name_or_id) - 1


def patient_record(row):
    """All 62 CSV columns of a row as a dict (the pandas Series the reference passes around)."""
    cols, rows = _rows('vpatient_params')
    return dict(zip(cols, rows[row]))


def cohort_table():
    """(params [30, 32], x0 [30, 13]) float64."""
    cols, rows = _rows('vpatient_params')
    idx = [cols.index(c) for c in HOT_PARAMS]
    tbl = np.array([[r[i] for i in idx] for r in rows], dtype=np.float64)
    x0 = np.array([r[2:15] for r in rows], dtype=np.float64)
    return tbl, x0


def quest_table():
    """[30, 4] float64: CR, CF, Age, TDI in vpatient_params row order (Quest.csv has the same order)."""
    cols, rows = _rows('Quest')
    names = patient_names()
    by_name = {r[0]: r for r in rows}
    return np.array([[by_name[n][cols.index(c)] for c in ('CR', 'CF', 'Age', 'TDI')] for n in names], dtype=np.float64)


def quest_record(name):
    cols, rows = _rows('Quest')
    for r in rows:
        if r[0] == name:
            return dict(zip(cols, r))
    return None


def sensor_params(name='Dexcom'):
    cols, rows = _rows('sensor_params')
    for r in rows:
        if r[0] == name:
            d = dict(zip(cols, r))
            d['sample_time'] = int(d['sample_time'])      # an integer column in sensor_params.csv
            return d
    raise KeyError(name)


def pump_params(name='Insulet'):
    cols, rows = _rows('pump_params')
    for r in rows:
        if r[0] == name:
            return dict(zip(cols, r))
    raise KeyError(name)


def sensor_names():
    return [r[0] for r in _rows('sensor_params')[1]]


def pump_names():
    return [r[0] for r in _rows('pump_params')[1]]


def patient_series(row):
    """The pandas Series the reference hands to T1DPatient (one row of vpatient_params.csv, all 62
    columns in file order, so that params.iloc[2:15] is x0 and params.BW etc. work)."""
    import pandas as pd
    cols, rows = _rows('vpatient_params')
    return pd.Series(rows[row], index=cols, name=row)
