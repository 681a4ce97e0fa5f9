"""Import the UVa/Padova parameter TABLES (data, not code) from the reference.

Reads simglucose/params/{vpatient_params,Quest,pump_params,sensor_params}.csv
of the reference checkout with pandas -- the same parser the reference uses
(patient/t1dpatient.py:54,66; controller/basal_bolus_ctrller.py:22-23;
actuator/pump.py:21; sensor/cgm.py:23) so every float64 is bit-identical to
what the reference computes with -- and writes them as one JSON document with
repr()-exact floats to simglucose_b200/params/tables.json.

Run in the build container only:  python tools/import_params.py
"""
import json
import os
import sys

import pandas as pd

REF = os.environ.get('SIMGLUCOSE_REFERENCE', '/root/reference')
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), '..',
                   'simglucose_b200', 'params', 'tables.json')


def table(name):
    df = pd.read_csv(os.path.join(REF, 'simglucose', 'params', name))
    rows = []
    for _, r in df.iterrows():
        rows.append([v if isinstance(v, str) else float(v) for v in r.tolist()])
    return {'columns': [str(c) for c in df.columns], 'rows': rows}


def main():
    doc = {
        'source': 'Sawyerbatch/simglucose simglucose/params/*.csv (UVa/Padova 2008 virtual cohort)',
        'vpatient_params': table('vpatient_params.csv'),
        'Quest': table('Quest.csv'),
        'pump_params': table('pump_params.csv'),
        'sensor_params': table('sensor_params.csv'),
    }
    with open(OUT, 'w') as f:
        json.dump(doc, f, indent=0)
    back = json.load(open(OUT))
    assert back == doc
    print('wrote', os.path.normpath(OUT), {k: len(v['rows']) for k, v in doc.items() if isinstance(v, dict)})


if __name__ == '__main__':
    sys.exit(main())
