#!/usr/bin/env python
"""profiles/run_kernel_traffic.json from an ncu capture of the run_kernel launches of ONE pass of the default bench
workload (ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -k regex:run_kernel
-s <launches per pass> -c <launches per pass> --csv --log-file X.csv python bench.py --steps 1 --warmup 1 ...):
    python tools/traffic_from_ncu.py X.csv <groups> <days> <coherence> <name of the committed copy under profiles/>"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

src, groups, days, coh, name = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), sys.argv[5]
rows = [r for r in csv.reader(open(src)) if len(r) > 10 and r[0] != 'ID' and 'run_kernel' in r[4]]
tot = {}
ids = set()
for r in rows:
    tot[r[12]] = tot.get(r[12], 0) + int(float(r[14].replace(',', '')))
    ids.add(r[0])
n_steps = days * 480
expect = (n_steps + coh - 1) // coh if coh > 0 else 1
assert len(ids) == expect, 'captured %d launches, one pass has %d' % (len(ids), expect)
t = {'dram_bytes_per_pass': tot['dram__bytes_read.sum'] + tot['dram__bytes_write.sum'], 'read': tot['dram__bytes_read.sum'],
     'write': tot['dram__bytes_write.sum'], 'kernel_ns_under_ncu': tot['gpu__time_duration.sum'],
     'launches_per_pass': len(ids), 'groups': groups, 'days': days, 'coherence_steps': coh,
     'kernel_sha16': bench.kernel_sha16(), 'source': 'profiles/' + name}
json.dump(t, open(os.path.join(ROOT, 'profiles', 'run_kernel_traffic.json'), 'w'), indent=1)
print(json.dumps(t))
