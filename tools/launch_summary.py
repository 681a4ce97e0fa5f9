#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (ncu --metrics gpu__time_duration.sum --csv --log-file launches.csv ...):
    python tools/launch_summary.py launches.csv [tail_fraction]
tail_fraction: only the last part of the list (the timed loop of a probe), default all."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi = hdr.index('Kernel Name'), hdr.index('Metric Value')
data = rows[1:]
frac = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
data = data[-max(1, int(len(data) * frac)):]
tot, cnt = collections.OrderedDict(), collections.Counter()
for r in data:
    k = r[ki][:70]
    tot[k] = tot.get(k, 0.0) + float(r[vi].replace(',', ''))
    cnt[k] += 1
for k, v in sorted(tot.items(), key=lambda x: -x[1]):
    print('%-72s n=%4d total %9.1f us  avg %8.1f us' % (k, cnt[k], v / 1e3, v / 1e3 / cnt[k]))
