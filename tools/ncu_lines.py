#!/usr/bin/env python
"""Per-source-line view of an ncu report captured with --import-source on (kernels built with -lineinfo):

    ncu -i prof.ncu-rep --page source --print-source cuda,sass --csv > src.csv
    python tools/ncu_lines.py src.csv [top]

Prints the share of warp-stall samples and of executed warp instructions per CUDA source line."""
import collections
import csv
import sys


def num(s):
    try:
        return int(s)
    except ValueError:
        return 0


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 50
    cur, agg = None, []
    for r in rows:
        if len(r) == 2 and r[0] == 'File Path':
            cur = r[1].split('/')[-1]
        elif r and r[0].isdigit():
            agg.append((cur, int(r[0]), r[1].strip()[:100], num(r[4]), num(r[7])))
    tot_s = sum(a[3] for a in agg) or 1
    tot_i = sum(a[4] for a in agg) or 1
    print('samples %d, warp instructions executed %d' % (tot_s, tot_i))
    byfile = collections.Counter()
    for a in agg:
        byfile[a[0]] += a[3]
    print(', '.join('%s %.1f%%' % (k, 100.0 * v / tot_s) for k, v in byfile.most_common()))
    for a in sorted(agg, key=lambda a: -a[3])[:top]:
        print('%-16s %4d  smp %5.2f%%  inst %5.2f%%  %s' % (a[0], a[1], 100.0 * a[3] / tot_s, 100.0 * a[4] / tot_i, a[2]))


if __name__ == '__main__':
    main()
