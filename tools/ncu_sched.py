#!/usr/bin/env python
"""Join an ncu SASS source page (ncu -i rep --page source --csv) with the static schedule of the same
binary (tools/sass_sched.py): dynamic issue cycles implied by the stall counts, per opcode and per
execution-frequency region.

    python tools/ncu_sched.py src.csv <object.o> <mangled kernel>
"""
import collections
import csv
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sass_sched as S  # noqa: E402


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    ia, ism, isrc = ix['Instructions Executed'], ix['# Samples'], ix['Source']
    body = [r for r in rows[2:] if len(r) > ia]
    ins = S.load(sys.argv[2], sys.argv[3])
    assert len(ins) == len(body), (len(ins), len(body))
    bad = sum(1 for a, b in zip(ins, body) if a['text'].split()[0] != b[isrc].strip().split()[0])
    assert bad == 0, 'binary differs from the profiled one (%d mismatches)' % bad
    tot_i = tot_c = tot_s = 0
    by = collections.defaultdict(lambda: [0, 0, 0])
    for a, b in zip(ins, body):
        n, s = int(b[ia] or 0), int(b[ism] or 0)
        c = n * max(1, a['stall'])
        op = S.opcode(a['text'])
        by[op][0] += n; by[op][1] += c; by[op][2] += s
        tot_i += n; tot_c += c; tot_s += s
    print('warp instructions %d, stall-count cycles %d (%.2f per instruction), samples %d' % (tot_i, tot_c, tot_c / tot_i, tot_s))
    for op, (n, c, s) in sorted(by.items(), key=lambda kv: -kv[1][1])[:24]:
        print('  %-10s %5.1f%% inst  %5.1f%% stall-cycles (%.2f/inst)  %5.1f%% samples' % (op, 100 * n / tot_i, 100 * c / tot_c, c / max(n, 1), 100 * s / tot_s))
    return ins, body


if __name__ == '__main__':
    main()
