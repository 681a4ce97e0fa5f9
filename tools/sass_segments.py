#!/usr/bin/env python
"""Segments of the integrator loop of run_kernel between unconditional branches (RHS + S2, S3, .., S7 + control, ...):
instructions, FP64-pipe instructions and issue cycles by the static stall counts.
    python tools/sass_segments.py [object] [kernel]"""
import collections
import re
import sys

sys.path.insert(0, __file__.rsplit('/', 1)[0])
import sass_sched as S

obj = sys.argv[1] if len(sys.argv) > 1 else 'simglucose_b200/lib/obj/sg_kernels_strict.o'
fun = sys.argv[2] if len(sys.argv) > 2 else '_ZN9sg_strict10run_kernelEN3sgk11EpisodeArgsE'
ins = S.load(obj, fun)
lo, hi = S.loops(ins)[0]
sel = [i for i in ins if lo <= i['addr'] <= hi]


def op(t):
    w = t.split()
    return (w[1] if w[0].startswith('@') else w[0]).split('.')[0]


seg, cur = [], []
for i in sel:
    cur.append(i)
    if 'BRA' in i['text'] and not i['text'].startswith('@'):
        seg.append(cur)
        cur = []
if cur:
    seg.append(cur)
print('loop %#x..%#x %d instructions' % (lo, hi, len(sel)))
for sg in seg:
    c = collections.Counter(op(i['text']) for i in sg)
    fp = sum(c[k] for k in ('DADD', 'DMUL', 'DFMA', 'DSETP'))
    print('%06x..%06x %4d instr  fp64 %3d  cycles %4d | %s' % (sg[0]['addr'], sg[-1]['addr'], len(sg), fp,
          sum(max(1, i['stall']) for i in sg), dict(c.most_common(7))))
