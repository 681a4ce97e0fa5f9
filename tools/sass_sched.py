#!/usr/bin/env python
"""Static schedule of a kernel: decode the control word of every SASS instruction (stall count, yield,
write/read barrier, wait mask) from `cuobjdump -sass` and summarise issue cycles per opcode class.

    python tools/sass_sched.py <object.o> <mangled kernel name> [lo hi] [--list]

stall = cycles the scheduler holds the warp after issuing the instruction (fixed-latency dependencies);
a wait mask means the instruction also waits on a scoreboard (variable latency: LDS, MUFU, ...).
"""
import collections
import re
import subprocess
import sys


def load(obj, fun):
    """Instructions of kernel `fun` (mangled name) with control fields and source line (needs -lineinfo)."""
    import os
    import tempfile
    tmp = tempfile.mkdtemp()
    subprocess.run(['cuobjdump', '-xelf', 'all', os.path.abspath(obj)], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
    cubin = [f for f in os.listdir(tmp) if f.endswith('.cubin')][0]
    txt = subprocess.run(['nvdisasm', '-g', '-hex', os.path.join(tmp, cubin)], capture_output=True, text=True, check=True).stdout
    ins, pend, cur, loc = [], None, False, ('?', 0)
    labels, waiting = {}, []
    load.labels = labels
    for line in txt.split('\n'):
        m = re.match(r'\s*(\.L_x_\d+):', line)
        if m and cur:
            waiting.append(m.group(1))
            continue
        m = re.match(r'\s*\.section\s+\.text\.(\S+?),', line)
        if m:
            cur = (m.group(1) == fun)
            continue
        if not cur:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', line)
        if m:
            loc = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r'\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);\s+/\* 0x([0-9a-f]{16}) \*/', line)
        if m:
            pend = (int(m.group(1), 16), m.group(2).strip(), int(m.group(3), 16))
            for w in waiting:
                labels[w] = pend[0]
            waiting = []
            continue
        m = re.match(r'\s+/\* 0x([0-9a-f]{16}) \*/', line)
        if m and pend:
            hi = int(m.group(1), 16)
            ctrl = hi >> 41                                # bits 105.. of the 128-bit instruction
            ins.append(dict(addr=pend[0], text=pend[1], stall=ctrl & 0xf, yield_=(ctrl >> 4) & 1,
                            wbar=(ctrl >> 5) & 7, rbar=(ctrl >> 8) & 7, wait=(ctrl >> 11) & 0x3f, loc=loc))
            pend = None
    return ins


def loops(ins):
    """(head, tail) address pairs of backward branches spanning more than 4 KB."""
    out = []
    for i in ins:
        m = re.search(r'BRA\S*\s+(?:\S+,\s+)*`?\(?(\.L_x_\d+|0x[0-9a-f]+)', i['text'])
        if m:
            t = m.group(1)
            tgt = int(t, 16) if t.startswith('0x') else getattr(load, 'labels', {}).get(t)
            if tgt is not None and tgt < i['addr'] and i['addr'] - tgt > 0x1000:
                out.append((tgt, i['addr']))
    return out


def opcode(t):
    w = t.split()
    return (w[1] if w[0].startswith('@') else w[0]).split('.')[0]


def main():
    args = [a for a in sys.argv[1:] if not a.startswith('--')]
    ins = load(args[0], args[1])
    lo = int(args[2], 16) if len(args) > 2 else 0
    hi = int(args[3], 16) if len(args) > 3 else 1 << 30
    for a, b in loops(ins):
        print('loop %#x..%#x  %d instructions (%.1f KB)' % (a, b, (b - a) // 16 + 1, (b - a + 16) / 1024))
    if '--hot' in sys.argv and loops(ins):                 # the first (innermost, hottest) long loop
        lo, hi = loops(ins)[0]
    sel = [i for i in ins if lo <= i['addr'] <= hi]
    cyc = sum(max(1, i['stall']) for i in sel)
    fp64 = [i for i in sel if opcode(i['text']) in ('DADD', 'DMUL', 'DFMA', 'DSETP')]
    print('%d instructions in range, %d issue cycles by stall counts (%.2f per instruction), %d FP64-pipe (x2 = %d cycles)'
          % (len(sel), cyc, cyc / max(1, len(sel)), len(fp64), 2 * len(fp64)))
    by = collections.Counter()
    n = collections.Counter()
    for i in sel:
        by[opcode(i['text'])] += max(1, i['stall'])
        n[opcode(i['text'])] += 1
    for k, v in by.most_common(20):
        print('  %-10s %5d instr %6d cycles  (%.2f)' % (k, n[k], v, v / n[k]))
    hist = collections.Counter(i['stall'] for i in sel)
    print('stall histogram:', dict(sorted(hist.items())))
    if '--lines' in sys.argv:
        bl = collections.defaultdict(lambda: [0, 0])
        for i in sel:
            bl[i['loc']][0] += 1
            bl[i['loc']][1] += max(1, i['stall'])
        for loc, (k, c) in sorted(bl.items(), key=lambda kv: -kv[1][1])[:40]:
            print('  %-20s %4d instr %5d cycles (%.2f)' % ('%s:%d' % loc, k, c, c / k))
    if '--list' in sys.argv:
        for i in sel:
            print('%06x  s%-2d %s w%d r%d m%02x  %s' % (i['addr'], i['stall'], 'Y' if i['yield_'] else ' ', i['wbar'], i['rbar'],
                                                   i['wait'], i['text']))


if __name__ == '__main__':
    main()
