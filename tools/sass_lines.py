#!/usr/bin/env python
"""Static SASS accounting per source line (needs -lineinfo): which lines of the CUDA sources own the
instructions of a kernel, optionally restricted to an address range (e.g. the hot loop).

    python tools/sass_lines.py <object.o> <kernel substring> [lo hi]     (addresses in hex)

Prints the loop spans (backward branches), then instruction counts per file:line split into FP64-pipe
instructions (DADD/DMUL/DFMA/DSETP) and everything else.
"""
import collections
import os
import re
import subprocess
import sys
import tempfile


def disassemble(obj):
    tmp = tempfile.mkdtemp()
    subprocess.run(['cuobjdump', '-xelf', 'all', os.path.abspath(obj)], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
    cubin = [f for f in os.listdir(tmp) if f.endswith('.cubin')][0]
    return subprocess.run(['nvdisasm', '-g', '-c', os.path.join(tmp, cubin)], check=True, capture_output=True,
                          text=True).stdout


def parse(dis, key):
    ins, cur, loc = [], False, ('?', 0)
    for line in dis.split('\n'):
        m = re.match(r'\s*\.section\s+\.text\.(\S+?),', line)
        if m:
            cur = key in m.group(1)
            continue
        if not cur:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', line)
        if m:
            loc = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r'\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);', line)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip(), loc))
    return ins


def opcode(t):
    w = t.split()
    return (w[1] if w[0].startswith('@') else w[0]).split('.')[0]


def main():
    obj, key = sys.argv[1], sys.argv[2]
    ins = parse(disassemble(obj), key)
    print('%d instructions, %.1f KB' % (len(ins), len(ins) * 16 / 1024))
    for a, t, _ in ins:
        m = re.search(r'BRA\S*\s+(?:\S+,\s+)*`?\(?(\.L_x_\d+|0x[0-9a-f]+)', t)
        if m and m.group(1).startswith('0x'):
            tgt = int(m.group(1), 16)
            if tgt < a and a - tgt > 0x400:
                print('  loop %#x..%#x  %d instructions (%.1f KB)' % (tgt, a, (a - tgt) // 16 + 1, (a - tgt + 16) / 1024))
    lo = int(sys.argv[3], 16) if len(sys.argv) > 3 else 0
    hi = int(sys.argv[4], 16) if len(sys.argv) > 4 else 1 << 30
    fp, other, ops = collections.Counter(), collections.Counter(), collections.Counter()
    for a, t, loc in ins:
        if lo <= a <= hi:
            op = opcode(t)
            ops[op] += 1
            (fp if op in ('DADD', 'DMUL', 'DFMA', 'DSETP') else other)[loc] += 1
    print('range %#x..%#x: %d FP64-pipe, %d other' % (lo, hi, sum(fp.values()), sum(other.values())))
    print('opcodes:', ', '.join('%s %d' % kv for kv in ops.most_common(25)))
    for loc in sorted(set(fp) | set(other)):
        print('%-22s %5d fp64 %5d other' % ('%s:%d' % loc, fp[loc], other[loc]))


if __name__ == '__main__':
    main()
