"""Check that the configs[3] population does not depend on how many GPUs simulated it.

    python tools/check_scale_vectors.py SCALE_r03.json            # a driver record, or
    python tools/check_scale_vectors.py line_1gpu.json line_2gpu.json ...   # bench.py lines, one file per N

Every bench.py line carries configs['configs[3]']['population_vector_hex']: the aggregate statistics vector [NSTAT]
after the NCCL all-reduce, bit for bit.  Its integer entries (history rows, the five time-in-range counters, episodes
done, DOPRI5 steps and rejects) and its extrema (min / max BG) must be IDENTICAL at every N -- an episode does not
know its shard; the floating-point sums may differ in their last bits (summation order) and are held to 1e-12.
Exit status 0 = consistent.
"""
import json
import sys

STAT_COLS = ['rows', 'sum_bg', 'min_bg', 'max_bg', 'sum_lbgi', 'sum_hbgi', 'sum_risk', 'sum_reward', 'lt50',
             'lt70', 'inrange', 'gt180', 'gt250', 'first_done', 'dopri_steps', 'dopri_rejects']
EXACT = [0, 2, 3, 8, 9, 10, 11, 12, 13, 14, 15]
SUMS = [1, 4, 5, 6, 7]


def bench_lines(obj):
    """Every dict below `obj` that looks like a bench.py line with a configs[3] entry."""
    if isinstance(obj, dict):
        c = obj.get('configs')
        if isinstance(c, dict) and isinstance(c.get('configs[3]'), dict) and 'n_gpus' in obj:
            yield obj
        for v in obj.values():
            yield from bench_lines(v)
    elif isinstance(obj, list):
        for v in obj:
            yield from bench_lines(v)
    elif isinstance(obj, str) and '"configs[3]"' in obj:
        for ln in obj.splitlines():
            ln = ln.strip()
            if ln.startswith('{'):
                try:
                    yield from bench_lines(json.loads(ln))
                except ValueError:
                    pass


def check(lines):
    seen = {}
    for ln in lines:
        c3 = ln['configs']['configs[3]']
        key = (int(ln['n_gpus']), c3.get('episodes'), c3.get('days'))
        seen[key] = [float.fromhex(h) for h in c3['population_vector_hex']]
    if len(seen) < 2:
        print('need configs[3] vectors from at least two GPU counts; found', sorted(seen))
        return 2
    sizes = {(k[1], k[2]) for k in seen}
    if len(sizes) != 1:
        print('the lines ran different configs[3] sizes:', sorted(seen))
        return 2
    keys = sorted(seen)
    base = seen[keys[0]]
    bad = 0
    for k in keys[1:]:
        v = seen[k]
        for i in EXACT:
            if v[i] != base[i]:
                print('N=%d: %s differs: %r vs %r at N=%d' % (k[0], STAT_COLS[i], v[i], base[i], keys[0][0]))
                bad += 1
        for i in SUMS:
            if abs(v[i] - base[i]) > 1e-12 * max(1.0, abs(base[i])):
                print('N=%d: %s differs beyond 1e-12: %r vs %r' % (k[0], STAT_COLS[i], v[i], base[i]))
                bad += 1
    print('configs[3] %s episodes x %s days at N = %s: %s' % (keys[0][1], keys[0][2], [k[0] for k in keys],
                                                              'consistent' if not bad else '%d mismatches' % bad))
    return 1 if bad else 0


if __name__ == '__main__':
    found = []
    for path in sys.argv[1:]:
        with open(path) as f:
            text = f.read()
        try:
            found.extend(bench_lines(json.loads(text)))
        except ValueError:
            found.extend(bench_lines(text))
    sys.exit(check(found))
