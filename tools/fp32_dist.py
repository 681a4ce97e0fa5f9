import sys, numpy as np, torch
sys.path.insert(0, '.')
sys.path.insert(0, 'tests')
import test_gpu_fp32 as T
for fp32 in (False, True):
    r = T._run(fp32, 'bb', E=960, n_steps=480, seed=7)
    c = r['counters'][:, 0]          # [minutes, E]
    print('fp32' if fp32 else 'fp64', 'mean', c.mean(), 'hist', np.bincount(c.ravel().clip(0, 12), minlength=13))
    w = c.reshape(c.shape[0], -1, 32)
    print('   mean over warps of max-in-warp', w.max(axis=2).mean(), 'lane eff', (2 + 6 * w.mean()) / (2 + 6 * w.max(axis=2).mean()))
    rej = r['counters'][:, 1]
    print('   rejects per minute', rej.mean())
r = T._run(True, 'bb', E=960, n_steps=480, seed=7)
c = r['counters'][:, 0]
mins, eps = np.nonzero(c >= 12)
print('heavy lane-minutes', len(mins), 'patients', np.bincount(eps % 30, minlength=30))
print('minute-of-day hist (per 2 h)', np.bincount(mins // 120, minlength=12))
r64 = T._run(False, 'bb', E=960, n_steps=480, seed=7)
c64 = r64['counters'][:, 0]
print('fp64 steps at those minutes: mean', c64[mins, eps].mean(), 'max', c64[mins, eps].max())
tr = r['trace']
from simglucose_b200 import _capi
Tn = _capi.TRACE_COLS.index
step = mins // 3
print('BG at heavy minutes: min %.1f mean %.1f max %.1f' % (tr[step, Tn('BG'), eps].min(), tr[step, Tn('BG'), eps].mean(), tr[step, Tn('BG'), eps].max()))
print('CHO at heavy minutes nonzero frac', (tr[step, Tn('CHO'), eps] > 0).mean())
