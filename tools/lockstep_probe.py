import sys, time, torch
sys.path.insert(0, '/root/repo')
from simglucose_b200 import streams as S
from simglucose_b200.engine import BatchEngine
E, n_steps = 1048576, 60
dev = torch.device('cuda', 0)
import itertools
for group, cap in itertools.product((1, 30), (0, 5, 8)):
    rows = (torch.arange(E, device=dev, dtype=torch.int64) % 30).to(torch.int32)
    noise = S.device_noise(E, n_steps + 2, 'Dexcom', seed=3, group_size=group, device=dev)
    mm, mg = S.device_scenarios(E, n_steps * 3, start_minute=6*60, seed=3, group_size=group, device=dev)
    acts = torch.stack([S.device_uniform(E, 0.0, 0.05, seed=0, offset=k * E) for k in range(n_steps)])
    zeros = torch.zeros(E, dtype=torch.float64, device=dev)
    full = torch.stack([acts, torch.zeros_like(acts)], dim=1).contiguous()
    eng = BatchEngine(E)
    eng.set_step_cap(cap)
    def fused():
        eng.reset(rows, noise, mm, mg, row0=False)
        eng.run(n_steps, controller='actions', actions=full, trace=False)
    out = eng.empty(12)
    def lock():
        eng.reset(rows, noise, mm, mg, row0=False)
        for k in range(n_steps):
            eng.step(acts[k], zeros, out=out)
    for name, fn in (('fused', fused), ('lockstep', lock)):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); e0.record(); fn(); e1.record(); t1 = time.perf_counter(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print('cap %d group %d %-9s %.1f ms  (%.3f ms/env-step, %.3g patient-min/s)  host launch time %.1f ms' % (cap, group, name, ms, ms / n_steps, E * n_steps * 3 / ms * 1e3, (t1 - t0) * 1e3))
    eng.close()
