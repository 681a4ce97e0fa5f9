"""GPU: the warp-pair kernel (SG_FLAG_PAIR: two threads per episode, csrc/sg_pair.cuh) against the single-thread
kernel.  Every state is integrated by one thread with the same operation order, so in strict arithmetic the two
organisations must agree BIT FOR BIT -- traces, DOPRI5 counters, resident state, statistics -- and the
single-thread kernel is itself pinned to the oracle by test_gpu_parity.py.  One oracle comparison is repeated
here directly so that the pair kernel does not only lean on that chain."""
import numpy as np
import pytest

import helpers as H
from helpers import O

pytestmark = pytest.mark.gpu


def _inputs(E, n_steps, seed, group=30):
    import torch
    from simglucose_b200 import streams as S
    dev = torch.device('cuda', 0)
    rows = (torch.arange(E, device=dev, dtype=torch.int64) % 30).to(torch.int32)
    noise = S.device_noise(E, n_steps + 2, 'Dexcom', seed=seed, group_size=group, device=dev)
    mm, mg = S.device_scenarios(E, n_steps * 3, start_minute=0, seed=seed, group_size=group, device=dev)
    return rows, noise, mm, mg


def _both(E, **kw):
    from simglucose_b200.engine import BatchEngine
    return BatchEngine(E, pair=False, **kw), BatchEngine(E, pair=True, **kw)


def _same(a, b):
    import torch
    # NaN-aware bitwise comparison
    return torch.equal(a.view(torch.int64) if a.dtype == torch.float64 else a,
                       b.view(torch.int64) if b.dtype == torch.float64 else b)


@pytest.mark.parametrize('controller', ['bb', 'pid'])
@pytest.mark.parametrize('carry_h', [False, True])
def test_pair_equals_single_thread_kernel_bitwise(controller, carry_h):
    import torch
    from simglucose_b200 import params as P
    E, n_steps = 2999, 200                           # ragged: the last block has 23 idle lanes
    rows, noise, mm, mg = _inputs(E, n_steps, seed=3)
    q = torch.from_numpy(P.quest_table()).cuda()[rows.long()]
    cr = torch.stack([torch.full_like(q[:, 0], 140.0), q[:, 0], q[:, 1], torch.zeros_like(q[:, 0])])
    outs = []
    for eng in _both(E, carry_h=carry_h):
        eng.reset(rows, noise, mm, mg)
        if controller == 'bb':
            o = eng.run(n_steps, controller='bb', ctrl_rows=cr, trace=True, counters=True)
        else:
            o = eng.run(n_steps, controller='pid', ctrl=(1e-3, 1e-5, 1e-3, 140.0), trace=True, counters=True)
        d, i = eng.episode_state()
        outs.append((o['trace'], o['counters'], d, i, eng.stats(), eng.status()))
        eng.close()
    for a, b in zip(*outs):
        assert _same(a, b)
    assert int(outs[1][1][:, 0].sum()) > 0          # DOPRI5 steps were counted


def test_pair_actions_in_two_launches_and_independent_meals():
    """Pre-drawn actions, meal schedules of their own per episode (group 1), the run split into two launches
    (state round trip through HBM between them)."""
    import torch
    E, n_steps = 1500, 120
    rows, noise, mm, mg = _inputs(E, n_steps, seed=11, group=1)
    g = torch.Generator(device='cuda').manual_seed(5)
    act = torch.zeros(n_steps, 2, E, dtype=torch.float64, device='cuda')
    act[:, 0] = torch.rand(n_steps, E, generator=g, device='cuda', dtype=torch.float64) * 0.05
    act[::7, 1] = torch.rand(len(range(0, n_steps, 7)), E, generator=g, device='cuda', dtype=torch.float64) * 0.3
    outs = []
    for eng in _both(E):
        eng.reset(rows, noise, mm, mg)
        o1 = eng.run(50, controller='actions', actions=act[:50].contiguous(), trace=True)
        o2 = eng.run(n_steps - 50, controller='actions', actions=act[50:].contiguous(), trace=True)
        outs.append((torch.cat([o1['trace'], o2['trace']]), eng.state(), eng.stats()))
        eng.close()
    for a, b in zip(*outs):
        assert _same(a, b)


def test_pair_perturbed_parameters_and_lockstep_steps():
    """Per-episode parameter rows (configs[3]) and sg_step (one env-step per launch)."""
    import torch
    from simglucose_b200 import streams as S
    E, n_steps = 1024, 40
    rows, noise, mm, mg = _inputs(E, n_steps, seed=2, group=1)
    params = S.device_perturbed_params(rows, amplitude=0.1, seed=1234)
    g = torch.Generator(device='cuda').manual_seed(9)
    outs = []
    for eng in _both(E):
        eng.reset(rows, noise, mm, mg, params=params)
        g.manual_seed(9)
        rowsout = []
        for _ in range(n_steps):
            basal = torch.rand(E, generator=g, device='cuda', dtype=torch.float64) * 0.05
            rowsout.append(eng.step(basal, torch.zeros_like(basal)).clone())
        outs.append((torch.stack(rowsout), eng.state(), eng.stats()))
        eng.close()
    for a, b in zip(*outs):
        assert _same(a, b)


def test_pair_against_oracle():
    """30 patients x 1 day, BBController, reference-compatible streams: the pair kernel against the CPU oracle."""
    import torch
    from simglucose_b200 import params as P, streams as S, _capi
    from simglucose_b200.engine import BatchEngine
    n_steps, E = 480, 30
    noise = S.cgm_noise_stream('Dexcom', 2, n_steps + 2)
    mm, mg = S.random_scenario_table(2, 0, n_steps * 3)
    tbl, x0 = P.cohort_table()
    q = P.quest_table()
    rows = np.arange(E, dtype=np.int32)
    cr = np.stack([np.full(E, 140.0), q[:, 0], q[:, 1], np.zeros(E)])
    eng = BatchEngine(E, pair=True)
    mmE, mgE = S.pad_tables([(mm, mg)] * E)
    row0 = eng.reset(rows, np.repeat(noise[:, None], E, 1), mmE, mgE)
    out = eng.run(n_steps, controller='bb', ctrl_rows=cr, trace=True, counters=True)
    T = _capi.TRACE_COLS
    bg = torch.cat([row0[None], out['trace']], 0)[:, T.index('BG'), :].cpu().numpy()
    cgm = torch.cat([row0[None], out['trace']], 0)[:, T.index('CGM'), :].cpu().numpy()
    ins = out['trace'][:, T.index('insulin'), :].cpu().numpy()
    cho = out['trace'][:, T.index('CHO'), :].cpu().numpy()
    worst = 0.0
    for e in range(E):
        o = O.run_episode(tbl[e], x0[e], n_steps, mm, mg, noise, controller='bb', ctrl=(140.0, q[e, 0], q[e, 1]))
        worst = max(worst, float(np.max(np.abs(bg[:, e] - o['BG']) / np.maximum(np.abs(o['BG']), 1.0))))
        worst = max(worst, float(np.max(np.abs(cgm[:, e] - o['CGM']) / np.maximum(np.abs(o['CGM']), 1.0))))
        assert np.array_equal(ins[:, e], o['insulin'][:-1])      # pump quantisation: exact
        assert np.array_equal(cho[:, e], o['CHO'][:-1])          # meal timing and grams: exact
    assert worst <= 1e-6, worst                                   # north-star tolerance
    assert (eng.status().cpu().numpy() == 1).all()
    eng.close()
