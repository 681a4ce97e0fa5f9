"""GPU: BASELINE.json-sized populations checked through size-independent properties (the oracle cannot follow
at these sizes): determinism, independence of episodes from their batch and shard, fused == lockstep,
population aggregate == reduction of the per-episode statistics, oracle spot checks on sampled episodes.

Sizes: configs[2] 1,048,576 lockstep episodes (a few env-steps), configs[4] 30 x 10,000 seeds x 1 day (fused),
configs[3] perturbed parameters (262,144 episodes, half a day)."""
import numpy as np
import pytest

import helpers as H
from helpers import O

pytestmark = pytest.mark.gpu


def _cohort_inputs(E, n_steps, seed, offset=0, group=30, dev=None):
    import torch
    from simglucose_b200 import streams as S
    dev = dev or torch.device('cuda', 0)
    rows = ((torch.arange(E, device=dev, dtype=torch.int64) + offset) % 30).to(torch.int32)
    noise = S.device_noise(E, n_steps + 2, 'Dexcom', seed=seed, episode_offset=offset, group_size=group, device=dev)
    mm, mg = S.device_scenarios(E, n_steps * 3, start_minute=0, seed=seed, episode_offset=offset, group_size=group, device=dev)
    return rows, noise, mm, mg


def test_config4_size_pid_population_properties():
    """30 patients x 10,000 seeds x 1 day, PID fused (BASELINE.json configs[4])."""
    import torch
    from simglucose_b200.engine import BatchEngine
    E, n_steps = 300000, 480
    rows, noise, mm, mg = _cohort_inputs(E, n_steps, seed=0)
    eng = BatchEngine(E)
    pid = (1e-3, 1e-5, 1e-3, 140.0)
    eng.reset(rows, noise, mm, mg, row0=False)
    eng.run(n_steps, controller='pid', ctrl=pid, trace=False)
    st = eng.stats().clone()
    agg = eng.reduce_stats().cpu().numpy()
    status = eng.status()
    assert int((status != 1).sum()) == 0
    s = st.cpu().numpy()
    assert np.all(np.isfinite(s))
    assert agg[0] == E * (n_steps + 1) == s[0].sum()
    assert agg[2] == s[2].min() and agg[3] == s[3].max() and agg[13] == (s[13] >= 0).sum()
    for k in (8, 9, 10, 11, 12, 14, 15):                         # integer counters: exact sums
        assert agg[k] == s[k].sum()
    for k in (1, 4, 5, 6, 7):
        assert abs(agg[k] - s[k].sum()) <= 1e-9 * max(1.0, abs(s[k].sum()))
    # determinism: the same launch again gives identical bits
    eng.reset(rows, noise, mm, mg, row0=False)
    eng.run(n_steps, controller='pid', ctrl=pid, trace=False)
    assert torch.equal(eng.stats(), st)
    # shard independence: episodes [lo, hi) simulated alone (another rank's view) == the same episodes in the big batch
    lo, hi = 123450, 123450 + 6000
    r2, n2, m2, g2 = _cohort_inputs(hi - lo, n_steps, seed=0, offset=lo)
    small = BatchEngine(hi - lo)
    small.reset(r2, n2, m2, g2, row0=False)
    out = small.run(n_steps, controller='pid', ctrl=pid, trace=True)
    assert torch.equal(small.stats(), st[:, lo:hi])
    # oracle spot checks on sampled episodes of the shard
    tr = out['trace'].cpu().numpy()
    nz, mmh, mgh = n2.cpu().numpy(), m2.cpu().numpy(), g2.cpu().numpy()
    for j in (0, 1777, 5999):
        row = int(r2[j])
        p, x0 = O.patient_params(row)
        k = mmh[:, j] >= 0
        o = O.run_episode(p, x0, n_steps, mmh[k, j], mgh[k, j], nz[:, j], controller='pid', ctrl=pid)
        assert H.rel_err(tr[:, 0, j], o['BG'][1:]) <= H.REL_TOL
        assert np.array_equal(tr[:, 9, j] != 0, o['done'][1:] != 0)
    small.close()
    eng.close()


def test_config2_size_lockstep_equals_fused():
    """1,048,576 environments stepped in lockstep (one launch per env-step) == the fused episode kernel."""
    import torch
    from simglucose_b200 import streams as S
    from simglucose_b200.engine import BatchEngine
    E, n_steps = 1048576, 6
    rows, noise, mm, mg = _cohort_inputs(E, n_steps, seed=3, group=1)
    acts = torch.stack([S.device_uniform(E, 0.0, 0.05, seed=0, offset=k * E) for k in range(n_steps)])
    zeros = torch.zeros_like(acts[0])
    a = BatchEngine(E)
    a.reset(rows, noise, mm, mg, row0=False)
    last = None
    for k in range(n_steps):
        last = a.step(acts[k], zeros)
    b = BatchEngine(E)
    b.reset(rows, noise, mm, mg, row0=False)
    full = torch.stack([acts, torch.zeros_like(acts)], dim=1).contiguous()
    out = b.run(n_steps, controller='actions', actions=full, trace=True)['trace']
    assert torch.equal(out[-1], last)
    assert torch.equal(a.state(), b.state()) and torch.equal(a.stats(), b.stats())
    assert float(acts.min()) >= 0.0 and float(acts.max()) <= 0.05 and abs(float(acts.mean()) - 0.025) < 1e-4
    a.close(); b.close()


def test_config3_size_perturbed_population():
    """+-10 % parameter perturbation (BASELINE.json configs[3]): bounds, determinism, table rows unperturbed at
    amplitude 0, every solver status ok over half a day with BBController."""
    import torch
    from simglucose_b200 import params as P, streams as S
    from simglucose_b200.engine import BatchEngine
    E, n_steps = 262144, 240
    rows, noise, mm, mg = _cohort_inputs(E, n_steps, seed=1234, group=1)
    tbl = torch.from_numpy(P.cohort_table()[0]).cuda()
    pp = S.device_perturbed_params(rows, amplitude=0.1, seed=1234)
    base = tbl[rows.long()].T
    ratio = (pp / base).cpu().numpy()
    assert ratio.min() >= 0.9 - 1e-12 and ratio.max() <= 1.1 + 1e-12 and abs(ratio.mean() - 1.0) < 1e-3
    assert torch.equal(S.device_perturbed_params(rows, amplitude=0.1, seed=1234), pp)
    assert torch.equal(S.device_perturbed_params(rows, amplitude=0.0, seed=1234), base.contiguous())
    q = torch.from_numpy(P.quest_table()).cuda()[rows.long()]
    ctrl_rows = torch.stack([torch.full_like(q[:, 0], 140.0), q[:, 0], q[:, 1], torch.zeros_like(q[:, 0])]).contiguous()
    eng = BatchEngine(E)
    eng.reset(rows, noise, mm, mg, params=pp, row0=False)
    out = eng.run(n_steps, controller='bb', ctrl_rows=ctrl_rows, trace=False)
    assert int((eng.status() != 1).sum()) == 0
    st = eng.stats().cpu().numpy()
    assert np.all(np.isfinite(st)) and st[0].min() == n_steps + 1
    # oracle spot check with the perturbed parameter rows
    pph, nz, mmh, mgh = pp.cpu().numpy(), noise.cpu().numpy(), mm.cpu().numpy(), mg.cpu().numpy()
    x0 = P.cohort_table()[1]
    qh = P.quest_table()
    for j in (5, 131071, 262143):
        r = int(rows[j])
        k = mmh[:, j] >= 0
        o = O.run_episode(pph[:, j], x0[r], n_steps, mmh[k, j], mgh[k, j], nz[:, j], controller='bb',
                          ctrl=(140.0, qh[r, 0], qh[r, 1]))
        ref = np.array([o['stats'][c] for c in O.STAT_COLS])
        assert np.array_equal(st[[0, 8, 9, 10, 11, 12, 13, 14, 15], j], ref[[0, 8, 9, 10, 11, 12, 13, 14, 15]])
        assert abs(st[1, j] - ref[1]) <= 1e-6 * abs(ref[1])
    eng.close()


@pytest.mark.parametrize('controller', ['bb', 'actions'])
def test_step_cap_compaction_is_bit_identical(controller):
    """sg_set_step_cap: heavy minutes finished by the second launch give the same bits as the single kernel
    (trace, per-minute DOPRI5 counters, state, statistics), fused and lockstep, independent schedules per lane."""
    import torch
    from simglucose_b200 import params as P, streams as S
    from simglucose_b200.engine import BatchEngine
    E, n_steps = 20000, 200                               # 10 h from 06:00: breakfast and lunch onsets
    dev = torch.device('cuda', 0)
    rows = (torch.arange(E, device=dev, dtype=torch.int64) % 30).to(torch.int32)
    noise = S.device_noise(E, n_steps + 2, 'Dexcom', seed=9, group_size=1, device=dev)
    mm, mg = S.device_scenarios(E, n_steps * 3, start_minute=360, seed=9, group_size=1, device=dev)
    q = torch.from_numpy(P.quest_table()).cuda()[rows.long()]
    ctrl_rows = torch.stack([torch.full_like(q[:, 0], 140.0), q[:, 0], q[:, 1], torch.zeros_like(q[:, 0])]).contiguous()
    acts = None
    if controller == 'actions':
        acts = torch.stack([S.device_uniform(2 * E, 0.0, 0.04, seed=1, offset=k * 2 * E).reshape(2, E) for k in range(n_steps)])
        acts[:, 1] = torch.where(acts[:, 1] > 0.039, acts[:, 1] * 20, torch.zeros_like(acts[:, 1]))    # occasional boluses

    def run(cap, lockstep=False):
        eng = BatchEngine(E)
        eng.set_step_cap(cap)
        eng.reset(rows, noise, mm, mg, row0=False)
        if lockstep:
            tr = torch.stack([eng.step(acts[k, 0], acts[k, 1]).clone() for k in range(n_steps)])
            cn = None
        else:
            out = eng.run(n_steps, controller=controller, ctrl_rows=ctrl_rows if controller == 'bb' else None,
                          actions=acts, trace=True, counters=True)
            tr, cn = out['trace'], out['counters']
        res = (tr, cn, eng.state(), eng.stats(), eng.status(), eng.launches)
        eng.close()
        return res
    base = run(0)
    capped = run(5)
    assert torch.equal(base[0], capped[0], ) or torch.equal(torch.nan_to_num(base[0]), torch.nan_to_num(capped[0]))
    assert torch.equal(base[1], capped[1]) and torch.equal(base[2], capped[2])
    assert torch.equal(base[3], capped[3]) and torch.equal(base[4], capped[4])
    assert capped[5] == 1 + 2 * n_steps and base[5] == 2                     # reset + (A, B) per env-step
    heavy = int((base[1][:, 0, :] > 5).sum())
    assert heavy > 100                                                        # the cap was exercised
    if controller == 'actions':
        lock = run(6, lockstep=True)
        assert torch.equal(torch.nan_to_num(lock[0]), torch.nan_to_num(base[0]))
        assert torch.equal(lock[2], base[2]) and torch.equal(lock[3], base[3])


def test_run_shards_streams_host_population_identically():
    """engine.run_shards (host-resident shards, upload of shard k+1 overlapped with the kernel of shard k on a second
    stream) == one resident batch, bit for bit, and its aggregate == the reduction of the per-episode statistics."""
    import torch
    from simglucose_b200 import params as P
    from simglucose_b200.engine import BatchEngine, run_shards
    E, n_steps = 30 * 500, 96
    rows, noise, mm, mg = _cohort_inputs(E, n_steps, seed=4)
    q = torch.from_numpy(P.quest_table()).cuda()[rows.long()]
    cr = torch.stack([torch.full_like(q[:, 0], 140.0), q[:, 0], q[:, 1], torch.zeros_like(q[:, 0])])
    eng = BatchEngine(E)
    eng.reset(rows, noise, mm, mg, row0=False)
    eng.run(n_steps, controller='bb', ctrl_rows=cr, trace=False)
    want = eng.stats().cpu()
    agg_want = eng.reduce_stats().cpu().numpy()
    eng.close()
    cuts = [0, 30 * 100, 30 * 101, 30 * 333, E]                      # ragged shards, one of a single seed group
    shards = []
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        sh = {'patient_row': rows[lo:hi], 'noise': noise[:, lo:hi], 'meal_minute': mm[:, lo:hi],
              'meal_grams': mg[:, lo:hi], 'ctrl_rows': cr[:, lo:hi]}
        shards.append({k: v.cpu().contiguous().pin_memory() for k, v in sh.items()})
    stats, agg = run_shards(shards, n_steps, controller='bb', slots=2)
    got = torch.cat(stats, dim=1)
    assert torch.equal(got, want)
    assert agg[0] == agg_want[0] and agg[2] == agg_want[2] and agg[3] == agg_want[3] and agg[13] == agg_want[13]
    for k in (8, 9, 10, 11, 12, 14, 15):
        assert agg[k] == agg_want[k]
    for k in (1, 4, 5, 6, 7):
        assert abs(agg[k] - agg_want[k]) <= 1e-9 * max(1.0, abs(agg_want[k]))
    # trajectories out: the selected history columns of every shard equal the resident batch's trace
    from simglucose_b200 import _capi
    from simglucose_b200.engine import ShardRunner
    eng = BatchEngine(E)
    eng.reset(rows, noise, mm, mg, row0=False)
    trace = eng.run(n_steps, controller='bb', ctrl_rows=cr, trace=True)['trace']
    cols = ('BG', 'CGM', 'CHO', 'insulin')
    want_tr = trace[:, [_capi.TRACE_COLS.index(c) for c in cols], :].cpu()
    eng.close()
    runner = ShardRunner(max(hi - lo for lo, hi in zip(cuts[:-1], cuts[1:])), slots=2)
    tr_out = [torch.empty(n_steps, len(cols), hi - lo, dtype=torch.float64).pin_memory() for lo, hi in zip(cuts[:-1], cuts[1:])]
    stats2, agg2 = runner.run(shards, n_steps, controller='bb', traces_out=tr_out, trace_cols=cols)
    torch.cuda.synchronize()
    # the same runner asked for OTHER columns (same count, then more): the selection belongs to each call
    for cols2 in (('Risk', 'BG', 'LBGI', 'HBGI'), ('BG', 'CGM', 'CHO', 'insulin', 'reward', 'done')):
        tr2 = [torch.empty(n_steps, len(cols2), hi - lo, dtype=torch.float64).pin_memory()
               for lo, hi in zip(cuts[:-1], cuts[1:])]
        runner.run(shards, n_steps, controller='bb', traces_out=tr2, trace_cols=cols2)
        torch.cuda.synchronize()
        want2 = trace[:, [_capi.TRACE_COLS.index(c) for c in cols2], :].cpu()
        assert torch.equal(torch.nan_to_num(torch.cat(tr2, dim=2)), torch.nan_to_num(want2)), cols2
    runner.close()
    assert torch.equal(torch.cat(tr_out, dim=2), want_tr)
    assert torch.equal(torch.cat(stats2, dim=1), want)


def test_config3_shard_halves_reduce_to_the_single_engine_aggregate():
    """The multi-GPU claim of configs[3] on one GPU: a 2^18-episode perturbed population (parameters and streams
    generated on device, keyed by GLOBAL episode index) simulated as ONE batch and as the two contiguous halves
    distributed.shard_range(E, r, 2) gives the two ranks.  Per-episode statistics are bit-identical (an episode does
    not know its shard), the counters of the aggregate are bitwise equal, the floating-point sums agree to 1e-12
    relative (only their summation order differs), min / max exactly."""
    import torch
    import bench
    from simglucose_b200 import _capi, streams as S
    from simglucose_b200.distributed import shard_range
    from simglucose_b200.engine import BatchEngine
    E, n_steps, seed = 262144, 96, 1234
    dev = torch.device('cuda', 0)

    def run(lo, hi):
        n = hi - lo
        rows = ((torch.arange(n, device=dev, dtype=torch.int64) + lo) % 30).to(torch.int32)
        pp = S.device_perturbed_params(rows, amplitude=0.1, seed=seed, episode_offset=lo)
        noise = S.device_noise(n, n_steps + 2, 'Dexcom', seed=seed, episode_offset=lo, group_size=1, device=dev)
        mm, mg = S.device_scenarios(n, n_steps * 3, start_minute=7 * 60, seed=seed, episode_offset=lo, group_size=1, device=dev)
        eng = BatchEngine(n, coherence=40)
        eng.reset(rows, noise, mm, mg, params=pp, row0=False)
        eng.run(n_steps, controller='bb', ctrl_rows=bench.controller_rows_device(rows), trace=False)
        assert int((eng.status() != 1).sum()) == 0
        st, agg = eng.stats().clone(), eng.reduce_stats().clone()
        eng.close()
        return st, agg, mm.shape[0]

    whole, agg_whole, _ = run(0, E)
    halves = [run(*shard_range(E, r, 2)) for r in range(2)]
    assert shard_range(E, 0, 2) == (0, E // 2) and shard_range(E, 1, 2) == (E // 2, E)
    assert torch.equal(torch.cat([h[0] for h in halves], dim=1), whole)
    a, b, w = halves[0][1].cpu().numpy(), halves[1][1].cpu().numpy(), agg_whole.cpu().numpy()
    names = _capi.STAT_COLS
    for k, name in enumerate(names):
        if name == 'min_bg':
            assert min(a[k], b[k]) == w[k]
        elif name == 'max_bg':
            assert max(a[k], b[k]) == w[k]
        elif name in ('rows', 'lt50', 'lt70', 'inrange', 'gt180', 'gt250', 'first_done', 'dopri_steps', 'dopri_rejects'):
            assert a[k] + b[k] == w[k], name                      # integer-valued doubles: exact in any order
        else:
            assert abs(a[k] + b[k] - w[k]) <= 1e-12 * max(1.0, abs(w[k])), name
    assert w[names.index('rows')] == E * (n_steps + 1)
