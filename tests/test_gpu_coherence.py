"""GPU: launch-order and reset machinery of round 2, each checked bit for bit against the plain path.

  * sg_set_coherence: episodes sorted by their next meal event run in a different thread -> per-episode results must
    not move by a bit (trace, per-minute DOPRI5 counters, resident state, statistics, status), fused and lockstep,
    FP64 and FP32, and the launch order must be a permutation;
  * sg_reset_masked: the masked episodes restart exactly like a fresh sg_reset, the others do not notice;
  * SG_CTRL_RANDOM: the action drawn inside the kernel is the one the trace reports, and feeding those actions back
    through SG_CTRL_ACTIONS gives the same episode;
  * BBController basal by name (ctrl[3]) for perturbed parameter rows;
  * status -5 when an episode runs past its noise stream;
  * handles on two devices in one process (skipped on a one-GPU box).
"""
import numpy as np
import pytest

import helpers as H
from helpers import O

pytestmark = pytest.mark.gpu


def _population(E, n_steps, seed, start_minute=360, dev=None):
    import torch
    from simglucose_b200 import params as P, streams as S
    dev = dev or torch.device('cuda', 0)
    rows = (torch.arange(E, device=dev, dtype=torch.int64) % 30).to(torch.int32)
    noise = S.device_noise(E, n_steps + 2, 'Dexcom', seed=seed, group_size=1, device=dev)
    mm, mg = S.device_scenarios(E, n_steps * 3, start_minute=start_minute, seed=seed, group_size=1, device=dev)
    q = torch.from_numpy(P.quest_table()).to(dev)[rows.long()]
    tbl = P.cohort_table()[0]
    basal = torch.from_numpy(tbl[:, 1] * tbl[:, 0] / 6000.0).to(dev)[rows.long()]
    ctrl_rows = torch.stack([torch.full_like(q[:, 0], 140.0), q[:, 0], q[:, 1], basal]).contiguous()
    return rows, noise, mm, mg, ctrl_rows


def _snapshot(eng, out):
    import torch
    d, i = eng.episode_state()
    return dict(trace=torch.nan_to_num(out['trace']) if out.get('trace') is not None else None,
                counters=out.get('counters'), d=d, i=i, stats=eng.stats(), status=eng.status())


def _same(a, b):
    import torch
    for k in a:
        if a[k] is None:
            assert b[k] is None
        else:
            assert torch.equal(a[k], b[k]), k


@pytest.mark.parametrize('fp32', [False, True])
@pytest.mark.parametrize('controller', ['bb', 'pid'])
def test_coherent_launch_order_is_bit_identical(controller, fp32):
    import torch
    from simglucose_b200.engine import BatchEngine
    E, n_steps = 20000, 200                                # 10 h from 06:00: breakfast and lunch, independent schedules
    rows, noise, mm, mg, ctrl_rows = _population(E, n_steps, seed=17)
    kw = dict(ctrl_rows=ctrl_rows) if controller == 'bb' else dict(ctrl=(1e-3, 1e-5, 1e-3, 140.0))

    def run(coherence):
        eng = BatchEngine(E, fp32=fp32, coherence=coherence)
        eng.reset(rows, noise, mm, mg, row0=False)
        out = eng.run(n_steps, controller=controller, trace=True, counters=True, **kw)
        snap = _snapshot(eng, out)
        perm = eng.launch_order().cpu().numpy() if coherence else None
        launches = eng.launches
        eng.close()
        return snap, perm, launches
    base, _, l0 = run(0)
    for seg in (25, 64):
        got, perm, l1 = run(seg)
        _same(base, got)
        assert np.array_equal(np.sort(perm), np.arange(E))
        n_seg = -(-n_steps // seg)
        assert l1 == 1 + n_seg * 4 and l0 == 2             # reset + (3 sort launches + 1 episode launch) per segment
    heavy = int((base['counters'][:, 0, :] > 5).sum())
    assert heavy > 100                                     # there were expensive minutes to line up
    # the adaptive order (sg_set_sort_mode 1: recent DOPRI5 steps + meals announced inside the segment), re-sorted every
    # few env-steps as a lockstep vector environment does: the same results, a permutation, expensive episodes together
    # (segments of 1 and 2 env-steps have fewer than 1,024 keys and take the shared-memory histogram path of the sort)
    for seg, adaptive in ((4, True), (1, True), (2, False)):
        eng = BatchEngine(E, fp32=fp32)
        eng.set_coherence(seg, adaptive=adaptive)
        eng.reset(rows, noise, mm, mg, row0=False)
        out = eng.run(n_steps if seg > 1 else 60, controller=controller, trace=True, counters=True, **kw)
        snap = _snapshot(eng, out)
        if seg > 1:
            _same(base, snap)
        else:                                              # 60 one-step segments: against the first 60 steps of the unsorted run
            assert torch.equal(snap['trace'], base['trace'][:60]) and torch.equal(snap['counters'], base['counters'][:180])
        perm = eng.launch_order().cpu().numpy()
        assert np.array_equal(np.sort(perm), np.arange(E))
        eng.close()


def test_coherent_order_groups_meal_onsets():
    """The sort puts episodes whose next meal starts in the same minute next to each other: within a warp-sized run of
    the launch order the spread of next-meal minutes is (far) smaller than in index order."""
    import torch
    from simglucose_b200.engine import BatchEngine
    E, n_steps = 32768, 40
    rows, noise, mm, mg, ctrl_rows = _population(E, 200, seed=23)
    eng = BatchEngine(E, coherence=n_steps)
    eng.reset(rows, noise, mm, mg, row0=False)
    eng.run(1, controller='bb', ctrl_rows=ctrl_rows, trace=False)
    perm = eng.launch_order().cpu().numpy()
    first = mm[0].cpu().numpy().astype(np.int64)           # first meal of every episode (start 06:00: breakfast or later)
    first = np.where((first < 0) | (first >= n_steps * 3), 10 ** 6, first)

    def warp_spread(order):
        w = first[order].reshape(-1, 32)
        has = (w < 10 ** 6)
        lo = np.where(has, w, 10 ** 7).min(axis=1)
        hi = np.where(has, w, -1).max(axis=1)
        mixed = has.any(axis=1) & ~has.all(axis=1)
        return float(np.where(has.all(axis=1), hi - lo, 0).mean()), float(mixed.mean())
    s_sorted, mixed_sorted = warp_spread(perm)
    s_index, mixed_index = warp_spread(np.arange(E))
    assert s_sorted <= 1.0 and mixed_sorted <= 0.01        # warps are (almost) all-meal with one onset minute, or no-meal
    assert mixed_index > 0.5
    eng.close()


def test_coherent_lockstep_steps_equal_fused_run():
    import torch
    from simglucose_b200 import streams as S
    from simglucose_b200.engine import BatchEngine
    E, n_steps = 8192, 90
    rows, noise, mm, mg, _ = _population(E, n_steps, seed=5)
    acts = torch.stack([S.device_uniform(2 * E, 0.0, 0.04, seed=2, offset=k * 2 * E).reshape(2, E) for k in range(n_steps)])
    acts[:, 1] = torch.where(acts[:, 1] > 0.039, acts[:, 1] * 20, torch.zeros_like(acts[:, 1]))
    a = BatchEngine(E)
    a.reset(rows, noise, mm, mg, row0=False)
    want = _snapshot(a, a.run(n_steps, controller='actions', actions=acts, trace=True))
    a.close()
    b = BatchEngine(E, coherence=16)
    b.reset(rows, noise, mm, mg, row0=False)
    rows_out = torch.stack([b.step(acts[k, 0], acts[k, 1]).clone() for k in range(n_steps)])
    got = _snapshot(b, dict(trace=rows_out))
    assert b.launches == 1 + n_steps + 3 * -(-n_steps // 16)
    b.close()
    _same(want, got)


def test_sensor_switch_after_set_coherence_resizes_the_sort():
    """The sort keys count MINUTES ahead, so their number grows with segment_steps * sample_time.  Switching the sensor
    (Dexcom, 3 min -> GuardianRT, 5 min) after set_coherence used to leave the histogram sized for the old period and
    the sort wrote past it.  Now sg_set_hardware re-sizes it (or refuses a segment that no longer fits 4096 minutes),
    and the results stay those of the unsorted order."""
    import torch
    from simglucose_b200 import _capi, params as P, streams as S
    from simglucose_b200.engine import BatchEngine
    E, n_steps = 9000, 72                                  # 6 h at 5-minute env-steps, from 06:00
    dev = torch.device('cuda', 0)
    rows = (torch.arange(E, device=dev, dtype=torch.int64) % 30).to(torch.int32)
    noise = S.device_noise(E, n_steps + 2, 'GuardianRT', seed=23, group_size=1, device=dev)
    mm, mg = S.device_scenarios(E, n_steps * 5, start_minute=360, seed=23, group_size=1, device=dev)

    def run(eng):
        eng.reset(rows, noise, mm, mg, row0=False)
        out = eng.run(n_steps, controller='pid', ctrl=(1e-3, 1e-5, 1e-3, 140.0), trace=True, counters=True)
        return _snapshot(eng, out)
    plain = BatchEngine(E, sensor='GuardianRT')
    base = run(plain)
    plain.close()
    eng = BatchEngine(E, coherence=20)                      # Dexcom: 60-minute window, 7,809 keys
    eng.set_hardware('GuardianRT')                          # 100-minute window, 12,929 keys
    got = run(eng)
    perm = eng.launch_order().cpu().numpy()
    _same(base, got)
    assert np.array_equal(np.sort(perm), np.arange(E))
    # a per-step sort through the shared-memory histogram (small key space) after the same switch
    eng.set_coherence(1, adaptive=True)                     # GuardianRT: 5-minute window, 6 x 128 + 1 keys
    eng.set_hardware('Navigator')                           # 1-minute sensor: 2 x 128 + 1 keys
    eng.set_hardware('GuardianRT')
    _same(base, run(eng))
    # a segment that fits at 3 minutes but not at 5 is refused, and the handle keeps its sensor
    eng.set_hardware('Dexcom')
    eng.set_coherence(1000)                                 # 3,000 minutes
    with pytest.raises(_capi.SgError, match='4096 minutes'):
        eng.set_hardware('GuardianRT')                      # 5,000 minutes
    assert eng.sample_time == 3
    eng.close()


def test_masked_reset_restarts_only_the_masked_episodes():
    import torch
    from simglucose_b200 import streams as S
    from simglucose_b200.engine import BatchEngine
    E, n_steps = 3000, 60
    rows, noise, mm, mg, ctrl_rows = _population(E, 2 * n_steps, seed=31)
    noise, mm, mg = noise.clone(), mm.clone(), mg.clone()
    eng = BatchEngine(E)
    eng.reset(rows, noise, mm, mg, row0=False)
    eng.run(n_steps, controller='bb', ctrl_rows=ctrl_rows, trace=False)
    mask = (torch.arange(E, device=rows.device) % 3 == 0).to(torch.int32)
    # the masked episodes get new streams (keyed generators write the masked columns in place) and new initial states
    key = torch.arange(E, device=rows.device, dtype=torch.int64) + (7 << 24)
    old_noise = noise.clone()
    S.device_noise_keyed(noise, key, 'Dexcom', seed=99, mask=mask)
    start = torch.full((E,), 600, dtype=torch.int32, device=rows.device)
    S.device_scenarios_keyed(mm, mg, start, 2 * n_steps * 3, key, seed=99, mask=mask)
    _, x0 = S.device_episode_starts(key, rows, n_draws=2, seed=99, mask=None)
    keep = mask == 0
    assert torch.equal(noise[:, keep], old_noise[:, keep]) and not torch.equal(noise[:, ~keep], old_noise[:, ~keep])
    row0 = torch.zeros(12, E, dtype=torch.float64, device=rows.device)
    eng.reset_masked(mask, x0=x0, row0=row0)
    out = eng.run(n_steps, controller='bb', ctrl_rows=ctrl_rows, trace=True)
    got = _snapshot(eng, out)
    eng.close()
    # (a) the untouched episodes == one uninterrupted run of 2 * n_steps on the old streams
    a = BatchEngine(E)
    a.reset(rows, old_noise, *S.device_scenarios(E, 2 * n_steps * 3, start_minute=360, seed=31, group_size=1), row0=False)
    wa = _snapshot(a, dict(trace=a.run(2 * n_steps, controller='bb', ctrl_rows=ctrl_rows, trace=True)['trace'][n_steps:]))
    a.close()
    # (b) the masked episodes == a fresh reset on the new streams and states
    b = BatchEngine(E)
    r0b = b.reset(rows, noise, mm, mg, x0=x0)
    wb = _snapshot(b, b.run(n_steps, controller='bb', ctrl_rows=ctrl_rows, trace=True))
    b.close()
    m = mask.bool()
    for k in ('trace', 'd', 'i', 'stats', 'status'):
        assert torch.equal(got[k][..., keep], wa[k][..., keep]), k
        assert torch.equal(got[k][..., m], wb[k][..., m]), k
    assert torch.equal(torch.nan_to_num(row0[:, m]), torch.nan_to_num(r0b[:, m])) and float(row0[:, keep].abs().sum()) == 0.0


def test_random_action_controller():
    import torch
    from simglucose_b200 import _capi
    from simglucose_b200.engine import BatchEngine
    E, n_steps = 65536, 20
    rows, noise, mm, mg, _ = _population(E, n_steps, seed=3)
    T = _capi.TRACE_COLS.index
    a = BatchEngine(E)
    a.reset(rows, noise, mm, mg, row0=False)
    tr = a.run(n_steps, controller='random', ctrl=(0.0, 0.05, 11.0, 1000.0), trace=True)['trace']
    sa = _snapshot(a, dict(trace=tr))
    basal = tr[:, T('basal')]
    assert float(basal.min()) >= 0.0 and float(basal.max()) <= 0.05 and float(tr[:, T('bolus')].abs().max()) == 0.0
    assert abs(float(basal.mean()) - 0.025) < 5e-5 and abs(float(basal.var()) - 0.05 ** 2 / 12) < 2e-6
    assert abs(float(np.corrcoef(basal[:-1].flatten().cpu().numpy(), basal[1:].flatten().cpu().numpy())[0, 1])) < 5e-3
    # stepping one env-step at a time continues the same action stream (counter = env-step index of the handle) ...
    b = BatchEngine(E)
    b.reset(rows, noise, mm, mg, row0=False)
    rows_b = torch.stack([b.run(1, controller='random', ctrl=(0.0, 0.05, 11.0, 1000.0), trace=True)['trace'][0]
                          for _ in range(n_steps)])
    _same(sa, _snapshot(b, dict(trace=rows_b)))
    # ... a shard that holds episodes [lo, hi) of the population draws the same actions for them ...
    lo, hi = 4096, 8192
    c = BatchEngine(hi - lo)
    c.reset(rows[lo:hi], noise[:, lo:hi].contiguous(), mm[:, lo:hi].contiguous(), mg[:, lo:hi].contiguous(), row0=False)
    trc = c.run(n_steps, controller='random', ctrl=(0.0, 0.05, 11.0, 1000.0 + lo), trace=True)['trace']
    assert torch.equal(torch.nan_to_num(trc), torch.nan_to_num(tr[:, :, lo:hi]))
    # ... and the reported actions fed back through SG_CTRL_ACTIONS reproduce the episode
    acts = torch.stack([basal, tr[:, T('bolus')]], dim=1).contiguous()
    d = BatchEngine(E)
    d.reset(rows, noise, mm, mg, row0=False)
    _same(sa, _snapshot(d, dict(trace=d.run(n_steps, controller='actions', actions=acts, trace=True)['trace'])))
    # a REUSED handle (second wave of a sweep, a ShardRunner slot) draws what a fresh one does: the env-step index of
    # the Philox counter restarts at sg_reset, whatever ran through the handle before (here: a bb warm-up and a run)
    d.reset(rows, noise, mm, mg, row0=False)
    d.run(7, controller='bb', ctrl=(140.0, 10.0, 30.0), trace=False)
    d.reset(rows, noise, mm, mg, row0=False)
    _same(sa, _snapshot(d, dict(trace=d.run(n_steps, controller='random', ctrl=(0.0, 0.05, 11.0, 1000.0), trace=True)['trace'])))
    # seed and offset are cast to unsigned integers in the kernel: negative or fractional values are refused
    for bad in ((0.0, 0.05, -1.0, 0.0), (0.0, 0.05, 11.5, 0.0), (0.0, 0.05, 11.0, -4.0), (0.0, 0.05, float('nan'), 0.0)):
        with pytest.raises(_capi.SgError, match='whole numbers'):
            d.run(1, controller='random', ctrl=bad, trace=False)
    for e in (a, b, c, d):
        e.close()


def test_bb_basal_by_name_for_perturbed_rows():
    """basal_bolus_ctrller.py:52-64 looks u2ss and BW up by patient NAME: a perturbed parameter row keeps the nominal
    basal when ctrl[3] carries it, and the oracle with the same nominal basal agrees."""
    import torch
    from simglucose_b200 import params as P, streams as S, _capi
    from simglucose_b200.engine import BatchEngine
    E, n_steps = 600, 100
    rows, noise, mm, mg, ctrl_rows = _population(E, n_steps, seed=41)
    pp = S.device_perturbed_params(rows, amplitude=0.1, seed=7)
    T = _capi.TRACE_COLS.index
    eng = BatchEngine(E)
    eng.reset(rows, noise, mm, mg, params=pp, row0=False)
    named = eng.run(n_steps, controller='bb', ctrl_rows=ctrl_rows, trace=True)['trace']
    own = ctrl_rows.clone(); own[3] = 0.0
    eng.reset(rows, noise, mm, mg, params=pp, row0=False)
    own_tr = eng.run(n_steps, controller='bb', ctrl_rows=own, trace=True)['trace']
    assert torch.equal(named[:, T('basal')], ctrl_rows[3].expand(n_steps, E))
    pph = pp.cpu().numpy()                                  # numpy: a true division (torch multiplies by 1 / 6000 on CUDA)
    want_own = torch.from_numpy(pph[1] * pph[0] / 6000.0).to(pp.device).expand(n_steps, E)
    assert torch.equal(own_tr[:, T('basal')], want_own) and not torch.equal(named[:, T('basal')], own_tr[:, T('basal')])
    # unperturbed rows: both conventions coincide bit for bit
    eng.reset(rows, noise, mm, mg, row0=False)
    t1 = eng.run(n_steps, controller='bb', ctrl_rows=ctrl_rows, trace=True)['trace']
    eng.reset(rows, noise, mm, mg, row0=False)
    t2 = eng.run(n_steps, controller='bb', ctrl_rows=own, trace=True)['trace']
    assert torch.equal(torch.nan_to_num(t1), torch.nan_to_num(t2))
    eng.close()


def test_noise_stream_exhaustion_is_reported_per_episode():
    import torch
    from simglucose_b200 import _capi
    from simglucose_b200.engine import BatchEngine
    E, n_steps = 256, 30
    rows, noise, mm, mg, ctrl_rows = _population(E, n_steps, seed=2)
    eng = BatchEngine(E)
    eng.reset(rows, noise, mm, mg, row0=False)
    eng.run(20, controller='bb', ctrl_rows=ctrl_rows, trace=False)
    mask = (torch.arange(E, device=rows.device) < 100).to(torch.int32)
    eng.reset_masked(mask)                                  # episodes 0..99 restart, 100.. are 20 steps in
    eng.run(20, controller='bb', ctrl_rows=ctrl_rows, trace=False)        # 40 > 30 for the untouched ones
    st = eng.status().cpu().numpy()
    assert (st[:100] == 1).all() and (st[100:] == _capi.STATUS_STREAM_EXHAUSTED).all()
    # without a masked reset the host-side check refuses the call instead
    eng.reset(rows, noise, mm, mg, row0=False)
    with pytest.raises(_capi.SgError):
        eng.run(n_steps + 1, controller='bb', ctrl_rows=ctrl_rows, trace=False)
    eng.close()


def test_step_cap_refuses_fp32_and_unknown_flags():
    import ctypes as C
    from simglucose_b200 import _capi
    from simglucose_b200.engine import BatchEngine
    eng = BatchEngine(64, fp32=True)
    with pytest.raises(_capi.SgError):
        eng.set_step_cap(5)
    eng.close()
    h = C.c_void_p()
    assert _capi.load().sg_create(0, 16, 4, C.byref(h)) == -1       # the removed SG_FLAG_PAIR bit


def test_handles_on_two_devices_in_one_process():
    """cudaFuncAttributeMaxDynamicSharedMemorySize is per device: the 51 KB launch must work on every GPU a process
    creates handles on.  The smoke batch on device 0 and device 1, same bits."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('one GPU')
    from simglucose_b200.engine import BatchEngine
    from simglucose_b200.patient.t1dpatient import T1DPatient
    E, n_steps = 30, 60
    res = []
    for dev in (0, 1):
        d = torch.device('cuda', dev)
        with torch.cuda.device(d):
            rows, noise, mm, mg, ctrl_rows = _population(E, n_steps, seed=1, dev=d)
            if res:
                rows, noise, mm, mg, ctrl_rows = [t.to(d) for t in res[0][1]]
            eng = BatchEngine(E, device=dev)
            eng.reset(rows, noise, mm, mg, row0=False)
            tr = eng.run(n_steps, controller='bb', ctrl_rows=ctrl_rows, trace=True)['trace']
            res.append((tr.cpu(), (rows, noise, mm, mg, ctrl_rows)))
            p = T1DPatient.withID(1, device=dev)            # patient_step_kernel on this device
            p.step(type('A', (), {'CHO': 0.0, 'insulin': 0.01})())
            eng.close()
    assert torch.equal(torch.nan_to_num(res[0][0]), torch.nan_to_num(res[1][0]))
