"""GPU parity: the CUDA path (through the C-ABI) against (a) the golden fixtures recorded from the
unmodified reference and (b) the CPU oracle on the same inputs.

Tolerance (BASELINE.json north_star / SURVEY.md 8d): max |dBG| / max(|BG|, 1 mg/dL) <= 1e-6 in FP64,
same rule for CGM; integer/index work (announced CHO per step, pump-quantised insulin, done flags,
solver status) exact.
"""
import numpy as np
import pytest

import helpers as H
from helpers import O

pytestmark = pytest.mark.gpu

TOL = H.REL_TOL          # the north-star contract: 1e-6
# What the strict FP64 path actually delivers against the reference (measured 2e-13 .. 3e-11 over days; the kernel's exp /
# log / pow are within 1 ulp of libm and every accept / reject decision of DOPRI5 is the reference's): asserted beside
# the contract so that a regression of several orders of magnitude cannot hide under it.
MEASURED = 1e-9


def _engine(n, **kw):
    from simglucose_b200.engine import BatchEngine
    return BatchEngine(n, **kw)


def _run_fixture(d, idx, carry_h=False, fma=False, counters=True):
    """Run episodes idx (same n_steps / controller kind / hardware) of fixture d in one batch."""
    idx = list(idx)
    n_steps = int(d['n_steps'][idx[0]])
    kind, _ = H.controller_of(d, idx[0])
    sn, pn = H.hardware_of(d, idx[0])
    eng = _engine(len(idx), carry_h=carry_h, fma=fma, sensor=sn, pump=pn)
    ctrl_rows = np.zeros((4, len(idx)))
    for j, i in enumerate(idx):
        k, c = H.controller_of(d, i)
        assert k == kind and int(d['n_steps'][i]) == n_steps
        ctrl_rows[:len(c), j] = c
    params = np.ascontiguousarray(d['params'][idx].T) if 'params' in d.files else None     # perturbed rows [32, E]
    row0 = eng.reset(d['patient_row'][idx], d['noise'][idx].T, d['meal_minute'][idx].T, d['meal_grams'][idx].T,
                     x0=d['x0'][idx].T, params=params)
    actions = np.ascontiguousarray(d['actions'][idx].transpose(1, 2, 0)) if kind == 'actions' else None
    out = eng.run(n_steps, controller=kind, ctrl_rows=ctrl_rows, actions=actions, trace=True, counters=counters)
    tr = np.concatenate([row0.cpu().numpy()[None], out['trace'].cpu().numpy()], 0)   # [n_steps+1, NTRACE, E]
    res = dict(trace=tr, status=eng.status().cpu().numpy(), state=eng.state().cpu().numpy(),
               stats=eng.stats().cpu().numpy(), agg=eng.reduce_stats().cpu().numpy(),
               counters=out['counters'].cpu().numpy() if counters else None)
    eng.close()
    return res


def _col(res, name):
    from simglucose_b200 import _capi
    return res['trace'][:, _capi.TRACE_COLS.index(name), :].T      # [E, n_steps+1]


def _check_against_reference(d, idx, res, tol=TOL, strict=True):
    """Golden columns recorded from the reference.  CHO / insulin of reference row r describe the step
    r -> r+1 (env.py:88-97); the library stores them on the row the step produced.
    strict: the default arithmetic (no FMA contraction) is also held to the measured level; SG_FLAG_FMA, whose step
    decisions may differ from scipy's, only to the contract."""
    idx = list(idx)
    for name in ('BG', 'CGM', 'LBGI', 'HBGI', 'Risk'):
        got, ref = _col(res, name), d[name][idx]
        if name in ('BG', 'CGM'):
            assert H.rel_err(got, ref) <= tol, name
            if strict:
                assert H.rel_err(got, ref) <= MEASURED, (name, H.rel_err(got, ref))
        else:
            # risk is ~20x steeper than BG around 110 mg/dL: 1e-8 is the bound the oracle holds (test_oracle_golden.py)
            assert H.rel_err(got, ref) <= (1e-8 if strict else 2e-5), (name, H.rel_err(got, ref))
    cho, ins = _col(res, 'CHO')[:, 1:], _col(res, 'insulin')[:, 1:]
    assert np.array_equal(cho, d['CHO'][idx][:, :-1]), 'announced CHO per step must be exact'
    assert np.max(np.abs(ins - d['insulin'][idx][:, :-1])) <= 1e-15, 'pump-quantised insulin'
    assert H.rel_err(_col(res, 'obs'), d['obs'][idx]) <= (MEASURED if strict else tol)
    if 'reward' in d.files:
        assert H.rel_err(_col(res, 'reward'), d['reward'][idx]) <= (1e-8 if strict else 1e-4)
    assert np.array_equal(_col(res, 'done') != 0, d['done'][idx])
    assert np.all(res['status'] == 1)
    assert H.rel_err(res['state'].T, d['final_state'][idx]) <= tol
    if strict:
        # Hidden states are held to the measured level for every patient whose glucose stays physiological.  A patient
        # that the reference drives to BG ~ 0 (child#004 under the example PID gains) sits on the
        # (x >= 0) * dxdt masks: x3, x4 are pinned at +-1e-10 by a zero crossing whose overshoot is pure rounding, and
        # x4 later regrows EXPONENTIALLY from that seed (Vmx * x6 < 0), so its final value carries the seed's relative
        # rounding difference (measured 8.8e-8; BG itself agrees to 1e-13 there).  Contract bound only.
        alive = np.array([np.nanmin(d['BG'][i]) >= 1.0 for i in idx])
        if alive.any():
            assert H.rel_err(res['state'].T[alive], d['final_state'][idx][alive]) <= MEASURED


def test_golden_config_matches_sim_results_csv():
    """tests/test_sim_engine.py:87-113 of the reference: adolescent#001, 2 days, BB, seeds 1.
    carry_h=1 is the pinned scipy 1.6.3 behaviour that produced tests/sim_results.csv."""
    d = H.load('golden_config')
    csv = d['sim_results_csv']                        # BG, CGM, CHO, insulin, LBGI, HBGI, Risk
    res = _run_fixture(d, [0], carry_h=True)
    assert H.rel_err(_col(res, 'BG')[0], csv[:, 0]) <= 1e-9
    assert H.rel_err(_col(res, 'CGM')[0], csv[:, 1]) <= 1e-9
    assert np.max(np.abs(_col(res, 'CHO')[0, 1:] - csv[:-1, 2])) <= 1e-12
    assert np.max(np.abs(_col(res, 'insulin')[0, 1:] - csv[:-1, 3])) <= 1e-15
    # the reference's own criterion: assert_frame_equal rtol 1e-5 on every column
    for c, name in enumerate(['BG', 'CGM', None, None, 'LBGI', 'HBGI', 'Risk']):
        if name:
            np.testing.assert_allclose(_col(res, name)[0], csv[:, c], rtol=1e-5, atol=1e-8)
    # and the scipy >= 1.7 mode reproduces the reference run recorded in this image
    res0 = _run_fixture(d, [0], carry_h=False)
    _check_against_reference(d, [0], res0)
    np.testing.assert_allclose(_col(res0, 'BG')[0], csv[:, 0], rtol=1e-5, atol=1e-8)


@pytest.mark.parametrize('fixture', ['custom_scenario', 'open_loop_actions', 'hardware_0', 'hardware_1', 'hardware_2'])
def test_small_fixtures(fixture):
    d = H.load(fixture)
    n = d['patient_row'].shape[0]
    res = _run_fixture(d, range(n))
    _check_against_reference(d, range(n), res)
    # per-minute DOPRI5 step counts vs scipy's iwork (diagnostic with a hard floor)
    same = (res['counters'][:, 0, :].T == d['dop_counters'][:, :, 1])
    assert same.mean() >= 0.999, same.mean()


def test_cohort_bb_3day_all_30_patients():
    """BASELINE.json configs[1]: all 30 virtual patients x 3-day RandomScenario, BBController."""
    d = H.load('cohort_bb_3day')
    res = _run_fixture(d, range(30))
    _check_against_reference(d, range(30), res)
    same = (res['counters'][:, 0, :].T == d['dop_counters'][:, :, 1])
    same_rej = (res['counters'][:, 1, :].T == d['dop_counters'][:, :, 3]) if res['counters'].shape[1] > 1 else same
    # the count itself goes on record (pytest -s; profiles/r03_*): how many of the 129,600 patient-minutes take a
    # different number of DOPRI5 steps / rejections than scipy did (norms and step controller are <= 2 ulp, not IEEE
    # quotients: a decision can flip only when errsum lands within that of the threshold)
    print('cohort_bb_3day: %d of %d patient-minutes differ from scipy in DOPRI5 steps, %d in rejections'
          % (int((~same).sum()), same.size, int((~same_rej).sum())))
    # measured on a B200 (profiles/r03_c_gpu_all.log): 0 of 129,600 differ, in steps and in rejections -- held to that
    assert same.all() and same_rej.all(), (int((~same).sum()), int((~same_rej).sum()))


@pytest.mark.parametrize('fixture', ['perturbed_1day', 'perturbed_7day'])
def test_perturbed_parameter_rows_against_reference(fixture):
    """BASELINE.json configs[3] anchored in the reference: episodes of the +-10 % perturbed population (global indices
    0..59 for a day; five for a week, three of which collapse to BG ~ 0) simulated by the UNMODIFIED reference on the
    perturbed parameter Series, against the CUDA path through params_soa + ctrl_rows[3] (the nominal basal the
    reference's BBController finds by patient name, basal_bolus_ctrller.py:52-64).  Same bounds as every other
    reference fixture.  And the device generator hands the simulator exactly the rows the fixture was recorded on."""
    import torch
    from simglucose_b200 import streams as S
    d = H.load(fixture)
    n = d['patient_row'].shape[0]
    res = _run_fixture(d, range(n))
    _check_against_reference(d, range(n), res)
    same = (res['counters'][:, 0, :].T == d['dop_counters'][:, :, 1])
    print('%s: %d of %d patient-minutes differ from scipy in their DOPRI5 step count' % (fixture, int((~same).sum()), same.size))
    assert same.all(), int((~same).sum())          # measured: none (profiles/r03_c_gpu_all.log)
    if fixture == 'perturbed_7day':
        assert 2 <= int((np.nanmin(d['BG'], axis=1) < 1.0).sum()) < n      # collapsed and physiological rows both present
    # sg_gen_perturbed_params at these global indices == the rows the reference patients were built on, bit for bit
    for i in range(n):
        g = int(d['global_index'][i])
        row = torch.tensor([g % 30], dtype=torch.int32, device='cuda:0')
        got = S.device_perturbed_params(row, amplitude=float(d['perturb_amplitude']), seed=int(d['perturb_seed']),
                                        episode_offset=g)
        assert np.array_equal(got.cpu().numpy()[:, 0], d['params'][i]), g


def test_cohort_pid_1day():
    """BASELINE.json configs[4] subset: 30 patients x seeds 0..3, PID fused on device."""
    d = H.load('cohort_pid_1day')
    res = _run_fixture(d, range(120))
    _check_against_reference(d, range(120), res)


def test_oracle_bitwise_integer_work_and_stats():
    """CUDA path vs the CPU oracle on the same inputs: statistics, counters, trace."""
    d = H.load('custom_scenario')
    n = d['patient_row'].shape[0]
    res = _run_fixture(d, range(n))
    for i in range(n):
        o = H.oracle_episode(d, i)
        assert H.rel_err(_col(res, 'BG')[i], o['BG']) <= 1e-9
        assert H.rel_err(_col(res, 'reward')[i], o['reward']) <= 1e-9
        assert np.array_equal(res['counters'][:, 0, i], o['counters'][:, 1])
        st = dict(zip(O.STAT_COLS, res['stats'][:, i]))
        for k in ('rows', 'lt50', 'lt70', 'inrange', 'gt180', 'gt250', 'first_done', 'dopri_steps', 'dopri_rejects'):
            assert st[k] == o['stats'][k], k
        for k in ('sum_bg', 'min_bg', 'max_bg', 'sum_lbgi', 'sum_hbgi', 'sum_risk'):
            assert abs(st[k] - o['stats'][k]) <= 1e-6 * max(1.0, abs(o['stats'][k])), k
    # population aggregate == fixed-order reduction of the per-episode statistics
    agg, s = res['agg'], res['stats']
    assert agg[0] == s[0].sum() and agg[2] == s[2].min() and agg[3] == s[3].max()
    assert abs(agg[1] - s[1].sum()) <= 1e-9 * abs(s[1].sum())
    assert agg[13] == (s[13] >= 0).sum()


def test_lockstep_step_equals_fused_run():
    """sg_step (one launch per env-step, state round-trips through HBM) == sg_run (fused)."""
    import torch
    d = H.load('open_loop_actions')
    idx = list(range(d['patient_row'].shape[0]))
    fused = _run_fixture(d, idx, counters=False)
    eng = _engine(len(idx))
    eng.reset(d['patient_row'][idx], d['noise'][idx].T, d['meal_minute'][idx].T, d['meal_grams'][idx].T,
              x0=d['x0'][idx].T)
    acts = torch.from_numpy(np.ascontiguousarray(d['actions'][idx].transpose(1, 2, 0))).cuda()
    rows = [eng.step(acts[k, 0], acts[k, 1]).cpu().numpy() for k in range(acts.shape[0])]
    assert np.array_equal(np.stack(rows), fused['trace'][1:], equal_nan=True)
    assert np.array_equal(eng.state().cpu().numpy(), fused['state'])
    assert np.array_equal(eng.stats().cpu().numpy(), fused['stats'])
    eng.close()


def test_fma_mode_within_tolerance():
    """SG_FLAG_FMA (contraction allowed) stays inside the FP64 tolerance on the 3-day cohort."""
    d = H.load('cohort_bb_3day')
    res = _run_fixture(d, range(30), fma=True)
    _check_against_reference(d, range(30), res, strict=False)


def test_missing_noise_is_an_error():
    from simglucose_b200 import _capi
    d = H.load('custom_scenario')
    eng = _engine(1)
    eng.reset(d['patient_row'][:1], d['noise'][:1, :10].T, d['meal_minute'][:1].T, d['meal_grams'][:1].T)
    with pytest.raises(_capi.SgError):
        eng.run(20, controller='bb', ctrl=(140.0, 10.0, 30.0))
    eng.close()
