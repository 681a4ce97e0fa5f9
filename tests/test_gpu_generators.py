"""GPU: the device (Philox) generators of the per-episode input streams against the reference's distributions.

Every throughput number of bench.py runs on these streams, so they are pinned here against
  * the host MT19937 replay streams.cgm_noise_stream, itself pinned bit-for-bit to the reference's CGMNoise by
    tests/test_host_logic.py and tests/golden/generators.npz; 24,000 day schedules drawn by the reference's own
    RandomScenario in the build container (tests/golden/generator_samples.npz, oracle/make_golden.py), and
  * the analytic form of each draw in the reference source:
      CGMNoise / noise15_iter              simglucose/sensor/noise_gen.py:72-97
      RandomScenario.create_scenario       simglucose/simulation/scenario_gen.py:33-60
      +-10 % parameter jitter              BASELINE.json configs[3]
Thresholds are 4-5 sigma of the sampling error at the sizes used (stated next to each).
"""
import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def _acf(x, lag):
    """Mean within-stream autocorrelation at `lag`; x [n_streams, length]."""
    a = x - x.mean()
    return float((a[:, :-lag] * a[:, lag:]).mean() / a.var())


@pytest.mark.parametrize('sensor', ['Dexcom', 'GuardianRT', 'Navigator'])
def test_device_noise_matches_cgmnoise_distribution(sensor):
    import torch
    from scipy import stats
    from simglucose_b200 import params as P, streams as S
    L, E_dev, E_host = 256, 4096, 384                       # 2^20 device samples; 98,304 host samples per sensor
    dev = S.device_noise(E_dev, L, sensor, seed=12345, group_size=1).cpu().numpy().T          # [E, L]
    host = np.stack([S.cgm_noise_stream(sensor, 1000 + s, L) for s in range(E_host)])
    assert np.all(np.isfinite(dev))
    # the interpolated stream is strongly autocorrelated: the effective sample size is the number of 15-minute knots
    st = S.P.sensor_params(sensor)['sample_time']
    knots_host = E_host * L * st / 15.0
    se = 1.0 / np.sqrt(knots_host)
    sd = host.std()
    assert abs(dev.mean() - host.mean()) < 5 * sd * se
    assert abs(dev.std() / sd - 1.0) < 5 * se * 1.5          # heavy-tailed (Johnson SU): kurtosis inflates the error
    for lag in (1, 5):
        assert abs(_acf(dev, lag) - _acf(host, lag)) < 0.02, lag
    for q in (1, 50, 99):                                    # the host percentile sits at the same level of the device sample
        ph = np.percentile(host, q)
        level = float((dev <= ph).mean())
        assert abs(level - q / 100.0) < 5 * np.sqrt(q / 100.0 * (1 - q / 100.0)) * se * 1.1, (q, ph, level)
    # two-sample Kolmogorov-Smirnov on one sample per 15-minute knot (approximately independent draws)
    step = max(1, int(round(15.0 / st)))
    ks = stats.ks_2samp(dev[:, ::step].ravel(), host[:, ::step].ravel()).statistic
    n, m = dev[:, ::step].size, host[:, ::step].size
    assert ks < 2.0 * np.sqrt((n + m) / (n * m)), ks         # 1.36 is the 5 % point, 1.95 the 0.1 % point
    # the first sample of a stream (noise_gen.py:24,84-85): Johnson transform of a standard normal
    first_d, first_h = dev[:, 0], np.array([S.cgm_noise_stream(sensor, 5000 + s, 1)[0] for s in range(2000)])
    assert stats.ks_2samp(first_d, first_h).statistic < 2.0 * np.sqrt((4096 + 2000) / (4096 * 2000))
    # streams of different episodes are independent; a seed group shares one
    assert abs(np.corrcoef(dev[0::2, 40], dev[1::2, 40])[0, 1]) < 5 / np.sqrt(E_dev / 2)
    grp = S.device_noise(60, L, sensor, seed=7, group_size=30).cpu().numpy()
    assert np.array_equal(grp[:, :30], np.repeat(grp[:, :1], 30, 1)) and not np.array_equal(grp[:, 0], grp[:, 30])
    # keyed streams: the warp-cooperative kernel that rewrites a few masked columns of a resident buffer gives the same
    # bits as the one-thread-per-episode kernel, and leaves the other columns alone
    key = torch.arange(1000, device='cuda', dtype=torch.int64) * 3 + 11
    full = S.device_noise_keyed(torch.zeros(L + 3, 1000, dtype=torch.float64, device='cuda'), key, sensor, seed=3)
    mask = (torch.arange(1000, device='cuda') % 7 == 2).to(torch.int32)
    part = S.device_noise_keyed(torch.full((L + 3, 1000), -1.0, dtype=torch.float64, device='cuda'), key, sensor, seed=3, mask=mask)
    m = mask.bool()
    assert torch.equal(part[:, m], full[:, m]) and bool((part[:, ~m] == -1.0).all())
    ones = torch.ones(1000, dtype=torch.int32, device='cuda')
    assert torch.equal(S.device_noise_keyed(torch.zeros_like(full), key, sensor, seed=3, mask=ones), full)
    # ... and so does the kernel that works through a gathered LIST of episodes (any order, count read on the device)
    idx = m.nonzero().reshape(-1).to(torch.int32)
    lst = torch.zeros(1000, dtype=torch.int32, device='cuda')
    lst[:idx.numel()] = idx[torch.randperm(idx.numel(), device='cuda')]
    cnt = torch.tensor([idx.numel()], dtype=torch.int32, device='cuda')
    W = torch.from_numpy(S.interp_matrix(P.sensor_params(sensor)['sample_time'])).cuda()
    lp = S.device_noise_keyed(torch.full((L + 3, 1000), -1.0, dtype=torch.float64, device='cuda'), key, sensor, seed=3, W=W,
                              listed=(lst, cnt))
    assert torch.equal(lp[:, m], full[:, m]) and bool((lp[:, ~m] == -1.0).all())
    none = S.device_noise_keyed(torch.full((L + 3, 1000), -1.0, dtype=torch.float64, device='cuda'), key, sensor, seed=3, W=W,
                                listed=(lst, torch.zeros(1, dtype=torch.int32, device='cuda')))
    assert bool((none == -1.0).all())


_PROB = [0.95, 0.3, 0.95, 0.3, 0.95, 0.3]                   # scenario_gen.py:38-44
_LB = np.array([5, 9, 10, 14, 16, 20]) * 60
_UB = np.array([9, 10, 14, 16, 20, 23]) * 60
_MU = np.array([7, 9.5, 12, 15, 18, 21.5]) * 60
_SD = np.array([60, 30, 60, 30, 60, 30])
_AMU = [45, 10, 70, 10, 80, 10]
_ASD = [10, 5, 10, 5, 10, 5]


def _slot_of(minute_of_day):
    """Meal slot of a minute (windows touch at their ends: the boundary minute goes to the earlier slot)."""
    s = np.searchsorted(_UB, minute_of_day, side='left')
    return np.where((minute_of_day >= _LB[0]) & (minute_of_day <= _UB[-1]), s, -1)


def test_device_scenarios_match_random_scenario_distribution():
    from scipy import stats
    from simglucose_b200 import streams as S
    E, days = 65536, 2                                      # 131,072 day schedules on the device
    mm, mg = S.device_scenarios(E, days * 1440, start_minute=0, seed=777, group_size=1)
    mm, mg = mm.cpu().numpy(), mg.cpu().numpy()
    assert mm.shape[0] == 6 * days
    valid = mm >= 0
    # table shape: firing order, -1 padding only at the end, whole positive grams, one meal per minute
    assert np.all((np.diff(np.where(valid, mm, 10 ** 6), axis=0) > 0) | ~valid[1:])
    assert np.all(valid[:-1] | ~valid[1:])
    assert np.all(mg[valid] > 0) and np.all(mg[valid] == np.round(mg[valid])) and np.all(mg[~valid] == 0)
    tod = mm[valid] % 1440
    slot = _slot_of(tod)
    assert np.all(slot >= 0) and np.all(tod >= _LB[slot]) and np.all(tod <= _UB[slot])      # inside [lb, ub]
    n_days = E * days
    # 24,000 day schedules drawn by the REFERENCE's own RandomScenario.create_scenario in the build container
    # (oracle/make_golden.py generator_samples -> tests/golden/generator_samples.npz)
    ref = H.load('generator_samples')
    H_DAYS = ref['day_minute'].shape[0]
    host_t = [ref['day_minute'][:, k][(ref['day_minute'][:, k] >= 0) & (ref['day_grams'][:, k] > 0)].astype(float) for k in range(6)]
    host_g = [ref['day_grams'][:, k][(ref['day_minute'][:, k] >= 0) & (ref['day_grams'][:, k] > 0)].astype(float) for k in range(6)]
    for i in range(6):
        sel = slot == i
        n_i = int(sel.sum())
        # meal probability: p * P(round(N(mu, sd)) > 0) (a zero-gram draw is "no meal" in get_action); the boundary
        # minutes shared with the neighbouring window carry ~1e-3 of the mass
        p_zero = stats.norm.cdf((0.5 - _AMU[i]) / _ASD[i])
        p = _PROB[i] * (1.0 - p_zero)
        assert abs(n_i / n_days - p) < 4 * np.sqrt(p * (1 - p) / n_days) + 2e-3, (i, n_i / n_days, p)
        assert abs(len(host_t[i]) / H_DAYS - p) < 4 * np.sqrt(p * (1 - p) / H_DAYS) + 2e-3
        # meal time: truncated normal rounded to the minute -- analytic CDF and the host replay
        t_dev = tod[sel].astype(float)
        a, b = (_LB[i] - _MU[i]) / _SD[i], (_UB[i] - _MU[i]) / _SD[i]
        tn = stats.truncnorm(a, b, loc=_MU[i], scale=_SD[i])
        assert abs(t_dev.mean() - tn.mean()) < 5 * tn.std() / np.sqrt(n_i) + 0.05
        assert abs(t_dev.std() / tn.std() - 1.0) < 5 / np.sqrt(2 * n_i) + 2e-3
        grid = np.arange(_LB[i], _UB[i] + 1)
        ecdf = np.searchsorted(np.sort(t_dev), grid + 0.5) / n_i
        assert np.abs(ecdf - tn.cdf(grid + 0.5)).max() < 2.0 / np.sqrt(n_i) + 1e-3
        assert stats.ks_2samp(t_dev, np.array(host_t[i])).statistic < 2.0 * np.sqrt(1 / n_i + 1 / len(host_t[i])) + 2e-3
        # grams: max(round(normal), 0) -- analytic probability mass and the host replay
        g_dev = mg[valid][sel]
        ks_vals = np.arange(1, int(_AMU[i] + 6 * _ASD[i]))
        pm = stats.norm.cdf((ks_vals + 0.5 - _AMU[i]) / _ASD[i]) - stats.norm.cdf((ks_vals - 0.5 - _AMU[i]) / _ASD[i])
        pm = pm / pm.sum()
        emp = np.array([(g_dev == k).mean() for k in ks_vals])
        assert np.abs(emp - pm).max() < 4 * np.sqrt(0.25 / n_i) / 2 + 1e-3
        assert abs(g_dev.mean() - np.mean(host_g[i])) < 5 * _ASD[i] * np.sqrt(1 / n_i + 1 / len(host_g[i]))
        assert stats.ks_2samp(g_dev, np.array(host_g[i], dtype=float)).statistic < 2.0 * np.sqrt(1 / n_i + 1 / len(host_g[i])) + 5e-3
    # presence of the six meals of a day is independent (one rand() each): breakfast vs lunch
    day0 = [((mm >= 0) & (mm < 1440) & (_slot_of(mm % 1440) == i)).any(axis=0) for i in (0, 2)]
    assert abs(np.corrcoef(day0[0], day0[1])[0, 1]) < 5 / np.sqrt(E)
    # a later start loses the meals before it, as RandomScenario does on its first day (scenario_gen.py:20-29)
    mm6, _ = S.device_scenarios(4096, 1440, start_minute=600, seed=5, group_size=1)
    m6 = mm6.cpu().numpy()
    assert m6[m6 >= 0].min() >= 0 and ((m6[m6 >= 0] + 600) % 1440 >= 0).all() and m6.max() < 1440
    # keyed schedules: the warp-cooperative kernel that rewrites a few masked columns of resident tables (one candidate
    # meal per lane, random access into the Philox stream) gives the same tables as the one-thread-per-episode kernel,
    # for every start hour and horizons of one to four calendar days (up to 24 candidates), and leaves the rest alone
    import torch
    n = 3000
    key = torch.arange(n, device='cuda', dtype=torch.int64) * 5 + 3
    for horizon in (480, 1440, 3 * 1440 + 77):
        starts = ((torch.arange(n, device='cuda') * 37) % 1440).to(torch.int32)
        M = 6 * ((1439 + horizon + 1439) // 1440)
        full = S.device_scenarios_keyed(torch.full((M, n), -7, dtype=torch.int32, device='cuda'),
                                        torch.full((M, n), -7.0, dtype=torch.float64, device='cuda'), starts, horizon, key, seed=21)
        mask = (torch.arange(n, device='cuda') % 11 == 4).to(torch.int32)
        part = S.device_scenarios_keyed(torch.full((M, n), -7, dtype=torch.int32, device='cuda'),
                                        torch.full((M, n), -7.0, dtype=torch.float64, device='cuda'), starts, horizon, key, seed=21, mask=mask)
        mb = mask.bool()
        assert torch.equal(part[0][:, mb], full[0][:, mb]) and torch.equal(part[1][:, mb], full[1][:, mb])
        assert bool((part[0][:, ~mb] == -7).all()) and bool((part[1][:, ~mb] == -7.0).all())
        ones = torch.ones(n, dtype=torch.int32, device='cuda')
        allm = S.device_scenarios_keyed(torch.zeros_like(full[0]), torch.zeros_like(full[1]), starts, horizon, key, seed=21, mask=ones)
        assert torch.equal(allm[0], full[0]) and torch.equal(allm[1], full[1])
        assert int((full[0] >= 0).sum()) > n * (horizon // 480)                                # the tables are not empty
        idx = mb.nonzero().reshape(-1).to(torch.int32)
        lst = torch.zeros(n, dtype=torch.int32, device='cuda')
        lst[:idx.numel()] = idx[torch.randperm(idx.numel(), device='cuda')]
        cnt = torch.tensor([idx.numel()], dtype=torch.int32, device='cuda')
        lp = S.device_scenarios_keyed(torch.full((M, n), -7, dtype=torch.int32, device='cuda'),
                                      torch.full((M, n), -7.0, dtype=torch.float64, device='cuda'), starts, horizon, key, seed=21,
                                      listed=(lst, cnt))
        assert torch.equal(lp[0][:, mb], full[0][:, mb]) and torch.equal(lp[1][:, mb], full[1][:, mb])
        assert bool((lp[0][:, ~mb] == -7).all()) and bool((lp[1][:, ~mb] == -7.0).all())
    # the host table builder and the device kernel agree on how a schedule becomes a table (start 00:00, 1 day)
    hm, hg = S.random_scenario_table(3, 0, 1440)
    assert np.all(np.diff(hm) > 0) and np.all(hg > 0)


def test_device_perturbed_params_and_uniform():
    import torch
    from scipy import stats
    from simglucose_b200 import params as P, streams as S
    E = 262144
    rows = (torch.arange(E, device='cuda', dtype=torch.int64) % 30).to(torch.int32)
    pp = S.device_perturbed_params(rows, amplitude=0.1, seed=1234)
    base = torch.from_numpy(P.cohort_table()[0]).cuda()[rows.long()].T
    u = ((pp / base - 1.0) / 0.1).cpu().numpy()                            # ~ U(-1, 1) i.i.d. per parameter
    thr = 2.0 / np.sqrt(E)
    for i in range(32):
        assert u[i].min() >= -1.0 - 1e-12 and u[i].max() <= 1.0 + 1e-12
        assert stats.kstest(u[i], 'uniform', args=(-1.0, 2.0)).statistic < thr, i
    c = np.corrcoef(u)
    off = c - np.eye(32)
    assert np.abs(off).max() < 5.5 / np.sqrt(E)                            # 496 pairs: independent across parameters
    assert abs(np.corrcoef(u[0, :-1], u[0, 1:])[0, 1]) < 5 / np.sqrt(E)    # and across episodes
    # a shard draws the same parameters for its episodes (Philox keyed by global episode index)
    lo = 100000
    sh = S.device_perturbed_params(rows[lo:lo + 5000].contiguous(), amplitude=0.1, seed=1234, episode_offset=lo)
    assert torch.equal(sh, pp[:, lo:lo + 5000])
    un = S.device_uniform(1 << 20, 0.0, 0.05, seed=3, offset=17).cpu().numpy()
    assert stats.kstest(un, 'uniform', args=(0.0, 0.05)).statistic < 2.0 / np.sqrt(un.size)
    assert abs(np.corrcoef(un[:-1], un[1:])[0, 1]) < 5 / np.sqrt(un.size)
    assert np.array_equal(S.device_uniform(1000, 0.0, 0.05, seed=3, offset=17 + 500).cpu().numpy()[:500], un[500:1000])
