"""GPU: analysis/report.py statistics (percent_stats, risk_index_trace, CVGA) computed by sg_report_stats /
sg_cvga_zones against (a) the numbers the reference's own report.py produced on the cohort fixtures
(tests/golden/report_stats.npz, oracle/make_golden.py g_report) and (b) numpy on adversarial traces
(ties, constant columns, tiny / negative BG, ragged last chunk)."""
import numpy as np
import pandas as pd
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def _names(n=30):
    from simglucose_b200 import params as P
    return P.patient_names()[:n]


@pytest.mark.parametrize('tag', ['bb', 'pid'])
def test_report_functions_match_reference(tag):
    from simglucose_b200.analysis import report as R
    d = H.load('report_stats')
    BG = pd.DataFrame(d[tag + '_BG'], columns=_names())
    pstats, _, _ = R.percent_stats(BG)
    assert list(pstats.columns) == [str(c) for c in d[tag + '_percent_cols']] and list(pstats.index) == _names()
    assert np.array_equal(pstats.to_numpy(), d[tag + '_percent'])                 # integer counts / n * 100: exact
    ri_per_hour, ri_mean, _, _ = R.risk_index_trace(BG)
    assert ri_per_hour.shape == d[tag + '_ri_per_hour'].shape
    assert list(ri_mean.columns) == [str(c) for c in d[tag + '_ri_mean_cols']]
    ref = d[tag + '_ri_per_hour']
    got = ri_per_hour.to_numpy()
    assert np.array_equal(np.isnan(got), np.isnan(ref))
    assert np.nanmax(np.abs(got - ref) / np.maximum(np.abs(ref), 1.0)) <= 1e-10
    gm, rm = ri_mean.to_numpy(), d[tag + '_ri_mean']
    assert np.array_equal(np.isnan(gm), np.isnan(rm))
    assert np.nanmax(np.abs(gm - rm) / np.maximum(np.abs(rm), 1.0)) <= 1e-10
    mn, mx, A, B, C_, D_, E = R.CVGA_analysis(BG)
    assert np.max(np.abs(mn - d[tag + '_cvga_min'])) <= 1e-12 and np.max(np.abs(mx - d[tag + '_cvga_max'])) <= 1e-12
    assert np.allclose([A, B, C_, D_, E], d[tag + '_zones'], rtol=0, atol=1e-15)
    zs, _, _ = R.CVGA([BG, BG], label=['a', 'b'])
    assert zs.shape == (2, 5) and np.allclose(zs.to_numpy()[1], d[tag + '_zones'], atol=1e-15)


def test_report_writes_the_three_csv_files(tmp_path):
    from simglucose_b200.analysis import report as R
    d = H.load('report_stats')
    names = _names()
    t = pd.date_range('2018-01-01', periods=d['bb_BG'].shape[0], freq='3min')
    frames = [pd.DataFrame({'BG': d['bb_BG'][:, i], 'CGM': d['bb_BG'][:, i]}, index=pd.Index(t, name='Time')) for i in range(30)]
    df = pd.concat(frames, keys=names)
    results, ri_per_hour, zone_stats, figs, axes = R.report(df, str(tmp_path))
    assert list(results.columns) == R.PERCENT_COLS + R.RI_COLS and list(results.index) == names
    assert np.array_equal(results[R.PERCENT_COLS].to_numpy(), d['bb_percent'])
    for f in ('performance_stats.csv', 'risk_trace.csv', 'CVGA_stats.csv'):
        assert (tmp_path / f).exists()
    back = pd.read_csv(tmp_path / 'CVGA_stats.csv', index_col=0)
    assert list(back.columns) == ['A', 'B', 'C', 'D', 'E'] and np.allclose(back.to_numpy()[0], d['bb_zones'])


def _np_reference(bg, chunk):
    """numpy restatement of report.py:74-92, 95-133 (pandas 1.2.4 semantics: column means skip NaN), 198-204."""
    n = bg.shape[0]
    pct = np.stack([((bg >= 70) & (bg <= 180)).sum(0), (bg > 180).sum(0), (bg < 70).sum(0), (bg > 250).sum(0),
                    (bg < 50).sum(0)]) / n * 100
    ri = []
    with np.errstate(all='ignore'):
        for r0 in range(0, n, chunk):
            c = bg[r0:r0 + chunk]
            f = 1.509 * (np.log(np.where(c > 0, c, np.nan)) ** 1.084 - 5.381)
            m = np.array([np.nan if np.all(np.isnan(f[:, j])) else np.nanmean(f[:, j]) for j in range(f.shape[1])])
            lb, hb = 10 * (m * (m < 0)) ** 2, 10 * (m * (m > 0)) ** 2
            ri.append(np.stack([lb, hb, lb + hb]))
    mn, mx = np.percentile(bg, 2.5, axis=0), np.percentile(bg, 97.5, axis=0)
    return pct, np.stack(ri, 1), np.clip(mn, 50, 400), np.clip(mx, 50, 400)


@pytest.mark.parametrize('n_rows,chunk', [(1, 60), (2, 60), (61, 60), (481, 60), (1000, 7)])
def test_device_report_adversarial_traces(n_rows, chunk):
    import torch
    from simglucose_b200.analysis.report import device_report
    rng = np.random.RandomState(n_rows)
    E = 257
    bg = np.exp(rng.uniform(np.log(20.0), np.log(500.0), size=(n_rows, E)))
    bg[:, 0] = 112.5                                    # constant column: every order statistic ties
    bg[:, 1] = np.round(bg[:, 1] / 25.0) * 25.0         # heavy ties
    bg[:, 2] = rng.choice([70.0, 180.0, 250.0, 50.0], n_rows)      # exactly on the thresholds
    bg[:, 3] = np.where(rng.rand(n_rows) < 0.5, -1e-12, 0.5)       # dead patient: BG <= 0 or log(BG) < 0 -> NaN chunks
    bg[:, 4] = np.where(rng.rand(n_rows) < 0.3, 0.0, bg[:, 4])
    bg[::3, 5] = 400.0
    # embed as one column of a wider trace to exercise the row stride
    wide = np.full((n_rows, 3, E), np.nan)
    wide[:, 1, :] = bg
    tw = torch.from_numpy(wide).cuda()
    out = device_report(tw[:, 1, :], chunk_rows=chunk)
    pct, ri, mn, mx = _np_reference(bg, chunk)
    assert np.array_equal(out['percent'].cpu().numpy(), pct)
    got = out['ri'].cpu().numpy()
    assert got.shape == ri.shape and np.array_equal(np.isnan(got), np.isnan(ri))
    assert np.nanmax(np.abs(got - ri) / np.maximum(np.abs(ri), 1.0), initial=0.0) <= 1e-10
    assert np.array_equal(out['cvga'].cpu().numpy(), np.stack([mn, mx]))      # exact order statistics + numpy's lerp
    z = out['zones'].cpu().numpy()
    A = ((mn > 90) & (mn <= 110) & (mx >= 110) & (mx < 180)).sum()
    Ee = ((mn <= 70) & (mx >= 300)).sum()
    assert z[0] == A and z[4] == Ee and z.sum() <= 2 * E
