"""Post-hoc statistics of simulation results -- the numbers of simglucose/analysis/report.py (percent_stats
:74-92, risk_index_trace :95-133, CVGA_analysis :198-217, CVGA :220-243, report :246-268) computed by the
CUDA library (sg_report_stats, sg_cvga_zones).  The reference draws matplotlib figures next to the numbers;
this build is headless: functions keep the reference's return arity with None in place of figures / axes,
and report() writes the same three CSV files (performance_stats.csv, risk_trace.csv, CVGA_stats.csv).

`device_report` works directly on a device BG trace [rows, E] (e.g. the BG column of BatchEngine.run's
trace) for populations that never visit the host.
"""
import ctypes as C
import os

import numpy as np
import pandas as pd
import torch

from .. import _capi

PERCENT_COLS = ['70<=BG<=180', 'BG>180', 'BG<70', 'BG>250', 'BG<50']
RI_COLS = ['LBGI', 'HBGI', 'Risk Index']
CHUNK_ROWS = 60                      # risk_index_trace chunks 60 ROWS (3 h at Dexcom's 3-minute rate), report.py:96


def device_report(bg, row_stride=None, n_rows=None, n_episodes=None, chunk_rows=CHUNK_ROWS, percent=True, risk=True,
                  cvga=True):
    """bg: device tensor holding element (row r, episode e) at offset r * row_stride + e from its first element
    (a contiguous [rows, E] tensor by default).  Returns a dict of device tensors: percent [5, E],
    ri [3, n_chunks, E], ri_mean [3, E], cvga [2, E], zones [5] (counts: A, A-or-B, C, D, E)."""
    if not (isinstance(bg, torch.Tensor) and bg.is_cuda):
        raise _capi.SgError('device_report needs a CUDA tensor (there is no CPU fallback)')
    lib = _capi.load()
    if n_rows is None:
        n_rows, n_episodes = bg.shape[0], bg.shape[-1]
    if row_stride is None:
        row_stride = bg.stride(0)
    dev = bg.device
    n_chunks = -(-int(n_rows) // int(chunk_rows))
    out = {}
    if percent:
        out['percent'] = torch.empty(5, n_episodes, dtype=torch.float64, device=dev)
    if risk:
        out['ri'] = torch.empty(3, n_chunks, n_episodes, dtype=torch.float64, device=dev)
        out['ri_mean'] = torch.empty(3, n_episodes, dtype=torch.float64, device=dev)
    if cvga:
        out['cvga'] = torch.empty(2, n_episodes, dtype=torch.float64, device=dev)
        out['zones'] = torch.empty(5, dtype=torch.float64, device=dev)

    def ptr(name):
        return C.c_void_p(out[name].data_ptr()) if name in out else None
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    with torch.cuda.device(dev):
        _capi.check(lib.sg_report_stats(C.c_void_p(bg.data_ptr()), int(row_stride), int(n_rows), int(n_episodes),
                                        int(chunk_rows), ptr('percent'), ptr('ri'), ptr('ri_mean'), ptr('cvga'), stream))
        if cvga:
            _capi.check(lib.sg_cvga_zones(ptr('cvga'), int(n_episodes), ptr('zones'), stream))
    return out


def _to_device(BG):
    """[time x patient] DataFrame / array -> contiguous device tensor [rows, E] and the column labels."""
    if isinstance(BG, pd.Series):
        BG = BG.to_frame()
    cols = list(BG.columns) if isinstance(BG, pd.DataFrame) else list(range(np.asarray(BG).shape[1]))
    arr = np.ascontiguousarray(np.asarray(BG, dtype=np.float64))
    if not torch.cuda.is_available():
        raise _capi.SgError('simglucose_b200 needs a CUDA device (there is no CPU fallback)')
    return torch.from_numpy(arr).cuda(), cols


def percent_stats(BG, ax=None):
    """Percent of rows in range per patient.  Returns (DataFrame [patient x 5], None, None)."""
    t, cols = _to_device(BG)
    p = device_report(t, risk=False, cvga=False)['percent'].cpu().numpy()
    return pd.DataFrame(p.T, index=cols, columns=PERCENT_COLS), None, ax


def risk_index_trace(df_BG, visualize=False):
    """LBGI / HBGI / RI of every 60-row chunk (from the chunk mean of f(BG)) and their means.
    Returns (ri_per_hour [(metric, patient) x chunk], ri_mean [patient x 3], None, [])."""
    t, cols = _to_device(df_BG)
    r = device_report(t, percent=False, cvga=False)
    ri = r['ri'].cpu().numpy()                                        # [3, n_chunks, E]
    idx = pd.MultiIndex.from_product([RI_COLS, cols])
    ri_per_hour = pd.DataFrame(ri.transpose(0, 2, 1).reshape(3 * len(cols), -1), index=idx)
    ri_mean = pd.DataFrame(r['ri_mean'].cpu().numpy().T, index=cols, columns=RI_COLS)
    return ri_per_hour, ri_mean, None, []


def CVGA_analysis(BG):
    """(BG_min, BG_max, perA, perB, perC, perD, perE): 2.5 / 97.5 percentiles per patient clipped to [50, 400]
    and the fractions of patients per control-variability grid zone."""
    t, cols = _to_device(BG)
    r = device_report(t, percent=False, risk=False)
    mn, mx = r['cvga'].cpu().numpy()
    cA, cAB, cC, cD, cE = r['zones'].cpu().numpy() / float(len(cols))
    return mn, mx, cA, cAB - cA, cC, cD, cE


def CVGA(BG_list, label=None):
    """Zone statistics for one or several populations.  Returns (DataFrame [population x A..E], None, None)."""
    if not isinstance(BG_list, list):
        BG_list = [BG_list]
    zone_stats = [CVGA_analysis(BG)[2:] for BG in BG_list]
    return pd.DataFrame(zone_stats, columns=['A', 'B', 'C', 'D', 'E']), None, None


def report(df, save_path=None):
    """df: the concatenated results of batch_sim, `pd.concat(results, keys=[names])` (user_interface.py:381).
    Returns (results, ri_per_hour, zone_stats, [], []) and, with save_path, writes performance_stats.csv,
    risk_trace.csv and CVGA_stats.csv there."""
    BG = df.unstack(level=0).BG
    pstats, _, _ = percent_stats(BG)
    ri_per_hour, ri_mean, _, _ = risk_index_trace(BG, visualize=False)
    zone_stats, _, _ = CVGA(BG, label='')
    results = pd.concat([pstats, ri_mean], axis=1)
    if save_path is not None:
        os.makedirs(save_path, exist_ok=True)
        results.to_csv(os.path.join(save_path, 'performance_stats.csv'))
        ri_per_hour.to_csv(os.path.join(save_path, 'risk_trace.csv'))
        zone_stats.to_csv(os.path.join(save_path, 'CVGA_stats.csv'))
    return results, ri_per_hour, zone_stats, [], []
