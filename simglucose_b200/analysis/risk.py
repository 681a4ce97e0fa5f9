"""risk_index -- same call as simglucose/analysis/risk.py:5-17, evaluated by the CUDA library."""
import numpy as np

from .. import ops


def risk_index(BG, horizon):
    """(LBGI, HBGI, RI) of the last `horizon` samples of BG (mg/dL)."""
    window = np.asarray(BG, dtype=np.float64).reshape(-1)[-int(horizon):]
    if window.size == 0:
        return (0.0, 0.0, 0.0)
    lbgi, hbgi, ri = ops.risk_index_window(window).cpu().tolist()
    return (lbgi, hbgi, ri)
