"""Elementwise pieces of the closed loop as device operations (no host arithmetic): the pump
quantiser, the CGM sampling branch and the risk index, each a CUDA kernel behind the C-ABI.

Reference: simglucose/actuator/pump.py:23-39, simglucose/sensor/cgm.py:28-33,
simglucose/analysis/risk.py:5-17.
"""
import ctypes as C

import numpy as np
import torch

from . import _capi


def _dev(a, device):
    t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.atleast_1d(np.asarray(a, dtype=np.float64)))
    return t.to(device=device, dtype=torch.float64).contiguous().reshape(-1)


def _device(device=None):
    if not torch.cuda.is_available():
        raise _capi.SgError('simglucose_b200 needs a CUDA device (there is no CPU fallback)')
    return torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def pump_quantise(amount, vmin, vmax, inc, device=None):
    """InsulinPump.basal / .bolus on a tensor of amounts (U/min) -> device tensor."""
    dev = _device(device)
    a = _dev(amount, dev)
    out = torch.empty_like(a)
    with torch.cuda.device(dev):
        _capi.check(_capi.load().sg_pump(C.c_void_p(a.data_ptr()), C.c_void_p(out.data_ptr()), a.numel(),
                                         float(vmin), float(vmax), float(inc), _stream()))
    return out


def cgm_measure(bg, noise, sensor_min, sensor_max, device=None):
    """min(max(BG + noise, sensor_min), sensor_max), the sampling branch of CGMSensor.measure."""
    dev = _device(device)
    b, z = _dev(bg, dev), _dev(noise, dev)
    assert b.numel() == z.numel()
    out = torch.empty_like(b)
    with torch.cuda.device(dev):
        _capi.check(_capi.load().sg_cgm_measure(C.c_void_p(b.data_ptr()), C.c_void_p(z.data_ptr()),
                                                C.c_void_p(out.data_ptr()), b.numel(), float(sensor_min),
                                                float(sensor_max), _stream()))
    return out


def risk_index_elementwise(bg, device=None):
    """risk_index([BG], 1) for every element: [3, n] device tensor (LBGI, HBGI, RI)."""
    dev = _device(device)
    b = _dev(bg, dev)
    out = torch.empty(3, b.numel(), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _capi.check(_capi.load().sg_risk_index(C.c_void_p(b.data_ptr()), C.c_void_p(out.data_ptr()), b.numel(), _stream()))
    return out


def risk_index_window(bg, device=None):
    """risk_index(BG, len(BG)): [3] device tensor (LBGI, HBGI, RI) over the whole window."""
    dev = _device(device)
    b = _dev(bg, dev)
    out = torch.empty(3, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _capi.check(_capi.load().sg_risk_window(C.c_void_p(b.data_ptr()), C.c_void_p(out.data_ptr()), b.numel(), _stream()))
    return out
