"""TEST INFRASTRUCTURE ONLY -- host restatement of the device parameter jitter of BASELINE.json configs[3].

Restates, value for value, what `perturb_kernel` (simglucose_b200/csrc/sg_generators.cu, behind
sg_gen_perturbed_params) computes with the counter-based Philox4x32-10 of csrc/sg_philox.cuh, so that the
fixtures recorded from the reference (oracle/make_golden.py, group `perturbed`) use exactly the parameter
rows the device generator hands the simulator for the same global episode indices.  The reference itself
has no such generator: T1DPatient.__init__ takes any parameter Series (patient/t1dpatient.py:22-42).

The generator kernels are compiled with FMA contraction; the SASS of perturb_kernel shows
    u = (bits + 0.5) * 2^-53;  v = fma(u, 2, -1);  w = fma(v, amp, 1);  out = t * w
`fma(v, amp, 1)` is reproduced exactly with rational arithmetic (one rounding).
"""
from fractions import Fraction

import numpy as np

M0, M1 = 0xD2511F53, 0xCD9E8D57
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = 0xFFFFFFFF
TAG_PARA = 0x70617261        # 'para'


def philox4x32_10(ctr, key):
    """One Philox4x32-10 block: ctr and key are tuples of 32-bit words; returns the four output words."""
    c = list(ctr)
    k = list(key)
    for _ in range(10):
        p0, p1 = M0 * c[0], M1 * c[2]
        c = [(p1 >> 32) ^ c[1] ^ k[0], p1 & MASK, (p0 >> 32) ^ c[3] ^ k[1], p0 & MASK]
        k = [(k[0] + W0) & MASK, (k[1] + W1) & MASK]
    return c


def uniforms(seed, stream, substream, n):
    """The first n values a sequential caller of Philox::uniform() sees: block j // 2, words (3, 2) then (1, 0)."""
    key = (seed & MASK, (seed >> 32) & MASK)
    out = []
    for blk in range((n + 1) // 2):
        w = philox4x32_10((blk, substream, stream & MASK, (stream >> 32) & MASK), key)
        for a, b in ((w[3], w[2]), (w[1], w[0])):
            bits = ((a << 32) | b) >> 11
            out.append((np.float64(bits) + np.float64(0.5)) * np.float64(1.0 / 9007199254740992.0))
    return np.array(out[:n], dtype=np.float64)


def _fma(a, b, c):
    return float(Fraction(float(a)) * Fraction(float(b)) + Fraction(float(c)))


def perturbed_row(table_row, amplitude, seed, global_index):
    """table_row [32] (hot parameter order) -> the row episode `global_index` of the perturbed population gets."""
    u = uniforms(int(seed), int(global_index), TAG_PARA, len(table_row))
    out = np.empty(len(table_row), dtype=np.float64)
    for i, t in enumerate(table_row):
        v = _fma(u[i], 2.0, -1.0)
        w = _fma(v, amplitude, 1.0)
        out[i] = np.float64(t) * np.float64(w)
    return out
