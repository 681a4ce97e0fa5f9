"""The kernels' own FP64 helpers (csrc/sg_math.cuh) against CUDA's libm, on the GPU, through the C-ABI test
hook sg_math_probe.  div_nr must reproduce IEEE division bit for bit in the range the kernels use it."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _probe(kind, a, b):
    import torch
    from simglucose_b200 import _capi
    lib = _capi.load()
    ta = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).cuda()
    tb = torch.from_numpy(np.ascontiguousarray(b, dtype=np.float64)).cuda()
    out = torch.empty_like(ta)
    _capi.check(lib.sg_math_probe(kind, C.c_void_p(ta.data_ptr()), C.c_void_p(tb.data_ptr()),
                                  C.c_void_p(out.data_ptr()), ta.numel(),
                                  C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_div_nr_is_ieee_division_bit_for_bit():
    rng = np.random.RandomState(0)
    n = 1 << 22
    # numerators over 500 decades (both signs, exact zeros), denominators like the kernels': tolerances
    # (1e-12 .. 1e-3), body weights / volumes (1 .. 1e3), Km0 + x (1e2 .. 1e4)
    a = np.exp(rng.uniform(-600.0, 300.0, n)) * rng.choice([-1.0, 1.0], n)
    a[::97] = 0.0
    b = np.exp(rng.uniform(np.log(1e-13), np.log(1e5), n))
    got, ref = _probe(0, a, b), _probe(1, a, b)
    big = np.abs(a) >= 2.0 ** -960
    assert np.array_equal(got[big].view(np.int64), ref[big].view(np.int64))
    assert np.array_equal(got[a == 0.0], ref[a == 0.0])
    small = ~big & (a != 0.0)
    assert np.all(np.abs(got[small] - ref[small]) <= np.abs(ref[small]) * 2.3e-16 + 5e-324)
    # mantissa corner cases
    m = np.array([1.0, 1.0 + 2.0 ** -52, 2.0 - 2.0 ** -52, 1.5, 4.0 / 3.0, 1.0 / 3.0, 0.1, 0.05, 6000.0, 1e-12, 1e-6])
    aa, bb = np.meshgrid(m, m)
    assert np.array_equal(_probe(0, aa.ravel(), bb.ravel()).view(np.int64), _probe(1, aa.ravel(), bb.ravel()).view(np.int64))


def test_exp_log_within_one_ulp_of_libm():
    rng = np.random.RandomState(1)
    n = 1 << 20
    t = rng.uniform(-160.0, 160.0, n)
    got, ref = _probe(6, t, t), _probe(7, t, t)
    assert np.max(np.abs(got - ref) / ref) <= 3.4e-16
    x = np.exp(rng.uniform(-690.0, 690.0, n))
    got, ref = _probe(8, x, x), _probe(9, x, x)
    ok = np.abs(ref) > 1e-3
    assert np.max(np.abs(got[ok] - ref[ok]) / np.abs(ref[ok])) <= 3.4e-16
    assert np.max(np.abs(got - ref)) <= 1e-13


def test_tanh_diff_absolute_error():
    rng = np.random.RandomState(2)
    n = 1 << 20
    u = rng.uniform(-60.0, 5.0, n)      # aa * (qsto - b*Dbar): -2.5 b/(1-b) .. 2.5
    v = rng.uniform(-5.0, 30.0, n)      # cc * (qsto - d*Dbar): -2.5 .. 2.5 (1-d)/d
    got, ref = _probe(2, u, v), _probe(3, u, v)
    assert np.max(np.abs(got - ref)) <= 1.2e-15
    z = np.zeros(4)
    assert np.array_equal(_probe(2, z, z), z)
    big = np.array([1e300, -1e300, 700.0, -700.0])
    assert np.array_equal(_probe(2, big, -big), np.array([2.0, -2.0, 2.0, -2.0]))


def test_tanh_sum_of_the_gastric_emptying_law():
    """sg_math.cuh tanh_sum_half (what rhs() evaluates for t1dpatient.py:138-142): tanh(u) - tanh(v) + 2 as one quotient
    of two positive polynomials in e^2u, e^2v, the exponentials through the nine-operation exp_tanh.  Relative error
    <= 2 ulp of the O(1) value against CUDA's libm, also in saturation."""
    rng = np.random.RandomState(12)
    n = 1 << 21
    u = np.concatenate([rng.uniform(-60.0, 5.0, n), rng.uniform(-3.0, 3.0, n), rng.uniform(-45.0, 45.0, n)])
    v = np.concatenate([rng.uniform(-5.0, 30.0, n), rng.uniform(-3.0, 3.0, n), rng.uniform(-45.0, 45.0, n)])
    got, ref = _probe(10, u, v), _probe(11, u, v)
    assert np.max(np.abs(got - ref) / np.maximum(ref, 1.0)) <= 6.7e-16
    assert np.max(np.abs(got - ref)) <= 2.0e-15
    z = np.zeros(4)
    assert np.array_equal(_probe(10, z, z), z + 2.0)
    big = np.array([1e300, -1e300, 700.0, -700.0, np.inf, -np.inf])
    sat = _probe(10, big, -big)                                     # saturation: 4 and 0 to the last place or two
    assert np.all(np.abs(sat - np.array([4.0, 0.0, 4.0, 0.0, 4.0, 0.0])) <= 1.8e-15) and np.all(sat[1::2] >= 0.0)
    assert np.all(np.abs(_probe(10, big, big) - 2.0) <= 9e-16)
    # the exponential on its own: the one-part reduction costs |t| * 46 * 1.7e-18 relative, by design
    t = rng.uniform(-80.0, 80.0, n)
    got, ref = _probe(14, t, t), _probe(7, t, t)
    assert np.max(np.abs(got - ref) / ref) <= 1.2e-14
    small = np.abs(t) <= 5.0
    assert np.max(np.abs(got[small] - ref[small]) / ref[small]) <= 1.0e-15


def test_rcp_ulp_and_integer_sign_tests():
    """rcp_ulp (error norm, HINIT norms): within 1.5 ulp of 1 / x for the scales the integrator divides by.  is_neg /
    is_pos (the RHS's masks): exactly x < 0 and x > 0 for every non-NaN double, -0.0 and denormals included."""
    rng = np.random.RandomState(13)
    x = np.exp(rng.uniform(np.log(1e-13), np.log(1e6), 1 << 20))
    got = _probe(12, x, x)
    assert np.max(np.abs(got * x - 1.0)) <= 3.4e-16
    tiny = np.array([5e-324, 2.2250738585072014e-308, 1e-300])
    e = np.concatenate([[0.0, -0.0, np.inf, -np.inf, 1.0, -1.0], tiny, -tiny, rng.standard_normal(1000)])
    got = _probe(13, e, e)
    assert np.array_equal(got, (e < 0) * 1.0 + (e > 0) * 2.0)


def test_pow_pos_step_controller_exponents():
    rng = np.random.RandomState(3)
    n = 1 << 20
    for y in (0.2 - 0.04 * 0.75, 0.04, 0.2):
        x = np.exp(rng.uniform(-14.0, 14.0, n))      # err, facold, 0.01/der12 seen by the controller
        got, ref = _probe(4, x, np.full(n, y)), _probe(5, x, np.full(n, y))
        assert np.max(np.abs(got - ref) / ref) <= 6e-16
    # clamps: 0 and inf behave like any value beyond the controller's limits (facc1 = 5, facc2 = 0.1)
    edge = _probe(4, np.array([0.0, np.inf]), np.array([0.17, 0.17]))
    assert edge[0] < 1e-3 and edge[1] > 1e3


def test_risk_index_against_oracle_over_the_whole_range_and_edge_cases():
    """risk.py:5-17 through sg_risk_index (the function the episode kernels call, sg_model.cuh risk_f: in-tree log /
    exp instead of CUDA's pow) against the oracle's libm evaluation: 1e-12 relative over BG 1..1e6 mg/dL -- the slope
    of 10 f^2 around f = 0 (BG 112.5) amplifies one ulp of pow to ~1e-14 absolute -- and exactly the reference's
    nan_to_num results for BG < 1, BG == 1, zero (log -> -inf, pow -> +inf: HBGI = DBL_MAX), negative, NaN and +inf."""
    import torch
    from helpers import O
    from simglucose_b200 import ops
    rng = np.random.default_rng(7)
    bg = np.concatenate([np.exp(rng.uniform(0.0, np.log(1e6), 20000)), np.linspace(100.0, 125.0, 5001),
                         [1.0, np.nextafter(1.0, 2.0), 1.0000001, 39.0, 600.0, 1e300, 1.7e308]])
    out = [t.cpu().numpy() for t in ops.risk_index_elementwise(torch.from_numpy(bg).cuda())]
    ref = np.array([O.risk(v) for v in bg]).T
    for got, want in zip(out, ref):
        assert np.all(np.abs(got - want) <= 1e-12 * np.maximum(np.abs(want), 1.0) + 1e-13)
    edge = np.array([0.5, 0.999999, 0.0, -0.0, -3.5e-12, -100.0, np.nan, np.inf, 1.0])
    out = [t.cpu().numpy() for t in ops.risk_index_elementwise(torch.from_numpy(edge).cuda())]
    ref = np.array([O.risk(v) for v in edge]).T
    for got, want in zip(out, ref):
        assert np.array_equal(got[:8], want[:8])                  # zeros, and DBL_MAX for BG == +-0 and +inf
        assert abs(got[8] - want[8]) <= 1e-12 * max(abs(want[8]), 1.0)
    big = np.finfo(float).max
    assert out[0][8] > 650.0 and out[1][2] == big and out[1][3] == big and out[1][7] == big    # BG == 1; +-0; +inf
