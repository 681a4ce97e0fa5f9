"""GPU: sg_reduce_stats_static, the population aggregate for callers that are not allowed to allocate device memory
(a plugin that only receives device pointers from its host application and may not call cudaMalloc, so it cannot
own an sg_handle).  Its partial sums live in an array built into the library instead of a cudaMalloc'ed buffer.

What is checked, always through the C-ABI: its output equals sg_reduce_stats' output on the same per-episode
statistics BIT FOR BIT (the tolerance the other aggregate tests use for counters and extrema; here the summation
order is the same too, so the sums are held to it as well), on calls made back to back that share the built-in
array -- a large population, then a small one that leaves most of the array holding the large one's rows, then the
large one again -- and that a population beyond the stated limit is refused with a readable message.

This file sorts last on purpose: `pytest -x` stops at the first failure, and a failure here must not hide the
parity tests.  It makes no assumption about what ran before it (the first call on the device does its own set-up).
"""
import ctypes as C

import pytest

pytestmark = pytest.mark.gpu


def _simulated_population(E, n_steps, seed):
    """An engine whose E episodes (30-patient cohort x seeds, basal-bolus controller) have run n_steps env-steps."""
    import torch
    from simglucose_b200 import streams as S
    from simglucose_b200.engine import BatchEngine
    dev = torch.device('cuda', 0)
    rows = (torch.arange(E, device=dev, dtype=torch.int64) % 30).to(torch.int32)
    noise = S.device_noise(E, n_steps + 2, 'Dexcom', seed=seed, episode_offset=0, group_size=30, device=dev)
    mm, mg = S.device_scenarios(E, n_steps * 3, start_minute=0, seed=seed, episode_offset=0, group_size=30, device=dev)
    eng = BatchEngine(E)
    eng.reset(rows, noise, mm, mg, row0=False)
    eng.run(n_steps, controller='bb', ctrl=(140.0, 10.0, 30.0), trace=False)
    return eng


def _static_reduce(lib, stats, n, out):
    """One call of the entry point under test.  `stats` and `out` were allocated by the caller beforehand."""
    import torch
    from simglucose_b200 import _capi
    assert stats.is_contiguous() and stats.dtype == torch.float64 and stats.shape[0] == _capi.NSTAT
    _capi.check(lib.sg_reduce_stats_static(C.c_void_p(stats.data_ptr()), n, stats.shape[1], C.c_void_p(out.data_ptr()),
                                           C.c_void_p(torch.cuda.current_stream().cuda_stream)))


def test_static_reduce_matches_the_allocating_entry_point_on_consecutive_calls():
    import torch
    from simglucose_b200 import _capi
    lib = _capi.load()
    torch.cuda.set_device(0)
    big_n, small_n, n_steps = 30 * 700, 30 * 40, 48       # 83 and 5 partial rows of the built-in array
    big = _simulated_population(big_n, n_steps, seed=11)
    small = _simulated_population(small_n, n_steps, seed=12)
    try:
        want_big, want_small = big.reduce_stats().clone(), small.reduce_stats().clone()     # existing entry point
        stats_big, stats_small = big.stats().contiguous(), small.stats().contiguous()
        out = [torch.full((_capi.NSTAT,), float('nan'), dtype=torch.float64, device='cuda:0') for _ in range(4)]
        _static_reduce(lib, stats_big, big_n, out[0])       # back to back on one stream: no sync, no clearing
        _static_reduce(lib, stats_small, small_n, out[1])   # reads 5 rows; rows 5..82 still hold the big population
        _static_reduce(lib, stats_big, big_n, out[2])
        # a prefix of a wider array (capacity > n_episodes): the first small_n episodes of the big population are
        # the same episodes a small_n-episode engine of that seed simulates (streams are keyed by global index)
        prefix = _simulated_population(small_n, n_steps, seed=11)
        want_prefix = prefix.reduce_stats().clone()
        prefix.close()
        _static_reduce(lib, stats_big, small_n, out[3])
        torch.cuda.synchronize()
        rows = n_steps + 1                                  # history rows per episode: the reset row and one per env-step
        assert want_big[0].item() == big_n * rows and want_small[0].item() == small_n * rows
        assert torch.equal(out[0], want_big), (out[0] - want_big).abs().max().item()
        assert torch.equal(out[1], want_small), (out[1] - want_small).abs().max().item()
        assert torch.equal(out[2], want_big)
        assert torch.equal(out[3], want_prefix)
        assert not torch.equal(want_big, want_small)        # the two populations do differ
    finally:
        big.close()
        small.close()


def test_static_reduce_refuses_what_does_not_fit():
    import torch
    from simglucose_b200 import _capi
    lib = _capi.load()
    torch.cuda.set_device(0)
    stats = torch.zeros(_capi.NSTAT, 64, dtype=torch.float64, device='cuda:0')
    out = torch.zeros(_capi.NSTAT, dtype=torch.float64, device='cuda:0')
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    too_many = _capi.REDUCE_STATIC_MAX_EPISODES + 1
    # refused on the size alone, before anything is launched: the 64-column buffer is never touched
    rc = lib.sg_reduce_stats_static(C.c_void_p(stats.data_ptr()), too_many, too_many, C.c_void_p(out.data_ptr()), stream)
    assert rc == -1
    msg = lib.sg_last_error().decode()
    assert 'exceed the limit' in msg and str(_capi.REDUCE_STATIC_MAX_EPISODES) in msg
    assert lib.sg_reduce_stats_static(C.c_void_p(stats.data_ptr()), 65, 64, C.c_void_p(out.data_ptr()), stream) == -1
    assert lib.sg_reduce_stats_static(None, 64, 64, C.c_void_p(out.data_ptr()), stream) == -1
    _static_reduce(lib, stats, 64, out)                     # and a call that fits still works afterwards
    torch.cuda.synchronize()
    assert out[2].item() == 0.0 and out[13].item() == 64.0  # min over zeros; every zero counts as "reported done"
