"""CPU: host-side logic of the package -- stream generators against fixtures recorded from the reference,
scenario classes, gym seeding chain, controllers' host policies, the C-ABI library's symbol table, and the
multi-process statistics reduction over gloo (world_size 2)."""
import ctypes
import os
import re
import subprocess
import sys
from datetime import datetime, timedelta

import numpy as np
import pytest

import helpers as H

ROOT = H.ROOT


def test_cabi_library_exports_every_declared_symbol():
    from simglucose_b200 import _capi, build
    assert os.path.exists(build.LIB_PATH), 'run __graft_entry__.build() first'
    header = open(os.path.join(ROOT, 'include', 'simglucose_b200.h')).read()
    declared = set(re.findall(r'^\s*(?:int|int64_t|const char \*)\s*\*?\s*(sg_\w+)\s*\(', header, flags=re.M))
    assert len(declared) >= 30
    assert declared == set(_capi.SYMBOLS), declared ^ set(_capi.SYMBOLS)
    lib = ctypes.CDLL(build.LIB_PATH)                     # loads without a GPU; no compute call is made
    for name in declared:
        assert hasattr(lib, name), name
    lib.sg_version.restype = ctypes.c_int
    assert lib.sg_version() == _capi.SG_ABI_VERSION
    h = ctypes.c_void_p()
    lib.sg_create.argtypes = [ctypes.c_int, ctypes.c_int64, ctypes.c_uint32, ctypes.POINTER(ctypes.c_void_p)]
    import torch
    if not torch.cuda.is_available():
        assert lib.sg_create(0, 16, 0, ctypes.byref(h)) < 0          # fails loudly: there is no CPU fallback
        lib.sg_last_error.restype = ctypes.c_char_p
        assert b'CUDA' in lib.sg_last_error()


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under simglucose_b200/ may reference it."""
    pkg = os.path.join(ROOT, 'simglucose_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                src = open(os.path.join(dirpath, f)).read()
                assert 'sg_oracle' not in src and 'import oracle' not in src and "'oracle'" not in src, f


def test_engine_refuses_to_run_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    from simglucose_b200 import _capi
    from simglucose_b200.engine import BatchEngine
    from simglucose_b200.patient.t1dpatient import T1DPatient
    with pytest.raises(_capi.SgError):
        BatchEngine(4)
    with pytest.raises(_capi.SgError):
        T1DPatient.withID(1)


def test_cgm_noise_streams_match_reference():
    from simglucose_b200 import streams as S
    from simglucose_b200.sensor.cgm import CGMSensor
    d = H.load('generators')
    for sensor in ('Dexcom', 'GuardianRT', 'Navigator'):
        for seed in range(6):
            got = S.cgm_noise_stream(sensor, seed, 400)
            assert np.max(np.abs(got - d['noise_' + sensor][seed])) <= 1e-12
        W = S.interp_matrix(CGMSensor.withName(sensor).sample_time)      # the constant-matrix form (device generator)
        assert W.shape == (S.noise_window_samples(CGMSensor.withName(sensor).sample_time), 11)
        assert np.allclose(W.sum(axis=1), 1.0, atol=1e-12)
    s = CGMSensor.withName('Dexcom', seed=2)
    it = [next(s._noise_generator) for _ in range(130)]
    assert np.max(np.abs(np.array(it) - d['noise_Dexcom'][2][:130])) <= 1e-12
    s.reset()
    assert next(s._noise_generator) == it[0]


def test_random_scenario_matches_reference():
    from simglucose_b200 import streams as S
    from simglucose_b200.simulation.scenario_gen import RandomScenario
    d = H.load('generators')
    for k, (seed, hour) in enumerate(d['sched_seed_hour']):
        mm, mg = S.random_scenario_table(int(seed), int(hour) * 60, 3 * 1440)
        n = int((d['sched_minute'][k] >= 0).sum())
        assert np.array_equal(mm[:24], d['sched_minute'][k][:n]) and np.array_equal(mg[:24], d['sched_grams'][k][:n])
    start = datetime(2018, 1, 1, 6)
    sc = RandomScenario(start_time=start, seed=3)
    got = [(m, sc.get_action(start + timedelta(minutes=m)).meal) for m in range(3 * 1440)]
    got = [(m, g) for m, g in got if g != 0]
    k = [tuple(x) for x in d['sched_seed_hour'].tolist()].index((3, 6))
    n = int((d['sched_minute'][k] >= 0).sum())
    assert [m for m, _ in got][:n] == d['sched_minute'][k][:n].tolist()
    sc.reset()
    first = dict(sc.scenario['meal'])
    sc.seed = 3
    assert sc.scenario['meal'] == first


def test_custom_scenario_time_semantics():
    """A number is hours AFTER start_time (scenario.py:49-50), a timedelta an offset, a datetime absolute."""
    from simglucose_b200 import streams as S
    from simglucose_b200.simulation.scenario import CustomScenario, parseTime
    start = datetime(2018, 1, 1, 6)
    sc = CustomScenario(start_time=start, scenario=[(7, 45), (timedelta(hours=1, minutes=30), 20),
                                                    (datetime(2018, 1, 1, 8, 0), 30), (7, 99)])
    assert sc.get_action(datetime(2018, 1, 1, 13, 0)).meal == 45         # first match wins
    assert sc.get_action(datetime(2018, 1, 1, 7, 30)).meal == 20
    assert sc.get_action(datetime(2018, 1, 1, 8, 0)).meal == 30
    assert sc.get_action(datetime(2018, 1, 1, 7, 0)).meal == 0
    assert parseTime(0.5, start) == datetime(2018, 1, 1, 6, 30)
    with pytest.raises(ValueError):
        parseTime('noon', start)
    assert CustomScenario(start, []).get_action(start).meal == 0
    mm, mg = S.custom_scenario_table(sc.scenario, start, 1440)
    assert mm.tolist() == [90, 120, 420] and mg.tolist() == [20.0, 30.0, 45.0]


def test_gym_seeding_chain_known_answers():
    """gym==0.9.4 seeding restated (third party); pinned by the reference's tests/test_seed.py and the fixture."""
    from simglucose_b200 import streams as S
    d = H.load('gym_seeded')
    for i in range(d['seed'].shape[0]):
        hours = []
        for _ in range(1):
            rng, s1 = S.np_random(int(d['seed'][i]))
            chain = []
            for _call in range(2):                        # _seed() builds an env, _reset() builds the next
                s2 = S.hash_seed(rng.randint(0, 1000)) % 2 ** 31
                s3 = S.hash_seed(s2 + 1) % 2 ** 31
                s4 = S.hash_seed(s3 + 1) % 2 ** 31
                hours.append(rng.randint(low=0.0, high=24.0))
                chain.append([s1, s2, s3, s4])
        assert chain[0] == [int(v) for v in d['seeds'][i]]
        assert hours[1] == int(d['start_hour'][i])
        assert chain[1][1] == int(d['sensor_seed'][i]) and chain[1][2] == int(d['scenario_seed'][i])
        assert chain[1][3] == int(d['patient_seed'][i])
    rng, _ = S.np_random(0)
    assert S.hash_seed(0) == S.hash_seed(0) and S.create_seed(2 ** 64 + 5) == 5
    with pytest.raises(ValueError):
        S.np_random(-1)


def test_random_init_state_matches_fixture():
    from simglucose_b200 import streams as S
    d = H.load('open_loop_actions')            # random_init_bg=True, patient_seed = 200 + i (oracle/make_golden.py)
    for i in range(d['patient_row'].shape[0]):
        _, x0 = H.O.patient_params(int(d['patient_row'][i]))
        # the reference harness resets twice (constructor + env.reset): second draw around the first
        x1 = S.random_init_state(x0, 200 + i)
        x2 = S.random_init_state(x1, 200 + i)
        assert np.max(np.abs(x2 - d['x0'][i])) <= 1e-9 * np.max(np.abs(x0))


def test_controller_host_policies():
    from collections import namedtuple
    from simglucose_b200.controller.basal_bolus_ctrller import BBController
    from simglucose_b200.controller.pid_ctrller import PIDController
    Obs = namedtuple('Observation', ['CGM'])
    bb = BBController(target=140)
    p, _ = H.O.patient_params(0)
    cr, cf = H.O.quest(0)
    a = bb.policy(Obs(CGM=180.0), 0, False, patient_name='adolescent#001', meal=15.0, sample_time=3)
    assert a.basal == p[1] * p[0] / 6000
    assert a.bolus == ((15.0 * 3) / cr + (180.0 - 140) / cf) / 3
    assert bb.policy(Obs(CGM=120.0), 0, False, patient_name='adolescent#001', meal=0, sample_time=3).bolus == 0
    assert bb.policy(Obs(CGM=120.0), 0, False, patient_name='adolescent#001', meal=3.0, sample_time=3).bolus == 9.0 / cr / 3
    unk = bb.policy(Obs(CGM=100.0), 0, False, patient_name='someone', meal=0, sample_time=3)
    assert unk.basal == 1.43 * 57.0 / 6000 and bb.device_spec('someone') is None
    assert bb.device_spec('adolescent#001') == ('bb', (140.0, cr, cf, p[1] * p[0] / 6000))
    pid = PIDController(P=0.001, I=0.00001, D=0.001, target=140)
    a1 = pid.policy(Obs(CGM=150.0), 0, False, sample_time=3)
    assert a1.basal == 0.001 * 10.0 + 0.00001 * 0 + 0.001 * 150.0 / 3 and a1.bolus == 0      # derivative kick
    a2 = pid.policy(Obs(CGM=153.0), 0, False, sample_time=3)
    assert a2.basal == 0.001 * 13.0 + 0.00001 * 30.0 + 0.001 * 3.0 / 3
    assert pid.device_spec() is None                       # the kernel starts from a reset controller
    pid.reset()
    assert pid.device_spec() == ('pid', (0.001, 0.00001, 0.001, 140.0))


def test_shard_ranges():
    from simglucose_b200.distributed import shard_range
    for n in (1, 7, 30, 16777216):
        for w in (1, 2, 4, 8):
            r = [shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(w - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1


_GLOO_WORKER = r'''
import os, sys
sys.path.insert(0, %(root)r)
import torch, torch.distributed as dist
from simglucose_b200 import _capi
from simglucose_b200.distributed import allreduce_stats, shard_range, summarize
dist.init_process_group('gloo')
rank, world = dist.get_rank(), dist.get_world_size()
n_total = 101
lo, hi = shard_range(n_total, rank, world)
# per-episode statistics of a fake population: episode e has rows = 10, sum_bg = 100 + e, min = e, max = 2e
e = torch.arange(lo, hi, dtype=torch.float64)
agg = torch.zeros(_capi.NSTAT, dtype=torch.float64)
agg[0] = 10.0 * e.numel(); agg[1] = (100 + e).sum(); agg[2] = e.min(); agg[3] = (2 * e).max()
agg[10] = 7.0 * e.numel(); agg[13] = float((e %% 2 == 0).sum())
allreduce_stats(agg)
s = summarize(agg)
assert agg[0] == 10.0 * n_total and agg[2] == 0 and agg[3] == 200 and agg[13] == 51, agg
assert abs(agg[1] - (100 * n_total + n_total * (n_total - 1) / 2)) < 1e-9
assert abs(s['pct_in_range'] - 70.0) < 1e-12 and s['episodes_done'] == 51
dist.destroy_process_group()
sys.stdout.write('rank %%d ok\n' %% rank)        # one write() per rank so the lines cannot interleave
'''


def test_stats_allreduce_over_gloo_world_size_2(tmp_path):
    script = tmp_path / 'worker.py'
    script.write_text(_GLOO_WORKER % {'root': ROOT})
    import socket
    with socket.socket() as sk:                           # a free rendezvous port on the loopback interface
        sk.bind(('127.0.0.1', 0))
        port = sk.getsockname()[1]
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node=2', '--master-addr', '127.0.0.1',
           '--master-port', str(port), str(script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert 'rank 0 ok' in r.stdout and 'rank 1 ok' in r.stdout


def test_scenario_meal_tables_equal_the_per_minute_replay():
    """batch_sim's fused path asks this package's scenario classes for their meal table (one loop turn per calendar day)
    instead of calling get_action once per simulated minute as env.mini_step does (env.py:50): same meals, same
    generator state and current-day schedule afterwards (scenario_gen.py:15-31, scenario.py:33-42)."""
    from datetime import datetime, timedelta
    from simglucose_b200.simulation.scenario import CustomScenario
    from simglucose_b200.simulation.scenario_gen import RandomScenario

    def replay(sc, t0, n):
        mm, gg = [], []
        for m in range(n):
            g = sc.get_action(sc.start_time + timedelta(minutes=t0 + m)).meal
            if g != 0:
                mm.append(m + t0)
                gg.append(float(g))
        return np.array(mm, dtype=np.int32), np.array(gg)

    for seed in range(3):
        for start in (datetime(2018, 1, 1, 0, 0, 0), datetime(2018, 1, 1, 6, 30, 0), datetime(2018, 1, 1, 23, 59, 0)):
            for t0, n in ((0, 4320), (5, 1440), (1000, 3000)):
                a, b = RandomScenario(start_time=start, seed=seed), RandomScenario(start_time=start, seed=seed)
                want, got = replay(a, t0, n), b.meal_table(t0, n)
                assert np.array_equal(want[0], got[0]) and np.array_equal(want[1], got[1])
                assert a.scenario == b.scenario and a.random_gen.randint(1 << 30) == b.random_gen.randint(1 << 30)
    assert RandomScenario(start_time=datetime(2018, 1, 1, 0, 0, 30), seed=1).meal_table(0, 10) is None
    start = datetime(2018, 1, 1, 0, 0, 0)
    entries = [(7, 45), (12, 70), (16, 15), (18, 80), (23, 10), (7, 99), (timedelta(hours=30), 33),
               (datetime(2018, 1, 1, 9, 30), 5), (1.004, 8), (2, 0)]
    c = CustomScenario(start_time=start, scenario=entries)
    for t0, n in ((0, 1440), (0, 2880), (600, 500)):
        want, got = replay(c, t0, n), c.meal_table(t0, n)
        assert np.array_equal(want[0], got[0]) and np.array_equal(want[1], got[1])
