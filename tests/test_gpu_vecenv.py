"""GPU: the vector environment as "the gym env, batched" (reference
envs/simglucose_gym_env.py:27-73, patient/t1dpatient.py:256-270).

  * gym_seeding + reference streams: N environments of a T1DSimVecEnv reproduce N single `envs.T1DSimEnv` gym
    wrappers (which are themselves pinned to the reference by tests/test_gpu_facade.py) bit for bit -- seeds, start
    hours, randomised initial states, every step, THROUGH per-environment resets on `done`;
  * gym_seeding + device streams: the same semantics drawn from Philox on the GPU, checked through properties
    (determinism, shard independence, per-environment restart, distributions of start hour and initial state).
"""
import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def test_gym_seeded_vector_env_equals_single_gym_envs_through_resets():
    import torch
    from simglucose_b200.envs import T1DSimEnv, T1DSimVecEnv
    names = ['child#004', 'adolescent#002', 'adult#007', 'child#008', 'adult#001']
    seed, E, horizon, n_steps = 50, len(names), 48, 110
    venv = T1DSimVecEnv(E, patient_names=names, seed=seed, horizon_steps=horizon, stream_mode='reference',
                        gym_seeding=True, auto_reset=True)
    singles = [T1DSimEnv(patient_name=n, seed=seed + e) for e, n in enumerate(names)]
    obs = venv.reset().cpu().numpy()
    first = [s.reset() for s in singles]
    assert [f.CGM for f in first] == obs.tolist()
    for e, s in enumerate(singles):
        s2, s3, s4, hour = venv.episode_seeds[e]
        assert (s.env.sensor.seed, s.env.scenario.seed, s.env.patient.seed) == (s2, s3, s4)
        assert s.env.scenario.start_time.hour == hour
    assert np.array_equal(venv.state.cpu().numpy().T, np.stack([s.env.patient.state for s in singles]))
    rng = np.random.RandomState(4)
    # basal rates that push the two children out of range quickly (done), the others reach the horizon (truncation)
    acts = rng.uniform(0.0, 0.05, size=(n_steps, E)) * np.array([4.0, 1.0, 1.0, 4.0, 1.0])
    steps_in_episode = [0] * E
    n_done = n_trunc = 0
    for k in range(n_steps):
        o, r, dn, info = venv.step(torch.from_numpy(acts[k]))
        o, r, dn = o.cpu().numpy(), r.cpu().numpy(), dn.cpu().numpy()
        bg, term = info['bg'].cpu().numpy(), info['terminal_obs'].cpu().numpy()
        for e, s in enumerate(singles):
            so, sr, sd, si = s.step(acts[k, e])
            steps_in_episode[e] += 1
            assert (sr, bool(sd), si['bg']) == (r[e], bool(dn[e]), bg[e]) and so.CGM == term[e]
            if sd or steps_in_episode[e] >= horizon:                  # what an RL loop does: reset this environment
                n_done += bool(sd); n_trunc += (not sd)
                so = s.reset()
                steps_in_episode[e] = 0
                assert venv.episode_seeds[e][3] == s.env.scenario.start_time.hour
                assert venv.episode_seeds[e][0] == s.env.sensor.seed
            assert so.CGM == o[e]
        assert np.array_equal(venv.episode_steps.cpu().numpy(), steps_in_episode)
    assert n_done >= 2 and n_trunc >= 2                               # both reasons for a restart were exercised
    assert np.array_equal(venv.state.cpu().numpy().T, np.stack([s.env.patient.state for s in singles]))
    venv.close()


def test_device_seeded_vector_env_properties():
    import torch
    from simglucose_b200 import params as P
    from simglucose_b200.envs import T1DSimVecEnv
    E, horizon = 30000, 40
    kw = dict(seed=9, horizon_steps=horizon, gym_seeding=True, auto_reset=True)
    a = T1DSimVecEnv(E, **kw)
    oa = a.reset().clone()
    # start hours ~ randint(0, 24): every hour occurs, frequencies within 5 sigma of 1/24
    hours = (a.start_minutes // 60).cpu().numpy()
    assert a.start_minutes.remainder(60).abs().sum().item() == 0 and hours.min() == 0 and hours.max() == 23
    freq = np.bincount(hours, minlength=24) / E
    assert np.abs(freq - 1 / 24).max() < 5 * np.sqrt((1 / 24) * (23 / 24) / E)
    # random_init_bg twice: x' = m + sqrt(0.1 m) z, x'' = x' + sqrt(0.1 x') z on x[3], x[4], x[12]; the rest is the table
    x = a.state.cpu().numpy()
    tx = P.cohort_table()[1][a.patient_row.cpu().numpy()].T
    rest = [i for i in range(13) if i not in (3, 4, 12)]
    assert np.array_equal(x[rest], tx[rest])
    for i in (3, 4, 12):
        z = (x[i] - tx[i]) / (2.0 * np.sqrt(0.1 * tx[i]))             # ~ N(0, 1) up to O(z^2 / sqrt(m)) terms
        assert abs(z.mean()) < 0.03 and abs(z.std() - 1.0) < 0.03
    assert abs(np.corrcoef(x[3] - tx[3], x[4] - tx[4])[0, 1]) < 0.03   # diagonal covariance: independent components
    # determinism, and a shard of the population sees the same environments (streams keyed by global env index)
    lo, hi = 12000, 12600
    b = T1DSimVecEnv(hi - lo, episode_offset=lo, patient_names=[P.patient_names()[(lo + e) % 30] for e in range(hi - lo)], **kw)
    ob = b.reset()
    assert torch.equal(oa[lo:hi], ob)
    basal = torch.full((E,), 0.1, dtype=torch.float64, device=a.device)       # several times basal: episodes end early
    restarted = torch.zeros(E, dtype=torch.bool, device=a.device)
    n_done = 0
    for k in range(horizon + 5):
        prev_started = a.episodes_started.clone()
        o, r, dn, info = a.step(basal)
        o2, r2, dn2, info2 = b.step(basal[lo:hi])
        assert torch.equal(o[lo:hi], o2) and torch.equal(r[lo:hi], r2) and torch.equal(dn[lo:hi], dn2)
        fin = dn | info['truncated']
        assert torch.equal(a.episodes_started - prev_started, fin.to(torch.int64))
        assert torch.equal(a.episode_steps == 0, fin)
        # a restarted environment reports the first observation of its new episode, the others this step's
        assert torch.equal(o[~fin], info['terminal_obs'][~fin])
        restarted |= fin
        n_done += int(dn.sum())
        if k == horizon - 1:
            assert bool(restarted.all())                                       # done or truncated by now
    assert int((a.engine.status() != 1).sum()) == 0
    st = a.statistics(reduce=False).cpu().numpy()
    assert np.array_equal(st[0], a.episode_steps.cpu().numpy() + 1)            # rows of the CURRENT episodes
    assert n_done > 0 and int(a.episodes_started.max()) > 2                     # environments that ended on `done`
    # a fused rollout after per-environment restarts is bounded by the OLDEST running episode, not by the env-steps
    # taken since reset() (horizon + 5 by now)
    oldest = int(a.episode_steps.max())
    assert a.steps == horizon + 5 and oldest < horizon
    with pytest.raises(RuntimeError, match='past the episode horizon'):
        a.rollout(horizon - oldest + 1, controller='bb')
    a.rollout(horizon - oldest, controller='bb')
    assert int((a.engine.status() != 1).sum()) == 0                            # no stream ran out
    a.close(); b.close()


def test_vector_env_step_random_is_the_random_agent():
    import torch
    from simglucose_b200.envs import T1DSimVecEnv
    E = 4096
    a = T1DSimVecEnv(E, seed=1, horizon_steps=30)
    b = T1DSimVecEnv(E, seed=1, horizon_steps=30)
    a.reset(); b.reset()
    for k in range(10):
        o, r, dn, info = a.step_random(0.0, 0.05, seed=3)
        basal = a._row[10].clone()
        assert float(basal.min()) >= 0 and float(basal.max()) <= 0.05
        o2, r2, dn2, _ = b.step(basal)
        assert torch.equal(o, o2) and torch.equal(r, r2) and torch.equal(dn, dn2)
    a.close(); b.close()
