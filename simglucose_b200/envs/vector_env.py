"""T1DSimVecEnv -- lockstep vector environment: num_envs closed loops resident on one GPU, one kernel
launch per env-step, observations / rewards / dones as device tensors (no host round trip).

Semantics per environment are exactly those of the gym wrapper's inner T1DSimEnv (simulation/env.py:66-155
of the reference: 3-minute Dexcom steps, risk_diff reward on the CGM history, done = BG < 70 or BG > 350,
two sensor samples consumed at reset); the batch dimension is new.

Two ways to say what an environment is:

  gym_seeding=False  every environment starts at `start_hour` from the table state of its patient, sensor and
                     scenario seeded with seed + env index (the reference's SimObj set-up, batched);
  gym_seeding=True   every environment IS the gym wrapper `T1DSimEnv(patient_name, seed=seed + env index)`
                     (envs/simglucose_gym_env.py:27-73 of the reference): each reset draws, from that environment's
                     own generator, the sensor / scenario / patient seeds and the start hour, and the patient starts
                     from a randomised state (random_init_bg, patient/t1dpatient.py:256-270).

stream_mode
  'device'     everything random is drawn on the GPU (Philox; the same distributions as CGMNoise, RandomScenario,
               randint(0, 24), multivariate_normal), any population size, nothing crosses the PCIe bus;
  'reference'  the reference's own MT19937 streams, replayed on the host (streams.py): small populations, bit-level
               parity with N single-environment facades (tests/test_gpu_facade.py).

auto_reset=True restarts every environment ON ITS OWN when it reports `done` or reaches `horizon_steps` (what an RL
loop does with `if done: env.reset()`): a masked reset kernel (sg_reset_masked) after the step, new streams written
into the masked columns of the resident buffers.  The observation returned for such an environment is the first
observation of its new episode; reward, done and info['bg'] describe the step that ended the old one, and
info['terminal_obs'] is that step's observation.  Without auto_reset the environments keep simulating through
`done`, like SimObj.simulate (simulation/sim_engine.py:33).
"""
import numpy as np
import torch

from .. import _capi
from .. import params as P
from .. import streams as S
from ..engine import BatchEngine

_T = {name: i for i, name in enumerate(_capi.TRACE_COLS)}
_KEY_SHIFT = 24          # Philox stream of an episode: (global env index << 24) | episodes started in that slot


class T1DSimVecEnv(object):
    def __init__(self, num_envs, patient_names=None, seed=0, start_hour=0, horizon_steps=480, sensor='Dexcom',
                 pump='Insulet', stream_mode='device', group_size=1, params=None, x0=None, auto_reset=False,
                 device=0, carry_h=False, fma=False, episode_offset=0, step_cap=None, reward_fun=None, fp32=False,
                 gym_seeding=False, coherence=None, adaptive_sort=None):
        """patient_names: list of names / 1-based ids cycled over the environments (default: the 30 virtual
        patients in table order).  params [32, E] / x0 [13, E]: per-environment overrides (perturbed
        populations; x0 is ignored with gym_seeding, which randomises the table state).  group_size: environments
        per shared stream (the reference shares RandomScenario(seed) and CGMSensor(seed) across the patients of one
        batch_sim); 1 with gym_seeding.
        coherence / adaptive_sort: env-steps between two sorts of the launch order and what it is sorted by
        (BatchEngine.set_coherence).  Defaults for populations that fill the GPU (>= 65,536 environments with schedules of
        their own): with auto_reset every env-step by the DOPRI5 steps every environment just took (measured on 1,048,576
        gym environments: 3.64 -> 2.57 ms per env-step, profiles/r02_t_*), otherwise every 40 env-steps (two hours) by the
        next meal event; smaller populations: off.  step_cap: the older heavy-minute compaction (two launches per step), default off.
        fp32: the optional single-precision integrator (SG_FLAG_FP32; relative BG error <= 1e-3 per day, see
        csrc/sg_fp32.cuh) -- for RL populations.
        reward_fun: None = risk_diff fused into the step kernel (env.py:27-33), or a callable
        reward_fun(BG_last_hour) -> [E] tensor, the reference's custom-reward hook (env.py:66, 100-102) batched:
        BG_last_hour is the device tensor CGM_hist[-int(60 / sample_time):], oldest first (the reset sample included,
        as in the reference), a view of `self.cgm_window` valid until the next step.  Without auto_reset its shape is
        [n <= window, E]; with auto_reset environments are at different points of their episodes, the shape is always
        [window, E] and samples an episode does not have yet are NaN."""
        self.num_envs = int(num_envs)
        self.engine = BatchEngine(self.num_envs, device=device, carry_h=carry_h, fma=fma, sensor=sensor, pump=pump,
                                  fp32=fp32)
        self.device = self.engine.device
        self.sample_time = self.engine.sample_time
        self.gym_seeding = bool(gym_seeding)
        self.group_size = 1 if self.gym_seeding else int(group_size)
        own_schedules = self.group_size == 1
        self.step_cap = 0 if step_cap is None else int(step_cap)
        self.engine.set_step_cap(self.step_cap)
        big = own_schedules and self.num_envs >= 65536 and not self.step_cap
        # environments that restart one by one (a fresh start costs 8 - 9 DOPRI5 steps in its first minute against 3):
        # re-sort every env-step, by what every episode just cost; otherwise by meal events, every two hours
        restarts = bool(auto_reset)
        auto = (1 if restarts else 40) if big else 0
        self.coherence = auto if coherence is None else int(coherence)
        self.adaptive_sort = (restarts and big) if adaptive_sort is None else bool(adaptive_sort)
        self.engine.set_coherence(self.coherence, adaptive=self.adaptive_sort)
        names = P.patient_names() if patient_names is None else list(patient_names)
        rows = np.array([P.patient_row(n) for n in names], dtype=np.int32)
        self._rows_host = rows[np.arange(self.num_envs) % len(rows)]
        self.patient_row = torch.from_numpy(self._rows_host).to(self.device)
        self.patient_names = [P.patient_names()[r] for r in self._rows_host.tolist()] if self.num_envs <= 4096 else None
        self.seed = int(seed)
        self.start_minute = int(start_hour) * 60
        self.horizon_steps = int(horizon_steps)
        if stream_mode not in ('device', 'reference'):
            raise ValueError("stream_mode must be 'device' or 'reference'")
        self.stream_mode = stream_mode
        self.episode_offset = int(episode_offset)
        self._params, self._x0 = params, x0
        self.auto_reset = bool(auto_reset)
        E, dev = self.num_envs, self.device
        self._zeros = torch.zeros(E, dtype=torch.float64, device=dev)
        self._row = torch.empty(_capi.NTRACE, E, dtype=torch.float64, device=dev)
        self._row0 = torch.zeros(_capi.NTRACE, E, dtype=torch.float64, device=dev)
        self._obs = torch.empty(E, dtype=torch.float64, device=dev)
        self.steps = None                                   # env-steps since the last full reset()
        self.max_basal = float(self.engine.pump['max_basal'])
        self.reward_fun = reward_fun
        self.window_size = int(60 / self.sample_time)                      # env.py:100
        self.cgm_window = (torch.zeros(self.window_size, E, dtype=torch.float64, device=dev)
                           if reward_fun is not None else None)
        self._window_n = 0
        # resident stream buffers (the engine borrows them; masked resets rewrite columns in place)
        n_min = self.horizon_steps * self.sample_time
        self._noise = torch.zeros(self.horizon_steps + 2, E, dtype=torch.float64, device=dev)
        self._max_meals = 6 * ((1439 + n_min + 1439) // 1440)              # any start minute of the day
        self._mm = torch.full((self._max_meals, E), -1, dtype=torch.int32, device=dev)
        self._mg = torch.zeros(self._max_meals, E, dtype=torch.float64, device=dev)
        self._x0_buf = torch.zeros(13, E, dtype=torch.float64, device=dev)
        self._table_x0 = torch.from_numpy(P.cohort_table()[1]).to(dev).contiguous()
        self._W = torch.from_numpy(S.interp_matrix(self.sample_time)).to(dev).contiguous()
        # per-environment episode bookkeeping
        self.episodes_started = torch.zeros(E, dtype=torch.int64, device=dev)     # in each slot
        self.episode_steps = torch.zeros(E, dtype=torch.int32, device=dev)        # env-steps of the current episode
        self.start_minutes = torch.full((E,), self.start_minute, dtype=torch.int32, device=dev)
        self._env_index = torch.arange(E, dtype=torch.int64, device=dev) + self.episode_offset
        # outputs of the book-keeping kernels around per-environment resets (sg_vec_finish / sg_vec_restart)
        self._mask = torch.zeros(E, dtype=torch.int32, device=dev)
        self._key = torch.zeros(E, dtype=torch.int64, device=dev)
        self._terminal_obs = torch.zeros(E, dtype=torch.float64, device=dev)
        self._done = torch.zeros(E, dtype=torch.bool, device=dev)
        self._truncated = torch.zeros(E, dtype=torch.bool, device=dev)
        self._list = torch.zeros(E, dtype=torch.int32, device=dev)          # the environments that ended, gathered
        self._count = torch.zeros(1, dtype=torch.int32, device=dev)
        if self.stream_mode == 'reference' and self.gym_seeding:
            # one gym np_random per environment (simglucose_gym_env.py:38); the wrapper's constructor already builds
            # one env from it (:39), which consumes the first chain of draws
            self._np_random = [S.np_random(self.seed + self.episode_offset + e)[0] for e in range(E)]
            for rng in self._np_random:
                self._gym_chain(rng)
        self.episode_seeds = None             # reference + gym_seeding: [(sensor, scenario, patient seed, hour)] per env

    # -- streams of new episodes -----------------------------------------------------------------------------------
    @staticmethod
    def _gym_chain(rng):
        """_create_env_from_random_state, simglucose_gym_env.py:58-73: (sensor seed, scenario seed, patient seed, hour)."""
        seed2 = S.hash_seed(rng.randint(0, 1000)) % 2 ** 31
        seed3 = S.hash_seed(seed2 + 1) % 2 ** 31
        seed4 = S.hash_seed(seed3 + 1) % 2 ** 31
        hour = rng.randint(low=0.0, high=24.0)
        return int(seed2), int(seed3), int(seed4), int(hour)

    def _new_episodes_device(self, mask, key=None, listed=None):
        """Philox streams for the environments in mask (int32 [E] device tensor, None = all), no host involvement.
        key / listed: the stream keys and the gathered list of the masked environments when the caller already has
        them (sg_vec_finish)."""
        if key is None:
            key = (self._env_index // self.group_size << _KEY_SHIFT) | self.episodes_started
        x0 = None
        if self.gym_seeding:
            S.device_episode_starts(key, self.patient_row, self._table_x0, n_draws=2, seed=self.seed, mask=mask,
                                    start_minute=self.start_minutes, x0=self._x0_buf)
            x0 = self._x0_buf
        S.device_noise_keyed(self._noise, key, self.engine.sensor, seed=self.seed, mask=mask, W=self._W, listed=listed)
        S.device_scenarios_keyed(self._mm, self._mg, self.start_minutes, self.horizon_steps * self.sample_time, key,
                                 seed=self.seed, mask=mask, listed=listed)
        return x0

    def _new_episodes_reference(self, idx):
        """The reference's MT19937 streams for the environments idx (host list), written into the device buffers."""
        n_noise, n_min = self.horizon_steps + 2, self.horizon_steps * self.sample_time
        started = self.episodes_started.cpu().numpy()
        x0_rows = P.cohort_table()[1]
        cols_noise, cols_tab, cols_x0, hours = [], [], [], []
        if self.episode_seeds is None:
            self.episode_seeds = [None] * self.num_envs
        cache = {}
        for e in idx:
            if self.gym_seeding:
                s2, s3, s4, hour = self._gym_chain(self._np_random[e])
                # the wrapper builds the patient (constructor reset) and then resets the env: two randomisations
                # with the same RandomState(seed4), the second around the first (t1dpatient.py:252, 268-270)
                x = S.random_init_state(S.random_init_state(x0_rows[self._rows_host[e]], s4), s4)
                start = hour * 60
            else:
                g = (self.episode_offset + e) // self.group_size
                s2 = s3 = self.seed + g + 7919 * int(started[e])
                s4, hour, start = None, self.start_minute // 60, self.start_minute
                x = None
            self.episode_seeds[e] = (s2, s3, s4, hour)
            k = (s2, s3, start)
            if k not in cache:
                cache[k] = (S.cgm_noise_stream(self.engine.sensor, s2, n_noise), S.random_scenario_table(s3, start, n_min))
            cols_noise.append(cache[k][0])
            cols_tab.append(cache[k][1])
            cols_x0.append(x)
            hours.append(start)
        ix = torch.as_tensor(list(idx), dtype=torch.int64, device=self.device)
        self._noise.index_copy_(1, ix, torch.from_numpy(np.stack(cols_noise, axis=1)).to(self.device))
        mm, mg = S.pad_tables(cols_tab, min_width=self._max_meals)
        assert mm.shape[0] == self._max_meals, 'meal table wider than 6 meals per day'
        self._mm.index_copy_(1, ix, torch.from_numpy(mm).to(self.device))
        self._mg.index_copy_(1, ix, torch.from_numpy(mg).to(self.device))
        self.start_minutes.index_copy_(0, ix, torch.as_tensor(hours, dtype=torch.int32, device=self.device))
        if self.gym_seeding:
            self._x0_buf.index_copy_(1, ix, torch.from_numpy(np.stack(cols_x0, axis=1)).to(self.device))
            return self._x0_buf
        return None

    def _x0_for_reset(self, generated):
        if generated is not None:
            return generated
        return self._x0

    # -- gym-like API --------------------------------------------------------------------------------------------
    def reset(self):
        """Start num_envs episodes.  Returns the observation tensor [E] (CGM, mg/dL).  `self.info` then holds
        the reset row (bg, lbgi, hbgi, risk)."""
        if self.stream_mode == 'device':
            x0 = self._new_episodes_device(None)
        else:
            x0 = self._new_episodes_reference(range(self.num_envs))
        row0 = self.engine.reset(self.patient_row, self._noise, self._mm, self._mg, x0=self._x0_for_reset(x0),
                                 params=self._params)
        self.steps = 0
        self.episodes_started += 1
        self.episode_steps.zero_()
        self.info = self._info(row0, reset=True)
        if self.cgm_window is not None:
            if self.auto_reset:
                self.cgm_window.fill_(float('nan'))
            self.cgm_window[-1] = row0[_T['CGM']]                          # CGM_hist[0], env.py:129
            self._window_n = 1
        return row0[_T['obs']]

    def _info(self, row, reset=False):
        return {'bg': row[_T['BG']], 'cgm': row[_T['CGM']], 'lbgi': row[_T['LBGI']], 'hbgi': row[_T['HBGI']],
                'risk': row[_T['Risk']], 'meal': self._zeros if reset else row[_T['CHO']],
                'insulin': self._zeros if reset else row[_T['insulin']], 'sample_time': self.sample_time,
                'patient_row': self.patient_row, 'minute': (self.steps or 0) * self.sample_time,
                'start_minute': self.start_minutes, 'episode_steps': self.episode_steps}

    def _split_action(self, action):
        if isinstance(action, (tuple, list)):
            return action
        a = action if isinstance(action, torch.Tensor) else torch.as_tensor(np.asarray(action, dtype=np.float64))
        if a.ndim == 2 and a.shape[0] == 2:
            return a[0], a[1]
        return a.reshape(-1), self._zeros

    def step(self, action):
        """action: basal rates [E] (the gym wrapper's action), or a (basal, bolus) tuple / [2, E] tensor, U/min.
        Returns (obs [E], reward [E], done [E] bool, info dict of [E] tensors); all tensors live on the GPU and
        are overwritten by the next step."""
        self._before_step()
        basal, bolus = self._split_action(action)
        row = self.engine.step(basal, bolus, out=self._row)
        return self._after_step(row)

    def step_random(self, low=0.0, high=None, seed=None):
        """One env-step under Action(basal ~ U(low, high), bolus 0) drawn INSIDE the step kernel (SG_CTRL_RANDOM): the
        random-agent baseline `env.step(env.action_space.sample())` without an action tensor or a sampling launch.
        high defaults to the pump's max_basal (the gym wrapper's action_space, simglucose_gym_env.py:78-81)."""
        self._before_step()
        high = self.max_basal if high is None else float(high)
        seed = self.seed if seed is None else int(seed)
        self.engine.run(1, controller='random', ctrl=(float(low), high, float(seed), float(self.episode_offset)),
                        trace=self._row)
        return self._after_step(self._row)

    def _before_step(self):
        if self.steps is None:
            raise RuntimeError('call reset() first')
        if not self.auto_reset and self.steps >= self.horizon_steps:
            raise RuntimeError('episode horizon (%d steps) reached: call reset()' % self.horizon_steps)

    def _after_step(self, row):
        self.steps += 1
        self.info = self._info(row)
        reward = row[_T['reward']]
        if self.auto_reset and self.stream_mode == 'device':
            # one launch: episode_steps += 1, done, finished mask, truncation, terminal observation, keys of the new episodes
            lib, E = self.engine.lib, self.num_envs
            _capi.check(lib.sg_vec_finish(row.data_ptr(), E, _T['done'], _T['obs'], self.episode_steps.data_ptr(),
                                          self.horizon_steps, self.episodes_started.data_ptr(), self.episode_offset,
                                          self.group_size, _KEY_SHIFT, self._mask.data_ptr(), self._key.data_ptr(),
                                          self._terminal_obs.data_ptr(), self._done.data_ptr(), self._truncated.data_ptr(),
                                          self._list.data_ptr(), self._count.data_ptr(),
                                          torch.cuda.current_stream(self.device).cuda_stream))
            done = self._done
        else:
            self.episode_steps += 1
            done = row[_T['done']] != 0
        if self.cgm_window is not None:                                    # env.py:96, 100-102
            w = self.cgm_window
            w[:-1] = w[1:].clone()
            w[-1] = row[_T['CGM']]
            self._window_n = min(self._window_n + 1, self.window_size)
            reward = self.reward_fun(w if self.auto_reset else w[self.window_size - self._window_n:])
        obs = row[_T['obs']]
        if self.auto_reset:
            obs = self._reset_finished(row, done)
        return obs, reward, done, self.info

    def _reset_finished(self, row, done):
        """Per-environment reset of the episodes that reported done or reached the horizon."""
        if self.stream_mode == 'device':
            mask = self._mask                                              # filled by sg_vec_finish
            self.info['terminal_obs'] = self._terminal_obs
            self.info['truncated'] = self._truncated
            x0 = self._new_episodes_device(mask, key=self._key, listed=(self._list, self._count))
            self.engine.reset_masked(mask, x0=self._x0_for_reset(x0), row0=self._row0)
            if self.cgm_window is not None:
                finished = mask != 0
                self.cgm_window[:, finished] = float('nan')
                self.cgm_window[-1] = torch.where(finished, self._row0[_T['CGM']], self.cgm_window[-1])
            _capi.check(self.engine.lib.sg_vec_restart(mask.data_ptr(), self.num_envs, self.episodes_started.data_ptr(),
                                                       self.episode_steps.data_ptr(), self._row0[_T['obs']].data_ptr(),
                                                       row[_T['obs']].data_ptr(), self._obs.data_ptr(),
                                                       torch.cuda.current_stream(self.device).cuda_stream))
            return self._obs
        finished = done | (self.episode_steps >= self.horizon_steps)
        mask = finished.to(torch.int32)
        self.info['terminal_obs'] = row[_T['obs']].clone()
        self.info['truncated'] = finished & ~done
        idx = finished.nonzero().reshape(-1).cpu().tolist()                # host streams: small populations
        if not idx:
            return row[_T['obs']]
        x0 = self._new_episodes_reference(idx)
        self.engine.reset_masked(mask, x0=self._x0_for_reset(x0), row0=self._row0)
        self.episodes_started += mask.to(torch.int64)
        self.episode_steps.masked_fill_(finished, 0)
        if self.cgm_window is not None:
            self.cgm_window[:, finished] = float('nan')
            self.cgm_window[-1] = torch.where(finished, self._row0[_T['CGM']], self.cgm_window[-1])
        torch.where(finished, self._row0[_T['obs']], row[_T['obs']], out=self._obs)
        return self._obs

    def rollout(self, n_steps, controller='bb', ctrl=None, ctrl_rows=None, actions=None, trace=False):
        """n_steps env-steps in ONE launch with the controller evaluated on device ('bb', 'pid', 'random') or a
        pre-drawn action tensor [n_steps, 2, E] ('actions').  Returns the engine's dict (trace / counters).  No
        per-environment resets happen inside a rollout."""
        if self.steps is None:
            raise RuntimeError('call reset() first')
        # with auto_reset the environments restarted one by one: what bounds a rollout is the OLDEST running episode
        # (its noise stream ends at the horizon), not the env-steps taken since reset() -- one device read per rollout
        oldest = int(self.episode_steps.max()) if self.auto_reset else self.steps
        if oldest + n_steps > self.horizon_steps:
            raise RuntimeError('rollout past the episode horizon (oldest episode at env-step %d of %d, %d more asked)'
                               % (oldest, self.horizon_steps, n_steps))
        if controller == 'bb' and ctrl is None and ctrl_rows is None:
            q = torch.from_numpy(P.quest_table()).to(self.device)[self.patient_row.long()]
            tbl = P.cohort_table()[0]
            basal = torch.from_numpy(tbl[:, 1] * tbl[:, 0] / 6000.0).to(self.device)[self.patient_row.long()]
            ctrl_rows = torch.stack([torch.full_like(q[:, 0], 140.0), q[:, 0], q[:, 1], basal])
        if self.step_cap:
            self.engine.set_step_cap(0)                  # one fused launch; the cap is for lockstep stepping
        try:
            out = self.engine.run(n_steps, controller=controller, ctrl=ctrl, ctrl_rows=ctrl_rows, actions=actions,
                                  trace=trace)
        finally:
            if self.step_cap:
                self.engine.set_step_cap(self.step_cap)
        self.steps += n_steps
        self.episode_steps += n_steps
        return out

    def statistics(self, reduce=True):
        """Population aggregate [NSTAT] (reduce=True, one reduction launch) or per-episode [NSTAT, E] of the CURRENT
        episodes (a per-environment reset restarts that environment's statistics)."""
        return self.engine.reduce_stats() if reduce else self.engine.stats()

    @property
    def state(self):
        """Patient states [13, E]."""
        return self.engine.state()

    def close(self):
        self.engine.close()
