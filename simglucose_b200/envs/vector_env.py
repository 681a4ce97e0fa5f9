"""T1DSimVecEnv -- lockstep vector environment: num_envs closed loops resident on one GPU, one kernel
launch per env-step, observations / rewards / dones as device tensors (no host round trip).

Semantics per environment are exactly those of the gym wrapper's inner T1DSimEnv (simulation/env.py:66-155
of the reference: 3-minute Dexcom steps, risk_diff reward on the CGM history, done = BG < 70 or BG > 350,
two sensor samples consumed at reset); the batch dimension is new.  Episodes all have the same length
(`horizon_steps`); like SimObj.simulate the environment keeps simulating through `done` -- the flag is
reported, nothing is reset behind the caller's back unless auto_reset=True, which starts a fresh batch of
episodes (new streams) once the horizon is reached.

stream_mode
  'device'     CGM noise and meal schedules drawn on the GPU (Philox; same distributions as CGMNoise /
               RandomScenario), any population size;
  'reference'  the reference's own MT19937 streams for sensor seed = scenario seed = seed + env index
               (host-generated: small populations, parity with the single-env facade).
"""
import numpy as np
import torch

from .. import _capi
from .. import params as P
from .. import streams as S
from ..engine import BatchEngine

_T = {name: i for i, name in enumerate(_capi.TRACE_COLS)}


class T1DSimVecEnv(object):
    def __init__(self, num_envs, patient_names=None, seed=0, start_hour=0, horizon_steps=480, sensor='Dexcom',
                 pump='Insulet', stream_mode='device', group_size=1, params=None, x0=None, auto_reset=False,
                 device=0, carry_h=False, fma=False, episode_offset=0, step_cap=None, reward_fun=None, fp32=False):
        """patient_names: list of names / 1-based ids cycled over the environments (default: the 30 virtual
        patients in table order).  params [32, E] / x0 [13, E]: per-environment overrides (perturbed
        populations).  group_size: environments per shared stream (the reference shares RandomScenario(seed)
        and CGMSensor(seed) across the patients of one batch_sim).  step_cap: heavy-minute compaction
        (BatchEngine.set_step_cap) for step(); default 5 when every environment has a meal schedule of its own
        (group_size == 1) and the population is large enough for the second launch to pay (>= 524,288), else off.
        fp32: the optional single-precision integrator (SG_FLAG_FP32; relative BG error <= 1e-3 per day, see
        csrc/sg_fp32.cuh) -- for RL populations; the heavy-minute compaction does not apply to it.
        reward_fun: None = risk_diff fused into the step kernel (env.py:27-33), or a callable
        reward_fun(BG_last_hour) -> [E] tensor, the reference's custom-reward hook (env.py:66, 100-102) batched:
        BG_last_hour is the device tensor CGM_hist[-int(60 / sample_time):], shape [n <= window, E], oldest first
        (the reset sample included, as in the reference); it is a view of `self.cgm_window`, valid until the next
        step."""
        self.num_envs = int(num_envs)
        self.engine = BatchEngine(self.num_envs, device=device, carry_h=carry_h, fma=fma, sensor=sensor, pump=pump,
                                  fp32=fp32)
        self.device = self.engine.device
        self.sample_time = self.engine.sample_time
        auto = 5 if (int(group_size) == 1 and self.num_envs >= 524288 and not fp32) else 0
        self.step_cap = auto if step_cap is None else int(step_cap)
        self.engine.set_step_cap(self.step_cap)
        names = P.patient_names() if patient_names is None else list(patient_names)
        rows = np.array([P.patient_row(n) for n in names], dtype=np.int32)
        self.patient_row = torch.from_numpy(rows[np.arange(self.num_envs) % len(rows)]).to(self.device)
        self.patient_names = [P.patient_names()[r] for r in self.patient_row.cpu().tolist()] if self.num_envs <= 4096 else None
        self.seed = int(seed)
        self.start_minute = int(start_hour) * 60
        self.horizon_steps = int(horizon_steps)
        self.stream_mode = stream_mode
        self.group_size = int(group_size)
        self.episode_offset = int(episode_offset)
        self._params, self._x0 = params, x0
        self.auto_reset = bool(auto_reset)
        self._epoch = 0
        self._zeros = torch.zeros(self.num_envs, dtype=torch.float64, device=self.device)
        self._row = torch.empty(_capi.NTRACE, self.num_envs, dtype=torch.float64, device=self.device)
        self.steps = None
        self.max_basal = float(self.engine.pump['max_basal'])
        self.reward_fun = reward_fun
        self.window_size = int(60 / self.sample_time)                      # env.py:100
        self.cgm_window = (torch.zeros(self.window_size, self.num_envs, dtype=torch.float64, device=self.device)
                           if reward_fun is not None else None)
        self._window_n = 0

    # -- streams -----------------------------------------------------------------------------------------------
    def _make_streams(self):
        E, n_noise, n_min = self.num_envs, self.horizon_steps + 2, self.horizon_steps * self.sample_time
        seed = self.seed + 7919 * self._epoch
        if self.stream_mode == 'device':
            noise = S.device_noise(E, n_noise, self.engine.sensor, seed=seed, episode_offset=self.episode_offset,
                                   group_size=self.group_size, device=self.device)
            mm, mg = S.device_scenarios(E, n_min, start_minute=self.start_minute, seed=seed,
                                        episode_offset=self.episode_offset, group_size=self.group_size,
                                        device=self.device)
            return noise, mm, mg
        if self.stream_mode != 'reference':
            raise ValueError("stream_mode must be 'device' or 'reference'")
        groups = [(self.episode_offset + e) // self.group_size for e in range(E)]
        cache = {}
        for g in sorted(set(groups)):
            cache[g] = (S.cgm_noise_stream(self.engine.sensor, seed + g, n_noise),
                        S.random_scenario_table(seed + g, self.start_minute, n_min))
        noise = np.stack([cache[g][0] for g in groups], axis=1)
        mm, mg = S.pad_tables([cache[g][1] for g in groups])
        return noise, mm, mg

    # -- gym-like API --------------------------------------------------------------------------------------------
    def reset(self):
        """Start num_envs episodes.  Returns the observation tensor [E] (CGM, mg/dL).  `self.info` then holds
        the reset row (bg, lbgi, hbgi, risk)."""
        noise, mm, mg = self._make_streams()
        row0 = self.engine.reset(self.patient_row, noise, mm, mg, x0=self._x0, params=self._params)
        self.steps = 0
        self._epoch += 1
        self.info = self._info(row0, reset=True)
        if self.cgm_window is not None:
            self.cgm_window[-1] = row0[_T['CGM']]                          # CGM_hist[0], env.py:129
            self._window_n = 1
        return row0[_T['obs']]

    def _info(self, row, reset=False):
        return {'bg': row[_T['BG']], 'cgm': row[_T['CGM']], 'lbgi': row[_T['LBGI']], 'hbgi': row[_T['HBGI']],
                'risk': row[_T['Risk']], 'meal': self._zeros if reset else row[_T['CHO']],
                'insulin': self._zeros if reset else row[_T['insulin']], 'sample_time': self.sample_time,
                'patient_row': self.patient_row, 'minute': (self.steps or 0) * self.sample_time}

    def step(self, action):
        """action: basal rates [E] (the gym wrapper's action), or a pair / [2, E] tensor (basal, bolus), U/min.
        Returns (obs [E], reward [E], done [E] bool, info dict of [E] tensors); all tensors live on the GPU and
        are overwritten by the next step."""
        if self.steps is None:
            raise RuntimeError('call reset() first')
        if self.steps >= self.horizon_steps:
            if not self.auto_reset:
                raise RuntimeError('episode horizon (%d steps) reached: call reset()' % self.horizon_steps)
            self.reset()
        if isinstance(action, (tuple, list)):
            basal, bolus = action
        else:
            a = action if isinstance(action, torch.Tensor) else torch.as_tensor(np.asarray(action, dtype=np.float64))
            if a.ndim == 2 and a.shape[0] == 2:
                basal, bolus = a[0], a[1]
            else:
                basal, bolus = a.reshape(-1), self._zeros
        row = self.engine.step(basal, bolus, out=self._row)
        self.steps += 1
        self.info = self._info(row)
        reward = row[_T['reward']]
        if self.cgm_window is not None:                                    # env.py:96, 100-102
            w = self.cgm_window
            w[:-1] = w[1:].clone()
            w[-1] = row[_T['CGM']]
            self._window_n = min(self._window_n + 1, self.window_size)
            reward = self.reward_fun(w[self.window_size - self._window_n:])
        return row[_T['obs']], reward, row[_T['done']] != 0, self.info

    def rollout(self, n_steps, controller='bb', ctrl=None, ctrl_rows=None, actions=None, trace=False):
        """n_steps env-steps in ONE launch with the controller evaluated on device ('bb', 'pid') or a
        pre-drawn action tensor [n_steps, 2, E] ('actions').  Returns the engine's dict (trace / counters)."""
        if self.steps is None:
            raise RuntimeError('call reset() first')
        if self.steps + n_steps > self.horizon_steps:
            raise RuntimeError('rollout past the episode horizon')
        if controller == 'bb' and ctrl is None and ctrl_rows is None:
            q = torch.from_numpy(P.quest_table()).to(self.device)[self.patient_row.long()]
            ctrl_rows = torch.stack([torch.full_like(q[:, 0], 140.0), q[:, 0], q[:, 1], torch.zeros_like(q[:, 0])])
        if self.step_cap:
            self.engine.set_step_cap(0)                  # one fused launch; the cap is for lockstep stepping
        try:
            out = self.engine.run(n_steps, controller=controller, ctrl=ctrl, ctrl_rows=ctrl_rows, actions=actions,
                                  trace=trace)
        finally:
            if self.step_cap:
                self.engine.set_step_cap(self.step_cap)
        self.steps += n_steps
        return out

    def statistics(self, reduce=True):
        """Population aggregate [NSTAT] (reduce=True, one reduction launch) or per-episode [NSTAT, E]."""
        return self.engine.reduce_stats() if reduce else self.engine.stats()

    @property
    def state(self):
        """Patient states [13, E]."""
        return self.engine.state()

    def close(self):
        self.engine.close()
