"""T1DSimEnv -- the surface of simglucose/simulation/env.py:36-180 on top of the CUDA library.

With this package's T1DPatient + CGMSensor + InsulinPump the whole env-step (sample_time minutes of:
meal announcement, pump quantisation, meal state machine, DOPRI5 minute, CGM sample/hold; then the
averages, risk index, reward and done flag) is ONE launch of the step kernel on the patient's handle
(sg_step_meals); the scenario -- any Scenario subclass -- is evaluated on the host because it is a
user-extensible Python object (it only maps wall-clock time to grams).  Foreign sensor / pump objects
are honoured through the reference's own minute-by-minute choreography, with the patient minute
(T1DPatient.step) and the risk index still computed on the GPU.
"""
import logging
from collections import namedtuple
from datetime import timedelta

import numpy as np
import pandas as pd
import torch

from .. import _capi
from ..actuator.pump import InsulinPump
from ..analysis.risk import risk_index
from ..patient.t1dpatient import Action, T1DPatient
from ..sensor.cgm import CGMSensor

try:
    from rllab.envs.base import Step
except ImportError:
    _Step = namedtuple("Step", ["observation", "reward", "done", "info"])

    def Step(observation, reward, done, **kwargs):
        """Result of reset()/step(): a 4-tuple (observation, reward, done, info dict)."""
        return _Step(observation, reward, done, kwargs)

Observation = namedtuple('Observation', ['CGM'])
logger = logging.getLogger(__name__)

_T = {name: i for i, name in enumerate(_capi.TRACE_COLS)}


def risk_diff(BG_last_hour):
    """Default reward: risk(previous sample) - risk(current sample); 0 with fewer than two samples."""
    if len(BG_last_hour) < 2:
        return 0
    _, _, risk_current = risk_index([BG_last_hour[-1]], 1)
    _, _, risk_prev = risk_index([BG_last_hour[-2]], 1)
    return risk_prev - risk_current


def _scalar(v):
    return float(np.asarray(v, dtype=np.float64).reshape(-1)[0])


class T1DSimEnv(object):
    def __init__(self, patient, sensor, pump, scenario):
        self.patient = patient
        self.sensor = sensor
        self.pump = pump
        self.scenario = scenario
        self._reset()

    # the fused path needs the library's own plant pieces (subclasses may override their behaviour)
    @property
    def fused(self):
        return (type(self.patient) is T1DPatient and type(self.sensor) is CGMSensor
                and type(self.pump) is InsulinPump)

    @property
    def time(self):
        return self.scenario.start_time + timedelta(minutes=self.patient.t)

    # -- fused path ------------------------------------------------------------------------------------
    def _noise_tensor(self, n):
        eng = self.patient._engine
        arr = self.sensor._noise_generator.peek(n)
        return torch.from_numpy(np.ascontiguousarray(arr.reshape(-1, 1))).to(eng.device)

    def _fused_reset(self):
        """env._reset + the observation of env.reset (env.py:119-155) in one reset launch; returns row 0."""
        eng = self.patient._engine
        eng.set_hardware(dict(self.sensor._params), dict(self.pump._params))
        self._n_noise = 2 + 1024
        eng.set_t0(self.patient.t0)
        x0 = np.asarray(self.patient.init_state, dtype=np.float64).reshape(13, 1)
        row0 = eng.reset(self.patient._row, self._noise_tensor(self._n_noise), x0=x0).cpu().numpy()[:, 0]
        self.patient._status = 1
        self._sync_plant()
        return row0

    def _sync_plant(self):
        """Host mirrors of the patient and the sensor after the kernel advanced the resident episode."""
        d, i = self.patient._pull()
        self.sensor._last_CGM = float(d[18])                     # SG_D_LAST_CGM
        self.sensor._noise_generator.count = int(i[2])           # SG_I_NOISE_POS

    def _fused_step(self, action, reward_fun):
        eng = self.patient._engine
        st = int(self.sample_time)
        if eng.steps_done + 3 > self._n_noise:                  # open-ended episode: extend the stream
            self._n_noise = 2 * self._n_noise
            eng.rebind_streams(self._noise_tensor(self._n_noise))
        now = self.time
        meals = [self.scenario.get_action(now + timedelta(minutes=j)).meal for j in range(st)]
        if self.patient._status < 0:
            logger.error('ODE solver failed!!')
            raise RuntimeError('ODE solver failed (dopri5 idid = %d)' % self.patient._status)
        row = eng.step_meals(torch.tensor([_scalar(action.basal)], dtype=torch.float64),
                             torch.tensor([_scalar(action.bolus)], dtype=torch.float64),
                             torch.tensor(meals, dtype=torch.float64).reshape(st, 1)).cpu().numpy()[:, 0]
        self._sync_plant()
        return row

    # -- reference choreography for foreign components (env.py:48-64) ------------------------------------
    def mini_step(self, action):
        patient_action = self.scenario.get_action(self.time)
        basal = self.pump.basal(action.basal)
        bolus = self.pump.bolus(action.bolus)
        insulin = basal + bolus
        CHO = patient_action.meal
        self.patient.step(Action(insulin=insulin, CHO=CHO))
        BG = self.patient.observation.Gsub
        CGM = self.sensor.measure(self.patient)
        return CHO, insulin, BG, CGM

    def step(self, action, reward_fun=risk_diff):
        """action: namedtuple with .basal and .bolus (U/min), held for sample_time minutes."""
        if self.fused:
            row = self._fused_step(action, reward_fun)
            CHO, insulin, BG, CGM = row[_T['CHO']], row[_T['insulin']], row[_T['BG']], row[_T['CGM']]
            LBGI, HBGI, risk = row[_T['LBGI']], row[_T['HBGI']], row[_T['Risk']]
            CHO, insulin, BG, CGM, LBGI, HBGI, risk = map(float, (CHO, insulin, BG, CGM, LBGI, HBGI, risk))
        else:
            CHO = insulin = BG = CGM = 0.0
            for _ in range(int(self.sample_time)):
                tmp_CHO, tmp_insulin, tmp_BG, tmp_CGM = self.mini_step(action)
                CHO += tmp_CHO / self.sample_time
                insulin += tmp_insulin / self.sample_time
                BG += tmp_BG / self.sample_time
                CGM += tmp_CGM / self.sample_time
            LBGI, HBGI, risk = risk_index([BG], 1)
            row = None

        self.CHO_hist.append(CHO)
        self.insulin_hist.append(insulin)
        self.time_hist.append(self.time)
        self.BG_hist.append(BG)
        self.CGM_hist.append(CGM)
        self.risk_hist.append(risk)
        self.LBGI_hist.append(LBGI)
        self.HBGI_hist.append(HBGI)

        window_size = int(60 / self.sample_time)
        if row is not None and reward_fun is risk_diff:
            reward = float(row[_T['reward']])              # computed by the step kernel
        else:
            reward = reward_fun(self.CGM_hist[-window_size:])
        done = BG < 70 or BG > 350
        return Step(observation=Observation(CGM=CGM), reward=reward, done=done, sample_time=self.sample_time,
                    patient_name=self.patient.name, meal=CHO, patient_state=self.patient.state, time=self.time,
                    bg=BG, lbgi=LBGI, hbgi=HBGI, risk=risk)

    def _reset(self):
        self.sample_time = self.sensor.sample_time
        self.viewer = None
        if self.fused:
            self._row0 = self._fused_reset()
            BG, CGM = float(self._row0[_T['BG']]), float(self._row0[_T['CGM']])
            LBGI, HBGI, risk = (float(self._row0[_T[k]]) for k in ('LBGI', 'HBGI', 'Risk'))
        else:
            self._row0 = None
            BG = self.patient.observation.Gsub
            LBGI, HBGI, risk = risk_index([BG], 1)
            CGM = self.sensor.measure(self.patient)
        self.time_hist = [self.scenario.start_time]
        self.BG_hist = [BG]
        self.CGM_hist = [CGM]
        self.risk_hist = [risk]
        self.LBGI_hist = [LBGI]
        self.HBGI_hist = [HBGI]
        self.CHO_hist = []
        self.insulin_hist = []

    def reset(self):
        self.patient.reset()
        self.sensor.reset()
        self.pump.reset()
        self.scenario.reset()
        self._reset()
        if self._row0 is not None:
            CGM = float(self._row0[_T['obs']])             # second sensor sample at t = 0 (env.py:142)
        else:
            CGM = self.sensor.measure(self.patient)
        return Step(observation=Observation(CGM=CGM), reward=0, done=False, sample_time=self.sample_time,
                    patient_name=self.patient.name, meal=0, patient_state=self.patient.state, time=self.time,
                    bg=self.BG_hist[0], lbgi=self.LBGI_hist[0], hbgi=self.HBGI_hist[0], risk=self.risk_hist[0])

    def render(self, close=False):
        """The reference animates matplotlib figures here; this build is headless (no-op)."""
        return None

    def show_history(self):
        """DataFrame indexed by Time with columns BG, CGM, CHO, insulin, LBGI, HBGI, Risk.  CHO and insulin
        describe the step that STARTS at the row's time, so the last row has none (NaN)."""
        n = len(self.time_hist)

        def padded(values):
            return np.concatenate([np.asarray(values, dtype=np.float64), np.full(n - len(values), np.nan)])
        cols = [('BG', self.BG_hist), ('CGM', self.CGM_hist), ('CHO', self.CHO_hist), ('insulin', self.insulin_hist),
                ('LBGI', self.LBGI_hist), ('HBGI', self.HBGI_hist), ('Risk', self.risk_hist)]
        return pd.DataFrame({name: padded(v) for name, v in cols}, index=pd.Index(self.time_hist, name='Time'))

    # -- bulk hand-over used by SimObj.simulate / batch_sim -----------------------------------------------
    def _append_trace(self, trace):
        """trace [n_steps, NTRACE] (host): extend the histories as n_steps calls of step() would."""
        start, st = self.scenario.start_time, int(self.sample_time)
        k0 = len(self.BG_hist) - 1
        for k in range(trace.shape[0]):
            self.time_hist.append(start + timedelta(minutes=self.patient.t0 + (k0 + k + 1) * st))
        self.CHO_hist.extend(trace[:, _T['CHO']].tolist())
        self.insulin_hist.extend(trace[:, _T['insulin']].tolist())
        self.BG_hist.extend(trace[:, _T['BG']].tolist())
        self.CGM_hist.extend(trace[:, _T['CGM']].tolist())
        self.risk_hist.extend(trace[:, _T['Risk']].tolist())
        self.LBGI_hist.extend(trace[:, _T['LBGI']].tolist())
        self.HBGI_hist.extend(trace[:, _T['HBGI']].tolist())
