"""T1DPatient -- the surface of simglucose/patient/t1dpatient.py:18-281 on top of the CUDA library.

One patient object owns one C-ABI handle holding a single resident episode.  step() is one launch of
the patient-step kernel (meal state machine + one minute of DOPRI5, sg_patient_step); model() is one
launch of the RHS kernel (sg_model).  A T1DSimEnv built from this patient, a CGMSensor and an
InsulinPump drives the same handle through the fused env-step kernel instead (simulation/env.py).
There is no CPU integrator: without the CUDA library / a GPU the constructor raises.
"""
import logging
from collections import namedtuple

import numpy as np
import torch

from .. import _capi
from .. import params as P
from ..engine import BatchEngine
from .base import Patient

logger = logging.getLogger(__name__)

Action = namedtuple("patient_action", ['CHO', 'insulin'])
Observation = namedtuple("observation", ['Gsub'])

_D_PLANNED, _D_LAST_QSTO, _D_LAST_FOOD = 13, 15, 16     # include/simglucose_b200.h sg_state_d
_I_EATING, _I_MINUTE, _I_STATUS = 0, 1, 4               # sg_state_i


def _hot_row(params):
    """The 32 hot-path parameters of a patient record, in C-ABI order."""
    return np.array([float(params[c]) for c in P.HOT_PARAMS], dtype=np.float64)


class T1DPatient(Patient):
    SAMPLE_TIME = 1  # min
    EAT_RATE = 5  # g/min CHO

    def __init__(self, params, init_state=None, random_init_bg=False, seed=None, t0=0, device=0,
                 carry_h=False, fma=False):
        """params: one row of vpatient_params.csv (pandas Series or mapping).  init_state: 13-vector, default
        params.iloc[2:15].  t0: start minute (integral).  carry_h / fma: sg_create flags."""
        self._params = params
        if init_state is not None and not (isinstance(init_state, np.ndarray) and init_state.dtype == np.float64):
            init_state = np.array(init_state, dtype=np.float64)       # one persistent, mutable 13-vector
        self._init_state = init_state
        self.random_init_bg = random_init_bg
        self._seed = seed
        if int(t0) != t0 or t0 < 0:
            raise ValueError('t0 must be a non-negative whole number of minutes')
        self.t0 = int(t0)
        self._engine = BatchEngine(1, device=device, carry_h=carry_h, fma=fma)
        x0 = np.array([float(v) for v in list(params.values())[2:15]] if isinstance(params, dict)
                      else np.asarray(params.iloc[2:15], dtype=np.float64))
        # The reference mutates init_state in place when random_init_bg is set and, with the pinned pandas,
        # params.iloc[2:15] is a view: successive resets random-walk around the previous draw
        # (t1dpatient.py:252,268-270; the gym wrapper relies on it by resetting twice).  One persistent
        # array per patient reproduces that.
        self._x0_default = x0.copy()
        self._engine.load_table(_hot_row(params)[None, :], x0[None, :])
        self._row = torch.zeros(1, dtype=torch.int32, device=self._engine.device)
        self._dummy_noise = torch.zeros(2, 1, dtype=torch.float64, device=self._engine.device)
        self.reset()

    @classmethod
    def withID(cls, patient_id, **kwargs):
        """patient_id 1..30: adolescent#001-010, adult#001-010, child#001-010."""
        return cls(P.patient_series(int(patient_id) - 1), **kwargs)

    @classmethod
    def withName(cls, name, **kwargs):
        return cls(P.patient_series(P.patient_row(name)), **kwargs)

    # -- state mirrors (host copies of the resident episode) -----------------------------------------
    @property
    def state(self):
        return self._x.copy()

    @property
    def t(self):
        return float(self._minute)

    @property
    def sample_time(self):
        return self.SAMPLE_TIME

    @property
    def observation(self):
        """Subcutaneous glucose in mg/dL: x[12] / Vg (t1dpatient.py:210-220), as computed by the device."""
        return Observation(Gsub=self._gsub)

    def _pull(self):
        """Refresh the host mirrors from the resident episode (after the env stepped it)."""
        d, i = self._engine.episode_state()
        d, i = d.cpu().numpy()[:, 0], i.cpu().numpy()[:, 0]
        self._x = d[:13].copy()
        self.planned_meal = float(d[_D_PLANNED])
        self._last_Qsto = float(d[_D_LAST_QSTO])
        self._last_foodtaken = float(d[_D_LAST_FOOD])
        self._last_action = Action(CHO=float(d[14]), insulin=self._last_action.insulin)
        self.is_eating = bool(i[_I_EATING])
        self._minute = int(i[_I_MINUTE])
        self._status = int(i[_I_STATUS])
        # mirror of the device's x[12] / Vg (one correctly rounded IEEE division: same bits on both sides)
        self._gsub = float(self._x[12] / float(self._params['Vg']))
        return d, i

    def step(self, action):
        """One simulated minute under Action(CHO = grams announced now, insulin = U/min)."""
        if self._status < 0:
            # the reference notices a failed integration at the NEXT step (t1dpatient.py:112-116)
            logger.error('ODE solver failed!!')
            raise RuntimeError('ODE solver failed (dopri5 idid = %d)' % self._status)
        out = self._engine.patient_step(torch.tensor([float(action.CHO)], dtype=torch.float64),
                                        torch.tensor([float(action.insulin)], dtype=torch.float64))
        o = out.cpu().numpy()[:, 0]
        self._x = o[:13].copy()
        self._gsub = float(o[13])
        self._last_action = Action(CHO=float(o[14]), insulin=action.insulin)
        if o[15] < 0:
            self._status = int(o[15])
        self.planned_meal = float(o[17])
        self._last_Qsto = float(o[18])
        self._last_foodtaken = float(o[19])
        self.is_eating = bool(o[20])
        self._minute += 1

    _static_engine = None

    @staticmethod
    def model(t, x, action, params, last_Qsto, last_foodtaken):
        """dx/dt of the UVa/Padova-2008 model (t1dpatient.py:118-208), evaluated by the RHS kernel."""
        if T1DPatient._static_engine is None:
            T1DPatient._static_engine = BatchEngine(1)
        eng = T1DPatient._static_engine
        dx = eng.model(np.asarray(x, dtype=np.float64).reshape(13, 1), float(action.CHO), float(action.insulin),
                       float(last_Qsto), float(last_foodtaken), params=_hot_row(params).reshape(-1, 1))
        return dx.cpu().numpy()[:, 0]

    @property
    def seed(self):
        return self._seed

    @seed.setter
    def seed(self, seed):
        self._seed = seed
        self.reset()

    def reset(self):
        """Back to the initial state (t1dpatient.py:247-281)."""
        if self._init_state is None:
            self.init_state = self._x0_default
        else:
            self.init_state = self._init_state
        self.random_state = np.random.RandomState(self.seed)
        if self.random_init_bg:
            # only the glucose states x4, x5, x13 are randomised: N(mean, diag(0.1 * mean))
            mean = [1.0 * self.init_state[3], 1.0 * self.init_state[4], 1.0 * self.init_state[12]]
            cov = np.diag([0.1 * self.init_state[3], 0.1 * self.init_state[4], 0.1 * self.init_state[12]])
            bg_init = self.random_state.multivariate_normal(mean, cov)
            self.init_state[3] = 1.0 * bg_init[0]
            self.init_state[4] = 1.0 * bg_init[1]
            self.init_state[12] = 1.0 * bg_init[2]
        self.name = self._params['Name']
        self._last_action = Action(CHO=0, insulin=0)
        self._engine.set_t0(self.t0)
        x0 = np.asarray(self.init_state, dtype=np.float64).reshape(13, 1)
        self._engine.reset(self._row, self._dummy_noise, x0=x0, row0=False)
        self._status = 1
        self._pull()
