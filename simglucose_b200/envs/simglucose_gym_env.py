"""Gym environment 'simglucose-v0' -- the surface of simglucose/envs/simglucose_gym_env.py:18-85.

Old-gym hooks (_step/_reset/_seed/_render) and their modern names are both provided.  The seeding chain
(gym==0.9.4 utils.seeding, restated in streams.py and pinned by the reference's tests/test_seed.py) decides
the start hour, the sensor / scenario / patient seeds; every env-step is one launch of the step kernel.
"""
from datetime import datetime

import numpy as np

from .. import streams
from ..actuator.pump import InsulinPump
from ..controller.base import Action
from ..patient.t1dpatient import T1DPatient
from ..sensor.cgm import CGMSensor
from ..simulation.env import T1DSimEnv as _T1DSimEnv
from ..simulation.scenario_gen import RandomScenario
from . import spaces

try:                                                 # pragma: no cover - gym is not in the build image
    import gym
    _Base = gym.Env
except Exception:
    gym = None
    _Base = object


class T1DSimEnv(_Base):
    """A single closed loop behind the gym API; the action is the basal rate (U/min)."""
    metadata = {'render.modes': ['human']}

    SENSOR_HARDWARE = 'Dexcom'
    INSULIN_PUMP_HARDWARE = 'Insulet'

    def __init__(self, patient_name=None, custom_scenario=None, reward_fun=None, seed=None):
        """patient_name: 'adolescent#001'..'adolescent#010', 'adult#001'..'adult#010', 'child#001'..'child#010'."""
        self.patient_name = 'adolescent#001' if patient_name is None else patient_name
        self.reward_fun = reward_fun
        self.np_random, _ = streams.np_random(seed=seed)
        self.env, _, _, _ = self._create_env_from_random_state(custom_scenario)

    # -- gym 0.9.4 hooks ---------------------------------------------------------------------------------
    def _step(self, action):
        act = Action(basal=action, bolus=0)          # this wrapper only controls basal insulin
        if self.reward_fun is None:
            return self.env.step(act)
        return self.env.step(act, reward_fun=self.reward_fun)

    def _reset(self):
        self.env, _, _, _ = self._create_env_from_random_state()
        obs, _, _, _ = self.env.reset()
        return obs

    def _seed(self, seed=None):
        self.np_random, seed1 = streams.np_random(seed=seed)
        self.env, seed2, seed3, seed4 = self._create_env_from_random_state()
        return [seed1, seed2, seed3, seed4]

    def _render(self, mode='human', close=False):
        self.env.render(close=close)

    # -- the same under their modern names -----------------------------------------------------------------
    def step(self, action):
        return self._step(action)

    def reset(self):
        return self._reset()

    def seed(self, seed=None):
        return self._seed(seed)

    def render(self, mode='human', close=False):
        return self._render(mode=mode, close=close)

    def close(self):
        pass

    def _create_env_from_random_state(self, custom_scenario=None):
        # seeds stay below 2**31 (they are checked as ints further down the chain)
        seed2 = streams.hash_seed(self.np_random.randint(0, 1000)) % 2 ** 31
        seed3 = streams.hash_seed(seed2 + 1) % 2 ** 31
        seed4 = streams.hash_seed(seed3 + 1) % 2 ** 31
        hour = self.np_random.randint(low=0.0, high=24.0)
        start_time = datetime(2018, 1, 1, hour, 0, 0)
        patient = T1DPatient.withName(self.patient_name, random_init_bg=True, seed=seed4)
        sensor = CGMSensor.withName(self.SENSOR_HARDWARE, seed=seed2)
        scenario = RandomScenario(start_time=start_time, seed=seed3) if custom_scenario is None else custom_scenario
        pump = InsulinPump.withName(self.INSULIN_PUMP_HARDWARE)
        return _T1DSimEnv(patient, sensor, pump, scenario), seed2, seed3, seed4

    @property
    def action_space(self):
        return spaces.Box(low=0, high=self.env.pump._params['max_basal'], shape=(1,))

    @property
    def observation_space(self):
        return spaces.Box(low=0, high=np.inf, shape=(1,))


def register_envs():
    """register('simglucose-v0') when a gym registry is importable (simglucose/__init__.py:3-6)."""
    if gym is None:
        return False
    from gym.envs.registration import register
    register(id='simglucose-v0', entry_point='simglucose_b200.envs:T1DSimEnv')
    return True
