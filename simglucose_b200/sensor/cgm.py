"""CGMSensor -- the surface of simglucose/sensor/cgm.py:12-50.

Inside a fused T1DSimEnv step the sensor is part of the step kernel (noise add, clip, zero-order hold);
measure() below serves callers that drive a patient themselves, with the arithmetic on the GPU
(sg_cgm_measure)."""
import logging

import pandas as pd

from .. import ops
from .. import params as P
from .noise_gen import CGMNoise

logger = logging.getLogger(__name__)


class CGMSensor(object):
    def __init__(self, params, seed=None):
        self._params = params
        self.name = params['Name']
        self.sample_time = params['sample_time']
        self.seed = seed
        self._last_CGM = 0

    @classmethod
    def withName(cls, name, **kwargs):
        return cls(pd.Series(P.sensor_params(name)), **kwargs)

    def measure(self, patient):
        if patient.t % self.sample_time == 0:
            BG = patient.observation.Gsub
            CGM = float(ops.cgm_measure(BG, next(self._noise_generator), self._params['min'],
                                        self._params['max']).cpu()[0])
            self._last_CGM = CGM
            return CGM
        return self._last_CGM          # zero-order hold

    @property
    def seed(self):
        return self._seed

    @seed.setter
    def seed(self, seed):
        self._seed = seed
        self._noise_generator = CGMNoise(self._params, seed=seed)

    def reset(self):
        logger.debug('Resetting CGM sensor ...')
        self._noise_generator = CGMNoise(self._params, seed=self.seed)
        self._last_CGM = 0
