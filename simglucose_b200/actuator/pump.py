"""InsulinPump -- the surface of simglucose/actuator/pump.py:14-43; the quantiser runs on the GPU
(sg_pump), and inside a fused T1DSimEnv step it is part of the step kernel."""
import logging

import pandas as pd

from .. import ops
from .. import params as P

logger = logging.getLogger(__name__)


class InsulinPump(object):
    U2PMOL = 6000

    def __init__(self, params):
        self._params = params

    @classmethod
    def withName(cls, name):
        return cls(pd.Series(P.pump_params(name)))

    def _quantise(self, amount, kind):
        p = self._params
        q = ops.pump_quantise(amount, p['min_' + kind], p['max_' + kind], p['inc_' + kind])
        return float(q.cpu()[0])

    def bolus(self, amount):
        """U/min -> pmol/min, rounded to inc_bolus (half to even), back to U/min, clipped (pump.py:23-30)."""
        return self._quantise(amount, 'bolus')

    def basal(self, amount):
        """Same for the basal rate (pump.py:32-39)."""
        return self._quantise(amount, 'basal')

    def reset(self):
        logger.info('Resetting insulin pump ...')
