"""RandomScenario -- the surface of simglucose/simulation/scenario_gen.py:10-75.  Draw order and
distributions are those of the reference (numpy MT19937 + scipy truncnorm), implemented once in
streams._create_day; populations too large for the host use the device generator
(streams.device_scenarios, Philox, same distributions)."""
import logging
from datetime import datetime

import numpy as np

from .. import streams
from .scenario import Action, Scenario

logger = logging.getLogger(__name__)


class RandomScenario(Scenario):
    def __init__(self, start_time, seed=None):
        Scenario.__init__(self, start_time=start_time)
        self.seed = seed

    def get_action(self, t):
        """t: datetime.  A new one-day schedule is drawn whenever t is midnight (scenario_gen.py:20-22)."""
        t_sec = (t - datetime.combine(t.date(), datetime.min.time())).total_seconds()
        if t_sec < 1:
            logger.info('Creating new one day scenario ...')
            self.scenario = self.create_scenario()
        t_min = np.floor(t_sec / 60.0)
        times = self.scenario['meal']['time']
        if t_min in times:
            return Action(meal=self.scenario['meal']['amount'][times.index(t_min)])
        return Action(meal=0)

    def create_scenario(self):
        times, amounts = streams._create_day(self.random_gen)
        return {'meal': {'time': times, 'amount': amounts}}

    def reset(self):
        self.random_gen = np.random.RandomState(self.seed)
        self.scenario = self.create_scenario()

    @property
    def seed(self):
        return self._seed

    @seed.setter
    def seed(self, seed):
        self._seed = seed
        self.reset()
