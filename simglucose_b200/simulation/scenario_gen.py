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

    def meal_table(self, first_minute, n_minutes):
        """What get_action returns for the minutes first_minute .. first_minute + n_minutes - 1 after start_time, as
        (minutes [k] int32, grams [k]) of the non-zero meals -- the same draws in the same order as calling get_action
        once per minute (one create_scenario per midnight passed), one loop turn per calendar day instead of per minute.
        None if start_time is not on a whole minute (the per-minute replay handles that)."""
        if self.start_time.second or self.start_time.microsecond:
            return None
        from datetime import timedelta
        mm, gg = [], []
        m, end = int(first_minute), int(first_minute) + int(n_minutes)
        while m < end:
            t = self.start_time + timedelta(minutes=m)
            tod = t.hour * 60 + t.minute
            if tod == 0:
                logger.info('Creating new one day scenario ...')
                self.scenario = self.create_scenario()
            hi = min(end, m + 1440 - tod)
            first = {}
            for tm, amount in zip(self.scenario['meal']['time'], self.scenario['meal']['amount']):
                first.setdefault(float(tm), amount)           # times.index(t_min): the first entry wins
            for tm in sorted(first):
                if tod <= tm < tod + (hi - m) and tm == np.floor(tm) and first[tm] != 0:
                    mm.append(m + int(tm) - tod)
                    gg.append(float(first[tm]))
            m = hi
        return np.array(mm, dtype=np.int32), np.array(gg, dtype=np.float64)

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
