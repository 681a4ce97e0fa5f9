"""Scenario / CustomScenario -- the surface of simglucose/simulation/scenario.py:7-59 (host-side: a
scenario is a lookup from wall-clock minute to announced grams; the kernels consume its table)."""
import logging
from collections import namedtuple
from datetime import datetime, timedelta

logger = logging.getLogger(__name__)
Action = namedtuple('scenario_action', ['meal'])


class Scenario(object):
    def __init__(self, start_time):
        self.start_time = start_time

    def get_action(self, t):
        raise NotImplementedError

    def reset(self):
        raise NotImplementedError


def parseTime(time, start_time):
    """A number is HOURS AFTER start_time, a timedelta an offset from it (both rounded to the minute),
    a datetime is absolute (scenario.py:48-59)."""
    if isinstance(time, (int, float)):
        return start_time + timedelta(minutes=round(time * 60.0))
    if isinstance(time, timedelta):
        return start_time + timedelta(minutes=round(time.total_seconds() / 60.0))
    if isinstance(time, datetime):
        return time
    raise ValueError('Expect time to be int, float, timedelta, datetime')


class CustomScenario(Scenario):
    def __init__(self, start_time, scenario):
        """scenario: list of (time, grams) with time a datetime, a timedelta or a number of hours."""
        Scenario.__init__(self, start_time=start_time)
        self.scenario = scenario

    def get_action(self, t):
        for when, grams in (self.scenario or []):         # first match wins, like list.index
            if parseTime(when, self.start_time) == t:
                return Action(meal=grams)
        return Action(meal=0)

    def meal_table(self, first_minute, n_minutes):
        """get_action for the minutes first_minute .. first_minute + n_minutes - 1 after start_time as (minutes, grams)
        of the non-zero meals (first matching entry wins, like the loop above)."""
        import numpy as np
        first = {}
        for when, grams in (self.scenario or []):
            sec = (parseTime(when, self.start_time) - self.start_time).total_seconds()
            if sec == int(sec) and int(sec) % 60 == 0:         # get_action is asked on whole minutes after start_time
                first.setdefault(int(sec) // 60, grams)
        lo, hi = int(first_minute), int(first_minute) + int(n_minutes)
        keep = sorted(m for m, g in first.items() if lo <= m < hi and g != 0)
        return np.array(keep, dtype=np.int32), np.array([float(first[m]) for m in keep], dtype=np.float64)

    def reset(self):
        pass
