"""BBController -- the surface of simglucose/controller/basal_bolus_ctrller.py:15-83.

SimObj.simulate / batch_sim recognise this class and evaluate the policy inside the episode kernel
(SG_CTRL_BB; `device_spec`).  policy() is the same formula for callers that run their own loop."""
import logging

from .. import params as P
from .base import Action, Controller

logger = logging.getLogger(__name__)


class BBController(Controller):
    """Basal-bolus therapy: basal = u2ss * BW / 6000 U/min; at a meal, bolus = carbs / CR +
    (CGM > 150) * (CGM - target) / CF units, delivered over one sample."""

    def __init__(self, target=140):
        self.target = target

    def _lookup(self, name):
        """(u2ss, BW, CR, CF) by patient name; population averages for unknown names
        (basal_bolus_ctrller.py:52-62)."""
        q = P.quest_record(name) if isinstance(name, str) else None
        if q is None:
            return 1.43, 57.0, 1 / 15, 1 / 50
        rec = P.patient_record(P.patient_row(name))
        return rec['u2ss'], rec['BW'], q['CR'], q['CF']

    def device_spec(self, name):
        """('bb', (target, CR, CF, basal)) when the kernel's BB policy applies to this patient name, else None.
        basal = u2ss * BW / 6000 of the CSV row found by NAME, as _bb_policy computes it -- not of the parameter
        Series the patient object may carry (a user-modified row keeps the cohort basal, as in the reference)."""
        if isinstance(name, str) and P.quest_record(name) is not None:
            u2ss, BW, CR, CF = self._lookup(name)
            return 'bb', (float(self.target), float(CR), float(CF), float(u2ss * BW / 6000))
        return None

    def policy(self, observation, reward, done, **kwargs):
        sample_time = kwargs.get('sample_time', 1)
        return self._bb_policy(kwargs.get('patient_name'), kwargs.get('meal'), observation.CGM, sample_time)

    def _bb_policy(self, name, meal, glucose, env_sample_time):
        u2ss, BW, CR, CF = self._lookup(name)
        basal = u2ss * BW / 6000
        if meal > 0:
            logger.info('Calculating bolus ...')
            bolus = (meal * env_sample_time) / CR + (glucose > 150) * (glucose - self.target) / CF
        else:
            bolus = 0
        return Action(basal=basal, bolus=bolus / env_sample_time)

    def reset(self):
        pass
