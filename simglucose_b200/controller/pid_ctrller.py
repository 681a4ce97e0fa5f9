"""PIDController -- the surface of simglucose/controller/pid_ctrller.py:8-40 (fused on device as
SG_CTRL_PID by SimObj.simulate / batch_sim; policy() for callers running their own loop)."""
import logging

from .base import Action, Controller

logger = logging.getLogger(__name__)


class PIDController(Controller):
    def __init__(self, P=1, I=0, D=0, target=140):
        self.P = P
        self.I = I
        self.D = D
        self.target = target
        self.integrated_state = 0
        self.prev_state = 0

    def device_spec(self, name=None):
        if self.integrated_state != 0 or self.prev_state != 0:
            return None                       # the kernel starts from a reset controller
        return 'pid', (float(self.P), float(self.I), float(self.D), float(self.target))

    def policy(self, observation, reward, done, **kwargs):
        sample_time = kwargs.get('sample_time')
        bg = observation.CGM
        err = bg - self.target
        u = self.P * err + self.I * self.integrated_state + self.D * (bg - self.prev_state) / sample_time
        self.prev_state = bg
        self.integrated_state += err * sample_time
        return Action(basal=u, bolus=0)

    def reset(self):
        self.integrated_state = 0
        self.prev_state = 0
