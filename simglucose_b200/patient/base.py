"""Patient protocol (simglucose/patient/base.py:4-31)."""


class Patient(object):
    def step(self, action):
        """Advance the patient dynamics by one sample time under `action`."""
        raise NotImplementedError

    @staticmethod
    def model(t, state, action, params):
        """Right-hand side of the patient ODE."""
        raise NotImplementedError

    def reset(self):
        """Back to the initial state."""
        raise NotImplementedError
