from .simglucose_gym_env import T1DSimEnv  # noqa: F401
from .vector_env import T1DSimVecEnv  # noqa: F401
