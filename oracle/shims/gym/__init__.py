"""TEST INFRASTRUCTURE ONLY -- minimal stand-in for gym==0.9.4 so that the
reference package (which does `from gym.envs.registration import register` in
simglucose/__init__.py:1) can be imported in the build container to generate
golden vectors.  Never imported by the product package.

The seeding functions restate gym 0.9.4's `gym.utils.seeding` (third-party,
not under /root/reference; pinned by the reference's setup.py:14); the
restatement is pinned by tests/test_seed.py:19,23 of the reference (start
hours 23 and 14 for seeds 0 and 1000).
"""
from . import spaces  # noqa: F401
from .envs import registration as _registration


class Env(object):
    metadata = {}

    def step(self, action):
        return self._step(action)

    def reset(self):
        return self._reset()

    def seed(self, seed=None):
        return self._seed(seed)

    def render(self, mode='human', close=False):
        return self._render(mode=mode, close=close)


def make(id, **kw):
    return _registration.make(id, **kw)
