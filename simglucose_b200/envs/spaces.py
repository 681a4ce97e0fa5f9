"""Box space: gym's (or gymnasium's) when one is installed, else a minimal stand-in with the same
attributes the reference's wrapper exposes (low, high, shape, sample, contains)."""
import numpy as np

try:                                                 # pragma: no cover - neither is in the build image
    from gym.spaces import Box
except Exception:
    try:
        from gymnasium.spaces import Box
    except Exception:
        class Box(object):
            def __init__(self, low, high, shape=None, dtype=np.float32):
                self.shape = tuple(shape) if shape is not None else np.shape(low)
                self.dtype = np.dtype(dtype)
                self.low = np.full(self.shape, low, dtype=self.dtype)
                self.high = np.full(self.shape, high, dtype=self.dtype)
                self._rng = np.random.RandomState()

            def seed(self, seed=None):
                self._rng = np.random.RandomState(seed)
                return [seed]

            def sample(self):
                hi = np.where(np.isfinite(self.high), self.high, self.low + 1e6)
                return self._rng.uniform(self.low, hi).astype(self.dtype)

            def contains(self, x):
                x = np.asarray(x)
                return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

            def __repr__(self):
                return 'Box(%s, %s, %s)' % (self.low.min(), self.high.max(), self.shape)
