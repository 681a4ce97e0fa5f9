"""CGMNoise -- iterator with the stream of simglucose/sensor/noise_gen.py:15-69 for a given seed.

The stream itself (AR(1) innovations -> Johnson SU -> cubic interpolation from 15-minute knots to the
sensor grid) is produced array-at-a-time by streams.cgm_noise_stream; this class only hands it out
sample by sample.  A fused T1DSimEnv uploads the same array and the step kernel indexes it."""
import numpy as np

from .. import streams


def johnson_transform_SU(xi, lam, gamma, delta, x):
    return xi + lam * np.sinh((x - gamma) / delta)


class CGMNoise(object):
    PRECOMPUTE = streams.PRECOMPUTE
    MDL_SAMPLE_TIME = streams.MDL_SAMPLE_TIME

    def __init__(self, params, n=np.inf, seed=None):
        self._params = params
        self.seed = seed
        # seed=None means "fresh entropy" in the reference (RandomState(None)); fix one here so that the
        # stream can be regenerated at a longer length with the same prefix
        self._stream_seed = int(np.random.randint(0, 2 ** 31 - 1)) if seed is None else seed
        self.n = n
        self.count = 0
        self._buf = np.zeros(0)

    def peek(self, n):
        """First n samples of the stream as an array (does not advance the iterator)."""
        if self._buf.size < n:
            self._buf = streams.cgm_noise_stream(dict(self._params), self._stream_seed, max(int(n), 2 * self._buf.size, 64))
        return self._buf[:int(n)]

    def __iter__(self):
        return self

    def __next__(self):
        if self.count >= self.n:
            raise StopIteration()
        v = float(self.peek(self.count + 1)[self.count])
        self.count += 1
        return v
