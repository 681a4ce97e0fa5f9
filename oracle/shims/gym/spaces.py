import numpy as np


class Box(object):
    def __init__(self, low, high, shape=None):
        self.low = np.full(shape, low, dtype=float)
        self.high = np.full(shape, high, dtype=float)
        self.shape = shape
