import numpy as np


class Box(object):
    def __init__(self, low, high, shape=None, dtype=np.float32):
        self.low = np.asarray(low, dtype=dtype)
        self.high = np.asarray(high, dtype=dtype)
        self.shape = self.low.shape
        self.dtype = dtype
        self.np_random = np.random.RandomState(0)

    def sample(self):
        return self.np_random.uniform(self.low, self.high).astype(self.dtype)
