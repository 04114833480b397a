"""gym.spaces.Discrete / Box as far as the reference's env touches them (n, shape, sample, seed)."""
import numpy as np


class Discrete:
    def __init__(self, n):
        self.n = int(n)
        self._rng = np.random.default_rng()

    def seed(self, seed=None):
        self._rng = np.random.default_rng(seed)
        return [seed]

    def sample(self):
        return int(self._rng.integers(self.n))


class Box:
    def __init__(self, low, high, shape=None, dtype=float):
        self.low, self.high, self.shape, self.dtype = low, high, tuple(shape), dtype
