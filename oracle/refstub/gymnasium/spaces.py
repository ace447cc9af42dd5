from __future__ import annotations

from typing import Generic, TypeVar

import numpy as np

from .utils import seeding

T = TypeVar("T")


class Space(Generic[T]):
    def __init__(self, shape=None, dtype=None, seed=None):
        self._shape = None if shape is None else tuple(shape)
        self.dtype = dtype
        self._np_random = None
        if seed is not None:
            self.seed(seed)

    @property
    def shape(self):
        return self._shape

    @property
    def np_random(self):
        if self._np_random is None:
            self.seed()
        return self._np_random

    def seed(self, seed=None):
        self._np_random, seed = seeding.np_random(seed)
        return [seed]

    def sample(self):
        raise NotImplementedError

    def contains(self, x):
        raise NotImplementedError

    def __contains__(self, x):
        return self.contains(x)


class Discrete(Space):
    def __init__(self, n, seed=None, start=0):
        self.n = int(n)
        self.start = int(start)
        super().__init__((), np.int64, seed)

    def sample(self):
        return int(self.start + self.np_random.integers(self.n))

    def contains(self, x):
        try:
            return self.start <= int(x) < self.start + self.n
        except Exception:
            return False


class Box(Space):
    def __init__(self, low, high, shape=None, dtype=np.float32, seed=None):
        self.low, self.high = low, high
        super().__init__(shape, dtype, seed)

    def contains(self, x):
        x = np.asarray(x)
        return x.shape == self.shape and bool(np.all(x >= self.low)) and bool(np.all(x <= self.high))


class Dict(Space):
    def __init__(self, spaces=None, seed=None):
        self.spaces = dict(spaces or {})
        super().__init__(None, None, seed)

    def __getitem__(self, k):
        return self.spaces[k]

    def contains(self, x):
        return isinstance(x, dict) and all(k in x and self.spaces[k].contains(x[k]) for k in self.spaces)


class MultiDiscrete(Space):
    def __init__(self, nvec, dtype=np.int64, seed=None):
        self.nvec = np.asarray(nvec, dtype=dtype)
        super().__init__(self.nvec.shape, dtype, seed)
