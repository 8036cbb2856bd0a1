"""``Box`` and ``Dict`` with the attributes the reference reads (``shape``, ``low``, ``high``, ``dtype``,
``sample()``, ``seed()``; reference src/rlsp/envs/environment_limits.py:54-90, src/rlsp/agents/models.py:124-125)."""
from collections import OrderedDict

import numpy as np

from .utils import seeding


class Space:
    def __init__(self, shape=None, dtype=None, seed=None):
        self._shape = None if shape is None else tuple(shape)
        self.dtype = None if dtype is None else np.dtype(dtype)
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


class Box(Space):
    def __init__(self, low, high, shape=None, dtype=np.float32, seed=None):
        dtype = np.dtype(dtype)
        if shape is None:
            shape = np.shape(low) if np.ndim(low) else np.shape(high)
        shape = tuple(int(d) for d in shape)
        super().__init__(shape, dtype, seed)
        self.low = np.full(shape, low, dtype=dtype) if np.isscalar(low) else np.asarray(low, dtype=dtype)
        self.high = np.full(shape, high, dtype=dtype) if np.isscalar(high) else np.asarray(high, dtype=dtype)
        self.bounded_below = -np.inf < self.low
        self.bounded_above = np.inf > self.high

    def is_bounded(self, manner="both"):
        return bool(np.all(self.bounded_below) and np.all(self.bounded_above))

    def sample(self):
        # gym 0.26.2: bounded dimensions are drawn uniformly from [low, high) for float dtypes
        high = self.high if self.dtype.kind == "f" else self.high.astype("int64") + 1
        if self.dtype.kind == "f":
            out = self.np_random.uniform(low=self.low, high=high, size=self.shape)
        else:
            out = self.np_random.integers(low=self.low, high=high, size=self.shape)
        return out.astype(self.dtype)

    def contains(self, x):
        x = np.asarray(x)
        return bool(x.shape == self.shape and np.all(x >= self.low) and np.all(x <= self.high))

    def __repr__(self):
        return f"Box({self.low.min()}, {self.high.max()}, {self.shape}, {self.dtype})"


class Dict(Space):
    def __init__(self, spaces=None, seed=None, **kw):
        spaces = OrderedDict(spaces or {})
        spaces.update(kw)
        self.spaces = spaces
        super().__init__(None, None, seed)

    def seed(self, seed=None):
        out = super().seed(seed)
        for s in self.spaces.values():
            out += s.seed(None if seed is None else int(self.np_random.integers(2 ** 31)))
        return out

    def sample(self):
        return OrderedDict((k, s.sample()) for k, s in self.spaces.items())

    def contains(self, x):
        return isinstance(x, dict) and all(k in x and s.contains(x[k]) for k, s in self.spaces.items())

    def __getitem__(self, key):
        return self.spaces[key]

    def __iter__(self):
        return iter(self.spaces)

    def __len__(self):
        return len(self.spaces)

    def keys(self):
        return self.spaces.keys()

    def items(self):
        return self.spaces.items()

    def values(self):
        return self.spaces.values()
