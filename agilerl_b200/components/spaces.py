"""Shape helpers for observation / action spaces without importing gymnasium.

Restates ``get_obs_shape`` / ``get_num_actions`` (agilerl/utils/algo_utils.py:307-350) by
duck-typing on the space's class name, so real ``gymnasium.spaces`` objects work when gymnasium
is installed, and provides minimal stand-in classes for when it is not (the B200 image).
"""

from __future__ import annotations

from collections import OrderedDict

import numpy as np


class Space:
    shape: tuple | None = None


class Box(Space):
    def __init__(self, low=-np.inf, high=np.inf, shape=None, dtype=np.float32):
        if shape is None:
            shape = np.shape(low)
        self.low, self.high, self.shape, self.dtype = low, high, tuple(shape), dtype


class Discrete(Space):
    def __init__(self, n: int, start: int = 0):
        self.n, self.start, self.shape, self.dtype = int(n), start, (), np.int64


class MultiDiscrete(Space):
    def __init__(self, nvec):
        self.nvec = np.asarray(nvec, dtype=np.int64)
        self.shape, self.dtype = self.nvec.shape, np.int64


class MultiBinary(Space):
    def __init__(self, n):
        self.n = n
        self.shape = (n,) if np.isscalar(n) else tuple(n)
        self.dtype = np.int8


class Dict(Space):
    def __init__(self, spaces=None, **kw):
        self.spaces = OrderedDict(spaces or {})
        self.spaces.update(kw)

    def __getitem__(self, k):
        return self.spaces[k]

    def keys(self):
        return self.spaces.keys()

    def items(self):
        return self.spaces.items()


class Tuple(Space):
    def __init__(self, spaces):
        self.spaces = tuple(spaces)

    def __iter__(self):
        return iter(self.spaces)

    def __getitem__(self, i):
        return self.spaces[i]


def _kind(space) -> str:
    return type(space).__name__


def is_dict_space(space) -> bool:
    return _kind(space) == "Dict"


def get_obs_shape(space):
    """algo_utils.py:307-330."""
    k = _kind(space)
    if k == "Box":
        return tuple(space.shape)
    if k == "Discrete":
        return (1,)
    if k == "MultiDiscrete":
        return (len(space.nvec),)
    if k == "MultiBinary":
        return tuple(space.shape)
    if k == "Dict":
        return {key: get_obs_shape(sub) for key, sub in space.spaces.items()}
    if k == "Tuple":
        return tuple(get_obs_shape(sub) for sub in space)
    raise NotImplementedError(f"{space} observation space is not supported")


def get_num_actions(space) -> int:
    """algo_utils.py:333-350."""
    k = _kind(space)
    if k == "Box":
        return int(np.prod(space.shape))  # spaces.flatdim(Box)
    if k == "Discrete":
        return 1
    if k == "MultiDiscrete":
        return len(space.nvec)
    if k == "MultiBinary":
        return int(np.prod(space.shape)) if not np.isscalar(space.n) else int(space.n)
    raise NotImplementedError(f"{space} action space is not supported by AgileRL.")
