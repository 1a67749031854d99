"""Input formatting with the semantics of agilerl/components/data.py (Transition :68-93,
ReplayDataset :96-116) without depending on ``tensordict``'s ``@tensorclass``."""

from __future__ import annotations

import warnings
from collections import OrderedDict
from collections.abc import Iterator
from numbers import Number

import numpy as np
import torch
from torch.utils.data import IterableDataset

from .tensor_batch import TensorBatch


def to_tensordict(data, dtype: torch.dtype = torch.float32) -> TensorBatch:
    """tuple / dict observation -> nested batch (data.py:15-45)."""
    if isinstance(data, tuple):
        assert all(isinstance(el, (torch.Tensor, np.ndarray, Number)) for el in data), (
            "Expected all elements of the tuple to be torch.Tensor or np.ndarray.")
        data = OrderedDict((f"tuple_obs_{i}", el) for i, el in enumerate(data))
    elif isinstance(data, dict):
        assert all(isinstance(el, (torch.Tensor, np.ndarray, Number)) for el in data.values()), (
            "Expected all values of the dict to be torch.Tensor or np.ndarray.")
    return TensorBatch({k: torch.as_tensor(v).to(dtype) for k, v in data.items()})


def to_torch_tensor(data, dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """data.py:48-65."""
    if isinstance(data, torch.Tensor):
        return data.to(dtype=dtype)
    return torch.tensor(np.asarray(data) if isinstance(data, (list, tuple)) else data, dtype=dtype)


class Transition:
    """obs / action / next_obs / reward / done with the reference's casts: action, reward, done ->
    fp32; 0-d reward/done -> (1,); dict/tuple observations -> nested batch (data.py:68-93)."""

    def __init__(self, obs, action, next_obs, reward, done, batch_size=None, device=None) -> None:
        if isinstance(obs, (dict, tuple)):
            obs = to_tensordict(obs)
        if isinstance(next_obs, (dict, tuple)):
            next_obs = to_tensordict(next_obs)
        self.obs = obs if isinstance(obs, TensorBatch) else torch.as_tensor(obs)
        self.next_obs = next_obs if isinstance(next_obs, TensorBatch) else torch.as_tensor(next_obs)
        self.action = to_torch_tensor(action)
        self.done = to_torch_tensor(done)
        self.reward = to_torch_tensor(reward)
        if self.done.ndim == 0:
            self.done = self.done.unsqueeze(-1)
        if self.reward.ndim == 0:
            self.reward = self.reward.unsqueeze(-1)
        self.batch_size = [] if batch_size is None else list(batch_size)
        self.device = device

    def to_tensordict(self) -> TensorBatch:
        bs = self.batch_size
        out = TensorBatch({}, batch_size=bs)
        for k in ("obs", "action", "next_obs", "reward", "done"):
            v = getattr(self, k)
            if isinstance(v, TensorBatch):
                v = TensorBatch(dict(v.items()), batch_size=bs)
            out[k] = v
        return out.to(self.device) if self.device is not None else out

    def unsqueeze(self, dim: int) -> "Transition":
        td = self.to_tensordict().unsqueeze(dim)
        t = Transition.__new__(Transition)
        for k in ("obs", "action", "next_obs", "reward", "done"):
            setattr(t, k, td[k])
        t.batch_size, t.device = list(td.batch_size), self.device
        return t


class ReplayDataset(IterableDataset):
    """data.py:96-116."""

    def __init__(self, buffer, batch_size: int = 256) -> None:
        from .replay_buffer import ReplayBuffer

        if not isinstance(buffer, ReplayBuffer):
            warnings.warn("Buffer is not an agilerl ReplayBuffer.", stacklevel=2)
        assert batch_size > 0, "Batch size must be greater than zero."
        self.buffer = buffer
        self.batch_size = batch_size

    def __iter__(self) -> Iterator:
        yield self.buffer.sample(self.batch_size)
