"""TensorBatch — the minimal TensorDict-compatible container the experience path hands out.

The reference returns ``tensordict.TensorDict`` objects (agilerl/components/replay_buffer.py:126,
rollout_buffer.py:229,577).  ``tensordict`` is not installable on the B200 image, so the engine
returns this mapping, which supports what the reference's callers do with a batch: ``[key]``,
``.get``, ``.keys()/.items()/.values()``, ``in``, ``is_empty()``, row indexing by int / slice /
tensor / ndarray, ``.to()``, ``.clone()``, ``.shape`` / ``.batch_size``, nesting (Dict spaces).
When the real ``tensordict`` is importable, ``to_tensordict()`` converts.
"""

from __future__ import annotations

from collections import OrderedDict
from typing import Any, Iterator, Mapping

import numpy as np
import torch


def _is_batch(x) -> bool:
    return isinstance(x, TensorBatch)


_RINGS: dict = {}  # device -> PinnedRing for index vectors


class TensorBatch:
    __slots__ = ("_d", "batch_size", "_non_tensor")

    def __init__(self, source: Mapping[str, Any] | None = None, batch_size=None, device=None):
        self._d: "OrderedDict[str, Any]" = OrderedDict()
        self._non_tensor: dict[str, Any] = {}
        if batch_size is None:
            batch_size = []
        if isinstance(batch_size, int):
            batch_size = [batch_size]
        self.batch_size = torch.Size(batch_size)
        for k, v in (source or {}).items():
            self[k] = v
        if device is not None:
            moved = self.to(device)
            self._d = moved._d

    # -- conversion --------------------------------------------------------------------------
    @staticmethod
    def _wrap(v, batch_size):
        if isinstance(v, torch.Tensor) or isinstance(v, TensorBatch):  # the common cases first
            return v
        if isinstance(v, Mapping):
            return TensorBatch(v, batch_size=batch_size)
        return torch.as_tensor(v)

    # -- mapping -----------------------------------------------------------------------------
    @property
    def shape(self) -> torch.Size:
        return self.batch_size

    @property
    def device(self):
        for v in self._d.values():
            return v.device
        return None

    def keys(self):
        return self._d.keys()

    def values(self):
        return self._d.values()

    def items(self):
        return self._d.items()

    def __iter__(self) -> Iterator[str]:
        return iter(self._d)

    def __contains__(self, k) -> bool:
        return k in self._d

    def __len__(self) -> int:
        return self.batch_size[0] if len(self.batch_size) else 0

    def get(self, k, default=None):
        return self._d.get(k, default)

    def is_empty(self) -> bool:
        return len(self._d) == 0

    def set_non_tensor(self, k, v) -> None:
        self._non_tensor[k] = v

    def get_non_tensor(self, k, default=None):
        return self._non_tensor.get(k, default)

    def select(self, *keys) -> "TensorBatch":
        out = TensorBatch({}, batch_size=self.batch_size)
        for k in keys:
            if k in self._d:
                out._d[k] = self._d[k]
        return out

    def exclude(self, *keys) -> "TensorBatch":
        out = TensorBatch({}, batch_size=self.batch_size)
        for k, v in self._d.items():
            if k not in keys:
                out._d[k] = v
        out._non_tensor = dict(self._non_tensor)
        return out

    def _map(self, fn, batch_size=None) -> "TensorBatch":
        out = TensorBatch({}, batch_size=self.batch_size if batch_size is None else batch_size)
        for k, v in self._d.items():
            out._d[k] = v._map(fn, batch_size) if _is_batch(v) else fn(v)
        out._non_tensor = dict(self._non_tensor)
        return out

    def _index_batch_size(self, idx) -> torch.Size:
        probe = torch.empty(tuple(self.batch_size) + (0,), device="meta")
        if isinstance(idx, torch.Tensor) and idx.device.type != "meta":
            idx = torch.empty(idx.shape, dtype=idx.dtype, device="meta") if idx.dtype != torch.bool else idx
        if isinstance(idx, np.ndarray):
            idx = torch.empty(idx.shape, dtype=torch.int64, device="meta")
        try:
            return probe[idx].shape[:-1]
        except Exception:  # boolean masks etc.: fall back to a real probe on CPU
            real = torch.empty(tuple(self.batch_size) + (0,))
            return real[idx].shape[:-1]

    def _leaf_items(self, prefix=()):
        for k, v in self._d.items():
            if _is_batch(v):
                yield from v._leaf_items(prefix + (k,))
            else:
                yield prefix + (k,), v

    def _gather_rows(self, idx: torch.Tensor):
        """``batch[idx]`` for a host integer index vector over a CUDA-resident batch: every key in
        ONE ``xp_gather`` launch (the reference's PPO minibatch slicing, algorithms/ppo.py:846-848,
        indexes each key separately).  Returns None when the fast path does not apply."""
        if len(self.batch_size) != 1 or idx.dim() != 1 or idx.is_cuda or idx.dtype not in (torch.int64, torch.int32):
            return None
        n = self.batch_size[0]
        leaves = list(self._leaf_items())
        if not leaves or n == 0:
            return None
        dev = leaves[0][1].device
        if dev.type != "cuda":
            return None
        for _, t in leaves:
            if t.device != dev or not t.is_contiguous() or t.dim() < 1 or t.shape[0] != n:
                return None
        m = idx.numel()
        if m == 0:
            return None
        lo, hi = int(idx.min()), int(idx.max())
        if hi >= n or lo < -n:  # what CPU indexing raises; a device gather would only assert
            raise IndexError(f"index {hi if hi >= n else lo} is out of bounds for dimension 0 with size {n}")
        if lo < 0:
            return None  # negative (from-the-end) indices: torch's own indexing handles them
        from .. import _native as nv
        from .. import xp
        from .staging import PinnedRing

        ring = _RINGS.get(dev)
        if ring is None:
            ring = _RINGS[dev] = PinnedRing(dev)
        with nv.on_device(dev):
            idx_dev = ring.to_device(idx if idx.dtype == torch.int64 else idx.to(torch.int64))
            outs = [torch.empty((m, *t.shape[1:]), dtype=t.dtype, device=dev) for _, t in leaves]
            for k in range(0, len(leaves), 24):
                xp.gather([(t, o) for (_, t), o in zip(leaves[k:k + 24], outs[k:k + 24])], idx_dev, n)
        root = TensorBatch({}, batch_size=[m])
        root._non_tensor = dict(self._non_tensor)
        for (path, _), o in zip(leaves, outs):
            node = root
            for key in path[:-1]:
                if key not in node._d:
                    node._d[key] = TensorBatch({}, batch_size=[m])
                node = node._d[key]
            node._d[path[-1]] = o
        return root

    def __getitem__(self, idx):
        if isinstance(idx, str):
            if idx in self._d:
                return self._d[idx]
            if idx in self._non_tensor:
                return self._non_tensor[idx]
            raise KeyError(idx)
        if isinstance(idx, np.ndarray):
            idx = torch.from_numpy(idx)
        if isinstance(idx, torch.Tensor):
            fast = self._gather_rows(idx)
            if fast is not None:
                return fast
        new_bs = self._index_batch_size(idx)

        def take(t):
            i = idx.to(t.device) if isinstance(idx, torch.Tensor) and idx.device != t.device else idx
            return t[i]

        return self._map(take, new_bs)

    def __setitem__(self, idx, value) -> None:
        if isinstance(idx, str):
            self._d[idx] = self._wrap(value, self.batch_size)
            return
        if isinstance(idx, np.ndarray):
            idx = torch.from_numpy(idx)
        for k, v in self._d.items():
            if isinstance(value, (TensorBatch, Mapping)):
                if k not in value:
                    continue
                src = value[k]
            else:
                src = value
            if _is_batch(v):
                v[idx] = src
            else:
                v[idx] = torch.as_tensor(src).to(v.device) if not isinstance(src, torch.Tensor) else src.to(v.device)

    # -- tensor-like ---------------------------------------------------------------------------
    def to(self, *args, **kwargs) -> "TensorBatch":
        return self._map(lambda t: t.to(*args, **kwargs))

    def cpu(self) -> "TensorBatch":
        return self.to("cpu")

    def clone(self) -> "TensorBatch":
        return self._map(lambda t: t.clone())

    def detach(self) -> "TensorBatch":
        return self._map(lambda t: t.detach())

    def contiguous(self) -> "TensorBatch":
        return self._map(lambda t: t.contiguous())

    def expand(self, *shape) -> "TensorBatch":
        if len(shape) == 1 and not isinstance(shape[0], int):
            shape = tuple(shape[0])
        nb = len(self.batch_size)
        return self._map(lambda t: t.expand(*shape, *t.shape[nb:]), torch.Size(shape))

    def unsqueeze(self, dim: int) -> "TensorBatch":
        bs = list(self.batch_size)
        if dim < 0:
            dim += len(bs) + 1
        bs.insert(dim, 1)
        return self._map(lambda t: t.unsqueeze(dim), torch.Size(bs))

    def view(self, *shape) -> "TensorBatch":
        if len(shape) == 1 and not isinstance(shape[0], int):
            shape = tuple(shape[0])
        nb = len(self.batch_size)
        probe = torch.empty(tuple(self.batch_size), device="meta").view(*shape)
        return self._map(lambda t: t.reshape(*probe.shape, *t.shape[nb:]), probe.shape)

    reshape = view

    def numpy(self) -> dict:
        return {k: (v.numpy() if _is_batch(v) else v.detach().cpu().numpy()) for k, v in self._d.items()}

    def to_dict(self) -> dict:
        return {k: (v.to_dict() if _is_batch(v) else v) for k, v in self._d.items()}

    def to_tensordict(self):
        """Real ``tensordict.TensorDict`` when that package is importable."""
        from tensordict import TensorDict  # noqa: PLC0415 - optional dependency

        return TensorDict(self.to_dict(), batch_size=list(self.batch_size))

    def __repr__(self) -> str:
        body = ", ".join(
            f"{k}: {v!r}" if _is_batch(v) else f"{k}: Tensor({tuple(v.shape)}, {v.dtype}, {v.device})"
            for k, v in self._d.items()
        )
        return f"TensorBatch({{{body}}}, batch_size={list(self.batch_size)})"

    # pickling (RolloutBuffer deep-copies through __getstate__, rollout_buffer.py:924-970)
    def __getstate__(self):
        return {"d": self._d, "bs": tuple(self.batch_size), "nt": self._non_tensor}

    def __setstate__(self, st):
        self._d = st["d"]
        self.batch_size = torch.Size(st["bs"])
        self._non_tensor = st["nt"]


def is_tensor_collection(x) -> bool:
    if isinstance(x, TensorBatch):
        return True
    try:
        from tensordict import is_tensor_collection as _itc  # noqa: PLC0415

        return bool(_itc(x))
    except Exception:
        return False


def as_batch(data, batch_size=None) -> TensorBatch:
    """Accepts a TensorBatch, a real TensorDict, a tensorclass (``Transition``) or a dict."""
    if isinstance(data, TensorBatch):
        return data
    if hasattr(data, "to_tensordict") and not isinstance(data, Mapping):
        data = data.to_tensordict()
    if hasattr(data, "batch_size") and hasattr(data, "items"):
        return TensorBatch({k: v for k, v in data.items()}, batch_size=list(data.batch_size))
    if isinstance(data, Mapping):
        if batch_size is None:
            first = next(iter(data.values()))
            while isinstance(first, Mapping):
                first = next(iter(first.values()))
            batch_size = [torch.as_tensor(first).shape[0]]
        return TensorBatch(data, batch_size=batch_size)
    raise TypeError(f"cannot interpret {type(data).__name__} as a batch of transitions")
