"""Device-resident segment trees with the API of agilerl/components/segment_tree.py.

``SumSegmentTree(capacity)`` / ``MinSegmentTree(capacity)`` keep the reference surface —
``tree[idx] = v``, ``tree[idx]``, ``.sum()/.min(start, end)``, ``.retrieve(ub)``, ``.capacity``,
``.tree`` (heap-order, ``2*capacity`` values), ``.operation`` — but the nodes live in HBM as fp64
in heap order and every operation is a CUDA kernel behind the C-ABI (include/b200xp.h).  Batched
forms (``set_batch``, ``retrieve_batch``) are what the prioritized buffer uses.

Numerics are the reference's: fp64, ``operator.add`` / Python ``min`` per node, every ancestor
recomputed from its final children (segment_tree.py:89-95), strict ``>`` descent with unfused
subtraction (:147-156), recursive range association (:28-62).
"""

from __future__ import annotations

import operator
from typing import Callable

import torch

from .. import _native as nv
from .. import xp


class _TreeStorage:
    """Node arrays of one tree or an interleaved (sum, min) pair + the int32 scratch."""

    def __init__(self, capacity: int, device, paired: bool):
        assert capacity > 0, "capacity must be positive and a power of 2."
        assert capacity & (capacity - 1) == 0, "capacity must be positive and a power of 2."
        self.device = nv.require_cuda(device)
        self.capacity = capacity
        self.paired = paired
        with nv.on_device(self.device):
            if paired:
                self.nodes = torch.empty(2 * capacity, 2, dtype=torch.float64, device=self.device)
            else:
                self.nodes = torch.empty(2 * capacity, 1, dtype=torch.float64, device=self.device)
            self.aux = torch.empty(2 * capacity, dtype=torch.int32, device=self.device)
            self.err = torch.zeros(1, dtype=torch.int32, device=self.device)
            self.scratch = torch.zeros(2, dtype=torch.float64, device=self.device)

    def column(self, j: int) -> torch.Tensor:
        return self.nodes[:, j]

    def handle(self, which: str | None) -> nv.XpTree:
        """which: 'sum', 'min' or None (both; paired only)."""
        stride = self.nodes.shape[1]
        if self.paired:
            s = self.column(0) if which in (None, "sum") else None
            m = self.column(1) if which in (None, "min") else None
        else:
            s = self.column(0) if which == "sum" else None
            m = self.column(0) if which == "min" else None
        return xp.make_tree(s, m, stride, self.capacity, self.aux)


class _NodeView:
    """List-like view of one tree's ``2*capacity`` node values (reference: ``SegmentTree.tree``)."""

    def __init__(self, column: torch.Tensor):
        self._c = column

    def __len__(self) -> int:
        return self._c.shape[0]

    def __getitem__(self, i):
        if isinstance(i, slice):
            return self._c[i].cpu().tolist()
        return float(self._c[i].item())

    def __iter__(self):
        return iter(self._c.cpu().tolist())

    def __eq__(self, other):
        return list(self) == list(other)

    def tolist(self) -> list:
        return self._c.cpu().tolist()

    def tensor(self) -> torch.Tensor:
        """The live device column (strided view); writes bypass ancestor maintenance."""
        return self._c

    def __repr__(self) -> str:
        return f"NodeView({self.tolist()!r})"


class SegmentTree:
    """Reference: agilerl/components/segment_tree.py:5-108."""

    _kind = "sum"

    def __init__(self, capacity: int, operation: Callable, init_value: float, *,
                 device=None, _storage: _TreeStorage | None = None) -> None:
        assert capacity > 0, "capacity must be positive and a power of 2."
        assert capacity & (capacity - 1) == 0, "capacity must be positive and a power of 2."
        self.capacity = capacity
        self.operation = operation
        self.init_value = init_value
        if _storage is None:
            self._storage = _TreeStorage(capacity, device, paired=False)
            self._col = self._storage.column(0)
            self._t = self._storage.handle(self._kind)
            with nv.on_device(self._storage.device):
                xp.tree_init(self._t, torch.cuda.current_stream().cuda_stream)
        else:
            self._storage = _storage
            self._col = _storage.column(0 if self._kind == "sum" else 1)
            self._t = _storage.handle(self._kind)
        self.tree = _NodeView(self._col)

    # -- reference API -------------------------------------------------------------------------
    @property
    def device(self):
        return self._storage.device

    def operate(self, start: int = 0, end: int = 0) -> float:
        out = self._storage.scratch
        with nv.on_device(self.device):
            xp.tree_query(self._t, int(start), int(end), out)
        return float(out[0 if self._kind == "sum" else 1].item())

    def __setitem__(self, idx: int, val: float) -> None:
        i = torch.tensor([int(idx)], dtype=torch.int64, device=self.device)
        v = torch.tensor([float(val)], dtype=torch.float64, device=self.device)
        self.set_batch(i, v, check=True)

    def __getitem__(self, idx: int) -> float:
        assert 0 <= idx < self.capacity
        return float(self._col[self.capacity + int(idx)].item())

    # -- batched forms (engine) ------------------------------------------------------------------
    def set_batch(self, idx: torch.Tensor, val: torch.Tensor, check: bool = False) -> None:
        """Sequential ``tree[idx[i]] = val[i]`` for the whole batch in one pass (last writer wins)."""
        idx = idx.to(device=self.device, dtype=torch.int64).contiguous().view(-1)
        val = val.to(device=self.device, dtype=torch.float64).contiguous().view(-1)
        with nv.on_device(self.device):
            xp.tree_set_batch(self._t, idx, val, self._storage.err)
        if check and int(self._storage.err.item()) != 0:
            self._storage.err.zero_()
            raise IndexError("segment tree index out of range")  # list assignment in the reference

    def leaves(self) -> torch.Tensor:
        return self._col[self.capacity:]


class SumSegmentTree(SegmentTree):
    """Reference: agilerl/components/segment_tree.py:111-156."""

    _kind = "sum"

    def __init__(self, capacity: int, *, device=None, _storage=None) -> None:
        super().__init__(capacity=capacity, operation=operator.add, init_value=0.0,
                         device=device, _storage=_storage)

    def sum(self, start: int = 0, end: int = 0) -> float:
        return super().operate(start, end)

    def retrieve(self, upperbound: float) -> int:
        assert 0 <= upperbound <= self.sum() + 1e-5, f"upperbound: {upperbound}"
        ub = torch.tensor([float(upperbound)], dtype=torch.float64, device=self.device)
        return int(self.retrieve_batch(ub).item())

    def retrieve_batch(self, upperbound: torch.Tensor) -> torch.Tensor:
        ub = upperbound.to(device=self.device, dtype=torch.float64).contiguous().view(-1)
        with nv.on_device(self.device):
            return xp.tree_retrieve(self._t, ub)


class MinSegmentTree(SegmentTree):
    """Reference: agilerl/components/segment_tree.py:159-182."""

    _kind = "min"

    def __init__(self, capacity: int, *, device=None, _storage=None) -> None:
        super().__init__(capacity=capacity, operation=min, init_value=float("inf"),
                         device=device, _storage=_storage)

    def min(self, start: int = 0, end: int = 0) -> float:
        return super().operate(start, end)


def make_pair(capacity: int, device=None) -> tuple[SumSegmentTree, MinSegmentTree, _TreeStorage]:
    """An interleaved (sum, min) pair sharing one node array: the two children of a node in
    both trees are one aligned 32-byte sector (PrioritizedReplayBuffer, replay_buffer.py:292-294)."""
    st = _TreeStorage(capacity, device, paired=True)
    with nv.on_device(st.device):
        xp.tree_init(st.handle(None), torch.cuda.current_stream().cuda_stream)
    return SumSegmentTree(capacity, _storage=st), MinSegmentTree(capacity, _storage=st), st
