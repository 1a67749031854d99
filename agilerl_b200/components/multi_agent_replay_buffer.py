"""Multi-agent experience replay on the B200 engine.

Reference: agilerl/components/multi_agent_replay_buffer.py:16-242 — a ``deque`` of namedtuples whose
fields are ``{agent_id: value}`` dicts; ``sample`` draws ``random.sample(memory, k)`` and stacks every
(field, agent) with numpy, one Python loop per sample per agent per field.

Here every (field, agent[, sub-key]) is one ``[memory_size, *shape]`` device array (struct of
arrays, like the single-agent buffers); a save is one ``xp_ring_insert`` launch per 24 columns and a
sample is ``random.sample(range(len), k)`` on the host — the same draws from Python's generator and the
same positions as ``random.sample(memory, k)`` — followed by one ``xp_gather`` launch per 24 columns.
Returned tensors are fp32 (``obs_to_tensor`` casts everything with ``.float()``, utils/algo_utils.py:
746-773), binary fields go through uint8 first (:137-138), 1-D stacks become ``[B, 1]`` (:80-96).
"""

from __future__ import annotations

import random
from collections import OrderedDict, namedtuple
from typing import Any

import numpy as np
import torch

from .. import _native as nv
from .. import xp

_BINARY_FIELDS = ("done", "termination", "terminated", "truncation", "truncated")
_MAX_KEYS = 24  # XP_MAX_KEYS


def _leaves(value) -> "list[tuple[tuple, Any]]":
    """One agent's value -> [(sub-path, array-like)] (dict / tuple observations are nested)."""
    if isinstance(value, dict):
        return [(("d", k), v) for k, v in value.items()]
    if isinstance(value, tuple):
        return [(("t", i), v) for i, v in enumerate(value)]
    return [((), value)]


class _MemoryView:
    """``buffer.memory`` of the reference is the deque itself; here it is a read-only window on
    the device columns: ``maxlen``, ``len()``, ``memory[i]`` -> Experience of host numpy values."""

    def __init__(self, owner: "MultiAgentReplayBuffer") -> None:
        self._o = owner

    @property
    def maxlen(self) -> int:
        return self._o.memory_size

    def __len__(self) -> int:
        return self._o._size

    def __getitem__(self, i: int):
        o = self._o
        n = len(self)
        if not -n <= i < n:
            raise IndexError("deque index out of range")
        slot = o._slot_of(i % n if i < 0 else i)
        fields = []
        for f in o.field_names:
            per_agent = {}
            for a in o.agent_ids:
                per_agent[a] = o._rebuild(f, a, lambda t: t[slot].cpu().numpy())
            fields.append(per_agent)
        return o.experience(*fields)

    def __iter__(self):
        return (self[i] for i in range(len(self)))


class MultiAgentReplayBuffer:
    def __init__(self, memory_size: int, field_names: list[str], agent_ids: list[str],
                 device: str | None = None) -> None:
        assert memory_size > 0, "Memory size must be greater than zero."
        assert len(field_names) > 0, "Field names must contain at least one field name."
        assert len(agent_ids) > 0, "Agent ids must contain at least one agent id."
        self.memory_size = memory_size
        self.field_names = field_names
        self.experience = namedtuple("Experience", self.field_names)
        self.counter = 0
        self.device = device
        self.agent_ids = agent_ids
        self._sdev = nv.require_cuda(device)
        self._cols: "OrderedDict[tuple, torch.Tensor] | None" = None  # (field, agent, sub) -> column
        self._kinds: dict = {}   # (field, agent) -> "array" | "dict" | "tuple"
        self._scalar: dict = {}  # column path -> True when the stored value is 0-d (stacks to [B, 1])
        self._size = 0
        self._cursor = 0
        self.memory = _MemoryView(self)

    def __len__(self) -> int:
        return self._size

    # -- layout ------------------------------------------------------------------------------------
    def _slot_of(self, j):
        """deque position (0 = oldest) -> ring slot."""
        if self._size < self.memory_size:
            return j
        return (self._cursor + j) % self.memory_size

    def _rebuild(self, field: str, agent: str, take):
        kind = self._kinds[(field, agent)]
        cols = [(p, t) for p, t in self._cols.items() if p[0] == field and p[1] == agent]
        if kind == "dict":
            return {p[2][1]: take(t) for p, t in cols}
        if kind == "tuple":
            return tuple(take(t) for p, t in cols)
        return take(cols[0][1])

    def _init_columns(self, args, vect: bool) -> None:
        self._cols = OrderedDict()
        for f, arg in zip(self.field_names, args):
            for a in self.agent_ids:
                v = arg[a]
                self._kinds[(f, a)] = "dict" if isinstance(v, dict) else "tuple" if isinstance(v, tuple) else "array"
                for sub, leaf in _leaves(v):
                    x = np.asarray(leaf[0] if vect else leaf)
                    t = torch.as_tensor(x)
                    self._scalar[(f, a, sub)] = x.ndim == 0
                    self._cols[(f, a, sub)] = torch.zeros((self.memory_size, *x.shape), dtype=t.dtype,
                                                          device=self._sdev)

    def _insert(self, args, n: int, vect: bool) -> None:
        if len(args) != len(self.field_names):
            raise TypeError(f"Experience expects {len(self.field_names)} fields, got {len(args)}")
        if self._cols is None:
            self._init_columns(args, vect)
        if n > self.memory_size:  # only the last memory_size rows survive the deque
            args = self._tail(args, n - self.memory_size)
            self.counter += n - self.memory_size
            self._advance(n - self.memory_size)
            n = self.memory_size
        pairs = []
        for f, arg in zip(self.field_names, args):
            for a in self.agent_ids:
                for sub, leaf in _leaves(arg[a]):
                    col = self._cols[(f, a, sub)]
                    x = np.asarray(leaf)
                    src = torch.as_tensor(x).reshape(n, *col.shape[1:])
                    if src.dtype != col.dtype:
                        src = src.to(col.dtype)
                    pairs.append((src.to(self._sdev, non_blocking=True), col))
        with nv.on_device(self._sdev):
            for k in range(0, len(pairs), _MAX_KEYS):
                xp.ring_insert(pairs[k:k + _MAX_KEYS], n, self._cursor, self.memory_size)
        self._advance(n)
        self.counter += n

    def _advance(self, n: int) -> None:
        self._cursor = (self._cursor + n) % self.memory_size
        self._size = min(self._size + n, self.memory_size)

    @staticmethod
    def _tail(args, skip: int):
        def cut(v):
            if isinstance(v, dict):
                return {k: cut(x) for k, x in v.items()}
            if isinstance(v, tuple):
                return tuple(cut(x) for x in v)
            return np.asarray(v)[skip:]
        return tuple({a: cut(v) for a, v in arg.items()} for arg in args)

    # -- saving (reference :98-107, :164-242) -------------------------------------------------------------
    def _add(self, *args) -> None:
        self._insert(args, 1, vect=False)
        self.counter -= 1  # _add itself does not count (:98-106); the callers below do

    def save_to_memory_single_env(self, *args) -> None:
        self._insert(args, 1, vect=False)

    def save_to_memory_vect_envs(self, *args) -> None:
        # the reference sizes the loop by the FIRST agent's value of the FIRST field (:199)
        num_entries = len(next(iter(args[0].values())))
        if num_entries == 0:
            return
        self._insert(self._head(args, num_entries), num_entries, vect=True)

    @staticmethod
    def _head(args, n: int):
        def cut(v):
            if isinstance(v, dict):
                return {k: cut(x) for k, x in v.items()}
            if isinstance(v, tuple):
                return tuple(cut(x) for x in v)
            return np.asarray(v)[:n]
        return tuple({a: cut(v) for a, v in arg.items()} for arg in args)

    def save_to_memory(self, *args, is_vectorised: bool = False) -> None:
        if is_vectorised:
            self.save_to_memory_vect_envs(*args)
        else:
            self.save_to_memory_single_env(*args)

    # -- sampling (reference :150-162) ------------------------------------------------------------------
    def sample(self, batch_size: int, *args: Any) -> tuple:
        positions = random.sample(range(self._size), k=batch_size)  # same draws as random.sample(memory, k)
        if self._cols is None:
            return tuple({} for _ in self.field_names)
        slots = [self._slot_of(j) for j in positions]
        idx = torch.tensor(slots, dtype=torch.int64).to(self._sdev, non_blocking=True)
        out = OrderedDict((p, torch.empty((batch_size, *c.shape[1:]), dtype=c.dtype, device=self._sdev))
                          for p, c in self._cols.items())
        with nv.on_device(self._sdev):
            paths = list(self._cols.keys())
            for k in range(0, len(paths), _MAX_KEYS):
                chunk = paths[k:k + _MAX_KEYS]
                xp.gather([(self._cols[p], out[p]) for p in chunk], idx, self.memory_size)
        target = None
        if self.device is not None and torch.device(self.device).type != "cuda":
            target = torch.device(self.device)
        transition = {f: {} for f in self.field_names}
        for f in self.field_names:
            binary = f in _BINARY_FIELDS
            for a in self.agent_ids:
                kind = self._kinds[(f, a)]
                cols = [(p, out[p]) for p in out if p[0] == f and p[1] == a]
                conv = []
                for p, t in cols:
                    if self._scalar[p]:
                        t = t.unsqueeze(1)
                    if binary and kind == "array":
                        # ts.astype(np.uint8) unless the stack holds a NaN (:137-138), then .float()
                        if not (t.is_floating_point() and bool(torch.isnan(t).any())):
                            t = t.to(torch.uint8)
                    t = t.float()
                    conv.append((p, t if target is None else t.to(target)))
                if kind == "dict":
                    transition[f][a] = {p[2][1]: t for p, t in conv}
                elif kind == "tuple":
                    transition[f][a] = tuple(t for _, t in conv)
                else:
                    transition[f][a] = conv[0][1]
        return tuple(transition.values())

    # -- host-side helpers kept for API parity (reference :58-96, :108-148) ----------------------------------
    @staticmethod
    def stack_transitions(transitions: list):
        field_type = type(transitions[0])
        ts = []
        for item in transitions:
            if field_type is dict:
                ts.append({k: np.array(v) for k, v in item.items()})
            elif field_type is tuple:
                ts.append(tuple(np.array(v) for v in item))
            else:
                ts.append(item)
        if field_type is dict:
            out = {}
            for k in ts[0].keys():
                kts = np.array([t[k] for t in ts])
                out[k] = np.expand_dims(kts, axis=1) if kts.ndim == 1 else kts
            return out
        if field_type is tuple:
            out = []
            for i in range(len(ts[0])):
                its = np.array([t[i] for t in ts])
                out.append(np.expand_dims(its, axis=1) if its.ndim == 1 else its)
            return tuple(out)
        ts = np.array(ts)
        return np.expand_dims(ts, axis=1) if ts.ndim == 1 else ts

    def _process_transition(self, experiences: list, np_array: bool = False) -> dict:
        def to_tensor(x):
            if isinstance(x, dict):
                return {k: torch.as_tensor(v, device=self.device).float() for k, v in x.items()}
            if isinstance(x, tuple):
                return tuple(torch.as_tensor(v, device=self.device).float() for v in x)
            return torch.as_tensor(x, device=self.device).float()

        transition = {field: {} for field in self.field_names}
        kept = [e for e in experiences if e is not None]
        for field in self.field_names:
            binary = field in _BINARY_FIELDS
            for agent_id in self.agent_ids:
                ts = MultiAgentReplayBuffer.stack_transitions([getattr(e, field)[agent_id] for e in kept])
                if binary and not np.isnan(ts).any():
                    ts = ts.astype(np.uint8)
                transition[field][agent_id] = ts if np_array else to_tensor(ts)
        return transition
