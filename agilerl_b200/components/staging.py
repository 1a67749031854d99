"""Pinned host staging for the small per-call inputs of the experience path.

The reference hands the buffers host data on every call: ``torch.rand`` draws for PER sampling
(replay_buffer.py:377) and priorities as numpy arrays (algorithms/dqn_rainbow.py:487).  A plain
``.to(device)`` of pageable memory blocks the host until the copy is done; going through a small
ring of pinned buffers keeps every call asynchronous (the copy is enqueued on the current stream,
the host moves on).  Every slot owns a pinned host buffer AND a device buffer; a slot is reused
only after the copy that read its host side has completed, and its device side is only ever
touched by stream-ordered work (copy, then the kernel that consumes it), so reuse is safe as long
as the consumer runs on the same stream — which is how the buffers use it.
"""

from __future__ import annotations

import torch

_SLOT_BYTES = 1 << 16


class _Slot:
    __slots__ = ("host", "dev", "ev", "views")

    def __init__(self, nbytes: int, device: torch.device) -> None:
        self.host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        self.dev = torch.empty(nbytes, dtype=torch.uint8, device=device)
        self.ev = torch.cuda.Event()
        self.views: dict = {}


class PinnedRing:
    def __init__(self, device: torch.device, slots: int = 16) -> None:
        self.device = device
        self._n = slots
        self._slots: list[_Slot | None] = [None] * slots
        self._next = 0
        self._used = [False] * slots

    def _slot(self, nbytes: int) -> _Slot:
        i = self._next
        self._next = (i + 1) % self._n
        s = self._slots[i]
        if s is None or s.host.numel() < nbytes:
            if s is not None and self._used[i]:
                s.ev.synchronize()
            s = _Slot(max(nbytes, _SLOT_BYTES), self.device)
            self._slots[i] = s
            self._used[i] = False
        elif self._used[i] and not s.ev.query():
            s.ev.synchronize()  # normally long complete
        return s

    def views(self, n: int, dtype: torch.dtype):
        """(slot, pinned host view [n], device view [n]) — fill the host view, then ``push``."""
        s = self._slot(n * dtype.itemsize)
        v = s.views.get((n, dtype))
        if v is None:
            nb = n * dtype.itemsize
            v = (s.host[:nb].view(dtype), s.dev[:nb].view(dtype))
            s.views[(n, dtype)] = v
        return s, v[0], v[1]

    def push(self, slot: _Slot, host: torch.Tensor, dev: torch.Tensor) -> torch.Tensor:
        """Enqueue host -> device on the current stream; returns the device view (valid for
        stream-ordered consumers until the ring wraps around)."""
        dev.copy_(host, non_blocking=True)
        slot.ev.record()
        self._used[self._slots.index(slot)] = True
        return dev

    def to_device(self, t: torch.Tensor) -> torch.Tensor:
        """Asynchronous copy of a (small, flat) CPU tensor."""
        slot, host, dev = self.views(t.numel(), t.dtype)
        host.copy_(t.reshape(-1))
        return self.push(slot, host, dev)
