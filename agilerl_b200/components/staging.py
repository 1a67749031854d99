"""Pinned host staging for the small per-call inputs of the experience path.

The reference hands the buffers host data on every call: ``torch.rand`` draws for PER sampling
(replay_buffer.py:377) and priorities as numpy arrays (algorithms/dqn_rainbow.py:487).  A plain
``.to(device)`` of pageable memory blocks the host until the copy is done; going through a small
ring of pinned buffers keeps every call asynchronous (the copy is enqueued on the current stream,
the host moves on).  A slot is reused only after the copy that read it has completed.
"""

from __future__ import annotations

import torch


class PinnedRing:
    def __init__(self, device: torch.device, slots: int = 16) -> None:
        self.device = device
        self._slots: list[dict] = [{"buf": None, "ev": None} for _ in range(slots)]
        self._next = 0

    def _slot(self, nbytes: int) -> dict:
        s = self._slots[self._next]
        self._next = (self._next + 1) % len(self._slots)
        if s["ev"] is not None and not s["ev"].query():
            s["ev"].synchronize()  # normally long complete
        if s["buf"] is None or s["buf"].numel() < nbytes:
            s["buf"] = torch.empty(max(nbytes, 4096), dtype=torch.uint8).pin_memory()
        if s["ev"] is None:
            s["ev"] = torch.cuda.Event()
        return s

    def host_view(self, n: int, dtype: torch.dtype):
        """(slot, pinned tensor [n] of dtype) to be filled by the caller, then ``push``ed."""
        s = self._slot(n * dtype.itemsize)
        return s, s["buf"][: n * dtype.itemsize].view(dtype)

    def push(self, slot: dict, host: torch.Tensor) -> torch.Tensor:
        """Enqueue host -> device on the current stream; returns the device tensor."""
        out = torch.empty(host.shape, dtype=host.dtype, device=self.device)
        out.copy_(host, non_blocking=True)
        slot["ev"].record()
        return out

    def to_device(self, t: torch.Tensor) -> torch.Tensor:
        """Asynchronous copy of a (small) CPU tensor."""
        t = t.contiguous()
        slot, host = self.host_view(t.numel(), t.dtype)
        host.copy_(t.view(-1))
        return self.push(slot, host).view(t.shape)
