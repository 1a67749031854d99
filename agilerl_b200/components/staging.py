"""Pinned host staging for the small per-call inputs of the experience path.

The reference hands the buffers host data on every call: ``torch.rand`` draws for PER sampling
(replay_buffer.py:377) and priorities as numpy arrays (algorithms/dqn_rainbow.py:487).  A plain
``.to(device)`` of pageable memory blocks the host until the copy is done; going through a small
ring of pinned buffers keeps every call asynchronous (the copy is enqueued on the current stream
by ``xp_copy_to_device``, the host moves on).  Every slot owns a pinned host buffer AND a device
buffer.  The ring is cut into two halves; an event is recorded after the last copy of a half and
waited on (normally long complete) before that half is written again, so a slot's host side is
never overwritten while a copy may still read it.  The device side is only ever touched by
stream-ordered work (copy, then the kernel that consumes it), so reuse is safe as long as the
consumer runs on the same stream — which is how the buffers use it.
"""

from __future__ import annotations

import torch

from .. import _native as nv

_SLOT_BYTES = 1 << 16
_STREAMS: dict = {}


def _stream_object(index: int):
    """torch Stream object of the current stream (cached by raw handle: the generic
    ``torch.cuda.current_stream()`` costs ~15 us of Python per call)."""
    raw = torch._C._cuda_getCurrentRawStream(index)
    s = _STREAMS.get((index, raw))
    if s is None:
        s = torch.cuda.current_stream(index)
        _STREAMS[(index, raw)] = s
    return s


class _Slot:
    __slots__ = ("host", "dev", "views", "host_ptr", "dev_ptr", "nbytes")

    def __init__(self, nbytes: int, device: torch.device) -> None:
        self.host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        self.dev = torch.empty(nbytes, dtype=torch.uint8, device=device)
        self.host_ptr = self.host.data_ptr()
        self.dev_ptr = self.dev.data_ptr()
        self.nbytes = nbytes
        self.views: dict = {}


class PinnedRing:
    def __init__(self, device: torch.device, slots: int = 32) -> None:
        assert slots % 2 == 0
        self.device = device
        self._n = slots
        self._half = slots // 2
        self._slots: list[_Slot | None] = [None] * slots
        self._next = 0
        self._ev = [torch.cuda.Event(), torch.cuda.Event()]
        self._pending = [False, False]
        self._copy = nv.load().xp_copy_to_device
        self._last_raw = None

    def _slot(self, nbytes: int) -> _Slot:
        i = self._next
        if i % self._half == 0:
            h = i // self._half
            if self._pending[h]:  # copies issued from this half the last time round
                ev = self._ev[h]
                if not ev.query():
                    ev.synchronize()
                self._pending[h] = False
        s = self._slots[i]
        if s is None or s.nbytes < nbytes:
            if s is not None:
                # a copy from the old pinned block may still be queued (torch does not know
                # about copies issued through the C-ABI): drain before dropping it
                torch.cuda.current_stream(self.device).synchronize()
            s = _Slot(max(nbytes, _SLOT_BYTES), self.device)
            self._slots[i] = s
        return s

    def views(self, n: int, dtype: torch.dtype):
        """(slot, pinned host view [n], device view [n]) — fill the host view, then ``push``."""
        s = self._slot(n * dtype.itemsize)
        v = s.views.get((n, dtype))
        if v is None:
            nb = n * dtype.itemsize
            v = (s.host[:nb].view(dtype), s.dev[:nb].view(dtype), nb)
            s.views[(n, dtype)] = v
        return s, v[0], v[1]

    def push(self, slot: _Slot, host: torch.Tensor, dev: torch.Tensor) -> torch.Tensor:
        """Enqueue host -> device on the current stream; returns the device view (valid for
        stream-ordered consumers until the ring wraps around)."""
        index = self.device.index
        raw = torch._C._cuda_getCurrentRawStream(index)
        if raw != self._last_raw:
            if self._last_raw is not None:  # the half events only order ONE stream: drain on a switch
                torch.cuda.synchronize(index)
            self._last_raw = raw
        nv.check(self._copy(slot.dev_ptr, slot.host_ptr, host.numel() * host.element_size(), raw),
                 "xp_copy_to_device")
        i = self._next
        self._next = (i + 1) % self._n
        if (i + 1) % self._half == 0:  # last slot of a half: one event covers the whole half
            h = i // self._half
            self._ev[h].record(_stream_object(index))
            self._pending[h] = True
        return dev

    def to_device(self, t: torch.Tensor) -> torch.Tensor:
        """Asynchronous copy of a (small, flat) CPU tensor."""
        slot, host, dev = self.views(t.numel(), t.dtype)
        host.copy_(t.reshape(-1))
        return self.push(slot, host, dev)
