"""Host-side mirror of torch's CPU generator (mt19937), so the engine consumes the same raw draws
the reference does ("identical uniform draws").

``torch.randperm(n)`` (agilerl/components/replay_buffer.py:125, rollout_buffer.py:570) draws one
raw 32-bit word per position and ``torch.rand`` (:377) one word per fp32 value; torch exposes no
API for raw words, so ``TorchCpuStream`` reads the generator state (``torch.get_rng_state()``,
the legacy THGeneratorState layout), runs the C mt19937 of libb200xp on it and writes the advanced
state back.  Pure host logic — needs no GPU.
"""

from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from .. import _native as nv

# THGeneratorState (ATen/CPUGeneratorImpl.cpp, legacy layout kept by get_state()):
#   uint64 the_initial_seed; int32 left; int32 seeded; uint64 next; uint64 state[624]; ...
_OFF_LEFT, _OFF_NEXT, _OFF_STATE, _N = 8, 16, 24, 624
_MIN_STATE_BYTES = _OFF_STATE + 8 * _N


class TorchCpuStream:
    """Raw-word view of a torch CPU generator (default: the global one).

    >>> s = TorchCpuStream(); raw = s.draw_u32(4); s.commit()   # == 4 calls of generator.random()
    """

    def __init__(self, generator: torch.Generator | None = None) -> None:
        self._gen = generator
        st = (generator.get_state() if generator is not None else torch.get_rng_state()).numpy().copy()
        if st.size < _MIN_STATE_BYTES:
            raise RuntimeError("unexpected torch CPU generator state layout")
        self._raw_state = st
        left = int(st[_OFF_LEFT:_OFF_LEFT + 4].view(np.int32)[0])
        nxt = int(st[_OFF_NEXT:_OFF_NEXT + 8].view(np.uint64)[0])
        words = st[_OFF_STATE:_OFF_STATE + 8 * _N].view(np.uint64).astype(np.uint32)
        self._g = nv.XpMt19937()
        C.memmove(self._g.mt, words.ctypes.data, 4 * _N)
        # torch twists when --left hits 0; left == 1 means "twist before the next word"
        self._g.idx = _N if left == 1 else nxt
        self._lib = nv.load()

    def draw_u32(self, n: int) -> torch.Tensor:
        """The next ``n`` raw words as an int32 tensor (bit pattern of the uint32 values)."""
        out = np.empty(n, dtype=np.uint32)
        self._lib.xp_mt19937_draw(C.byref(self._g), out.ctypes.data, n)
        return torch.from_numpy(out.view(np.int32))

    def skip(self, n: int) -> None:
        self._lib.xp_mt19937_skip(C.byref(self._g), int(n))

    def rand_f32(self, n: int) -> torch.Tensor:
        """What ``torch.rand(n)`` returns from this point of the stream."""
        raw = self.draw_u32(n).numpy().view(np.uint32)
        return torch.from_numpy((raw & np.uint32(0xFFFFFF)).astype(np.float32) * np.float32(2.0**-24))

    def commit(self) -> None:
        """Write the advanced state back so torch continues after the words consumed here."""
        st = self._raw_state
        words = np.ctypeslib.as_array(self._g.mt).astype(np.uint64)
        st[_OFF_STATE:_OFF_STATE + 8 * _N] = words.view(np.uint8)
        idx = int(self._g.idx)
        if idx >= _N:
            left, nxt = 1, _N
        else:
            left, nxt = _N + 1 - idx, idx
        st[_OFF_LEFT:_OFF_LEFT + 4] = np.array([left], dtype=np.int32).view(np.uint8)
        st[_OFF_NEXT:_OFF_NEXT + 8] = np.array([nxt], dtype=np.uint64).view(np.uint8)
        t = torch.from_numpy(st)
        if self._gen is not None:
            self._gen.set_state(t)
        else:
            torch.set_rng_state(t)
