"""NVLink peer-memory exchange for the replicated prioritized replay (one process per GPU).

``PeerExchange`` owns, on every rank, a block of peer-mapped ("symmetric") memory that the other
ranks of the group can read directly, plus a small flag array the peers signal through.
``ring_insert()`` launches ONE kernel (``xp_peer_ring_insert``) that signals, waits and writes every
rank's new transitions straight into the local replica's ring storage (exchange + insert fused).
``allgather()`` launches ONE kernel (``xp_peer_allgather``) that signals, waits and copies every
rank's block into local rank-major arrays — the job of two NCCL all-gathers, in one launch and
without NCCL's per-collective latency.  The memory mapping itself comes from
``torch.distributed._symmetric_memory`` (plumbing); the data path is the library's own kernel.

Reference: the index broadcast / priority gather of the distributed off-policy loop,
agilerl/training/train_off_policy.py:363-381.
"""

from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist

from . import _native as nv


class XpPeerSeg(C.Structure):
    _fields_ = [("src_off", C.c_int64), ("dst", C.c_void_p), ("bytes", C.c_int64)]


class XpPeerRingSeg(C.Structure):
    _fields_ = [("src_off", C.c_int64), ("dst", C.c_void_p), ("row_bytes", C.c_int64)]


def available(group=None) -> bool:
    """True when the group's ranks are CUDA devices whose memory can be peer-mapped."""
    try:
        import torch.distributed._symmetric_memory  # noqa: F401
    except Exception:
        return False
    return dist.is_available() and dist.is_initialized() and dist.get_backend(group) == "nccl"


class PeerExchange:
    """Two alternating blocks of ``block_bytes`` per rank + flags.  ``segments`` describes what an
    ``allgather`` collects: ``[(offset_in_block, nbytes, dtype)]``; outputs are ``[world * n]``
    tensors of that dtype, rank-major."""

    def __init__(self, block_bytes: int, segments, device: torch.device, group=None) -> None:
        import torch.distributed._symmetric_memory as symm

        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        assert self.world <= 16, "XP_MAX_PEERS"
        self.device = device
        self.block_bytes = (block_bytes + 255) // 256 * 256
        # [2 blocks | flags u32[world] padded]
        total = 2 * self.block_bytes + 256
        self._mem = symm.empty(total, dtype=torch.uint8, device=device)
        self._mem.zero_()
        self._hdl = symm.rendezvous(self._mem, self.group.group_name)
        torch.cuda.synchronize(device)
        self._hdl.barrier()  # every rank's flags are zero before anyone signals
        torch.cuda.synchronize(device)
        ptrs = [int(p) for p in self._hdl.buffer_ptrs]
        self._flags = (C.c_void_p * self.world)(*[p + 2 * self.block_bytes for p in ptrs])
        self._bases = [(C.c_void_p * self.world)(*[p + k * self.block_bytes for p in ptrs]) for k in (0, 1)]
        self._segs_meta = list(segments)
        self._out = []
        self._segs = (XpPeerSeg * len(self._segs_meta))()
        self._fn_ring = nv.load().xp_peer_ring_insert
        for i, (off, nbytes, dtype) in enumerate(self._segs_meta):
            o = torch.empty(self.world * nbytes // dtype.itemsize, dtype=dtype, device=device)
            self._out.append(o)
            self._segs[i].src_off = off
            self._segs[i].dst = o.data_ptr()
            self._segs[i].bytes = nbytes
        self.err = torch.zeros(1, dtype=torch.int32, device=device)
        self.epoch = 0
        self._fn = nv.load().xp_peer_allgather

    def block(self, k: int) -> torch.Tensor:
        """This rank's block ``k`` (0/1) as a uint8 tensor: producers write their segments here."""
        return self._mem[k * self.block_bytes:(k + 1) * self.block_bytes]

    def view(self, k: int, offset: int, n: int, dtype: torch.dtype) -> torch.Tensor:
        return self.block(k)[offset:offset + n * dtype.itemsize].view(dtype)

    @property
    def current(self) -> int:
        """Index of the block the NEXT ``allgather`` will publish."""
        return self.epoch & 1

    def allgather(self):
        """Publish block ``current`` and collect every rank's; returns the rank-major outputs (valid
        for stream-ordered consumers until the next call)."""
        k = self.epoch & 1
        self.epoch += 1
        nv.check(self._fn(self._bases[k], self._flags, self.rank, self.world, self.epoch & 0xFFFFFFFF,
                          self._segs, len(self._segs_meta), self.err.data_ptr(),
                          nv.raw_stream(self.device.index)), "xp_peer_allgather")
        return self._out

    def ring_insert(self, keys, n_rows: int, cursor: int, ring_rows: int) -> None:
        """Publish block ``current`` and insert every rank's rows into the local ring in ONE kernel
        (``xp_peer_ring_insert``).  ``keys``: ``[(offset_in_block, storage [ring_rows, ...])]`` —
        this rank has written its ``n_rows`` rows of each key at that offset of block ``current``;
        row ``r * n_rows + i`` of the rank-major concatenation lands in slot
        ``(cursor + r * n_rows + i) % ring_rows`` of ``storage``."""
        k = self.epoch & 1
        self.epoch += 1
        segs = (XpPeerRingSeg * len(keys))()
        for i, (off, storage) in enumerate(keys):
            segs[i].src_off = off
            segs[i].dst = storage.data_ptr()
            segs[i].row_bytes = storage[0].numel() * storage.element_size()
        nv.check(self._fn_ring(self._bases[k], self._flags, self.rank, self.world, self.epoch & 0xFFFFFFFF,
                               segs, len(keys), n_rows, cursor, ring_rows, self.err.data_ptr(),
                               nv.raw_stream(self.device.index)), "xp_peer_ring_insert")

    def check(self) -> None:
        if int(self.err.item()) != 0:
            raise nv.NativeError("xp_peer_allgather: a peer did not arrive (timeout)")
