"""NVLink peer-memory exchange for the replicated prioritized replay (one process per GPU).

``PeerExchange`` owns, on every rank, a block of peer-mapped ("symmetric") memory that the other
ranks of the group can read directly, plus a small flag array the peers signal through.
``ring_insert()`` launches ONE kernel (``xp_peer_ring_insert``) that signals, waits and writes every
rank's new transitions straight into the local replica's ring storage (exchange + insert fused).
``allgather()`` launches ONE kernel (``xp_peer_allgather``) that signals, waits and copies every
rank's block into local rank-major arrays — the job of two NCCL all-gathers, in one launch and
without NCCL's per-collective latency.  The memory mapping itself comes from
``torch.distributed._symmetric_memory`` (plumbing); the data path is the library's own kernel.

The reference itself has no such exchange — one ``memory`` object serves the whole population in one
process (agilerl/training/train_off_policy.py:46,343-345,363-376).  It exists here because the
population is sharded one process per GPU with the shared memory replicated (BASELINE.json north_star);
what it must preserve is the reference's per-operation result: every replica ends up as if ONE buffer
had received every rank's ``add`` / ``update_priorities`` in rank-major order.

A peer that never arrives makes the kernel give up after a bounded wait (``set_timeout``; ~30 s by
default) and set ``err``; the exchanged data of that call is then invalid.  ``err`` is read back
without blocking after every call and the failure is raised by the next call (``check()`` forces it).
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
        # epoch kept on the device: the launch parameters never change, so a step that contains the
        # exchange can be captured in a CUDA graph (xp_peer_allgather_graph)
        self._epoch_dev = torch.zeros(2, dtype=torch.int32, device=device)
        self._fn_graph = nv.load().xp_peer_allgather_graph
        self._err_pin = torch.zeros(1, dtype=torch.int32).pin_memory()
        self._err_ev = None
        self._side = None

    def base_addrs(self, k: int) -> list[int]:
        """Peer-mapped address of block ``k`` on every rank (for native callers)."""
        return [int(p) + k * self.block_bytes for p in self._hdl.buffer_ptrs]

    def flag_addrs(self) -> list[int]:
        return [int(p) + 2 * self.block_bytes for p in self._hdl.buffer_ptrs]

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
        for stream-ordered consumers until the next call).  The epoch lives on the device, so the call
        may be captured in a CUDA graph; after replaying a graph that holds ``n`` exchanges call
        ``advance(n)`` so that ``current`` keeps naming the block the next exchange publishes."""
        capturing = torch.cuda.is_current_stream_capturing()
        if not capturing:
            self._raise_if_failed()
        self.epoch += 1
        nv.check(self._fn_graph(self._bases[0], self._bases[1], self._flags, self.rank, self.world,
                                self._epoch_dev.data_ptr(), self._segs, len(self._segs_meta),
                                self.err.data_ptr(), nv.raw_stream(self.device.index)),
                 "xp_peer_allgather_graph")
        if not capturing:
            self._poll_err_async()
        return self._out

    def advance(self, n: int) -> None:
        """Account for ``n`` exchanges issued by graph replays (the device epoch advanced by itself)."""
        self.epoch += n

    # -- failure detection (a peer that never arrived) ----------------------------------------------
    def _poll_err_async(self) -> None:
        """Queue a non-blocking read-back of the error flag behind the exchange just launched (on a
        side stream, so that the caller's stream is not delayed)."""
        if self._err_ev is not None and not self._err_ev.query():
            return  # the previous read-back is still in flight
        if self._side is None:
            self._side = torch.cuda.Stream(self.device)
            self._err_ev = torch.cuda.Event()
        self._side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self._side):
            self._err_pin.copy_(self.err, non_blocking=True)
            self._err_ev.record()

    def _raise_if_failed(self) -> None:
        if self._err_ev is not None and self._err_ev.query() and int(self._err_pin[0]) != 0:
            raise nv.NativeError("peer exchange: a peer did not arrive within the time-out; the replicas "
                                 "are no longer identical")

    @staticmethod
    def set_timeout(seconds: float, sm_ghz: float = 2.0) -> None:
        nv.load().xp_peer_set_timeout_cycles(int(seconds * sm_ghz * 1e9))

    def ring_insert(self, keys, n_rows: int, cursor: int, ring_rows: int) -> None:
        """Publish block ``current`` and insert every rank's rows into the local ring in ONE kernel
        (``xp_peer_ring_insert``).  ``keys``: ``[(offset_in_block, storage [ring_rows, ...])]`` —
        this rank has written its ``n_rows`` rows of each key at that offset of block ``current``;
        row ``r * n_rows + i`` of the rank-major concatenation lands in slot
        ``(cursor + r * n_rows + i) % ring_rows`` of ``storage``."""
        self._raise_if_failed()
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
        self._poll_err_async()

    def check(self) -> None:
        """Block until every queued exchange ran and raise if one of them timed out."""
        if int(self.err.item()) != 0:
            raise nv.NativeError("peer exchange: a peer did not arrive within the time-out; the replicas "
                                 "are no longer identical")
