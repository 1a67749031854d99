"""Population sharding across the GPUs of one box (one process per GPU, torch.distributed).

The reference trains the agents of a population sequentially on one device and shares ONE
replay ``memory`` between them (training/train_off_policy.py:46,249,343-345).  Here:

* each rank owns a contiguous slice of the agents and their RolloutBuffers entirely — GAE / add /
  minibatching need no communication;
* the shared off-policy memory is REPLICATED: every rank holds the full storage and both trees.
  Per env step each rank contributes its agents' new rows; an all-gather in fixed rank-major,
  agent-major order makes every replica insert the same rows at the same cursor.  PER priority
  updates are all-gathered as (idx, priority) pairs and applied on every replica in the same
  canonical order through the same last-writer-wins kernel, so replicas stay bit-identical.

The exchange is device-agnostic (NCCL on GPUs; the CPU tests run it over gloo, world_size 2).
"""

from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as dist

from .components.tensor_batch import TensorBatch, as_batch


@dataclass(frozen=True)
class ShardPlan:
    population: int
    world_size: int
    rank: int

    def owner(self, agent: int) -> int:
        """Rank owning ``agent``: contiguous blocks, the first ``population % world_size`` ranks
        hold one extra agent."""
        base, extra = divmod(self.population, self.world_size)
        cut = extra * (base + 1)
        return agent // (base + 1) if agent < cut else extra + (agent - cut) // max(base, 1)

    def agents_of(self, rank: int) -> range:
        base, extra = divmod(self.population, self.world_size)
        start = rank * base + min(rank, extra)
        return range(start, start + base + (1 if rank < extra else 0))

    @property
    def local_agents(self) -> range:
        return self.agents_of(self.rank)

    @property
    def max_local(self) -> int:
        return len(self.agents_of(0))


def _group_size(group=None) -> int:
    return dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1


def all_gather_rows(t: torch.Tensor, group=None) -> torch.Tensor:
    """[n, ...] on every rank -> [world*n, ...], rank-major.  Equal ``n`` on every rank."""
    world = _group_size(group)
    if world == 1:
        return t
    t = t.contiguous()
    out = torch.empty((world * t.shape[0], *t.shape[1:]), dtype=t.dtype, device=t.device)
    if t.dtype == torch.bool:  # NCCL has no bool; ship bytes
        dist.all_gather_into_tensor(out.view(torch.uint8), t.view(torch.uint8), group=group)
    else:
        dist.all_gather_into_tensor(out, t, group=group)
    return out


def all_gather_batch(batch: TensorBatch, group=None) -> TensorBatch:
    """Every key of a ``[n]`` batch gathered rank-major into a ``[world*n]`` batch."""
    world = _group_size(group)
    if world == 1:
        return batch
    n = batch.batch_size[0]
    out = batch._map(lambda t: all_gather_rows(t, group), torch.Size([world * n]))
    return out


class ReplicatedReplay:
    """A replay buffer replicated on every rank, kept identical by canonical-order exchange.

    ``memory`` is any of the engine's buffers (``ReplayBuffer``, ``PrioritizedReplayBuffer``;
    for n-step wrap the ``MultiStepReplayBuffer`` whose output feeds the PER buffer).  Each rank
    calls the methods below collectively with the rows / priorities of ITS agents.
    """

    def __init__(self, memory, group=None) -> None:
        self.memory = memory
        self.group = group
        self._px = None      # peer.PeerExchange, created on the first CUDA update
        self._px_n = 0
        self._peer_ok = None
        self._ix = None      # peer.PeerExchange of the insert path (blocks sized for one add)
        self._ix_sig = None
        self._ix_ok = None

    def add(self, local_rows) -> None:
        """``local_rows``: the new transitions of this rank's agents, agent-major ``[n]`` (equal
        ``n`` on every rank).  Every replica inserts all ranks' rows, rank-major."""
        local = as_batch(local_rows)
        dev = self.memory.storage_device if _is_cuda_backend(self.group) else None
        if dev is not None:
            local = local.to(dev)
        if dev is not None and _group_size(self.group) > 1 and self._ix_ok is not False \
                and type(self.memory) in _PEER_INSERT_TYPES and self._peer_add(local):
            return
        self.memory.add(all_gather_batch(local, self.group))

    def _peer_add(self, local) -> bool:
        """Exchange + insert in ONE kernel over NVLink peer memory (``xp_peer_ring_insert``): each
        rank stages its rows in its peer-mapped block, every replica pulls all ranks' rows straight
        into its ring at the rank-major slots.  Returns False when peer memory is unavailable (the
        NCCL all-gather then serves)."""
        mem = self.memory
        data, n = mem._prepare(local)
        if mem._storage is None:
            mem._init(data)
        world = _group_size(self.group)
        if world * n > mem.max_size:
            return False
        pairs = mem._pairs(data, mem._leaves)
        sig = (n, tuple((tuple(t.shape), t.dtype) for t, _ in pairs))
        try:
            if self._ix is None or self._ix_sig != sig:
                from . import peer

                if len(pairs) > 16 or not peer.available(self.group):
                    raise RuntimeError("peer insert unavailable")
                offs, off = [], 0
                for t, _ in pairs:
                    offs.append(off)
                    off += (t.numel() * t.element_size() + 255) // 256 * 256
                self._ix = peer.PeerExchange(off, [], torch.device(mem.storage_device), self.group)
                self._ix_sig, self._ix_offs, self._ix_ok = sig, offs, True
        except Exception:
            if self._ix_ok:
                raise
            self._ix_ok = False
            return False
        px, k = self._ix, self._ix.current
        for (t, _), off in zip(pairs, self._ix_offs):
            px.view(k, off, t.numel(), t.dtype).copy_(t.reshape(-1))
        keys = [(off, d) for (_, d), off in zip(pairs, self._ix_offs)]
        mem._add_rows_with(world * n, lambda cursor: px.ring_insert(keys, n, cursor, mem.max_size))
        return True

    def update_priorities(self, local_idx: torch.Tensor, local_prio: torch.Tensor) -> None:
        """(idx, priority) pairs of this rank's agents -> applied everywhere in rank-major order
        (one batched last-writer-wins update per replica)."""
        world = _group_size(self.group)
        if world > 1 and self._peer_ok is not False and self._native_update(local_idx, local_prio):
            return
        dev = self.memory.storage_device if _is_cuda_backend(self.group) else local_idx.device
        idx = torch.as_tensor(local_idx).to(device=dev, dtype=torch.int64).reshape(-1)
        pr = torch.as_tensor(local_prio).to(device=dev, dtype=torch.float32).reshape(-1)
        if world > 1 and _is_cuda_backend(self.group) and self._peer_ok is not False:
            # NVLink peer-memory exchange: one kernel signals, waits and pulls every rank's block
            n = idx.shape[0]
            try:
                if self._px is None or self._px_n != n:
                    from . import peer

                    if not peer.available(self.group):
                        raise RuntimeError("symmetric memory unavailable")
                    self._px = peer.PeerExchange(12 * n, [(0, 8 * n, torch.int64), (8 * n, 4 * n, torch.float32)],
                                                 torch.device(dev), self.group)
                    self._px_n = n
                    self._peer_ok = True
            except Exception:
                if self._peer_ok:  # it worked before: a real error, not an unsupported box
                    raise
                self._peer_ok = False
            if self._peer_ok:
                k = self._px.current
                self._px.view(k, 0, n, torch.int64).copy_(idx)
                self._px.view(k, 8 * n, n, torch.float32).copy_(pr)
                all_idx, all_pr = self._px.allgather()
                self.memory.update_priorities(all_idx, all_pr)
                return
        if world > 1 and _is_cuda_backend(self.group):
            # both all-gathers in ONE NCCL launch
            all_idx = torch.empty(world * idx.shape[0], dtype=idx.dtype, device=dev)
            all_pr = torch.empty(world * pr.shape[0], dtype=pr.dtype, device=dev)
            with dist._coalescing_manager(group=self.group, device=torch.device(dev)):
                dist.all_gather_into_tensor(all_idx, idx.contiguous(), group=self.group)
                dist.all_gather_into_tensor(all_pr, pr.contiguous(), group=self.group)
            self.memory.update_priorities(all_idx, all_pr)
            return
        self.memory.update_priorities(all_gather_rows(idx, self.group), all_gather_rows(pr, self.group))

    def _native_update(self, local_idx, local_prio) -> bool:
        """The whole replicated update in ONE native call (csrc/hostaccel.cpp ``FastPer.update_replicated``):
        indices / priorities into this rank's peer-mapped block, the NVLink exchange kernel, the rank-major
        update.  Taken when the indices are the canonical device tensor ``sample`` returned and the priorities
        an fp32 array (host or device); anything else goes through the generic path below.  Returns False when
        it did not run."""
        mem = self.memory
        if not (isinstance(local_idx, torch.Tensor) and local_idx.is_cuda and local_idx.dtype == torch.int64
                and local_idx.is_contiguous() and _is_cuda_backend(self.group)):
            return False
        fp = getattr(mem, "_fast_per", None)
        fp = fp() if fp is not None else None
        if fp is None or local_idx.device != mem.storage_device or torch._C._cuda_getDevice() != local_idx.device.index:
            return False
        pr = local_prio if isinstance(local_prio, torch.Tensor) else torch.from_numpy(local_prio)
        n = local_idx.numel()
        if pr.dtype != torch.float32 or not pr.is_contiguous() or pr.numel() != n or (pr.is_cuda and pr.device != local_idx.device):
            return False
        try:
            if self._px is None or self._px_n != n:
                from . import peer

                if not peer.available(self.group):
                    raise RuntimeError("symmetric memory unavailable")
                self._px = peer.PeerExchange(12 * n, [(0, 8 * n, torch.int64), (8 * n, 4 * n, torch.float32)],
                                             torch.device(mem.storage_device), self.group)
                self._px_n = n
                self._peer_ok = True
        except Exception:
            if self._peer_ok:
                raise
            self._peer_ok = False
            return False
        px = self._px
        if getattr(mem, "_xchg", None) is not px:  # bind the exchange to the buffer's native object once
            fp.set_exchange(px.base_addrs(0), px.base_addrs(1), px.flag_addrs(), px.rank, px._epoch_dev,
                            px._out[0], px._out[1], px.err, n)
            mem._xchg = px
        if mem._pending_err is not None:
            mem._raise_deferred()
        px._raise_if_failed()
        fp.update_replicated(px.current, local_idx, pr)
        px.epoch += 1
        self._since_poll = getattr(self, "_since_poll", 0) + 1
        if self._since_poll >= 32:  # a peer that never arrived / an index out of range: raised by a later call
            self._since_poll = 0
            px._poll_err_async()
            if mem._pending_err is None:
                mem._poll_err_async()
        return True

    def sample(self, *args, **kwargs):
        """Sampling and gathering are local to a replica."""
        return self.memory.sample(*args, **kwargs)


from .components.replay_buffer import PrioritizedReplayBuffer, ReplayBuffer  # noqa: E402

_PEER_INSERT_TYPES = (ReplayBuffer, PrioritizedReplayBuffer)


def _is_cuda_backend(group=None) -> bool:
    if not (dist.is_available() and dist.is_initialized()):
        return True
    return dist.get_backend(group) == "nccl"
