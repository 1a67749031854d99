"""Device-resident replay buffers with the API of agilerl/components/replay_buffer.py.

``ReplayBuffer`` / ``MultiStepReplayBuffer`` / ``PrioritizedReplayBuffer`` keep the reference's
names, inheritance, constructor signatures, attributes and returned keys/shapes, so ``Sampler``'s
``isinstance`` dispatch (sampler.py:71-72) and the training loops (train_off_policy.py:341-412)
work unchanged.  Storage is struct-of-arrays in HBM; add / sample / n-step fold / priority
update are CUDA kernels behind the C-ABI (include/b200xp.h).  There is no CPU path.

Random draws ("identical uniform draws"): with ``rng="torch"`` (default) the buffers consume
torch's global CPU mt19937 stream exactly where the reference does — ``torch.rand`` per stratum
for PER (:377) and the raw draws of ``torch.randperm`` for uniform sampling (:125) — so a seeded
run yields the reference's indices bit-for-bit.  ``rng="device"`` draws with Philox on the GPU.
"""

from __future__ import annotations

from collections import deque
from typing import Any

import numpy as np
import torch

from .. import _hostaccel
from .. import _native as nv
from .. import xp
from .rng import TorchCpuStream
from .staging import PinnedRing
from .tensor_batch import TensorBatch, as_batch

DataType = Any

_TORCH_RAND = torch.rand  # the builtin: a patched ``torch.rand`` (the reference's tests mock it, test_replay_buffer.py:956-1000) is honoured call by call


def _flatten(batch: TensorBatch, prefix=()) -> list[tuple[tuple, torch.Tensor]]:
    out = []
    for k, v in batch.items():
        if isinstance(v, TensorBatch):
            out.extend(_flatten(v, prefix + (k,)))
        else:
            out.append((prefix + (k,), v))
    return out


def _unflatten(leaves: dict[tuple, torch.Tensor], batch_size) -> TensorBatch:
    root = TensorBatch({}, batch_size=batch_size)
    for path, t in leaves.items():
        node = root
        for k in path[:-1]:
            if k not in node:
                node._d[k] = TensorBatch({}, batch_size=batch_size)
            node = node._d[k]
        node._d[path[-1]] = t
    return root


class ReplayBuffer:
    """A circular replay buffer (reference: replay_buffer.py:12-138), stored on the GPU.

    :param max_size: Maximum number of transitions to store
    :param device: Device sampled batches are returned on (as in the reference).  Storage
        lives on this device when it is a CUDA device, else on the current CUDA device.
    :param dtype: kept for signature compatibility (the reference ignores it too)
    :param rng: "torch" (reference's CPU mt19937 stream) or "device" (Philox on the GPU)
    :param exact_rng_advance: (default) also skip the ``size-1-batch`` words torch.randperm would
        have consumed, so that a whole seeded RUN — later ``sample`` calls and every other consumer
        of torch's global generator — stays aligned with the reference, not just the first call.
        O(size) host work at ~0.4 ns per word (0.4 ms at 1 M rows; the reference's own randperm is
        ~60 ms there).  ``False`` trades that alignment for throughput.
    """

    def __init__(self, max_size: int, device: str | torch.device = "cpu",
                 dtype: torch.dtype = torch.float32, *, rng: str = "torch",
                 exact_rng_advance: bool = True, seed: int = 0) -> None:
        self.max_size = max_size
        self.device = device
        self.dtype = dtype
        self.counter = 0
        self.initialized = False

        self._cursor = 0
        self._size = 0
        self._storage: TensorBatch | None = None
        self._leaves: dict[tuple, torch.Tensor] = {}
        self._sdev: torch.device | None = None
        assert rng in ("torch", "device")
        self.rng = rng
        self.exact_rng_advance = exact_rng_advance
        self._philox_seed = seed
        self._philox_offset = 0
        self._err: torch.Tensor | None = None
        self._ring: PinnedRing | None = None

    def _stage(self) -> PinnedRing:
        if self._ring is None:
            self._ring = PinnedRing(self.storage_device)
        return self._ring

    # -- properties (reference :39-58) ----------------------------------------------------------
    @property
    def storage(self) -> TensorBatch | None:
        return self._storage

    @property
    def size(self) -> int:
        return self._size

    @size.setter
    def size(self, value: int) -> None:
        self._size = value

    @property
    def is_full(self) -> bool:
        return len(self) == self.max_size

    def __len__(self) -> int:
        return self._size

    @property
    def storage_device(self) -> torch.device:
        if self._sdev is None:
            self._sdev = nv.require_cuda(self.device)
        return self._sdev

    def _out_is_storage(self) -> bool:
        same = getattr(self, "_out_same", None)
        if same is None:
            dev = torch.device(self.device)
            same = dev.type == "cuda" and (dev.index is None or dev == self.storage_device)
            self._out_same = same
        return same

    def _out(self, batch: TensorBatch) -> TensorBatch:
        """Move a sampled batch to ``self.device`` (no-op when that is the storage device)."""
        return batch if self._out_is_storage() else batch.to(torch.device(self.device))

    # -- add (reference :60-112) ---------------------------------------------------------------
    def _init(self, data: TensorBatch) -> None:
        self._leaves = {}
        for path, t in _flatten(data):
            self._leaves[path] = torch.zeros((self.max_size, *t.shape[1:]), dtype=t.dtype,
                                             device=self.storage_device)
        self._storage = _unflatten(self._leaves, [self.max_size])
        self._err = torch.zeros(1, dtype=torch.int32, device=self.storage_device)
        self.initialized = True

    def _prepare(self, data) -> tuple[TensorBatch, int]:
        """``.to(device)`` and the 1-D -> (n,1) reshape of :79-94 (which mutates ``data``)."""
        data = as_batch(data)
        # fast path (the per-env-step case): flat keys, every tensor already on the storage device — `.to` is
        # the identity then and only the reference's in-place 1-D -> (n,1) reshape is left
        d = data._d
        dev = self.storage_device
        for item in d.values():
            if not isinstance(item, torch.Tensor) or item.device != dev:
                break
        else:
            n = data.batch_size[0]
            for key, item in d.items():
                if item.ndim == 1:
                    d[key] = item.reshape(n, 1)
            return data, n
        data = data.to(self.storage_device)
        n = data.shape[0]
        for key, item in list(data.items()):
            if isinstance(item, TensorBatch):
                for k, v in list(item.items()):
                    if v.ndim == 1:
                        item[k] = v.reshape(n, 1)
            elif item.ndim == 1:
                item = item.reshape(n, 1)
            data[key] = item
        return data, n

    def _pairs(self, data: TensorBatch, dst: dict[tuple, torch.Tensor]):
        pairs = []
        for path, t in _flatten(data):
            if path not in dst:
                raise KeyError(f"key {'.'.join(path)} not in buffer storage")
            d = dst[path]
            if t.dtype != d.dtype:
                t = t.to(d.dtype)
            pairs.append((t.contiguous(), d))
        return pairs

    def _insert(self, data: TensorBatch, dst: dict[tuple, torch.Tensor], n: int, cursor: int, rows: int,
                slot: str) -> None:
        """Ring-write the ``n`` rows of ``data`` into the table ``dst`` at ``cursor`` (one launch for every key).
        The launch descriptor of a table — destination pointers, row sizes, dtypes — is built once and kept
        under ``slot``; per call only the source pointers change."""
        tab = self.__dict__.get(slot)
        if tab is None or tab[0] is not dst:
            paths = list(dst.keys())
            keys = (nv.XpKey * len(paths))()
            meta = []
            for i, p in enumerate(paths):
                t = dst[p]
                keys[i].dst = t.data_ptr()
                keys[i].row_bytes = xp.row_bytes(t)
                meta.append((p, p[0] if len(p) == 1 else None, t.dtype, tuple(t.shape[1:])))
            tab = (dst, keys, meta, nv.load().xp_ring_insert, len(paths))
            self.__dict__[slot] = tab
        _, keys, meta, fn, nk = tab
        d = data._d
        if len(d) != nk or n == 0:
            # nested / missing / extra keys, empty batch: the general path (same checks and errors as before)
            with nv.on_device(self.storage_device):
                xp.ring_insert(self._pairs(data, dst), n, cursor, rows)
            return
        keep = []
        for i, (path, flat_key, dtype, tail) in enumerate(meta):
            t = d.get(flat_key) if flat_key is not None else None
            if not isinstance(t, torch.Tensor) or tuple(t.shape[1:]) != tail:
                with nv.on_device(self.storage_device):
                    xp.ring_insert(self._pairs(data, dst), n, cursor, rows)
                return
            if t.dtype != dtype:
                t = t.to(dtype)
                keep.append(t)
            if not t.is_contiguous():
                t = t.contiguous()
                keep.append(t)
            keys[i].src = t.data_ptr()
        dev = self.storage_device
        with nv.on_device(dev):
            nv.check(fn(keys, nk, n, cursor, rows, nv.raw_stream(dev.index)), "xp_ring_insert")

    def add(self, data: DataType) -> None:
        data, n = self._prepare(data)
        if self._storage is None:
            self._init(data)
        self._insert(data, self._leaves, n, self._cursor, self.max_size, "_ins_storage")
        self._advance(n)

    def _add_rows_with(self, n: int, insert) -> None:
        """``add`` of ``n`` rows whose movement into the ring is done by ``insert(cursor)`` — the fused
        exchange + insert of the replicated shared memory (``population.ReplicatedReplay``)."""
        with nv.on_device(self.storage_device):
            insert(self._cursor)
        self._advance(n)

    def _advance(self, n: int) -> None:
        self._cursor = (self._cursor + n) % self.max_size
        self._size = min(self._size + n, self.max_size)
        self.counter += n

    # -- sample (reference :114-131) --------------------------------------------------------------
    def _uniform_indices(self, batch_size: int) -> torch.Tensor:
        """First ``batch_size`` entries of torch.randperm(self.size), computed on the device from
        the raw draws torch would consume (sparse forward Fisher-Yates, exact)."""
        n = self.size
        b = min(int(batch_size), n)
        dev = self.storage_device
        if b == 0:
            return torch.empty(0, dtype=torch.int64, device=dev)
        if self.rng == "torch":
            stream = TorchCpuStream()
            draws = min(b, n - 1)  # randperm draws once per i in [0, n-2]
            raw = torch.zeros(b, dtype=torch.int32)
            if draws:
                raw[:draws] = stream.draw_u32(draws)
            if self.exact_rng_advance and n - 1 > draws:
                stream.skip(n - 1 - draws)
            stream.commit()
            with nv.on_device(dev):
                raw = self._stage().to_device(raw)
        else:
            with nv.on_device(dev):
                raw = xp.philox_u32(self._philox_seed, self._philox_offset, b, dev)
            self._philox_offset += (b + 3) // 4
        with nv.on_device(dev):
            if b > 16384:  # beyond the kernel's shared-memory table: the chunked prefix
                from .rollout_buffer import chunked_randperm_prefix

                return chunked_randperm_prefix(raw, b, n)
            return xp.randperm_prefix(raw, b, 1, n)

    def _gather_plan(self):
        """Per-storage launch descriptor, built once: the ``xp_key`` array with the source
        pointers / row sizes filled in, and the keys grouped by (tail shape, dtype) — the outputs
        of a group are carved from ONE allocation (fewer allocator round trips per sample)."""
        plan = getattr(self, "_gplan", None)
        if plan is None or plan[0] is not self._leaves:
            paths = list(self._leaves.keys())
            keys = (nv.XpKey * len(paths))()
            groups: dict = {}
            for i, p in enumerate(paths):
                t = self._leaves[p]
                keys[i].src = t.data_ptr()
                keys[i].row_bytes = xp.row_bytes(t)
                groups.setdefault((tuple(t.shape[1:]), t.dtype), []).append((i, p, xp.row_bytes(t)))
            flat = all(len(p) == 1 for p in paths)
            plan = (self._leaves, keys, list(groups.items()), flat, nv.load().xp_gather, len(paths))
            self._gplan = plan
        return plan

    def _fast_table(self):
        """Native host fast path for ``storage[idx]`` (flat key sets on the current device), or
        None.  Rebuilt when the storage is re-allocated."""
        ft = getattr(self, "_ftab", None)
        if ft is None or ft[0] is not self._leaves:
            acc = _hostaccel.load()
            obj = None
            if (acc is not None and self._leaves and len(self._leaves) <= 24
                    and all(len(p) == 1 and t.is_cuda for p, t in self._leaves.items())):
                obj = acc.FastGather(list(self._leaves.values()), [p[0] for p in self._leaves],
                                     self.max_size, self._err)
            ft = (self._leaves, obj)
            self._ftab = ft
        return ft[1]

    def _gather(self, indices: torch.Tensor, lead: tuple | None = None) -> TensorBatch:
        """``storage[indices]`` as fresh copies, every key in one launch.  ``lead``: leading
        shape of the result (the shape of ``indices`` in the reference's ``storage[idxs]``)."""
        dev = self.storage_device
        idx = indices
        if not idx.is_cuda and idx.numel():
            # host indices: checked here like `storage[idx]` does (IndexError; Python negative indexing)
            lo, hi = int(idx.min()), int(idx.max())
            if lo < -self.max_size or hi >= self.max_size:
                raise IndexError(f"index {lo if lo < -self.max_size else hi} is out of bounds for dimension 0 "
                                 f"with size {self.max_size}")
            if lo < 0:
                idx = torch.where(idx < 0, idx + self.max_size, idx)
        elif idx.is_cuda:
            # device indices: the kernel zero-fills a row whose index is out of range and raises a sticky flag,
            # read back without blocking every 32 gathers and reported by a later call (check_errors() forces it)
            self._gathers_since_poll = getattr(self, "_gathers_since_poll", 0) + 1
            if getattr(self, "_gerr_pending", None) is not None:
                self._raise_gather_error()
            elif self._gathers_since_poll >= 32 and self._err is not None:
                self._gathers_since_poll = 0
                self._poll_gather_error()
        if idx.device != dev or idx.dtype != torch.int64:
            idx = idx.to(device=dev, dtype=torch.int64)
        if idx.dim() != 1 or not idx.is_contiguous():
            idx = idx.contiguous().view(-1)
        B = idx.shape[0]
        if lead is None:
            lead = (B,)
        fast = self._fast_table()
        if fast is not None and torch._C._cuda_getDevice() == dev.index:
            tb = TensorBatch.__new__(TensorBatch)
            tb._d = fast.gather(idx, lead)
            tb._non_tensor = {}
            tb.batch_size = torch.Size(lead)
            return tb
        _, keys, groups, flat, fn, n_keys = self._gather_plan()
        out = {}
        with nv.on_device(dev):
            for (tail, dt), members in groups:
                if len(members) == 1:
                    i, p, _ = members[0]
                    o = torch.empty((*lead, *tail), dtype=dt, device=dev)
                    keys[i].dst = o.data_ptr()
                    out[p] = o
                else:
                    slab = torch.empty((len(members), *lead, *tail), dtype=dt, device=dev)
                    base = slab.data_ptr()
                    for j, ((i, p, rb), o) in enumerate(zip(members, slab.unbind(0))):
                        keys[i].dst = base + j * B * rb
                        out[p] = o
            if B:
                nv.check(fn(keys, n_keys, idx.data_ptr(), B, self.max_size, self._err.data_ptr(),
                            nv.raw_stream(dev.index)), "xp_gather")
        if flat:  # no nested keys: build the batch directly (in storage key order)
            tb = TensorBatch.__new__(TensorBatch)
            tb._d = {p[0]: out[p] for p in self._leaves}
            tb._non_tensor = {}
            tb.batch_size = torch.Size(lead)
            return tb
        return _unflatten({p: out[p] for p in self._leaves}, list(lead))

    def _poll_gather_error(self) -> None:
        if getattr(self, "_gerr_pin", None) is None:
            self._gerr_pin = torch.zeros(1, dtype=torch.int32).pin_memory()
            self._gerr_ev = torch.cuda.Event()
        with nv.on_device(self.storage_device):
            self._gerr_pin.copy_(self._err, non_blocking=True)
            self._gerr_ev.record()
        self._gerr_pending = True

    def _raise_gather_error(self, wait: bool = False) -> None:
        if wait:
            self._gerr_ev.synchronize()
        if self._gerr_ev.query():
            self._gerr_pending = None
            if int(self._gerr_pin[0]) != 0:
                self._err.zero_()
                raise IndexError(f"a sampled index was out of bounds for dimension 0 with size {self.max_size} "
                                 "(device indices are checked by the gather kernel; the rows came back as zeros)")

    def check_gather_errors(self) -> None:
        """Block until every queued gather ran and raise its IndexError, if any (device-side indices)."""
        if self._err is None:
            return
        self._poll_gather_error()
        self._raise_gather_error(wait=True)

    def sample(self, batch_size: int, return_idx: bool = False) -> TensorBatch:
        indices = self._uniform_indices(batch_size)
        samples = self._gather(indices)
        if return_idx:
            samples["idxs"] = indices
        return self._out(samples)

    def clear(self) -> None:
        self._size = 0
        self._cursor = 0
        self._storage = None
        self._leaves = {}
        self.initialized = False


class MultiStepReplayBuffer(ReplayBuffer):
    """n-step returns (reference: replay_buffer.py:141-258).  The last ``n_step`` env steps live
    in a device stage ring; the fold (fp32, product then sum, "any env done" truncation) and the
    ring insert are one kernel."""

    def __init__(self, max_size: int, n_step: int = 3, gamma: float = 0.99,
                 device: str | torch.device = "cpu", dtype: torch.dtype = torch.float32,
                 **kw) -> None:
        super().__init__(max_size, device, dtype, **kw)
        self.n_step = n_step
        self.gamma = gamma
        self.n_step_buffer: deque[TensorBatch] = deque(maxlen=n_step)
        self.reward_key = "reward"
        self.done_key = None
        self.ns_key = "next_obs"
        self._nstage: dict[tuple, torch.Tensor] = {}
        self._nstage_n = 0
        self._nstage_count = 0
        self._prefetched = None  # (idxs tensor, its version, self.counter, {key: rows}) from a linked PER sample

    def _init_stage(self, data: TensorBatch, n: int) -> None:
        self._nstage = {p: torch.zeros((self.n_step * n, *t.shape[1:]), dtype=t.dtype,
                                      device=self.storage_device)
                       for p, t in _flatten(data)}
        self._nstage_n = n
        # the views of the stage ring handed out as n_step_buffer entries never change: one per slot
        self._slot_views = [_unflatten({p: t[k * n:(k + 1) * n] for p, t in self._nstage.items()}, [n])
                            for k in range(self.n_step)]

    def _detect_keys(self) -> None:
        first = self.n_step_buffer[0]
        assert self.reward_key in first, (
            f"Reward key not found in transition. Expected key: {self.reward_key}")
        assert self.ns_key in first, (
            f"Next observation key not found in transition. Expected key: {self.ns_key}")
        done_key = None
        expected_keys = ["done", "termination", "terminated"]
        for key in expected_keys:
            if key in first:
                done_key = key
                break
        assert done_key is not None, (
            f"No done/termination key found in transition. Expected keys: {expected_keys}")
        self.done_key = done_key

    def _fold_into(self, dst: dict[tuple, torch.Tensor], cursor: int, dst_rows: int) -> None:
        if not self.initialized or self.done_key is None:
            self._detect_keys()
        fd = self.__dict__.get("_fold_desc")
        if fd is not None and fd[0] is self._nstage and fd[1] is dst and fd[2] == (self.n_step, self.gamma, self.reward_key,
                                                                                   self.done_key, self.ns_key):
            # the launch descriptor of this (stage ring, destination table) pair, built by the general path below
            _, _, _, keys, roles, nk, done_idx, done_bytes, gpow, fn = fd
            dev = self.storage_device
            with nv.on_device(dev):
                nv.check(fn(keys, roles, nk, done_idx, done_bytes, self.n_step, self._nstage_count % self.n_step, gpow,
                            self._nstage_n, cursor, dst_rows, nv.raw_stream(dev.index)), "xp_nstep_fold_insert")
            return
        paths = list(self._nstage.keys())
        roles = []
        for p in paths:
            if p[0] == self.reward_key:
                roles.append(nv.XP_ROLE_REWARD)
            elif p[0] in (self.ns_key, self.done_key):
                roles.append(nv.XP_ROLE_LAST)
            else:
                roles.append(nv.XP_ROLE_FIRST)
        done_path = (self.done_key,)
        done_t = self._nstage[done_path]
        if done_t.dtype in (torch.bool, torch.uint8, torch.int8):
            done_bytes = 1
        elif done_t.dtype in (torch.float32, torch.int32):
            done_bytes = 4
        else:
            raise TypeError(f"unsupported done dtype {done_t.dtype}")
        if self._nstage[(self.reward_key,)].dtype != torch.float32:
            raise TypeError("n-step reward must be float32 (Transition casts it, data.py:85-87)")
        gpow = [float(np.float32(self.gamma**i)) for i in range(self.n_step)]
        slot0 = self._nstage_count % self.n_step  # oldest step (the ring is full here)
        with nv.on_device(self.storage_device):
            xp.nstep_fold_insert([(self._nstage[p], dst[p]) for p in paths], roles,
                                 paths.index(done_path), done_bytes, self.n_step, slot0, gpow,
                                 self._nstage_n, cursor, dst_rows)
        import ctypes as _C

        self._fold_desc = (self._nstage, dst, (self.n_step, self.gamma, self.reward_key, self.done_key, self.ns_key),
                           xp.make_keys([(self._nstage[p], dst[p]) for p in paths]), (_C.c_int * len(paths))(*roles),
                           len(paths), paths.index(done_path), done_bytes, (_C.c_float * len(gpow))(*gpow),
                           nv.load().xp_nstep_fold_insert)

    def add(self, data: DataType) -> TensorBatch | None:
        data, n = self._prepare(data)
        if not self._nstage:
            self._init_stage(data, n)
        if n != self._nstage_n:
            raise ValueError(f"n-step buffer was initialised with {self._nstage_n} envs, got {n}")
        slot = self._nstage_count % self.n_step
        self._insert(data, self._nstage, n, slot * n, self.n_step * n, "_ins_stage")
        self._nstage_count += 1
        view = self._slot_views[slot]
        self.n_step_buffer.append(view)

        if len(self.n_step_buffer) < self.n_step:
            return None

        if self._storage is None:
            if self.done_key is None:
                self._detect_keys()
            self._init(view)
        self._fold_into(self._leaves, self._cursor, self.max_size)
        self._advance(n)
        return self.n_step_buffer[0]

    def sample_from_indices(self, idxs: torch.Tensor) -> TensorBatch:
        pf = self._prefetched
        if pf is not None:
            # rows for exactly THESE indices were gathered together with the linked PER buffer's batch
            # (link_n_step): valid while it is the same, unmodified index tensor and nothing was added since
            self._prefetched = None
            if pf[0] is idxs and idxs._version == pf[1] and self.counter == pf[2]:
                tb = TensorBatch.__new__(TensorBatch)
                tb._d = pf[3]
                tb._non_tensor = {}
                tb.batch_size = idxs.shape
                return self._out(tb)
        if not isinstance(idxs, torch.Tensor):
            idxs = torch.as_tensor(idxs)
        # storage[idxs] with (B,1) indices gives [B,1,...] in the reference
        return self._out(self._gather(idxs, tuple(idxs.shape) if idxs.ndim > 1 else None))

    def _get_n_step_info(self) -> TensorBatch:
        """The folded n-step transition of the transitions currently in ``n_step_buffer``, oldest
        first (reference :206-258) — whatever put them there: ``add`` (device views of the stage
        ring) or a direct ``n_step_buffer.append`` as the reference's tests do.  Not the hot path:
        ``add`` folds straight out of the stage ring into the storage."""
        entries = [as_batch(e) for e in self.n_step_buffer]
        assert entries, "n_step_buffer is empty"
        if not self.initialized or self.done_key is None:
            self._detect_keys()
        dev = self.storage_device
        n = entries[0].shape[0] if len(entries[0].shape) else 1
        m = len(entries)
        flat = [dict(_flatten(e.to(dev))) for e in entries]
        paths = list(flat[0].keys())
        stage = {p: torch.cat([f[p].reshape(n, -1) for f in flat], 0).contiguous() for p in paths}
        tmp = {p: torch.empty_like(stage[p][:n]) for p in paths}
        saved = self._nstage, self._nstage_n, self._nstage_count, self.n_step
        try:
            self._nstage, self._nstage_n, self._nstage_count, self.n_step = stage, n, 0, m
            self._fold_into(tmp, 0, n)
        finally:
            self._nstage, self._nstage_n, self._nstage_count, self.n_step = saved
        out = {p: tmp[p].reshape(flat[0][p].shape) for p in paths}
        return self._out(_unflatten(out, list(entries[0].batch_size)))

    def clear(self) -> None:
        super().clear()


class PrioritizedReplayBuffer(ReplayBuffer):
    """Prioritized replay (reference: replay_buffer.py:261-428) over an interleaved (sum, min)
    device tree pair; ``max_priority`` is device-resident so add/update never synchronise."""

    def __init__(self, max_size: int, alpha: float = 0.6, device: str | torch.device = "cpu",
                 dtype: torch.dtype = torch.float32, **kw) -> None:
        super().__init__(max_size, device, dtype, **kw)
        self.alpha = alpha
        self.tree_ptr = 0
        tree_capacity = 1
        while tree_capacity < max_size:
            tree_capacity *= 2
        self._tree_capacity = tree_capacity
        self._trees = None
        self._pending_max_priority = 1.0
        self._sibling = None  # linked n-step buffer (link_n_step)
        self._pending_err = None
        self._err_pin = None
        self._err_ev = None
        self._updates_since_poll = 0

    # trees are created on first use so that constructing a buffer needs no GPU
    def _ensure_trees(self):
        if self._trees is None:
            from .segment_tree import make_pair

            s, m, st = make_pair(self._tree_capacity, self.storage_device)
            self._state = torch.tensor([self._pending_max_priority], dtype=torch.float64,
                                       device=self.storage_device)
            self._trees = (s, m, st, st.handle(None))
            import ctypes as _C

            self._lib = nv.load()
            self._href = _C.byref(self._trees[3])
        return self._trees

    @property
    def sum_tree(self):
        return self._ensure_trees()[0]

    @property
    def min_tree(self):
        return self._ensure_trees()[1]

    @property
    def max_priority(self) -> float:
        if self._trees is None:
            return self._pending_max_priority
        return float(self._state.item())

    @max_priority.setter
    def max_priority(self, v: float) -> None:
        if self._trees is None:
            self._pending_max_priority = float(v)
        else:
            self._state.fill_(float(v))

    def add(self, data: DataType) -> None:
        data, n = self._prepare(data)
        if self._storage is None:
            self._init(data)
        self._ensure_trees()
        self._insert(data, self._leaves, n, self._cursor, self.max_size, "_ins_storage")
        dev = self.storage_device
        with nv.on_device(dev):
            # leaves tree_ptr.. := max_priority ** alpha (:306-309), no host round trip
            nv.check(self._lib.xp_per_add(self._href, self._state.data_ptr(), self.tree_ptr, n, self.max_size,
                                          self.alpha, nv.raw_stream(dev.index)), "xp_per_add")
        self._advance(n)
        self.tree_ptr = (self.tree_ptr + n) % self.max_size

    def _add_rows_with(self, n: int, insert) -> None:
        _, _, st, h = self._ensure_trees()
        with nv.on_device(self.storage_device):
            insert(self._cursor)
            xp.per_add(h, self._state, self.tree_ptr, n, self.max_size, self.alpha)
        self._advance(n)
        self.tree_ptr = (self.tree_ptr + n) % self.max_size

    def _update_priority(self, idx: int, priority: float) -> None:
        assert 0 <= idx < self.max_size
        s, m, st, h = self._ensure_trees()
        v = torch.tensor([float(priority) ** self.alpha], dtype=torch.float64, device=self.storage_device)
        i = torch.tensor([int(idx)], dtype=torch.int64, device=self.storage_device)
        with nv.on_device(self.storage_device):
            xp.tree_set_batch(h, i, v, st.err)
        self.max_priority = max(self.max_priority, priority)

    def _draw_uniforms(self, batch_size: int) -> torch.Tensor:
        dev = self.storage_device
        if self.rng == "torch":
            # batch_size x torch.rand(1).item() == torch.rand(batch_size) on the same stream
            with nv.on_device(dev):
                ring = self._stage()
                slot, host, dv = ring.views(batch_size, torch.float32)
                if torch.rand is _TORCH_RAND:
                    torch.rand(batch_size, out=host)
                else:  # someone replaced torch.rand: make the reference's calls, one per stratum (:377)
                    for i in range(batch_size):
                        host[i] = float(torch.rand(1).item())
                return ring.push(slot, host, dv)
        with nv.on_device(dev):
            raw = xp.philox_u32(self._philox_seed, self._philox_offset, batch_size, dev)
        self._philox_offset += (batch_size + 3) // 4
        return (raw & 0xFFFFFF).to(torch.float32) * (2.0**-24)

    def _sample_proportional(self, batch_size: int, u: torch.Tensor | None = None) -> torch.Tensor:
        _, _, st, h = self._ensure_trees()
        if u is None:
            u = self._draw_uniforms(batch_size)
        u = u.to(device=self.storage_device, dtype=torch.float32).contiguous()
        idx = torch.empty(batch_size, dtype=torch.int64, device=self.storage_device)
        with nv.on_device(self.storage_device):
            xp.per_sample(h, u, batch_size, 1, self.size, 0.0, idx, None)
        return idx.cpu()  # the reference builds these indices on the CPU whatever `device` is (:365)

    def _calculate_weights(self, indices: torch.Tensor, beta: float) -> torch.Tensor:
        _, _, st, h = self._ensure_trees()
        idx = torch.as_tensor(indices).to(device=self.storage_device, dtype=torch.int64).contiguous().view(-1)
        w = torch.empty(idx.numel(), dtype=torch.float32, device=self.storage_device)
        with nv.on_device(self.storage_device):
            xp.per_weights(h, idx, self.size, float(beta), w)
        return w if self._out_is_storage() else w.to(torch.device(self.device))

    def sample(self, batch_size: int, beta: float = 0.4, u: torch.Tensor | None = None) -> TensorBatch:
        """``u`` (optional): the fp32 uniforms to use instead of drawing them."""
        _, _, st, h = self._ensure_trees()
        if self._pending_err is not None:
            self._raise_deferred()
        dev = self.storage_device
        fp = self._fast_per()
        if fp is not None and torch._C._cuda_getDevice() == dev.index:
            if u is None and (self.rng != "torch" or torch.rand is not _TORCH_RAND):
                u = self._draw_uniforms(batch_size)
            if u is None or (u.is_cuda and u.device == dev and u.dtype == torch.float32
                             and u.is_contiguous() and u.numel() == batch_size):
                d = fp.sample(batch_size, float(beta), self._size, u)
                sib_rows = d.pop("__sibling__", None)
                if sib_rows is not None:
                    ix = d["idxs"]
                    self._sibling._prefetched = (ix, ix._version, self._sibling.counter, sib_rows)
                tb = TensorBatch.__new__(TensorBatch)
                tb._d = d
                tb._non_tensor = {}
                tb.batch_size = torch.Size((batch_size,))
                return self._out(tb)
        if u is None:
            u = self._draw_uniforms(batch_size)
        if u.device != dev or u.dtype != torch.float32 or not u.is_contiguous():
            u = u.to(device=dev, dtype=torch.float32).contiguous()
        with nv.on_device(dev):
            idx = torch.empty((batch_size, 1), dtype=torch.int64, device=dev)
            w = torch.empty((batch_size, 1), dtype=torch.float32, device=dev)
            nv.check(self._lib.xp_per_sample(self._href, u.data_ptr(), batch_size, 1, self._size,
                                             float(beta), idx.data_ptr(), w.data_ptr(),
                                             nv.raw_stream(dev.index)), "xp_per_sample")
        samples = self._gather(idx.view(-1))
        samples._d["weights"] = w
        samples._d["idxs"] = idx
        return self._out(samples)

    def _fast_per(self):
        """Native host fast path of sample / update_priorities, or None.  Valid while the storage table, the
        tree pair and the linked sibling's table are the objects it was built for (three identity tests)."""
        fp = self.__dict__.get("_fper")
        sib = self._sibling
        sl = sib._leaves if sib is not None else None
        if fp is not None and fp[0] is self._leaves and fp[1] is self._trees and fp[3] is sl:
            return fp[2]
        tab = self._fast_table()
        if tab is None:
            return None
        stab = None
        if sib is not None and sib._leaves and sib.max_size == self.max_size and sib.storage_device == self.storage_device \
                and len(sib._leaves) + len(self._leaves) <= 24:
            stab = sib._fast_table()
        import ctypes as _C

        _, _, st, h = self._ensure_trees()
        obj = _hostaccel.load().FastPer(tab, _C.addressof(h), st.nodes, st.aux, self._state, st.err,
                                        self.max_size, float(self.alpha))
        obj.set_sibling(stab)
        self._fper = (self._leaves, self._trees, obj, sl)
        self._xchg = None  # a replicated-update exchange bound to the previous object is gone with it
        return obj

    def link_n_step(self, n_step_memory: "MultiStepReplayBuffer | None") -> None:
        """Optional: tell this buffer which ``MultiStepReplayBuffer`` is filled in lock-step with it
        (train_off_policy.py:341-343 adds one row batch to both per env step, so row i of the two is the
        same transition).  ``sample(batch_size, beta)`` then moves the rows of BOTH tables in one gather
        launch and the following ``n_step_memory.sample_from_indices(batch["idxs"])`` — the reference's own
        call, unchanged — is served from it.  Results are identical with or without the link; any other
        use of either buffer falls back to its own gather."""
        self._sibling = n_step_memory

    def _raise_deferred(self) -> None:
        """Out-of-range indices that arrived as DEVICE tensors are detected by the kernel (a sticky
        device flag); the flag is read back without blocking and the AssertionError surfaces on
        a later call into the buffer (``check_errors()`` forces it)."""
        pend = self._pending_err
        if pend is not None and pend[1].query():
            self._pending_err = None
            if int(pend[0][0]) != 0:
                self._trees[2].err.zero_()
                raise AssertionError("priority index out of range [0, max_size)")

    def _poll_err_async(self) -> None:
        if self._err_pin is None:
            self._err_pin = torch.zeros(1, dtype=torch.int32).pin_memory()
            self._err_ev = torch.cuda.Event()
        self._err_pin.copy_(self._trees[2].err, non_blocking=True)
        self._err_ev.record()
        self._pending_err = (self._err_pin, self._err_ev)

    def update_priorities(self, indices, priorities, check: bool = True) -> None:
        """Sequential semantics of :411-428 in one batched pass (last writer wins; every
        priority, overwritten or not, counts towards ``max_priority``).

        Never blocks the host: host-side indices (numpy / CPU tensors) are range-checked on the
        host (AssertionError as in :319); device-side indices are checked by the kernel and the
        error is raised on a later call into the buffer."""
        _, _, st, h = self._ensure_trees()
        dev = self.storage_device
        if self._pending_err is not None:
            self._raise_deferred()
        idx = indices if isinstance(indices, torch.Tensor) else torch.as_tensor(indices)
        on_host = not idx.is_cuda
        if on_host and check and idx.numel():
            assert 0 <= int(idx.min()) and int(idx.max()) < self.max_size, (
                "priority index out of range [0, max_size)")
        if not on_host and torch._C._cuda_getDevice() == dev.index:
            fp = self._fast_per()
            pr = priorities if isinstance(priorities, torch.Tensor) else torch.as_tensor(priorities)
            if (fp is not None and idx.device == dev and idx.dtype == torch.int64 and idx.is_contiguous()
                    and pr.dtype == torch.float32 and pr.is_contiguous()
                    and (not pr.is_cuda or pr.device == dev)):
                fp.update(idx, pr)
                if check:
                    self._updates_since_poll += 1
                    if self._updates_since_poll >= 32 and self._pending_err is None:
                        self._updates_since_poll = 0
                        self._poll_err_async()
                return
        with nv.on_device(dev):
            if on_host:
                idx = self._stage().to_device(idx.to(torch.int64).reshape(-1))
            else:
                if idx.device != dev or idx.dtype != torch.int64:
                    idx = idx.to(device=dev, dtype=torch.int64)
                idx = idx.view(-1) if idx.is_contiguous() else idx.contiguous().view(-1)
            pr = priorities if isinstance(priorities, torch.Tensor) else torch.as_tensor(priorities)
            if pr.dtype != torch.float32:
                # priority.item() of a float64 array keeps fp64; the kernel takes fp32 inputs,
                # which is what the algorithms produce (dqn_rainbow.py:487-488)
                pr = pr.to(torch.float32)
            if pr.is_cuda:
                pr = pr.to(device=dev).contiguous().view(-1)
            else:
                pr = self._stage().to_device(pr.reshape(-1))
            n = min(idx.shape[0], pr.shape[0])  # zip(..., strict=False)
            nv.check(self._lib.xp_per_update(self._href, self._state.data_ptr(), idx.data_ptr(),
                                             pr.data_ptr(), n, self.max_size, self.alpha,
                                             st.err.data_ptr(), nv.raw_stream(dev.index)),
                     "xp_per_update")
            if check and not on_host:
                self._updates_since_poll += 1
                if self._updates_since_poll >= 32 and self._pending_err is None:
                    self._updates_since_poll = 0
                    self._poll_err_async()

    def check_errors(self) -> None:
        """Block until every queued update ran and raise its index error, if any."""
        if self._trees is None:
            return
        with nv.on_device(self.storage_device):
            self._poll_err_async()
        self._pending_err[1].synchronize()
        self._raise_deferred()
