"""Device-resident on-policy rollout storage with the API of
agilerl/components/rollout_buffer.py (RolloutBuffer, :26-970).

The ``[capacity, num_envs]`` table lives in HBM (the reference keeps it on the CPU and re-uploads
it on every learn step, :229-233,:577).  ``add`` writes one time row for every key in one launch,
``compute_returns_and_advantages`` is the reverse-time scan kernel (GAE / Monte-Carlo with the
reference's exact mixed fp32/fp64 rounding sequence), ``get_tensor_batch`` returns views
(``batch_size=None``) or a kernel gather of a ``torch.randperm`` prefix.

``RolloutArena`` allocates the buffers of a whole population as one ``[P, capacity, num_envs]``
slab so that the returns of every agent are computed by ONE launch.
"""

from __future__ import annotations

import warnings
from collections import OrderedDict
from enum import Enum
from typing import Any

import numpy as np
import torch

from .. import _native as nv
from .. import xp
from .rng import TorchCpuStream
from .spaces import get_num_actions, get_obs_shape, is_dict_space
from .tensor_batch import TensorBatch


class BPTTSequenceType(Enum):
    """agilerl/typing.py:160-174."""

    CHUNKED = "chunked"
    MAXIMUM = "maximum"
    FIFTY_PERCENT_OVERLAP = "fifty_percent_overlap"


def _field_layout(observation_space, action_space, recurrent, hidden_state_architecture):
    """name -> (trailing shape, dtype) of every pre-allocated key (reference :137-233)."""
    obs_shape = get_obs_shape(observation_space)
    num_actions = get_num_actions(action_space)
    fields: "OrderedDict[tuple, tuple]" = OrderedDict()
    if is_dict_space(observation_space):
        for key, shape in obs_shape.items():
            fields[("observations", key)] = (tuple(shape), torch.float32)
        for key, shape in obs_shape.items():
            fields[("next_observations", key)] = (tuple(shape), torch.float32)
    else:
        fields[("observations",)] = (tuple(obs_shape), torch.float32)
        fields[("next_observations",)] = (tuple(obs_shape), torch.float32)
    fields[("actions",)] = ((num_actions,), torch.float32)
    fields[("rewards",)] = ((), torch.float32)
    fields[("dones",)] = ((), torch.bool)
    fields[("values",)] = ((), torch.float32)
    fields[("log_probs",)] = ((), torch.float32)
    fields[("advantages",)] = ((), torch.float32)
    fields[("returns",)] = ((), torch.float32)
    if recurrent:
        if hidden_state_architecture is None:
            raise ValueError("hidden_state_architecture must be provided if recurrent=True")
        for key, arch in hidden_state_architecture.items():
            fields[("hidden_states", key)] = ((arch[0], arch[2]), torch.float32)
    return fields


def _nest(leaves: "dict[tuple, torch.Tensor]", batch_size) -> TensorBatch:
    root = TensorBatch({}, batch_size=batch_size)
    for path, t in leaves.items():
        node = root
        for k in path[:-1]:
            if k not in node:
                node._d[k] = TensorBatch({}, batch_size=batch_size)
            node = node._d[k]
        node._d[path[-1]] = t
    return root


def _flat(batch: TensorBatch, prefix=()):
    for k, v in batch.items():
        if isinstance(v, TensorBatch):
            yield from _flat(v, prefix + (k,))
        else:
            yield prefix + (k,), v


class RolloutBuffer:
    """Reference: agilerl/components/rollout_buffer.py:26-970 (same constructor signature)."""

    def __init__(
        self,
        capacity: int,
        observation_space,
        action_space,
        num_envs: int = 1,
        device: str = "cpu",
        gae_lambda: float = 0.95,
        gamma: float = 0.99,
        recurrent: bool = False,
        hidden_state_architecture: dict[str, tuple[int, int, int]] | None = None,
        use_gae: bool = True,
        wrap_at_capacity: bool = False,
        max_seq_len: int | None = None,
        bptt_sequence_type: BPTTSequenceType = BPTTSequenceType.CHUNKED,
        *,
        _slab: "dict[tuple, torch.Tensor] | None" = None,
    ) -> None:
        self.capacity = capacity
        self.observation_space = observation_space
        self.action_space = action_space
        self.num_envs = num_envs
        self.device = device
        self.gamma = gamma
        self.gae_lambda = gae_lambda
        self.recurrent = recurrent
        self.hidden_state_architecture = hidden_state_architecture
        self.use_gae = use_gae
        self.wrap_at_capacity = wrap_at_capacity
        self.max_seq_len = max_seq_len
        self.bptt_sequence_type = bptt_sequence_type

        # keep torch's CPU generator where the reference leaves it after get_tensor_batch(batch_size)
        # (torch.randperm consumes total-1 words, :570); class-level default, an instance may set False
        self.exact_rng_advance = True
        self.pos = 0
        self.full = False
        self.num_sequences = None
        self.max_sequence_length = None
        self.unpadded_slices = None
        self.padded_data = None
        self.unpadded_data = None
        self._sdev = nv.require_cuda(device)
        self._initialize_buffers(_slab)

    # -- bookkeeping (reference :100-116) ---------------------------------------------------------
    def size(self) -> int:
        return (self.capacity if self.full else self.pos) * self.num_envs

    def reset(self) -> None:
        self.pos = 0
        self.full = False
        self.num_sequences = None
        self.max_sequence_length = None
        self.unpadded_slices = None
        self.padded_data = None
        self.unpadded_data = None

    @property
    def storage_device(self) -> torch.device:
        return self._sdev

    def _initialize_buffers(self, slab=None) -> None:
        fields = _field_layout(self.observation_space, self.action_space, self.recurrent,
                                self.hidden_state_architecture)
        self._leaves: "OrderedDict[tuple, torch.Tensor]" = OrderedDict()
        for path, (shape, dtype) in fields.items():
            if slab is not None:
                self._leaves[path] = slab[path]
            else:
                self._leaves[path] = torch.zeros((self.capacity, self.num_envs, *shape), dtype=dtype,
                                                 device=self._sdev)
        self.buffer = _nest(self._leaves, [self.capacity, self.num_envs])

    def _lazy_key(self, path: tuple, trailing: tuple, dtype, fill) -> torch.Tensor:
        if path not in self._leaves:
            self._leaves[path] = torch.full((self.capacity, self.num_envs, *trailing), fill,
                                            dtype=dtype, device=self._sdev)
            self.buffer = _nest(self._leaves, [self.capacity, self.num_envs])
        return self._leaves[path]

    # -- add (reference :235-411) -------------------------------------------------------------------
    def _row(self, value, dtype, trailing: tuple) -> torch.Tensor:
        t = torch.as_tensor(value)
        if t.device != self._sdev or t.dtype != dtype:
            t = t.to(device=self._sdev, dtype=dtype, non_blocking=True)
        return t.reshape(self.num_envs, *trailing).contiguous()

    def add(self, obs, action, reward, done, value, log_prob, next_obs=None, hidden_state=None,
            next_hidden_state=None, episode_start=None, action_mask=None) -> None:
        if self.pos == self.capacity:
            if self.wrap_at_capacity:
                self.pos = 0
            else:
                msg = (
                    f"Buffer has reached capacity ({self.capacity} transitions) but received more transitions. "
                    "Either increase buffer capacity or set wrap_at_capacity=True."
                )
                raise ValueError(msg)

        L = self._leaves
        pairs = []

        def put(path, value, dtype=None, trailing=None):
            dst = L[path]
            pairs.append((self._row(value, dst.dtype if dtype is None else dtype,
                                    tuple(dst.shape[2:]) if trailing is None else trailing), dst))

        if is_dict_space(self.observation_space):
            for key, item in obs.items():
                put(("observations", key), item)
            if next_obs is not None:
                for key, item in next_obs.items():
                    put(("next_observations", key), item)
        else:
            put(("observations",), obs)
            if next_obs is not None:
                put(("next_observations",), next_obs)
        put(("actions",), action)
        put(("rewards",), reward)
        put(("dones",), done)
        put(("values",), value)
        put(("log_probs",), log_prob)

        es = self._lazy_key(("episode_starts",), (), torch.bool, False)
        if episode_start is not None:
            pairs.append((self._row(episode_start, torch.bool, ()), es))
        else:
            pairs.append((torch.zeros(self.num_envs, dtype=torch.bool, device=self._sdev), es))

        if action_mask is not None:
            m = torch.as_tensor(action_mask).to(device=self._sdev, dtype=torch.bool).reshape(self.num_envs, -1)
            dst = self._lazy_key(("action_masks",), (m.shape[-1],), torch.bool, True)
            pairs.append((m.contiguous(), dst))

        if self.recurrent and hidden_state is not None:
            for key, t in hidden_state.items():
                # PPO hands (layers, num_envs, size); stored as (num_envs, layers, size) (:381-392)
                put(("hidden_states", key), torch.as_tensor(t).permute(1, 0, 2))

        with nv.on_device(self._sdev):
            xp.ring_insert(pairs_as_rows(pairs), 1, self.pos, self.capacity)

        self.pos += 1
        if self.pos == self.capacity:
            self.full = True

    # -- returns (reference :413-481) ---------------------------------------------------------------
    def compute_returns_and_advantages(self, last_value, last_done) -> None:
        buffer_size = self.capacity if self.full else self.pos
        lv = _as_f64(last_value, self.num_envs, self._sdev)
        ld = _as_f64(last_done, self.num_envs, self._sdev)
        L = self._leaves
        with nv.on_device(self._sdev):
            xp.gae(L[("rewards",)], L[("values",)], L[("dones",)], lv, ld, L[("advantages",)],
                   L[("returns",)], buffer_size, self.num_envs, 1, self.capacity * self.num_envs,
                   float(self.gamma), float(self.gae_lambda), bool(self.use_gae))

    # -- retrieval (reference :483-577) -------------------------------------------------------------
    def _flat_views(self, buffer_size: int) -> "OrderedDict[tuple, torch.Tensor]":
        """[buffer_size * num_envs, ...] views, t-major (flat = t * num_envs + env)."""
        n = buffer_size * self.num_envs
        return OrderedDict((p, t[:buffer_size].reshape(n, *t.shape[2:])) for p, t in self._leaves.items())

    def _gather_flat(self, flat, idx: torch.Tensor) -> "OrderedDict[tuple, torch.Tensor]":
        B = idx.numel()
        out = OrderedDict((p, torch.empty((B, *s.shape[1:]), dtype=s.dtype, device=s.device))
                          for p, s in flat.items())
        n_rows = next(iter(flat.values())).shape[0]
        with nv.on_device(self._sdev):
            xp.gather([(flat[p], out[p]) for p in flat], idx, n_rows)
        return out

    def _randperm_prefix(self, total: int, batch_size: int) -> torch.Tensor:
        """``torch.randperm(total)[:batch_size]`` (:570) from torch's own CPU stream, on device."""
        stream = TorchCpuStream()
        draws = min(batch_size, total - 1)
        raw = torch.zeros(batch_size, dtype=torch.int32)
        if draws:
            raw[:draws] = stream.draw_u32(draws)
        if getattr(self, "exact_rng_advance", True) and total - 1 > draws:
            stream.skip(total - 1 - draws)  # torch.randperm consumes total-1 words: stay aligned for the run
        stream.commit()
        with nv.on_device(self._sdev):
            return chunked_randperm_prefix(raw.to(self._sdev), batch_size, total)

    def get_tensor_batch(self, batch_size: int | None = None, device: str | None = None):
        target_device = device or self.device
        buffer_size = self.capacity if self.full else self.pos
        total_samples = buffer_size * self.num_envs
        if total_samples == 0:
            return TensorBatch({}, batch_size=[0])  # {} in the reference; PPO calls .is_empty() on it
        flat = self._flat_views(buffer_size)
        if batch_size is not None:
            if batch_size > total_samples:
                warnings.warn(
                    f"Batch size {batch_size} is larger than buffer_size {total_samples}. "
                    "Returning all data.", stacklevel=2)
                return self._to_target(_nest(flat, [total_samples]), target_device)
            idx = self._randperm_prefix(total_samples, batch_size)
            return self._to_target(_nest(self._gather_flat(flat, idx), [batch_size]), target_device)
        return self._to_target(_nest(flat, [total_samples]), target_device)

    def _to_target(self, batch: TensorBatch, target) -> TensorBatch:
        dev = torch.device(target)
        if dev.type == "cuda" and (dev.index is None or dev == self._sdev):
            return batch  # already resident: views, no H2D re-upload (the reference copies, :577)
        return batch.to(dev)

    def get(self, batch_size: int | None = None) -> dict:
        buffer_size = self.capacity if self.full else self.pos
        total_samples = buffer_size * self.num_envs
        if total_samples == 0:
            return {}
        flat = self._flat_views(buffer_size)
        if batch_size is not None:
            if batch_size > total_samples:
                warnings.warn(
                    f"Batch size {batch_size} is larger than buffer size {total_samples}. Returning all data.",
                    stacklevel=2)
                return self._convert_td_to_np_dict(_nest(flat, [total_samples]))
            indices = np.random.choice(total_samples, size=batch_size, replace=False)  # :519
            idx = torch.from_numpy(indices).to(self._sdev)
            return self._convert_td_to_np_dict(_nest(self._gather_flat(flat, idx), [batch_size]))
        return self._convert_td_to_np_dict(_nest(flat, [total_samples]))

    def _convert_td_to_np_dict(self, td: TensorBatch) -> dict:
        np_dict = {}
        for key, value in td.items():
            if isinstance(value, TensorBatch):
                np_dict[key] = {k: v.cpu().numpy() for k, v in value.items()}
            else:
                np_dict[key] = value.cpu().numpy()
        return np_dict

    # -- truncated-BPTT sequences (reference :627-922) ----------------------------------------------
    @staticmethod
    def _pad_sequences(sequences, target_length: int | None = None):
        """Zero-pad a list of [len_i, ...] tensors (or batches of them) to [B, T, ...] (:627-675).
        Host-level utility kept for API parity; ``prepare_sequence_tensors`` pads inside the row
        gather instead."""
        if isinstance(sequences[0], TensorBatch):
            keys = list(sequences[0].keys())
            return TensorBatch({k: RolloutBuffer._pad_sequences([s[k] for s in sequences], target_length)
                                for k in keys})
        if target_length is not None:
            out = []
            for seq in sequences:
                pad = target_length - seq.size(0)
                out.append(torch.nn.functional.pad(seq, [0, 0] * (seq.dim() - 1) + [0, pad]))
            return torch.stack(out)
        return torch.nn.utils.rnn.pad_sequence(sequences, batch_first=True, padding_value=0)

    def _get_complete_sequences(self, data, episode_done_indices):
        """Split ``data`` [T, num_envs, ...] into per-episode sequences (:677-722); API parity
        helper over plain indexing (the batched path is ``xp_bptt_plan``)."""
        msl = self.max_seq_len
        sequences, max_length = [], 1
        for env_idx in range(self.num_envs):
            start = 0
            for done_index in episode_done_indices[env_idx]:
                episode = data[start:done_index + 1, env_idx]
                n = len(episode)
                if msl is not None and n >= msl:
                    sequences.extend(_cut_episode(episode, msl, self.bptt_sequence_type))
                else:
                    sequences.append(episode)
                max_length = max(max_length, n)
                start = done_index + 1
        return sequences, max_length

    def prepare_sequence_tensors(self, device: str | None = None) -> None:
        """All sequences of the buffer, zero-padded to a common length (:724-846): fills
        ``padded_data`` (batch [num_sequences * max_sequence_length]: observations, actions,
        [action_masks], pad_mask + non-tensor ``initial_hidden_states``), ``unpadded_data``
        (log_probs, advantages, values, returns, flat t-major) and ``unpadded_slices``."""
        if not self.recurrent:
            raise ValueError("prepare_sequence_tensors() is only supported when recurrent=True.")
        if self.size() == 0:
            raise ValueError("Attempting to prepare sequences with empty buffer.")
        target_device = device or self.device
        buffer_size = self.capacity if self.full else self.pos
        N = self.num_envs
        msl = self.max_seq_len
        kind = BPTTSequenceType(self.bptt_sequence_type).value
        if msl is not None:
            assert msl > 0, "max_seq_len must be greater than 0"
            if kind == "fifty_percent_overlap" and msl == 1:
                raise ValueError("range() arg 3 must not be zero")
        flat = self._flat_views(buffer_size)
        with nv.on_device(self._sdev):
            desc, off, header = xp.bptt_plan(self._leaves[("dones",)], buffer_size, N, msl, kind)
            # output shapes depend on the data: one small read-back (the reference's
            # dones.nonzero().tolist() per env is the same dependency)
            S, L, max_ep, total = (int(x) for x in header.cpu())
            pad_idx, pad_mask, unp_idx, first = xp.bptt_indices(desc, off, S, L, N, total)
            self._seq_desc = desc[:, :S].cpu().numpy()  # host copy (t0, env, len) for the minibatch plans
            pad_paths = [p for p in flat if p[0] in ("observations", "actions")]
            pad_paths += [p for p in flat if p[0] == "action_masks"]
            n_rows = buffer_size * N
            padded = OrderedDict((p, torch.empty((S * L, *flat[p].shape[1:]), dtype=flat[p].dtype,
                                                 device=self._sdev)) for p in pad_paths)
            xp.gather([(flat[p], padded[p]) for p in pad_paths], pad_idx, n_rows)  # -1 rows -> zeros
            hid_paths = [p for p in flat if p[0] == "hidden_states"]
            init = OrderedDict((p, torch.empty((S, *flat[p].shape[1:]), dtype=flat[p].dtype, device=self._sdev))
                               for p in hid_paths)
            if hid_paths:
                xp.gather([(flat[p], init[p]) for p in hid_paths], first, n_rows)
            unp = OrderedDict((p, flat[p].clone()) for p in
                              (("log_probs",), ("advantages",), ("values",), ("returns",)))
        padded[("pad_mask",)] = pad_mask
        self.unpadded_slices = _LazySlices(self._seq_desc, N)
        self.max_sequence_length = min(max_ep, msl) if msl is not None else max_ep
        self.num_sequences = S
        pd = self._to_target(_nest(padded, [S * L]), target_device)
        if hid_paths:
            pd.set_non_tensor("initial_hidden_states",
                              OrderedDict((p[1], t.to(pd.device)) for p, t in init.items()))
        self.padded_data = pd
        self.unpadded_data = self._to_target(_nest(unp, [n_rows]), target_device)

    def get_minibatch_sequences(self, batch_size: int):
        """Yield (padded, unpadded) minibatches of whole sequences (:848-922); the sequence order
        is ``torch.randperm(num_sequences)`` from torch's CPU generator, as in the reference."""
        if self.unpadded_slices is None:
            raise ValueError("Attempting to fetch minibatches before preparing sequences. "
                             "Call prepare_sequence_tensors() first.")
        S, L, N = self.num_sequences, self.max_sequence_length, self.num_envs
        if batch_size > S:
            warnings.warn(f"Batch size {batch_size} is larger than the number of sequences "
                          f"({S}), using batch_size = {S}.", stacklevel=2)
            batch_size = S
        sequence_indices = torch.randperm(S)
        sel_host = sequence_indices.numpy()
        t0, env, ln = (self._seq_desc[i].astype(np.int64) for i in range(3))
        pdata, udata = self.padded_data, self.unpadded_data
        pdev = pdata.device if pdata.device is not None else self._sdev
        on_dev = torch.device(pdev).type == "cuda"
        p_leaves = OrderedDict(_flat(pdata))
        u_leaves = OrderedDict(_flat(udata))
        init = pdata.get_non_tensor("initial_hidden_states", None)
        sel_dev = sequence_indices.to(pdev) if on_dev else sequence_indices
        for start in range(0, S, batch_size):
            end = min(start + batch_size, S)
            sel = sel_host[start:end]
            k = end - start
            lens = ln[sel]
            # flat (t * N + env) steps of the selected sequences, concatenated (:889-900)
            first = np.repeat(t0[sel] * N + env[sel], lens)
            within = np.arange(int(lens.sum())) - np.repeat(np.cumsum(lens) - lens, lens)
            unp_idx = torch.from_numpy(first + within * N)
            if on_dev:
                with nv.on_device(self._sdev):
                    sd = sel_dev[start:end]
                    rows = OrderedDict((p, t.view(S, -1)) for p, t in p_leaves.items())
                    out = OrderedDict((p, torch.empty((k * L, *t.shape[1:]), dtype=t.dtype, device=t.device))
                                      for p, t in p_leaves.items())
                    xp.gather([(rows[p], out[p].view(k, -1)) for p in rows], sd, S)
                    padded = _nest(out, [k * L])
                    if init is not None:
                        hs = OrderedDict((h, torch.empty((k, *t.shape[1:]), dtype=t.dtype, device=t.device))
                                         for h, t in init.items())
                        xp.gather([(init[h], hs[h]) for h in init], sd, S)
                        padded.set_non_tensor("initial_hidden_states", dict(hs))
                    ui = unp_idx.to(self._sdev, non_blocking=True)
                    uo = OrderedDict((p, torch.empty((ui.numel(), *t.shape[1:]), dtype=t.dtype, device=t.device))
                                     for p, t in u_leaves.items())
                    xp.gather([(u_leaves[p], uo[p]) for p in u_leaves], ui, next(iter(u_leaves.values())).shape[0])
                    unpadded = _nest(uo, [ui.numel()])
            else:  # data was requested on the host: plain indexing of the host copies
                sq = sequence_indices[start:end]
                pidx = (sq.view(-1, 1) * L + torch.arange(L)).view(-1)
                padded = pdata[pidx]
                if init is not None:
                    padded.set_non_tensor("initial_hidden_states", {h: t[sq].clone() for h, t in init.items()})
                unpadded = udata[unp_idx]
            yield padded, unpadded

    # -- pickling (reference :924-970; clone() deep-copies the buffer) ------------------------------
    def __getstate__(self) -> dict[str, Any]:
        state = self.__dict__.copy()
        state["_leaves"] = OrderedDict((p, t.cpu()) for p, t in self._leaves.items())
        for k in ("buffer", "padded_data", "unpadded_data"):
            state[k] = None
        state["_sdev"] = str(self._sdev)
        return state

    def __setstate__(self, state: dict[str, Any]) -> None:
        """Unpickling with the reference's integrity checks (:940-970): a missing / invalid buffer is
        re-initialised with a warning (or left alone, with a second warning, when the spaces are
        missing too); a buffer whose leading shape is not ``[capacity, num_envs]`` is re-initialised."""
        self.__dict__.update(state)
        self._sdev = nv.require_cuda(state.get("_sdev"))
        leaves = state.get("_leaves")
        given = state.get("buffer", "missing")
        valid = (isinstance(leaves, dict) and len(leaves) > 0 and all(isinstance(t, torch.Tensor) for t in leaves.values())
                 and (given is None or isinstance(given, TensorBatch)))
        if not valid:
            warnings.warn("TensorDict buffer not found or invalid during unpickling. Re-initializing.", stacklevel=2)
            if not hasattr(self, "observation_space") or not hasattr(self, "action_space"):
                warnings.warn("Observation or action space missing during RolloutBuffer unpickling. "
                              "Buffer might be invalid.", stacklevel=2)
                return
            self._initialize_buffers()
            return
        expected = (self.capacity, self.num_envs)
        if any(tuple(t.shape[:2]) != expected for t in leaves.values()):
            got = tuple(next(iter(leaves.values())).shape[:2])
            warnings.warn(f"Loaded TensorDict has batch_size {torch.Size(got)}, expected {torch.Size(expected)}. "
                          "Re-initializing.", stacklevel=2)
            self._initialize_buffers()
            return
        self._leaves = OrderedDict((p, t.to(self._sdev)) for p, t in leaves.items())
        self.buffer = _nest(self._leaves, [self.capacity, self.num_envs])


def _cut_episode(episode, max_seq_len: int, sequence_type):
    """extract_sequences_from_episode (utils/algo_utils.py:214-265) on one episode."""
    n = len(episode)
    assert max_seq_len > 0, "max_seq_len must be greater than 0"
    assert n > 0, "episode must be non-empty"
    assert max_seq_len <= n, "max_seq_len must be less than or equal to the length of the episode"
    kind = BPTTSequenceType(sequence_type)
    if kind == BPTTSequenceType.CHUNKED:
        starts = range(0, max(1, n // max_seq_len) * max_seq_len, max_seq_len)
    elif kind == BPTTSequenceType.MAXIMUM:
        starts = range(n - max_seq_len + 1)
    else:
        starts = range(0, n - max_seq_len + 1, max_seq_len // 2)
    return [episode[s:s + max_seq_len] for s in starts]


class _LazySlices:
    """``unpadded_slices`` (:752-762): per sequence, the flat ``t * num_envs + env`` indices of its
    steps — materialised on demand from the (t0, env, len) descriptors."""

    def __init__(self, desc: np.ndarray, num_envs: int) -> None:
        self._d, self._n = desc, num_envs

    def __len__(self) -> int:
        return self._d.shape[1]

    def __getitem__(self, i: int) -> torch.Tensor:
        if not -len(self) <= i < len(self):
            raise IndexError(i)
        t0, env, ln = (int(x) for x in self._d[:, i])
        return torch.arange(t0, t0 + ln) * self._n + env

    def __iter__(self):
        return (self[i] for i in range(len(self)))


def pairs_as_rows(pairs):
    """(row [N,...], table [capacity, N, ...]) -> (src [1, N, ...], table) for the ring kernel."""
    return [(src.unsqueeze(0), dst) for src, dst in pairs]


def _as_f64(x, n: int, device) -> torch.Tensor:
    """last_value / last_done as exact fp64 [n] on the device (``.astype(float)``, :452-453)."""
    if isinstance(x, torch.Tensor):
        t = x.detach()
    else:
        t = torch.from_numpy(np.ascontiguousarray(np.asarray(x)))
    return t.to(device=device, dtype=torch.float64).reshape(n).contiguous()


_PREFIX_MAX = 16384


def chunked_randperm_prefix(raw: torch.Tensor, batch_size: int, total: int) -> torch.Tensor:
    """Exact randperm prefix; one kernel launch up to 16384 draws (the kernel's shared-memory
    sort limit), beyond that the swap map is replayed on the host (rare: PPO minibatches are
    drawn by index arrays, algorithms/ppo.py:838-848)."""
    if batch_size <= _PREFIX_MAX:
        return xp.randperm_prefix(raw, batch_size, 1, total)
    lib = nv.load()
    r = raw.cpu().numpy().view(np.uint32).copy()
    out = np.empty(batch_size, dtype=np.int64)
    disp: dict[int, int] = {}
    for i in range(batch_size):
        if i >= total - 1:
            out[i] = disp.get(i, i)
            continue
        j = i + int(r[i]) % (total - i)
        vi, vj = disp.get(i, i), disp.get(j, j)
        out[i] = vj
        disp[j] = vi
    del lib
    return torch.from_numpy(out).to(raw.device)


class RolloutArena:
    """The rollout buffers of a whole population as one ``[P, capacity, num_envs, ...]`` slab.

    ``arena.buffers[i]`` is a regular :class:`RolloutBuffer` whose tables are views of the slab;
    ``arena.compute_returns_and_advantages(last_values, last_dones)`` runs GAE / MC for every
    agent in ONE kernel launch (the reference walks agents sequentially,
    training/train_on_policy.py:230-245).
    """

    def __init__(self, population: int, capacity: int, observation_space, action_space,
                 num_envs: int = 1, device: str = "cpu", **kwargs) -> None:
        self.population = population
        self.capacity = capacity
        self.num_envs = num_envs
        self._sdev = nv.require_cuda(device)
        fields = _field_layout(observation_space, action_space, kwargs.get("recurrent", False),
                               kwargs.get("hidden_state_architecture"))
        self.slab = OrderedDict(
            (path, torch.zeros((population, capacity, num_envs, *shape), dtype=dtype, device=self._sdev))
            for path, (shape, dtype) in fields.items())
        self.buffers = [
            RolloutBuffer(capacity, observation_space, action_space, num_envs=num_envs, device=device,
                          _slab={p: t[i] for p, t in self.slab.items()}, **kwargs)
            for i in range(population)
        ]
        self.gamma = kwargs.get("gamma", 0.99)
        self.gae_lambda = kwargs.get("gae_lambda", 0.95)
        self.use_gae = kwargs.get("use_gae", True)
        self._foreign: set[int] = set()  # slots whose buffer does not live in the slab (adopt)

    # -- keeping the one-launch path alive through evolution -------------------------------------------------
    @classmethod
    def from_buffers(cls, buffers: "list[RolloutBuffer]") -> "RolloutArena":
        """Re-slab a population after a generation: an arena shaped like ``buffers[0]`` that ADOPTS every
        buffer of the list (``[agent.rollout_buffer for agent in pop]``).  Buffers whose capacity or layout
        differs (a mutated ``learn_step``, algorithms/ppo.py:336-339) stay standalone and get their own
        launch."""
        b0 = buffers[0]
        arena = cls(len(buffers), b0.capacity, b0.observation_space, b0.action_space, num_envs=b0.num_envs,
                    device=b0.device, gae_lambda=b0.gae_lambda, gamma=b0.gamma, recurrent=b0.recurrent,
                    hidden_state_architecture=b0.hidden_state_architecture, use_gae=b0.use_gae,
                    wrap_at_capacity=b0.wrap_at_capacity, max_seq_len=b0.max_seq_len,
                    bptt_sequence_type=b0.bptt_sequence_type)
        for i, b in enumerate(buffers):
            arena.adopt(i, b)
        return arena

    def _fits(self, buffer: "RolloutBuffer") -> bool:
        if buffer.capacity != self.capacity or buffer.num_envs != self.num_envs or buffer._sdev != self._sdev:
            return False
        return all(p in buffer._leaves and buffer._leaves[p].shape == t.shape[1:] and buffer._leaves[p].dtype == t.dtype
                   for p, t in self.slab.items())

    def adopt(self, index: int, buffer: "RolloutBuffer") -> "RolloutBuffer":
        """Population slot ``index`` is served by ``buffer`` from now on — the SAME object the agent holds:
        a deep copy made by ``agent.clone()`` (algorithms/core/base.py:444-503,871-937) or a fresh buffer from
        ``create_rollout_buffer`` after a mutation (algorithms/ppo.py:336-339,360-378).  Its stored steps are
        copied into the slab slot and its tables are re-bound to slab views, so one launch keeps covering the
        whole population.  A buffer of another capacity / layout is kept as it is and computed by a launch of
        its own.  Returns ``buffer``."""
        if not 0 <= index < self.population:
            raise IndexError(index)
        self.buffers[index] = buffer
        self._foreign.discard(index)
        if not self._fits(buffer):
            self._foreign.add(index)
            return buffer
        with nv.on_device(self._sdev):
            for path, t in self.slab.items():
                src = buffer._leaves[path]
                if src.data_ptr() != t[index].data_ptr():
                    t[index].copy_(src)
                buffer._leaves[path] = t[index]
        buffer.buffer = _nest(buffer._leaves, [buffer.capacity, buffer.num_envs])
        return buffer

    def compute_returns_and_advantages(self, last_values, last_dones) -> None:
        """last_values / last_dones: ``[P, num_envs]`` (tensor or array). All agents must hold the
        same number of steps (they do: one collection phase per generation)."""
        P, N = self.population, self.num_envs
        members = [i for i in range(P) if i not in self._foreign]
        lv = _as_f64(last_values, P * N, self._sdev)
        ld = _as_f64(last_dones, P * N, self._sdev)
        if members:
            sizes = {self.buffers[i].capacity if self.buffers[i].full else self.buffers[i].pos for i in members}
            if len(sizes) != 1:
                raise ValueError(f"agents hold different numbers of steps: {sorted(sizes)}")
            T = sizes.pop()
            S = self.slab
            # ONE launch over every slot of the slab (a slot whose buffer is standalone is computed too and
            # ignored: cheaper than a second launch)
            with nv.on_device(self._sdev):
                xp.gae(S[("rewards",)], S[("values",)], S[("dones",)], lv, ld, S[("advantages",)],
                       S[("returns",)], T, N, P, self.capacity * N, float(self.gamma),
                       float(self.gae_lambda), bool(self.use_gae))
        for i in sorted(self._foreign):
            self.buffers[i].compute_returns_and_advantages(lv.view(P, N)[i], ld.view(P, N)[i])
