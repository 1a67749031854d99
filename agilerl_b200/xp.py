"""Tensor-level wrappers over the C-ABI (include/b200xp.h).

Each function marshals torch CUDA tensors to raw device pointers, launches on the current CUDA
stream of the tensors' device and returns without synchronising.  No arithmetic happens here.
"""

from __future__ import annotations

import ctypes as C
from typing import Sequence

import torch

from . import _native as nv
from ._native import XpKey, XpTree, check, ptr


def _stream(t: torch.Tensor) -> int:
    return torch._C._cuda_getCurrentRawStream(t.device.index)


def _need(t: torch.Tensor, dtype, name: str) -> torch.Tensor:
    if t.dtype != dtype:
        raise TypeError(f"{name}: expected {dtype}, got {t.dtype}")
    if not t.is_cuda:
        raise nv.NativeError(f"{name}: tensor must live on a CUDA device (no CPU fallback)")
    if not t.is_contiguous():
        raise ValueError(f"{name}: tensor must be contiguous")
    return t


def row_bytes(t: torch.Tensor) -> int:
    n = t.element_size()
    for s in t.shape[1:]:
        n *= s
    return n


def make_keys(pairs: Sequence[tuple[torch.Tensor, torch.Tensor]]):
    """(src, dst) tensor pairs -> XpKey array; row size is taken from dst's trailing dims."""
    arr = (XpKey * len(pairs))()
    for i, (src, dst) in enumerate(pairs):
        rb = row_bytes(dst)
        if src.dtype != dst.dtype or row_bytes(src) != rb:
            raise ValueError(
                f"key {i}: src {tuple(src.shape)}/{src.dtype} does not match dst {tuple(dst.shape)}/{dst.dtype}"
            )
        arr[i].src = src.data_ptr()
        arr[i].dst = dst.data_ptr()
        arr[i].row_bytes = rb
    return arr


# ---- ring / gather ---------------------------------------------------------------------------
def ring_insert(pairs, n_rows: int, cursor: int, max_size: int) -> None:
    """pairs: [(rows [n,...], storage [max_size,...])]."""
    if not pairs or n_rows == 0:
        return
    lib = nv.load()
    keys = make_keys(pairs)
    check(lib.xp_ring_insert(keys, len(pairs), n_rows, cursor, max_size, _stream(pairs[0][1])),
          "xp_ring_insert")


def gather(pairs, idx: torch.Tensor, n_src_rows: int, err_flag: torch.Tensor | None = None) -> None:
    """pairs: [(storage [rows,...], out [B,...])]; idx int64[B] on the same device."""
    if not pairs or idx.numel() == 0:
        return
    lib = nv.load()
    _need(idx, torch.int64, "idx")
    keys = make_keys(pairs)
    check(lib.xp_gather(keys, len(pairs), idx.data_ptr(), idx.numel(), n_src_rows, ptr(err_flag),
                        _stream(idx)), "xp_gather")


def nstep_fold_insert(pairs, roles, done_key: int, done_elem_bytes: int, n_step: int, slot0: int,
                      gpow: Sequence[float], N: int, cursor: int, max_size: int) -> None:
    """pairs: [(stage [n_step*N,...], storage [max_size,...])]."""
    lib = nv.load()
    keys = make_keys(pairs)
    r = (C.c_int * len(pairs))(*roles)
    g = (C.c_float * len(gpow))(*gpow)
    check(lib.xp_nstep_fold_insert(keys, r, len(pairs), done_key, done_elem_bytes, n_step, slot0, g,
                                   N, cursor, max_size, _stream(pairs[0][1])), "xp_nstep_fold_insert")


# ---- uniform sampling --------------------------------------------------------------------------
def randperm_prefix(raw: torch.Tensor, B: int, P: int, n: int, out: torch.Tensor | None = None) -> torch.Tensor:
    """raw: int32/uint32-bit draws as an int32 tensor [P*B] (bit pattern); returns int64 [P*B]."""
    lib = nv.load()
    if raw.dtype not in (torch.int32, torch.uint32):
        raise TypeError("raw draws must be a 32-bit integer tensor")
    if out is None:
        out = torch.empty(P * B, dtype=torch.int64, device=raw.device)
    check(lib.xp_randperm_prefix(raw.data_ptr(), B, P, n, out.data_ptr(), _stream(raw)),
          "xp_randperm_prefix")
    return out


def philox_u32(seed: int, offset: int, n: int, device) -> torch.Tensor:
    lib = nv.load()
    out = torch.empty(n, dtype=torch.int32, device=device)
    check(lib.xp_philox_u32(seed & (2**64 - 1), offset & (2**64 - 1), out.data_ptr(), n, _stream(out)),
          "xp_philox_u32")
    return out


# ---- trees -------------------------------------------------------------------------------------
def make_tree(sum_t: torch.Tensor | None, min_t: torch.Tensor | None, stride: int, cap: int,
              aux: torch.Tensor | None) -> XpTree:
    t = XpTree()
    t.sum = ptr(sum_t)
    t.min = ptr(min_t)
    t.stride = stride
    t.cap = cap
    t.aux = ptr(aux)
    return t


def tree_init(tree: XpTree, stream: int) -> None:
    check(nv.load().xp_tree_init(C.byref(tree), stream), "xp_tree_init")


def tree_set_batch(tree: XpTree, idx: torch.Tensor, val: torch.Tensor, err_flag=None) -> None:
    _need(idx, torch.int64, "idx")
    _need(val, torch.float64, "val")
    check(nv.load().xp_tree_set_batch(C.byref(tree), idx.data_ptr(), val.data_ptr(), idx.numel(),
                                      ptr(err_flag), _stream(idx)), "xp_tree_set_batch")


def tree_query(tree: XpTree, start: int, end: int, out: torch.Tensor) -> None:
    _need(out, torch.float64, "out")
    check(nv.load().xp_tree_query(C.byref(tree), start, end, out.data_ptr(), _stream(out)),
          "xp_tree_query")


def tree_retrieve(tree: XpTree, ub: torch.Tensor) -> torch.Tensor:
    _need(ub, torch.float64, "ub")
    out = torch.empty(ub.numel(), dtype=torch.int64, device=ub.device)
    check(nv.load().xp_tree_retrieve(C.byref(tree), ub.data_ptr(), ub.numel(), out.data_ptr(),
                                     _stream(ub)), "xp_tree_retrieve")
    return out


def per_add(tree: XpTree, state: torch.Tensor, tree_ptr: int, n: int, max_size: int, alpha: float) -> None:
    check(nv.load().xp_per_add(C.byref(tree), state.data_ptr(), tree_ptr, n, max_size, alpha,
                               _stream(state)), "xp_per_add")


def per_sample(tree: XpTree, u: torch.Tensor, B: int, P: int, size: int, beta: float,
               idx_out: torch.Tensor, w_out: torch.Tensor | None) -> None:
    _need(u, torch.float32, "u")
    check(nv.load().xp_per_sample(C.byref(tree), u.data_ptr(), B, P, size, beta, idx_out.data_ptr(),
                                  ptr(w_out), _stream(u)), "xp_per_sample")


def per_weights(tree: XpTree, idx: torch.Tensor, size: int, beta: float, w_out: torch.Tensor) -> None:
    _need(idx, torch.int64, "idx")
    check(nv.load().xp_per_weights(C.byref(tree), idx.data_ptr(), idx.numel(), size, beta,
                                   w_out.data_ptr(), _stream(idx)), "xp_per_weights")


def per_update(tree: XpTree, state: torch.Tensor, idx: torch.Tensor, prio: torch.Tensor,
               max_size: int, alpha: float, err_flag=None) -> None:
    _need(idx, torch.int64, "idx")
    _need(prio, torch.float32, "prio")
    check(nv.load().xp_per_update(C.byref(tree), state.data_ptr(), idx.data_ptr(), prio.data_ptr(),
                                  idx.numel(), max_size, alpha, ptr(err_flag), _stream(idx)),
          "xp_per_update")


def per_sample_hostu(tree: XpTree, u_host, B: int, size: int, beta: float, idx_out: torch.Tensor,
                     w_out: torch.Tensor | None) -> None:
    """xp_per_sample for one agent with the fp32 uniforms still on the host (numpy array / CPU tensor)."""
    u = torch.as_tensor(u_host)
    if u.is_cuda or u.dtype != torch.float32 or not u.is_contiguous() or u.numel() < B:
        raise ValueError("u_host: contiguous fp32 host vector of at least B elements")
    check(nv.load().xp_per_sample_hostu(C.byref(tree), u.data_ptr(), B, size, beta, idx_out.data_ptr(),
                                        ptr(w_out), _stream(idx_out)), "xp_per_sample_hostu")


def per_update_hostp(tree: XpTree, state: torch.Tensor, idx: torch.Tensor, prio_host, max_size: int,
                     alpha: float, err_flag=None) -> None:
    """xp_per_update with the fp32 priorities still on the host (numpy array / CPU tensor), B <= 1024."""
    _need(idx, torch.int64, "idx")
    p = torch.as_tensor(prio_host)
    if p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous() or p.numel() < idx.numel():
        raise ValueError("prio_host: contiguous fp32 host vector of at least len(idx) elements")
    check(nv.load().xp_per_update_hostp(C.byref(tree), state.data_ptr(), idx.data_ptr(), p.data_ptr(),
                                        idx.numel(), max_size, alpha, ptr(err_flag), _stream(idx)),
          "xp_per_update_hostp")


def pow_batch(x: torch.Tensor, y: float) -> torch.Tensor:
    _need(x, torch.float64, "x")
    out = torch.empty_like(x)
    check(nv.load().xp_pow_batch(x.data_ptr(), y, out.data_ptr(), x.numel(), _stream(x)), "xp_pow_batch")
    return out


# ---- rollout returns ---------------------------------------------------------------------------
def gae(rewards, values, dones, last_value, last_done, adv, ret, T: int, N: int, P: int,
        agent_stride: int, gamma: float, gae_lambda: float, use_gae: bool = True) -> None:
    """rewards/values/adv/ret fp32, dones bool/uint8, last_* fp64 — all CUDA, layout [P][T][N]."""
    lib = nv.load()
    for name, t, dt in (("rewards", rewards, torch.float32), ("values", values, torch.float32),
                        ("adv", adv, torch.float32), ("ret", ret, torch.float32),
                        ("last_value", last_value, torch.float64), ("last_done", last_done, torch.float64)):
        _need(t, dt, name)
    if dones.dtype not in (torch.bool, torch.uint8) or not dones.is_cuda:
        raise TypeError("dones must be a CUDA bool/uint8 tensor")
    s = _stream(rewards)
    if use_gae:
        check(lib.xp_gae(rewards.data_ptr(), values.data_ptr(), dones.data_ptr(), last_value.data_ptr(),
                         last_done.data_ptr(), adv.data_ptr(), ret.data_ptr(), T, N, P, agent_stride,
                         gamma, gae_lambda, s), "xp_gae")
    else:
        check(lib.xp_mc_returns(rewards.data_ptr(), values.data_ptr(), dones.data_ptr(),
                                last_value.data_ptr(), last_done.data_ptr(), adv.data_ptr(),
                                ret.data_ptr(), T, N, P, agent_stride, gamma, s), "xp_mc_returns")


def gae_f32(rewards, values, dones, next_value, next_done, adv, ret, T: int, N: int, P: int,
            agent_stride: int, gamma: float, gae_lambda: float) -> None:
    """The all-fp32 GAE of the legacy experiences path: everything fp32 except dones (bool/uint8)."""
    for name, t in (("rewards", rewards), ("values", values), ("adv", adv), ("ret", ret),
                    ("next_value", next_value), ("next_done", next_done)):
        _need(t, torch.float32, name)
    if dones.dtype not in (torch.bool, torch.uint8) or not dones.is_cuda:
        raise TypeError("dones must be a CUDA bool/uint8 tensor")
    check(nv.load().xp_gae_f32(rewards.data_ptr(), values.data_ptr(), dones.data_ptr(),
                               next_value.data_ptr(), next_done.data_ptr(), adv.data_ptr(),
                               ret.data_ptr(), T, N, P, agent_stride, gamma, gae_lambda,
                               _stream(rewards)), "xp_gae_f32")


# ---- token- / turn-level GAE of the LLM PPO path ------------------------------------------------
def gae_tokens(rewards, values, mask, adv, ret, final_mask: bool, gamma: float, gae_lambda: float) -> None:
    """[B, S] fp32 CUDA tensors (time contiguous); see xp_gae_tokens."""
    for name, t in (("rewards", rewards), ("values", values), ("mask", mask), ("adv", adv), ("ret", ret)):
        _need(t, torch.float32, name)
    B, S = rewards.shape
    check(nv.load().xp_gae_tokens(rewards.data_ptr(), values.data_ptr(), mask.data_ptr(), adv.data_ptr(),
                                  ret.data_ptr(), B, S, gamma, gae_lambda, int(final_mask), _stream(rewards)),
          "xp_gae_tokens")


def gae_turns(rewards, values, turn_ids, turn_values, turn_adv, turn_mask, num_turns: int, final_value: bool,
              gamma: float, gae_lambda: float) -> None:
    for name, t in (("rewards", rewards), ("values", values), ("turn_values", turn_values),
                    ("turn_adv", turn_adv), ("turn_mask", turn_mask)):
        _need(t, torch.float32, name)
    _need(turn_ids, torch.int64, "turn_ids")
    B, S = rewards.shape
    check(nv.load().xp_gae_turns(rewards.data_ptr(), values.data_ptr(), turn_ids.data_ptr(), turn_values.data_ptr(),
                                 turn_adv.data_ptr(), turn_mask.data_ptr(), B, S, num_turns, int(final_value),
                                 gamma, gae_lambda, _stream(rewards)), "xp_gae_turns")


def gae_turn_broadcast(turn_ids, turn_values, turn_adv, turn_adv_pg, ret, adv, num_turns: int) -> None:
    B, S = turn_ids.shape
    check(nv.load().xp_gae_turn_broadcast(turn_ids.data_ptr(), turn_values.data_ptr(), turn_adv.data_ptr(),
                                          turn_adv_pg.data_ptr(), ret.data_ptr(), adv.data_ptr(), B, S, num_turns,
                                          _stream(turn_ids)), "xp_gae_turn_broadcast")


WHITEN_SCRATCH = 520  # XP_WHITEN_SCRATCH


def masked_whiten(x, mask, out, times_mask: bool, scratch, err) -> None:
    for name, t in (("x", x), ("mask", mask), ("out", out)):
        _need(t, torch.float32, name)
    _need(scratch, torch.float64, "scratch")
    check(nv.load().xp_masked_whiten(x.data_ptr(), mask.data_ptr(), out.data_ptr(), x.numel(), int(times_mask),
                                     scratch.data_ptr(), err.data_ptr(), _stream(x)), "xp_masked_whiten")


def standardize(x, out, scratch) -> None:
    """out = (x - x.mean()) / (x.std() + 1e-8) over all elements (fp32 CUDA, contiguous; out may be x)."""
    _need(x, torch.float32, "x")
    _need(out, torch.float32, "out")
    _need(scratch, torch.float64, "scratch")
    check(nv.load().xp_standardize(x.data_ptr(), out.data_ptr(), x.numel(), scratch.data_ptr(), _stream(x)),
          "xp_standardize")


# ---- truncated-BPTT sequence preparation ------------------------------------------------------------
BPTT_TYPES = {"chunked": 0, "maximum": 1, "fifty_percent_overlap": 2}


def bptt_plan(dones: torch.Tensor, T: int, N: int, max_seq_len: int | None, seq_type: str):
    """Cut the first ``T`` stored steps of ``dones`` [>=T, N] (CUDA bool/uint8) into sequences.
    Returns (desc int32 [3, T*N] = rows t0/env/len, seq_off int64 [T*N], header int64 [4] =
    {num_sequences, longest sequence, longest episode, total unpadded steps}) — all on device."""
    if dones.dtype not in (torch.bool, torch.uint8) or not dones.is_cuda or not dones.is_contiguous():
        raise TypeError("dones must be a contiguous CUDA bool/uint8 tensor")
    dev = dones.device
    n = T * N
    desc = torch.empty((3, n), dtype=torch.int32, device=dev)
    off = torch.empty(n, dtype=torch.int64, device=dev)
    scratch = torch.empty(2 * N + 2, dtype=torch.int32, device=dev)
    header = torch.empty(4, dtype=torch.int64, device=dev)
    check(nv.load().xp_bptt_plan(dones.data_ptr(), T, N, int(max_seq_len or 0), BPTT_TYPES[seq_type],
                                 scratch.data_ptr(), desc[0].data_ptr(), desc[1].data_ptr(),
                                 desc[2].data_ptr(), off.data_ptr(), header.data_ptr(), _stream(dones)),
          "xp_bptt_plan")
    return desc, off, header


def bptt_indices(desc: torch.Tensor, off: torch.Tensor, num_seq: int, pad_len: int, N: int, total: int):
    """(pad_idx int64 [S*L] with -1 padding, pad_mask bool [S*L], unp_idx int64 [total],
    first_idx int64 [S]) for the sequences described by ``bptt_plan``."""
    dev = desc.device
    pad_idx = torch.empty(num_seq * pad_len, dtype=torch.int64, device=dev)
    pad_mask = torch.empty(num_seq * pad_len, dtype=torch.bool, device=dev)
    unp_idx = torch.empty(total, dtype=torch.int64, device=dev)
    first = torch.empty(num_seq, dtype=torch.int64, device=dev)
    check(nv.load().xp_bptt_indices(desc[0].data_ptr(), desc[1].data_ptr(), desc[2].data_ptr(), off.data_ptr(),
                                    num_seq, pad_len, N, pad_idx.data_ptr(), pad_mask.data_ptr(),
                                    unp_idx.data_ptr(), first.data_ptr(), _stream(desc)), "xp_bptt_indices")
    return pad_idx, pad_mask, unp_idx, first
