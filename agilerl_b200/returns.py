"""Advantage / return estimation of the reference's sibling GAE sites, on the B200 engine.

The reference computes these with a Python ``for t in reversed(range(T))`` loop of torch ops
inside its algorithm classes (no standalone function), so the host mirror is one function per site,
with the argument meaning of the variables the reference loop reads:

  * :func:`gae_from_experiences` — ``PPO._deprecated_learn_from_experiences``
    (algorithms/ppo.py:686-713) and ``IPPO._learn_individual`` (algorithms/ippo.py:703-724);
  * :func:`normalize_advantages` — the global advantage normalisation of
    ``PPO._learn_from_rollout_buffer_flat`` (algorithms/ppo.py:829-833);
  * :func:`gae_returns_token` — ``PPO._compute_gae_returns_token`` of the LLM path
    (algorithms/ppo_llm.py:833-870);
  * :func:`gae_returns_turn` — ``PPO._compute_gae_returns`` (algorithms/ppo_llm.py:758-831) with
    ``pool_by_turns`` and ``masked_whiten`` (utils/llm_utils.py:215-312).

Every function launches hand-written kernels from ``libb200xp.so`` and fails loudly without it; the
results are bit-identical to the reference loop (all-fp32, one rounding per torch op).  Where the
reference reduces with ``tensor.sum()`` (turn pooling with ``"mean"``, whitening statistics) torch
fixes no summation order; those sums are taken in fp64 in a fixed order and rounded once, so the
results agree with the reference's to a few fp32 ulp (tests: rtol 2e-5) instead of bit for bit.
"""

from __future__ import annotations

import torch

from . import _native as nv
from . import xp


def _dev_f32(x, device) -> torch.Tensor:
    t = torch.as_tensor(x)
    if t.dtype != torch.float32:
        t = t.float()
    return t.to(device, non_blocking=True).contiguous()


def gae_from_experiences(rewards, values, dones, next_value, next_done, gamma: float,
                         gae_lambda: float, device=None) -> "tuple[torch.Tensor, torch.Tensor]":
    """``advantages, returns`` of the legacy experiences path.

    rewards, values: ``[num_steps, num_envs]`` (or ``[num_steps]`` for one env) fp32; dones the
    same shape, 0/1 (the reference's ``dones.long()``); next_value / next_done: ``num_envs`` values
    that bootstrap ``t = num_steps - 1`` (the reference's ``critic(next_obs).reshape(1, -1)`` and
    ``next_done``).  A leading population axis ``[P, num_steps, num_envs]`` is one launch for all
    agents.  Outputs are fp32 CUDA tensors shaped like ``rewards``; ``returns = advantages + values``.
    """
    dev = nv.require_cuda(device if device is not None else
                          (rewards.device if isinstance(rewards, torch.Tensor) and rewards.is_cuda else None))
    r = _dev_f32(rewards, dev)
    v = _dev_f32(values, dev)
    if v.shape != r.shape:
        raise ValueError(f"values {tuple(v.shape)} does not match rewards {tuple(r.shape)}")
    d = torch.as_tensor(dones).to(dev, non_blocking=True)
    if d.shape != r.shape:
        raise ValueError(f"dones {tuple(d.shape)} does not match rewards {tuple(r.shape)}")
    if d.dtype not in (torch.bool, torch.uint8):
        d = d != 0
    d = d.contiguous()
    if r.dim() == 1:
        P, T, N = 1, r.shape[0], 1
    elif r.dim() == 2:
        P, (T, N) = 1, r.shape
    elif r.dim() == 3:
        P, T, N = r.shape
    else:
        raise ValueError("rewards must be [T], [T, N] or [P, T, N]")
    nvl = _dev_f32(next_value, dev).reshape(-1)
    nd = _dev_f32(next_done, dev).reshape(-1)
    if nvl.numel() != P * N or nd.numel() != P * N:
        raise ValueError(f"next_value / next_done must hold {P * N} values")
    adv = torch.empty_like(r)
    ret = torch.empty_like(r)
    with nv.on_device(dev):
        xp.gae_f32(r, v, d, nvl, nd, adv, ret, T, N, P, T * N, float(gamma), float(gae_lambda))
    return adv, ret


_whiten_scratch: dict = {}


def _whiten(x: torch.Tensor, mask: torch.Tensor, times_mask: bool) -> torch.Tensor:
    """``masked_whiten(x, mask)`` (utils/llm_utils.py:245-254), optionally followed by ``* mask``.
    Raises the reference's ValueError when fewer than two values are unmasked."""
    dev = x.device
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    if key not in _whiten_scratch:
        _whiten_scratch[key] = (torch.zeros(xp.WHITEN_SCRATCH, dtype=torch.float64, device=dev),
                                torch.zeros(1, dtype=torch.int32, device=dev))
    scratch, err = _whiten_scratch[key]
    out = torch.empty_like(x)
    xp.masked_whiten(x, mask, out, times_mask, scratch, err)
    if int(err.item()) != 0:  # the reference's `if mask_sum <= 1` is a host read of the same kind
        err.zero_()
        raise ValueError("Unbiased masked variance requires at least 2 unmasked values; "
                         "increase `mini_batch_size` or `gradient_accumulation_steps`.")
    return out


def normalize_advantages(advantages, out=None, device=None) -> torch.Tensor:
    """``(adv - adv.mean()) / (adv.std() + 1e-8)`` over the whole rollout — the global advantage
    normalisation at the top of ``PPO._learn_from_rollout_buffer_flat`` (algorithms/ppo.py:829-833) —
    on the device, so a device-resident ``get_tensor_batch()`` never leaves HBM.  ``out`` may be the
    input itself (in place, what ``buffer_td["advantages"] = ...`` amounts to)."""
    dev = nv.require_cuda(device if device is not None else
                          (advantages.device if isinstance(advantages, torch.Tensor) and advantages.is_cuda else None))
    x = _dev_f32(advantages, dev)
    if out is None:
        out = torch.empty_like(x)
    elif out.dtype != torch.float32 or not out.is_cuda or not out.is_contiguous() or out.shape != x.shape:
        raise ValueError("out must be a contiguous fp32 CUDA tensor shaped like advantages")
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    if key not in _whiten_scratch:
        _whiten_scratch[key] = (torch.zeros(xp.WHITEN_SCRATCH, dtype=torch.float64, device=dev),
                                torch.zeros(1, dtype=torch.int32, device=dev))
    with nv.on_device(dev):
        xp.standardize(x, out, _whiten_scratch[key][0])
    return out


def _llm_inputs(rewards, values, device):
    dev = nv.require_cuda(device if device is not None else
                          (values.device if isinstance(values, torch.Tensor) and values.is_cuda else None))
    r = _dev_f32(rewards, dev)
    v = _dev_f32(values, dev)
    if r.dim() != 2 or v.shape != r.shape:
        raise ValueError(f"rewards {tuple(r.shape)} / values {tuple(v.shape)} must both be [batch, seq_len]")
    return dev, r, v


def gae_returns_token(rewards, values, action_mask, gamma: float, gae_lambda: float,
                      adv_whitening: bool = True, device=None) -> "tuple[torch.Tensor, torch.Tensor]":
    """``(token_returns, token_advantages)`` of ``PPO._compute_gae_returns_token``
    (algorithms/ppo_llm.py:833-870): rewards / values / action_mask ``[batch, seq_len]``; the mask may
    be bool or float (the reference takes ``action_mask.float()``)."""
    dev, r, v = _llm_inputs(rewards, values, device)
    m = _dev_f32(action_mask, dev)
    if m.shape != r.shape:
        raise ValueError(f"action_mask {tuple(m.shape)} does not match values {tuple(r.shape)}")
    adv, ret = torch.empty_like(r), torch.empty_like(r)
    with nv.on_device(dev):
        xp.gae_tokens(r, v, m, adv, ret, not adv_whitening, float(gamma), float(gae_lambda))
        if adv_whitening:
            adv = _whiten(adv, m, True)
    return ret, adv


def gae_returns_turn(rewards, values, action_mask, turn_ids, gamma: float, gae_lambda: float,
                     turn_value_reduction: str = "final_value", adv_whitening: bool = True,
                     device=None) -> "tuple[torch.Tensor, torch.Tensor]":
    """``(token_returns, token_advantages)`` of ``PPO._compute_gae_returns`` (algorithms/ppo_llm.py:758-831):
    each generation turn is one RL action.  ``turn_ids`` int ``[batch, seq_len]``, -1 outside action
    tokens; ``action_mask`` is accepted for signature parity (the reference does not read it)."""
    if turn_value_reduction not in ("mean", "final_value"):
        raise ValueError("turn_value_reduction must be one of ['final_value', 'mean'].")
    dev, r, v = _llm_inputs(rewards, values, device)
    tid = torch.as_tensor(turn_ids).to(dev, non_blocking=True).to(torch.int64).contiguous()
    if tid.shape != r.shape:
        raise ValueError(f"turn_ids {tuple(tid.shape)} does not match values {tuple(r.shape)}")
    B, S = r.shape
    num_turns = int(tid.max().item()) + 1  # `turn_ids.max().item() + 1`, the reference's own host read
    if num_turns < 1:
        raise ValueError("turn_ids holds no turn")
    tv = torch.empty(B, num_turns, dtype=torch.float32, device=dev)
    tadv, tmask = torch.empty_like(tv), torch.empty_like(tv)
    ret, adv = torch.empty_like(r), torch.empty_like(r)
    with nv.on_device(dev):
        xp.gae_turns(r, v, tid, tv, tadv, tmask, num_turns, turn_value_reduction == "final_value",
                     float(gamma), float(gae_lambda))
        tadv_pg = _whiten(tadv, tmask, False) if adv_whitening else tadv
        xp.gae_turn_broadcast(tid, tv, tadv, tadv_pg, ret, adv, num_turns)
    return ret, adv
