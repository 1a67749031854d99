"""Advantage / return estimation of the reference's sibling GAE sites, on the B200 engine.

The reference computes these with a Python ``for t in reversed(range(T))`` loop of torch ops
inside its algorithm classes (no standalone function), so the host mirror is one function per site,
with the argument meaning of the variables the reference loop reads:

  * :func:`gae_from_experiences` — ``PPO._deprecated_learn_from_experiences``
    (algorithms/ppo.py:686-713) and ``IPPO._learn_individual`` (algorithms/ippo.py:703-724).

Every function launches hand-written kernels from ``libb200xp.so`` and fails loudly without it; the
results are bit-identical to the reference loop (all-fp32, one rounding per torch op).
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
