"""TEST INFRASTRUCTURE — loads the UNMODIFIED reference (AgileRL) hot-path modules by file path.

Reads /root/reference where it exists (the build container) and otherwise the unmodified copy that
``oracle/stage_reference.py`` staged under git-ignored ``oracle/_ref`` (sha256-checked), which travels
to the GPU box.  Used by ``oracle/make_golden.py`` to generate the committed fixtures under
``tests/golden/``, by ``tests/test_oracle_vs_reference_cpu.py``, and by ``bench.py``'s reference arm /
``cpu_baseline`` legs.  Nothing in the product package imports this file.

``import agilerl`` itself fails here (tensordict / gymnasium / accelerate are not installed), so:
  * agilerl/components/segment_tree.py is stdlib-only        -> loaded as is;
  * agilerl/components/replay_buffer.py needs `tensordict`   -> a minimal shim is registered;
  * RolloutBuffer.compute_returns_and_advantages             -> AST-extracted and exec'd.
No reference source is copied into this repository.
"""

from __future__ import annotations

import ast
import importlib.util
import os
import sys
import textwrap
import types

import numpy as np
import torch

def _resolve_root() -> tuple[str, str]:
    """The reference tree itself where it exists (build container), else the byte-for-byte staged copy
    under oracle/_ref (oracle/stage_reference.py; sha256-checked) — e.g. on the GPU box."""
    live = os.environ.get("AGILERL_REFERENCE", "/root/reference")
    if os.path.isfile(os.path.join(live, "agilerl", "components", "replay_buffer.py")):
        return live, "live"
    from . import stage_reference as sr

    if sr.verify():
        return sr.STAGE_ROOT, "staged"
    return live, "absent"


REF_ROOT, REF_SOURCE = _resolve_root()
REF_PKG = os.path.join(REF_ROOT, "agilerl")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_PKG, "components", "replay_buffer.py"))


# --------------------------------------------------------------------------------------------
# Minimal TensorDict stand-in: just the operations replay_buffer.py performs on its storage.
# --------------------------------------------------------------------------------------------
class ShimTensorDict:
    def __init__(self, source=None, batch_size=None, device=None):
        self._d = {}
        source = source or {}
        for k, v in source.items():
            if isinstance(v, dict):
                v = ShimTensorDict(v, batch_size=batch_size)
            elif not isinstance(v, (torch.Tensor, ShimTensorDict)):
                v = torch.as_tensor(v)
            self._d[k] = v
        if batch_size is None:
            batch_size = []
        if isinstance(batch_size, int):
            batch_size = [batch_size]
        self.batch_size = torch.Size(batch_size)

    # -- mapping ---------------------------------------------------------------------------
    @property
    def shape(self):
        return self.batch_size

    def keys(self):
        return self._d.keys()

    def items(self):
        return list(self._d.items())

    def values(self):
        return self._d.values()

    def __contains__(self, k):
        return k in self._d

    def get(self, k, default=None):
        return self._d.get(k, default)

    def _map(self, fn, batch_size=None):
        out = ShimTensorDict({}, batch_size=self.batch_size if batch_size is None else batch_size)
        for k, v in self._d.items():
            out._d[k] = v._map(fn, batch_size) if isinstance(v, ShimTensorDict) else fn(v)
        return out

    def __getitem__(self, idx):
        if isinstance(idx, str):
            return self._d[idx]
        probe = torch.empty(self.batch_size + (1,), device="meta")[idx]
        new_bs = probe.shape[:-1]
        return self._map(lambda t: t[idx], new_bs)

    def __setitem__(self, idx, value):
        if isinstance(idx, str):
            self._d[idx] = value
            return
        for k, v in self._d.items():
            src = value[k] if not isinstance(value, torch.Tensor) else value
            if isinstance(v, ShimTensorDict):
                v[idx] = src
            else:
                v[idx] = src

    def to(self, *a, **kw):
        return self._map(lambda t: t.to(*a, **kw))

    def clone(self):
        return self._map(lambda t: t.clone())

    def expand(self, shape):
        shape = tuple(shape)
        nb = len(self.batch_size)
        return self._map(lambda t: t.expand(*shape, *t.shape[nb:]), torch.Size(shape))

    def unsqueeze(self, dim):
        bs = list(self.batch_size)
        bs.insert(dim, 1)
        return self._map(lambda t: t.unsqueeze(dim), torch.Size(bs))

    @classmethod
    def __torch_function__(cls, func, types_, args=(), kwargs=None):
        kwargs = kwargs or {}
        if func is torch.zeros_like:
            return args[0]._map(lambda t: torch.zeros_like(t))
        return NotImplemented


def _load(name: str, path: str):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


_cache: dict = {}


def load_reference():
    """Returns a namespace with the reference's SumSegmentTree, MinSegmentTree, ReplayBuffer,
    MultiStepReplayBuffer, PrioritizedReplayBuffer (unmodified code) and TensorDict (the shim)."""
    if "ns" in _cache:
        return _cache["ns"]
    if not available():
        raise RuntimeError(f"reference tree not found under {REF_ROOT}")
    saved = {k: sys.modules.get(k) for k in
             ("agilerl", "agilerl.components", "agilerl.typing", "tensordict",
              "agilerl.components.segment_tree", "agilerl.components.replay_buffer")}
    try:
        for n in ("agilerl", "agilerl.components"):
            pkg = types.ModuleType(n)
            pkg.__path__ = []
            sys.modules[n] = pkg
        seg = _load("agilerl.components.segment_tree",
                    os.path.join(REF_PKG, "components", "segment_tree.py"))
        td = types.ModuleType("tensordict")
        td.TensorDict = td.TensorDictBase = ShimTensorDict
        td.is_tensor_collection = lambda x: isinstance(x, ShimTensorDict)
        sys.modules["tensordict"] = td
        ty = types.ModuleType("agilerl.typing")
        ty.ArrayOrTensor = object
        sys.modules["agilerl.typing"] = ty
        rb = _load("agilerl.components.replay_buffer",
                   os.path.join(REF_PKG, "components", "replay_buffer.py"))
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    ns = types.SimpleNamespace(
        SumSegmentTree=seg.SumSegmentTree, MinSegmentTree=seg.MinSegmentTree,
        ReplayBuffer=rb.ReplayBuffer, MultiStepReplayBuffer=rb.MultiStepReplayBuffer,
        PrioritizedReplayBuffer=rb.PrioritizedReplayBuffer, TensorDict=ShimTensorDict,
    )
    _cache["ns"] = ns
    return ns


def load_reference_gae():
    """AST-extract RolloutBuffer.compute_returns_and_advantages (rollout_buffer.py:413-481) and
    return ``f(self_like, last_value, last_done)``; ``self_like`` needs num_envs, capacity, full,
    pos, gamma, gae_lambda, use_gae and ``buffer`` = dict of CPU torch tensors."""
    if "gae" in _cache:
        return _cache["gae"]
    path = os.path.join(REF_PKG, "components", "rollout_buffer.py")
    src = open(path).read()
    tree = ast.parse(src)
    fn_src = None
    for node in tree.body:
        if isinstance(node, ast.ClassDef) and node.name == "RolloutBuffer":
            for item in node.body:
                if isinstance(item, ast.FunctionDef) and item.name == "compute_returns_and_advantages":
                    fn_src = ast.get_source_segment(src, item)
    assert fn_src is not None
    g = {"np": np, "torch": torch, "ArrayOrTensor": object}
    exec(textwrap.dedent(fn_src), g)
    _cache["gae"] = g["compute_returns_and_advantages"]
    return _cache["gae"]


def reference_gae(rewards, values, dones, last_value, last_done, gamma, gae_lambda, use_gae=True):
    """Run the reference GAE/MC on numpy inputs [T,N]; returns (advantages, returns) fp32."""
    f = load_reference_gae()
    T, N = rewards.shape
    self_like = types.SimpleNamespace(
        num_envs=N, capacity=T, full=True, pos=0, gamma=gamma, gae_lambda=gae_lambda,
        use_gae=use_gae,
        buffer={
            "rewards": torch.from_numpy(np.ascontiguousarray(rewards)),
            "values": torch.from_numpy(np.ascontiguousarray(values)),
            "dones": torch.from_numpy(np.ascontiguousarray(dones)),
            "advantages": torch.zeros(T, N), "returns": torch.zeros(T, N),
        },
    )
    f(self_like, last_value, last_done)
    return self_like.buffer["advantages"].numpy().copy(), self_like.buffer["returns"].numpy().copy()


def _extract(path: str, names: list[str], cls: str | None = None) -> dict[str, str]:
    """Source text of top-level defs/classes (or methods of ``cls``) called ``names``."""
    src = open(path).read()
    tree = ast.parse(src)
    out = {}
    body = tree.body
    if cls is not None:
        body = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == cls).body
    for node in body:
        if isinstance(node, (ast.FunctionDef, ast.ClassDef)) and node.name in names:
            out[node.name] = textwrap.dedent(ast.get_source_segment(src, node))
    missing = set(names) - set(out)
    assert not missing, f"not found in {path}: {missing}"
    return out


def load_reference_bptt():
    """The UNMODIFIED sequence-cutting code of the reference, extracted by name:
    ``BPTTSequenceType`` (agilerl/typing.py:160), ``extract_sequences_from_episode``
    (utils/algo_utils.py:214-265), ``RolloutBuffer._get_complete_sequences`` (:677-722) and the
    tensor branch of ``RolloutBuffer._pad_sequences`` (:627-675).  Returns a namespace."""
    if "bptt" in _cache:
        return _cache["bptt"]
    import enum

    g = {"np": np, "torch": torch, "Enum": enum.Enum, "TensorDict": dict,
         "is_tensor_collection": lambda x: False}
    exec(_extract(os.path.join(REF_PKG, "typing.py"), ["BPTTSequenceType"])["BPTTSequenceType"], g)
    exec(_extract(os.path.join(REF_PKG, "utils", "algo_utils.py"),
                  ["extract_sequences_from_episode"])["extract_sequences_from_episode"], g)
    m = _extract(os.path.join(REF_PKG, "components", "rollout_buffer.py"),
                 ["_get_complete_sequences", "_pad_sequences"], cls="RolloutBuffer")
    exec(m["_get_complete_sequences"], g)
    exec(m["_pad_sequences"].replace("@staticmethod", ""), g)
    g["RolloutBuffer"] = types.SimpleNamespace(_pad_sequences=g["_pad_sequences"])
    _cache["bptt"] = types.SimpleNamespace(**{k: g[k] for k in (
        "BPTTSequenceType", "extract_sequences_from_episode", "_get_complete_sequences", "_pad_sequences")})
    return _cache["bptt"]


def reference_bptt(dones, data, max_seq_len, seq_type: str):
    """Drive the extracted reference functions the way prepare_sequence_tensors does
    (rollout_buffer.py:736-800) on ``dones`` bool [T,N] and ONE data field ``data`` [T,N,...].
    Returns (unpadded_slices list[np.ndarray], padded np [S,L,...], pad_mask np [S,L],
    max_sequence_length, num_sequences)."""
    ns = load_reference_bptt()
    T, N = dones.shape
    d = torch.from_numpy(np.ascontiguousarray(dones))
    self_like = types.SimpleNamespace(num_envs=N, max_seq_len=max_seq_len,
                                      bptt_sequence_type=ns.BPTTSequenceType(seq_type))
    # the done-index loop of prepare_sequence_tensors (:738-750), statement for statement
    episode_done_indices = []
    for env_idx in range(N):
        env_dones = d[:, env_idx]
        lst = env_dones.nonzero().squeeze(-1).tolist()
        episode_done_indices.append(lst)
        if len(episode_done_indices[env_idx]) == 0 or episode_done_indices[env_idx][-1] != T - 1:
            episode_done_indices[env_idx].append(T - 1)
    flat = torch.arange(T * N).reshape(T, N)
    slices, _ = ns._get_complete_sequences(self_like, flat, episode_done_indices)
    x = torch.from_numpy(np.ascontiguousarray(data))
    seqs, max_len = ns._get_complete_sequences(self_like, x, episode_done_indices)
    padded = ns._pad_sequences(seqs, target_length=None)
    ones = torch.ones(T, N, dtype=torch.bool)
    mseqs, _ = ns._get_complete_sequences(self_like, ones, episode_done_indices)
    mask = ns._pad_sequences(mseqs, target_length=None)
    msl = min(max_len, max_seq_len) if max_seq_len is not None else max_len
    return [s.numpy().copy() for s in slices], padded.numpy(), mask.numpy(), int(msl), len(seqs)


def load_reference_marb():
    """The UNMODIFIED ``MultiAgentReplayBuffer`` (components/multi_agent_replay_buffer.py), loaded
    by path with two stub modules: ``agilerl.typing`` (a type alias only) and
    ``agilerl.utils.algo_utils`` holding the reference's own ``obs_to_tensor`` (AST-extracted,
    utils/algo_utils.py:746-773)."""
    if "marb" in _cache:
        return _cache["marb"]
    from numbers import Number

    names = ("agilerl", "agilerl.components", "agilerl.typing", "agilerl.utils", "agilerl.utils.algo_utils",
             "agilerl.components.multi_agent_replay_buffer")
    saved = {k: sys.modules.get(k) for k in names}
    try:
        for n in ("agilerl", "agilerl.components", "agilerl.utils"):
            pkg = types.ModuleType(n)
            pkg.__path__ = []
            sys.modules[n] = pkg
        ty = types.ModuleType("agilerl.typing")
        ty.NumpyObsType = object
        sys.modules["agilerl.typing"] = ty
        au = types.ModuleType("agilerl.utils.algo_utils")
        g = {"np": np, "torch": torch, "Number": Number, "TensorDict": ShimTensorDict,
             "ObservationType": object, "TorchObsType": object}
        exec(_extract(os.path.join(REF_PKG, "utils", "algo_utils.py"), ["obs_to_tensor"])["obs_to_tensor"], g)
        au.obs_to_tensor = g["obs_to_tensor"]
        sys.modules["agilerl.utils.algo_utils"] = au
        mod = _load("agilerl.components.multi_agent_replay_buffer",
                    os.path.join(REF_PKG, "components", "multi_agent_replay_buffer.py"))
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    _cache["marb"] = mod.MultiAgentReplayBuffer
    return _cache["marb"]


def _reference_gae_loop(path: str, method: str, cls: str):
    """Source of the ``for t in reversed(range(num_steps))`` statement inside ``cls.method`` —
    the all-fp32 GAE loop of the legacy experiences path (algorithms/ppo.py:693-711,
    algorithms/ippo.py:705-720) — extracted unmodified."""
    src = open(path).read()
    tree = ast.parse(src)
    klass = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == cls)
    fn = next(n for n in klass.body if isinstance(n, ast.FunctionDef) and n.name == method)
    for node in ast.walk(fn):
        if (isinstance(node, ast.For) and isinstance(node.target, ast.Name) and node.target.id == "t"
                and "last_gae_lambda" in ast.get_source_segment(src, node)):
            return textwrap.dedent(" " * node.col_offset + ast.get_source_segment(src, node))
    raise AssertionError(f"GAE loop not found in {path}:{cls}.{method}")


def reference_gae_f32(rewards, values, dones, next_value, next_done, gamma, gae_lambda, site="ppo"):
    """Run the reference's own loop statement (site 'ppo' = PPO._deprecated_learn_from_experiences,
    'ippo' = IPPO._learn_individual) on fp32 torch CPU tensors, with the few assignments that
    precede it in the reference (ppo.py:688-692) restated.  rewards/values fp32 [T,N], dones 0/1
    [T,N], next_value fp32 [N], next_done fp32 [N].  Returns (advantages, returns) fp32 numpy."""
    key = "gae32_" + site
    if key not in _cache:
        if site == "ppo":
            _cache[key] = _reference_gae_loop(os.path.join(REF_PKG, "algorithms", "ppo.py"),
                                              "_deprecated_learn_from_experiences", "PPO")
        else:
            _cache[key] = _reference_gae_loop(os.path.join(REF_PKG, "algorithms", "ippo.py"),
                                              "_learn_individual", "IPPO")
    rewards = torch.from_numpy(np.ascontiguousarray(rewards, dtype=np.float32))
    values = torch.from_numpy(np.ascontiguousarray(values, dtype=np.float32))
    d = torch.from_numpy(np.ascontiguousarray(dones))
    g = {
        "torch": torch,
        "self": types.SimpleNamespace(gamma=gamma, gae_lambda=gae_lambda),
        "rewards": rewards, "values": values,
        "dones": d.long() if site == "ppo" else d.float(),
        "next_done": torch.from_numpy(np.ascontiguousarray(next_done, dtype=np.float32)).reshape(1, -1),
        "next_value": torch.from_numpy(np.ascontiguousarray(next_value, dtype=np.float32)).reshape(1, -1),
        "num_steps": rewards.size(0),
        "advantages": torch.zeros_like(rewards).float(),
        "last_gae_lambda": 0,
    }
    with torch.no_grad():
        exec(_cache[key], g)
    adv = g["advantages"]
    return adv.numpy().copy(), (adv + values).numpy().copy()


def load_reference_llm_gae():
    """The UNMODIFIED token- and turn-level GAE of the LLM PPO path, extracted by name:
    ``PPO._compute_gae_returns_token`` / ``PPO._compute_gae_returns`` (algorithms/ppo_llm.py:758-872)
    with ``masked_mean / masked_var / masked_whiten / pool_by_turns`` (utils/llm_utils.py:215-312)."""
    if "llm_gae" in _cache:
        return _cache["llm_gae"]
    g = {"torch": torch, "np": np}
    u = _extract(os.path.join(REF_PKG, "utils", "llm_utils.py"),
                 ["masked_mean", "masked_var", "masked_whiten", "pool_by_turns"])
    for name in ("masked_mean", "masked_var", "masked_whiten", "pool_by_turns"):
        exec("from __future__ import annotations\n" + u[name], g)
    path = os.path.join(REF_PKG, "algorithms", "ppo_llm.py")
    cls = next(n.name for n in ast.parse(open(path).read()).body
               if isinstance(n, ast.ClassDef) and any(isinstance(b, ast.FunctionDef) and b.name == "_compute_gae_returns_token"
                                                      for b in n.body))
    m = _extract(path, ["_compute_gae_returns_token", "_compute_gae_returns"], cls=cls)
    for name in ("_compute_gae_returns_token", "_compute_gae_returns"):
        exec("from __future__ import annotations\n" + m[name], g)
    _cache["llm_gae"] = types.SimpleNamespace(token=g["_compute_gae_returns_token"], turn=g["_compute_gae_returns"])
    return _cache["llm_gae"]


def reference_llm_gae(rewards, values, action_mask, gamma, gae_lambda, adv_whitening, turn_ids=None,
                      turn_value_reduction="final_value"):
    """Run the reference method on fp32 CPU tensors ``[batch, seq_len]``; ``turn_ids`` None selects the
    token-level variant.  Returns ``(token_returns, token_advantages)`` as fp32 numpy."""
    ref = load_reference_llm_gae()
    me = types.SimpleNamespace(gamma=gamma, gae_lambda=gae_lambda, adv_whitening=adv_whitening,
                               turn_value_reduction=turn_value_reduction)
    r = torch.from_numpy(np.ascontiguousarray(rewards, dtype=np.float32))
    v = torch.from_numpy(np.ascontiguousarray(values, dtype=np.float32))
    m = torch.from_numpy(np.ascontiguousarray(action_mask))
    with torch.no_grad():
        if turn_ids is None:
            ret, adv = ref.token(me, r, v, m)
        else:
            ret, adv = ref.turn(me, r, v, m, torch.from_numpy(np.ascontiguousarray(turn_ids, dtype=np.int64)))
    return ret.numpy().copy(), adv.numpy().copy()


def reference_normalize_advantages(advantages):
    """The reference's own statement ``normalized_advantages = (valid_advantages - valid_advantages.mean())
    / (valid_advantages.std() + 1e-8)`` (PPO._learn_from_rollout_buffer_flat, algorithms/ppo.py:829-833),
    extracted unmodified and run on an fp32 CPU tensor."""
    if "advnorm" not in _cache:
        path = os.path.join(REF_PKG, "algorithms", "ppo.py")
        src = open(path).read()
        tree = ast.parse(src)
        klass = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "PPO")
        fn = next(n for n in klass.body if isinstance(n, ast.FunctionDef) and n.name == "_learn_from_rollout_buffer_flat")
        stmt = next(n for n in ast.walk(fn) if isinstance(n, ast.Assign) and isinstance(n.targets[0], ast.Name)
                    and n.targets[0].id == "normalized_advantages")
        _cache["advnorm"] = textwrap.dedent(" " * stmt.col_offset + ast.get_source_segment(src, stmt))
    g = {"torch": torch, "valid_advantages": torch.from_numpy(np.ascontiguousarray(advantages, dtype=np.float32))}
    exec(_cache["advnorm"], g)
    return g["normalized_advantages"].numpy().copy()
