"""TEST INFRASTRUCTURE — CPU restatement (numpy / pure Python) of the AgileRL experience path.

This is the parity ORACLE, not product code: only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import it.  Each function cites the reference
file:line (relative to the AgileRL tree) it follows.  It deliberately keeps the reference's
data structures (Python lists of floats for the trees, per-sample Python loops, numpy time loop
for GAE) so that timing it is representative of the reference's own CPU path.

Parity pin: tests/test_oracle_golden.py checks every function against the fixtures in
tests/golden/ that oracle/make_golden.py produced by running the unmodified reference code
(oracle/ref_loader.py) in the build container, and against the reference's own known-answer
tests (tests/test_components/test_segment_tree.py, test_replay_buffer.py).
"""

from __future__ import annotations

import math
import operator

import numpy as np

# ---------------------------------------------------------------------------------------------
# segment trees — agilerl/components/segment_tree.py:5-182
# ---------------------------------------------------------------------------------------------


class SegTree:
    """Heap-layout tree over a Python list of floats (segment_tree.py:19-26)."""

    def __init__(self, capacity: int, kind: str):
        assert capacity > 0 and capacity & (capacity - 1) == 0, (
            "capacity must be positive and a power of 2."
        )
        self.capacity = capacity
        self.kind = kind
        if kind == "sum":
            self.operation, init = operator.add, 0.0
        else:
            self.operation, init = min, float("inf")
        self.tree = [init for _ in range(2 * capacity)]

    def _helper(self, start, end, node, ns, ne):  # segment_tree.py:28-62
        if start == ns and end == ne:
            return self.tree[node]
        mid = (ns + ne) // 2
        if end <= mid:
            return self._helper(start, end, 2 * node, ns, mid)
        if mid + 1 <= start:
            return self._helper(start, end, 2 * node + 1, mid + 1, ne)
        return self.operation(
            self._helper(start, mid, 2 * node, ns, mid),
            self._helper(mid + 1, end, 2 * node + 1, mid + 1, ne),
        )

    def operate(self, start=0, end=0):  # segment_tree.py:64-79
        if end <= 0:
            end += self.capacity
        end -= 1
        return self._helper(start, end, 1, 0, self.capacity - 1)

    def __setitem__(self, idx, val):  # segment_tree.py:81-95
        idx += self.capacity
        self.tree[idx] = val
        idx //= 2
        while idx >= 1:
            self.tree[idx] = self.operation(self.tree[2 * idx], self.tree[2 * idx + 1])
            idx //= 2

    def __getitem__(self, idx):  # segment_tree.py:97-108
        assert 0 <= idx < self.capacity
        return self.tree[self.capacity + idx]

    def retrieve(self, upperbound):  # segment_tree.py:136-156
        assert 0 <= upperbound <= self.operate() + 1e-5, f"upperbound: {upperbound}"
        idx = 1
        while idx < self.capacity:
            left = 2 * idx
            if self.tree[left] > upperbound:
                idx = left
            else:
                upperbound -= self.tree[left]
                idx = left + 1
        return idx - self.capacity

    def rebuild_from_leaves(self):
        """Bulk construction (not in the reference): every internal node = op(children).
        Yields exactly the tree the sequential root walks produce."""
        t = self.tree
        for i in range(self.capacity - 1, 0, -1):
            t[i] = self.operation(t[2 * i], t[2 * i + 1])


class PerOracle:
    """PrioritizedReplayBuffer's tree logic without storage (replay_buffer.py:261-428)."""

    def __init__(self, max_size: int, alpha: float = 0.6):
        self.max_size = max_size
        self.alpha = alpha
        self.max_priority = 1.0  # :284
        self.tree_ptr = 0
        self.size = 0
        cap = 1
        while cap < max_size:  # :288-290
            cap *= 2
        self.sum_tree = SegTree(cap, "sum")
        self.min_tree = SegTree(cap, "min")

    def _update_priority(self, idx: int, priority: float):  # :311-329
        assert 0 <= idx < self.max_size
        pa = priority**self.alpha
        self.sum_tree[idx] = pa
        self.min_tree[idx] = pa
        self.max_priority = max(self.max_priority, priority)

    def add(self, n: int):  # :296-309
        self.size = min(self.size + n, self.max_size)
        for _ in range(n):
            self._update_priority(self.tree_ptr, self.max_priority)
            self.tree_ptr = (self.tree_ptr + 1) % self.max_size

    def sample_proportional(self, batch_size: int, u) -> np.ndarray:  # :357-381
        """``u[i]`` is the value ``torch.rand(1).item()`` returns for stratum i (fp32 widened)."""
        indices = np.zeros(batch_size, dtype=np.int64)
        total = self.sum_tree.operate()
        segment = total / batch_size
        for i in range(batch_size):
            a = segment * i
            b = segment * (i + 1)
            upperbound = float(u[i]) * (b - a) + a
            indices[i] = self.sum_tree.retrieve(upperbound)
        return indices

    def calculate_weights(self, indices, beta: float) -> np.ndarray:  # :383-409
        w = np.zeros(len(indices), dtype=np.float32)
        p_min = self.min_tree.operate() / self.sum_tree.operate()
        max_weight = (p_min * self.size) ** -beta
        for i, idx in enumerate(indices):
            p_sample = self.sum_tree[int(idx)] / self.sum_tree.operate()
            weight = (p_sample * self.size) ** -beta
            w[i] = weight / max_weight
        return w

    def update_priorities(self, indices, priorities):  # :411-428
        for idx, p in zip(indices, priorities):
            self._update_priority(int(idx), max(float(p), 1e-5))


# ---------------------------------------------------------------------------------------------
# ring buffer — replay_buffer.py:72-131
# ---------------------------------------------------------------------------------------------


def ring_insert(storage: dict, data: dict, cursor: int, max_size: int) -> int:
    """ReplayBuffer.add's circular write (:99-110). ``storage[k]`` is [max_size, ...] numpy."""
    n = next(iter(data.values())).shape[0]
    end = cursor + n
    for k, v in data.items():
        if end > max_size:
            first = max_size - cursor
            storage[k][cursor:] = v[:first]
            storage[k][: n - first] = v[first:]
        else:
            storage[k][cursor:end] = v
    return end % max_size


def gather(storage: dict, idx) -> dict:
    """``storage[indices]`` (:126,:204,:345)."""
    return {k: v[np.asarray(idx)] for k, v in storage.items()}


def randperm_prefix(raw, batch: int, n: int) -> np.ndarray:
    """First ``batch`` entries of torch.randperm(n) given the raw mt19937 draws it consumes
    (replay_buffer.py:125; ATen randperm_cpu: ``z = random() % (n - i); swap(p[i], p[i+z])``)."""
    disp: dict[int, int] = {}
    out = np.empty(batch, dtype=np.int64)
    for i in range(batch):
        if i >= n - 1:  # the loop stops at n-2; the last element stays in place
            out[i] = disp.get(i, i)
            continue
        j = i + int(raw[i]) % (n - i)
        vi = disp.get(i, i)
        vj = disp.get(j, j)
        out[i] = vj
        disp[j] = vi
        disp[i] = vj
    return out


def mt19937_raw(seed: int, count: int) -> np.ndarray:
    """Raw 32-bit outputs of mt19937 seeded like torch.manual_seed(seed)."""
    rs = np.random.RandomState(seed & 0xFFFFFFFF)
    return np.frombuffer(rs.bytes(4 * count), dtype="<u4").copy()


def torch_rand_from_raw(raw) -> np.ndarray:
    """torch.rand fp32 values from raw draws: (x & (2^24-1)) * 2^-24."""
    return ((np.asarray(raw, dtype=np.uint32) & np.uint32(0xFFFFFF)).astype(np.float32)
            * np.float32(2.0**-24))


# ---------------------------------------------------------------------------------------------
# n-step fold — replay_buffer.py:206-258
# ---------------------------------------------------------------------------------------------


def nstep_fold(steps: list[dict], gamma: float, reward_key="reward", done_key="done",
               ns_key="next_obs") -> dict:
    """``steps`` = the deque contents, oldest first; values are numpy arrays [N, ...]."""
    first = {k: v.copy() for k, v in steps[0].items()}
    R = first[reward_key].copy()
    for i, tr in enumerate(steps[1:]):
        # fp32 tensor * Python float: the scalar is taken in fp32, product then sum (:244)
        R += tr[reward_key] * np.float32(gamma ** (i + 1))
        first[ns_key] = tr[ns_key].copy()
        first[done_key] = tr[done_key].copy()
        if tr[done_key].astype(bool).any():  # :252
            break
    first[reward_key] = R
    return first


# ---------------------------------------------------------------------------------------------
# rollout returns — rollout_buffer.py:413-481
# ---------------------------------------------------------------------------------------------


def gae(rewards, values, dones, last_value, last_done, gamma: float, gae_lambda: float):
    """GAE branch (:448-467). rewards/values fp32 [T,N], dones bool [T,N]."""
    T, N = rewards.shape
    last_value = np.asarray(last_value).reshape(N)
    last_done = np.asarray(last_done).reshape(N)
    adv = np.zeros((T, N), dtype=np.float32)
    last_gae_lambda = np.zeros(N, dtype=np.float32)
    for t in reversed(range(T)):
        if t == T - 1:
            next_non_terminal = 1.0 - last_done.astype(float)
            next_values = last_value.astype(float)
        else:
            next_non_terminal = 1.0 - dones[t + 1].astype(float)
            next_values = values[t + 1]
        delta = rewards[t] + gamma * next_values * next_non_terminal - values[t]
        adv[t] = last_gae_lambda = (
            delta + gamma * gae_lambda * next_non_terminal * last_gae_lambda
        )
    ret = adv + values
    return adv, ret


def mc_returns(rewards, values, dones, last_value, last_done, gamma: float):
    """Monte-Carlo branch (:468-477)."""
    T, N = rewards.shape
    last_value = np.asarray(last_value).reshape(N)
    last_done = np.asarray(last_done).reshape(N)
    ret = np.zeros((T, N), dtype=np.float32)
    last_returns = last_value.astype(float) * (1.0 - last_done.astype(float))
    for t in reversed(range(T)):
        ret[t] = last_returns = rewards[t] + gamma * last_returns * (1.0 - dones[t].astype(float))
    adv = ret - values
    return adv, ret


def gae_scalar(rewards, values, dones, last_value, last_done, gamma: float, gae_lambda: float):
    """Explicit per-element restatement of the rounding sequence; used to pin
    the semantics the CUDA kernel implements against the vectorised numpy form above."""
    T, N = rewards.shape
    adv = np.zeros((T, N), dtype=np.float32)
    g32 = np.float32(gamma)
    gl = gamma * gae_lambda
    for e in range(N):
        carry = 0.0
        for t in reversed(range(T)):
            if t == T - 1:
                nnt = 1.0 - float(last_done[e])
                g_nv = gamma * float(last_value[e])
            else:
                nnt = 1.0 - float(bool(dones[t + 1, e]))
                g_nv = float(np.float32(g32 * values[t + 1, e]))
            delta = (float(rewards[t, e]) + g_nv * nnt) - float(values[t, e])
            carry = delta + (gl * nnt) * carry
            adv[t, e] = np.float32(carry)
    return adv, adv + values


def gae_f32(rewards, values, dones, next_value, next_done, gamma: float, gae_lambda: float):
    """The all-fp32 GAE loop of the legacy experiences path (algorithms/ppo.py:693-713,
    algorithms/ippo.py:705-724).  torch evaluates it on fp32 tensors with Python-float scalars,
    i.e. every scalar enters as fp32 and every op rounds to fp32.  rewards/values fp32 [T,N],
    dones 0/1 [T,N], next_value/next_done fp32 [N]."""
    f = np.float32
    T, N = rewards.shape
    rewards = np.asarray(rewards, dtype=f)
    values = np.asarray(values, dtype=f)
    g32, gl32 = f(gamma), f(gamma * gae_lambda)
    adv = np.zeros((T, N), dtype=f)
    carry = np.zeros(N, dtype=f)
    for t in reversed(range(T)):
        if t == T - 1:
            nnt = f(1.0) - np.asarray(next_done, dtype=f).reshape(N)
            nv = np.asarray(next_value, dtype=f).reshape(N)
        else:
            nnt = f(1.0) - np.asarray(dones[t + 1]).astype(f)
            nv = values[t + 1]
        delta = (rewards[t] + (g32 * nv) * nnt) - values[t]
        adv[t] = carry = delta + (gl32 * nnt) * carry
    assert adv.dtype == f
    return adv, adv + values


# =================================================================================================
# token- and turn-level GAE of the LLM PPO path — algorithms/ppo_llm.py:758-872,
# utils/llm_utils.py:215-312.  [batch, seq_len] fp32; every torch op is one fp32 rounding.
# Sums (`tensor.sum()`) have no fixed order in torch: taken in fp64 here and rounded once.
# =================================================================================================
def masked_whiten(values, mask):
    """masked_whiten with shift_mean=True, unbiased variance (llm_utils.py:215-254)."""
    f = np.float32
    values = np.asarray(values, dtype=f)
    mask = np.asarray(mask, dtype=f)
    cnt = f(mask.astype(np.float64).sum())
    mean = f(f((values * mask).astype(np.float64).sum()) / cnt)               # masked_mean
    centered = values - mean
    var = f(f(((centered * centered) * mask).astype(np.float64).sum()) / cnt)   # masked_mean(centered**2)
    if cnt <= 1:
        raise ValueError("Unbiased masked variance requires at least 2 unmasked values")
    var = f(var * f(cnt / f(cnt - f(1.0))))
    rs = f(f(1.0) / np.sqrt(f(var + f(1e-8))))
    return (values - mean) * rs


def gae_tokens(rewards, values, action_mask, gamma: float, gae_lambda: float, adv_whitening: bool):
    """PPO._compute_gae_returns_token (ppo_llm.py:833-870) -> (token_returns, token_advantages)."""
    f = np.float32
    rewards = np.asarray(rewards, dtype=f)
    values = np.asarray(values, dtype=f)
    mask = np.asarray(action_mask).astype(f)
    B, S = values.shape
    g32, gl32 = f(gamma), f(gamma * gae_lambda)
    adv = np.zeros((B, S), dtype=f)
    last = np.zeros(B, dtype=f)
    for t in reversed(range(S)):
        if t == S - 1:
            nv, nm = np.zeros(B, dtype=f), np.zeros(B, dtype=f)
        else:
            nv, nm = values[:, t + 1], mask[:, t + 1]
        delta = (rewards[:, t] + (g32 * nv) * nm) - values[:, t]
        last = delta + (gl32 * last) * nm
        adv[:, t] = last * mask[:, t]
    ret = (adv + values) * mask
    if adv_whitening:
        adv = masked_whiten(adv, mask)
    return ret, adv * mask


def pool_by_turns(token_values, turn_ids, num_turns: int, reduction: str = "mean"):
    """llm_utils.py:256-312 ("mean" and "final_value")."""
    f = np.float32
    token_values = np.asarray(token_values, dtype=f)
    B, S = token_values.shape
    out = np.zeros((B, num_turns), dtype=f)
    for t in range(num_turns):
        mask_t = (turn_ids == t)
        if reduction == "mean":
            summed = (token_values * mask_t.astype(f)).astype(np.float64).sum(axis=1).astype(f)
            count = np.maximum(mask_t.sum(axis=1).astype(f), f(1.0))
            out[:, t] = summed / count
        else:
            pos = np.where(mask_t, np.arange(S)[None, :], -1).max(axis=1)
            out[:, t] = np.where(pos >= 0, token_values[np.arange(B), np.maximum(pos, 0)], f(0.0))
    return out


def gae_turns(rewards, values, turn_ids, gamma: float, gae_lambda: float, adv_whitening: bool,
              turn_value_reduction: str = "final_value"):
    """PPO._compute_gae_returns (ppo_llm.py:758-831) -> (token_returns, token_advantages)."""
    f = np.float32
    turn_ids = np.asarray(turn_ids).astype(np.int64)
    B, S = turn_ids.shape
    K = int(turn_ids.max()) + 1
    tv = pool_by_turns(values, turn_ids, K, turn_value_reduction)
    tr = pool_by_turns(rewards, turn_ids, K)
    g32, gl32 = f(gamma), f(gamma * gae_lambda)
    tadv = np.zeros((B, K), dtype=f)
    last = np.zeros(B, dtype=f)
    nb = turn_ids.max(axis=1) + 1
    for t in reversed(range(K)):
        is_last = t >= nb - 1
        nv = np.zeros(B, dtype=f) if t == K - 1 else tv[:, t + 1]
        nv = np.where(is_last, f(0.0), nv).astype(f)
        delta = (tr[:, t] + g32 * nv) - tv[:, t]
        has = (nb > t).astype(f)
        last = (delta + gl32 * last) * has
        tadv[:, t] = last
    turn_mask = (turn_ids[:, :, None] == np.arange(K)[None, None, :]).any(axis=1).astype(f)
    tadv_pg = masked_whiten(tadv, turn_mask) if adv_whitening else tadv
    valid = (turn_ids >= 0).astype(f)
    safe = np.maximum(turn_ids, 0)
    rows = np.arange(B)[:, None]
    return (tadv + tv)[rows, safe] * valid, tadv_pg[rows, safe] * valid


def normalize_advantages(advantages):
    """(adv - adv.mean()) / (adv.std() + 1e-8), std unbiased (algorithms/ppo.py:829-833); fp32 result,
    statistics in fp64 rounded once (torch fixes no summation order)."""
    f = np.float32
    a = np.asarray(advantages, dtype=f)
    mean = f(a.astype(np.float64).mean())
    sd = f(np.sqrt(((a.astype(np.float64) - np.float64(mean)) ** 2).sum() / (a.size - 1))) if a.size > 1 else f(np.nan)
    return (a - mean) / f(sd + f(1e-8))


def pow_cpython(x: float, y: float) -> float:
    """float ** float in CPython = libm pow (replay_buffer.py:322,401,406)."""
    return math.pow(x, y)


# =================================================================================================
# truncated-BPTT sequence preparation — rollout_buffer.py:627-800, utils/algo_utils.py:214-265
# =================================================================================================
def bptt_sequences(dones, max_seq_len, seq_type: str = "chunked"):
    """dones bool [T,N] -> ([(env, t0, len), ...] in reference order, longest episode).

    Episodes end at every done flag and at the last stored step (:738-750); an episode of
    len >= max_seq_len is cut by extract_sequences_from_episode (algo_utils.py:243-259), shorter
    ones (or max_seq_len None) are one sequence (:696-712)."""
    dones = np.asarray(dones).astype(bool)
    T, N = dones.shape
    seqs, max_ep = [], 1
    for env in range(N):
        ends = np.nonzero(dones[:, env])[0].tolist()
        if not ends or ends[-1] != T - 1:
            ends.append(T - 1)
        start = 0
        for e in ends:
            n = e - start + 1
            if max_seq_len is not None and n >= max_seq_len:
                if seq_type == "chunked":
                    for c in range(max(1, n // max_seq_len)):
                        seqs.append((env, start + c * max_seq_len, max_seq_len))
                elif seq_type == "maximum":
                    for s in range(n - max_seq_len + 1):
                        seqs.append((env, start + s, max_seq_len))
                elif seq_type == "fifty_percent_overlap":
                    for s in range(0, n - max_seq_len + 1, max_seq_len // 2):  # step 0 -> ValueError
                        seqs.append((env, start + s, max_seq_len))
                else:
                    raise NotImplementedError(seq_type)
            else:
                seqs.append((env, start, n))
            max_ep = max(max_ep, n)
            start = e + 1
    return seqs, max_ep


def bptt_prepare(dones, fields: dict, max_seq_len, seq_type: str = "chunked"):
    """prepare_sequence_tensors (:724-846) on numpy arrays: ``fields`` name -> [T,N,...].
    Returns dict(num_sequences, max_sequence_length, unpadded_slices (list of flat t*N+env index
    arrays), pad_mask [S,L], padded {name: [S,L,...]}, first_idx [S])."""
    dones = np.asarray(dones)
    T, N = dones.shape
    seqs, max_ep = bptt_sequences(dones, max_seq_len, seq_type)
    S = len(seqs)
    L = max(n for _, _, n in seqs)  # pad_sequence pads to the longest sequence (:669-675)
    slices = [(np.arange(t0, t0 + n) * N + env).astype(np.int64) for env, t0, n in seqs]
    mask = np.zeros((S, L), dtype=bool)
    padded = {k: np.zeros((S, L, *np.asarray(v).shape[2:]), dtype=np.asarray(v).dtype) for k, v in fields.items()}
    for s, (env, t0, n) in enumerate(seqs):
        mask[s, :n] = True
        for k, v in fields.items():
            padded[k][s, :n] = np.asarray(v)[t0:t0 + n, env]
    msl = min(max_ep, max_seq_len) if max_seq_len is not None else max_ep  # :806-810
    return {"num_sequences": S, "max_sequence_length": int(msl), "unpadded_slices": slices,
            "pad_mask": mask, "padded": padded,
            "first_idx": np.array([t0 * N + env for env, t0, _ in seqs], dtype=np.int64)}


# =================================================================================================
# MultiAgentReplayBuffer — components/multi_agent_replay_buffer.py:16-242
# =================================================================================================
class MarbOracle:
    """deque of per-step tuples of {agent: value}; sample = random.sample + numpy stacking +
    ``.float()`` (:108-162, utils/algo_utils.py:746-773).  Returns numpy fp32 arrays."""

    BINARY = ("done", "termination", "terminated", "truncation", "truncated")

    def __init__(self, memory_size: int, field_names, agent_ids):
        from collections import deque

        self.memory = deque(maxlen=memory_size)
        self.field_names, self.agent_ids = list(field_names), list(agent_ids)
        self.counter = 0

    def __len__(self):
        return len(self.memory)

    def save(self, *args, is_vectorised=False):
        if not is_vectorised:  # :164-173
            self.memory.append(args)
            self.counter += 1
            return
        n = len(next(iter(args[0].values())))  # :199

        def row(v, i):
            if isinstance(v, dict):
                return {k: np.asarray(x[i]) for k, x in v.items()}
            if isinstance(v, tuple):
                return tuple(np.asarray(x[i]) for x in v)
            return np.asarray(v[i])

        for i in range(n):  # :200-226
            self.memory.append(tuple({a: row(v, i) for a, v in arg.items()} for arg in args))
            self.counter += 1

    @staticmethod
    def _stack(items):  # stack_transitions :58-96
        def st(xs):
            a = np.array(xs)
            return a[:, None] if a.ndim == 1 else a

        if isinstance(items[0], dict):
            return {k: st([np.array(it[k]) for it in items]) for k in items[0]}
        if isinstance(items[0], tuple):
            return tuple(st([np.array(it[i]) for it in items]) for i in range(len(items[0])))
        return st(items)

    def sample(self, batch_size: int):
        import random

        exps = random.sample(self.memory, k=batch_size)  # :160

        def f32(x):
            return np.asarray(x).astype(np.float32)

        out = []
        for fi, f in enumerate(self.field_names):
            per = {}
            for a in self.agent_ids:
                ts = self._stack([e[fi][a] for e in exps])
                if f in self.BINARY and not isinstance(ts, (dict, tuple)) and not np.isnan(ts).any():
                    ts = ts.astype(np.uint8)  # :137-138
                if isinstance(ts, dict):
                    per[a] = {k: f32(v) for k, v in ts.items()}
                elif isinstance(ts, tuple):
                    per[a] = tuple(f32(v) for v in ts)
                else:
                    per[a] = f32(ts)
            out.append(per)
        return tuple(out)


# ---- array-based restatements for sizes the list trees cannot reach in seconds (C4: cap 2^24) --------
# Same arithmetic as SegTree / PerOracle above, on numpy fp64 arrays in heap order.  fp64 add / subtract /
# compare are single IEEE operations in numpy exactly as in CPython, so these are bit-identical to the
# list-based walk; tests/test_oracle_golden.py pins them against SegTree.retrieve and PerOracle at 2^16
# (and tests/test_oracle_vs_reference_cpu.py against the unmodified reference classes).
def np_trees_from_leaves(leaves: np.ndarray, cap: int):
    """(sum, min) heap arrays [2*cap] whose internal nodes are op(left, right) of their children — the
    tree the reference's sequential root walks leave behind (segment_tree.py:89-95).
    ``leaves`` [n <= cap] fp64; unfilled leaves are 0.0 / +inf (segment_tree.py:122,170)."""
    s = np.zeros(2 * cap, dtype=np.float64)
    m = np.full(2 * cap, np.inf, dtype=np.float64)
    s[cap:cap + len(leaves)] = leaves
    m[cap:cap + len(leaves)] = leaves
    lvl = cap >> 1
    while lvl >= 1:
        s[lvl:2 * lvl] = s[2 * lvl:4 * lvl:2] + s[2 * lvl + 1:4 * lvl:2]
        # builtin min(a, b) returns a on ties and b only when b < a (segment_tree.py:170)
        a, b = m[2 * lvl:4 * lvl:2], m[2 * lvl + 1:4 * lvl:2]
        m[lvl:2 * lvl] = np.where(b < a, b, a)
        lvl >>= 1
    return s, m


def np_retrieve(sum_tree: np.ndarray, cap: int, upperbound: np.ndarray) -> np.ndarray:
    """SumSegmentTree.retrieve (segment_tree.py:136-156) for a vector of upper bounds: strict '>' and one
    rounded subtraction per level, every element walking its own path."""
    ub = np.array(upperbound, dtype=np.float64, copy=True)
    assert np.all(ub >= 0) and np.all(ub <= sum_tree[1] + 1e-5)
    idx = np.ones(len(ub), dtype=np.int64)
    while idx[0] < cap:  # all paths have the same length
        left = 2 * idx
        lv = sum_tree[left]
        go_left = lv > ub
        ub = np.where(go_left, ub, ub - lv)
        idx = np.where(go_left, left, left + 1)
    return idx - cap


def np_sample_proportional(sum_tree: np.ndarray, cap: int, batch_size: int, u: np.ndarray) -> np.ndarray:
    """_sample_proportional (replay_buffer.py:357-381): u[i] = the fp32 torch.rand(1).item() of stratum i."""
    total = sum_tree[1]
    segment = total / batch_size
    i = np.arange(batch_size, dtype=np.float64)
    a = segment * i
    b = segment * (i + 1.0)
    ub = np.asarray(u, dtype=np.float32).astype(np.float64) * (b - a) + a  # four separately rounded fp64 ops
    return np_retrieve(sum_tree, cap, ub)


def np_calculate_weights(sum_tree: np.ndarray, min_tree: np.ndarray, cap: int, indices, size: int,
                         beta: float) -> np.ndarray:
    """_calculate_weights (replay_buffer.py:383-409); the two ``**`` stay CPython float pows (libm)."""
    total = float(sum_tree[1])
    p_min = float(min_tree[1]) / total
    max_weight = (p_min * size) ** -beta
    leaves = sum_tree[cap + np.asarray(indices, dtype=np.int64)]
    out = np.empty(len(leaves), dtype=np.float32)
    for k, lv in enumerate(leaves.tolist()):
        out[k] = ((lv / total) * size) ** -beta / max_weight
    return out
