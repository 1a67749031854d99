"""BASELINE INFRASTRUCTURE — times the reference's OWN CPU implementation of the experience path.

Imported only by ``bench.py``'s ``--impl reference`` arm and ``cpu_baseline`` legs (never by the
product).  Every timed call below is a call into the UNMODIFIED reference classes / methods loaded by
``oracle/ref_loader.py`` (from /root/reference, or from the byte-identical staged copy under
``oracle/_ref`` on the GPU box): ``PrioritizedReplayBuffer.sample / update_priorities``,
``MultiStepReplayBuffer.sample_from_indices``, ``ReplayBuffer.sample``,
``RolloutBuffer.compute_returns_and_advantages``.  ``kind`` is ``"reference"`` then.  (``TensorDict`` is
ref_loader's minimal stand-in — tensordict is not installed — whose indexing is per-key torch indexing,
which is what tensordict's own ``storage[idx]`` does.)  If neither copy of the reference is present the
legs fall back to the oracle port (``kind: "port"``) and say so.

Execution model: the reference trains the agents of a population SEQUENTIALLY in one Python process
against ONE shared ``memory`` (training/train_off_policy.py:46,249,343-345), so the headline CPU number
is one process (torch's intra-op threads are whatever the host gives it).  SETUP shortcuts, untimed and
stated in every ``sample`` string: trees are filled by direct leaf assignment + one bottom-up pass
(bit-identical to sequential ``_update_priority`` calls, SURVEY.md A.1 — 10^6..10^7 Python calls would
take minutes to hours) and storage rows are filled in bulk with random bytes.
"""

from __future__ import annotations

import os
import time

import numpy as np
import torch

from . import ref_loader
from . import xp_oracle as xo

ATARI = (84, 84, 4)


# ---- host facts -------------------------------------------------------------------------------------
def mem_available_bytes() -> int:
    try:
        with open("/proc/meminfo") as f:
            for ln in f:
                if ln.startswith("MemAvailable:"):
                    return int(ln.split()[1]) * 1024
    except OSError:
        pass
    return 16 << 30


def host_info() -> dict:
    model = ""
    try:
        with open("/proc/cpuinfo") as f:
            for ln in f:
                if ln.startswith("model name"):
                    model = ln.split(":", 1)[1].strip()
                    break
    except OSError:
        pass
    return {"host_cores": os.cpu_count(), "torch_threads": torch.get_num_threads(), "cpu_model": model,
            "mem_available_gb": round(mem_available_bytes() / 2**30, 1),
            "reference_source": ref_loader.REF_SOURCE}


def kind() -> str:
    return "reference" if ref_loader.available() else "port"


# ---- setup helpers (untimed) ---------------------------------------------------------------------------
def fill_rows(tensors, seconds: float, ram_bytes: float, seed: int = 0) -> int:
    """Write random bytes into rows [0, F) of every tensor in ``tensors`` (same leading dimension) until
    the time or RAM budget is reached; returns F.  Rows beyond F stay the untouched zero pages the
    reference's own ``torch.zeros_like`` allocation left (replay_buffer.py:68)."""
    rows = tensors[0].shape[0]
    per_row = sum(t[0].numel() * t.element_size() for t in tensors)
    max_rows = int(min(rows, ram_bytes // max(per_row, 1)))
    rng = np.random.default_rng(seed)
    block = torch.from_numpy(rng.integers(0, 256, (1 << 26) + 4096, dtype=np.uint8))
    t0 = time.perf_counter()
    done, k = 0, 0
    chunk = max(1, (1 << 28) // max(per_row, 1))
    while done < max_rows and time.perf_counter() - t0 < seconds:
        n = min(chunk, max_rows - done)
        for t in tensors:
            flat = t[done:done + n].view(-1).view(torch.uint8) if t.dtype != torch.uint8 else t[done:done + n].view(-1)
            off = 0
            while off < flat.numel():
                m = min(1 << 26, flat.numel() - off)
                s = (k * 1031) % 4096  # shifted so that rows are not periodic copies
                flat[off:off + m] = block[s:s + m]
                off += m
                k += 1
        done += n
    return done


def fill_tree_pair(sum_tree, min_tree, leaves: np.ndarray) -> None:
    """Direct leaf fill + one bottom-up pass over the reference's list trees (segment_tree.py:19-26).
    Yields exactly the tree the sequential ``__setitem__`` walks produce (every internal node is
    op(left, right) of final children); the levels are computed with numpy fp64 (the same
    single IEEE operations) and assigned to the reference's Python lists."""
    cap = sum_tree.capacity
    s, m = xo.np_trees_from_leaves(np.asarray(leaves, dtype=np.float64), cap)
    s[0], m[0] = sum_tree.tree[0], min_tree.tree[0]  # node 0 is unused; keep what the constructor put there
    sum_tree.tree[:] = s.tolist()
    min_tree.tree[:] = m.tolist()


def atari_transition(td_cls, n: int, g: torch.Generator):
    return td_cls({
        "obs": torch.randint(0, 256, (n, *ATARI), dtype=torch.uint8, generator=g),
        "action": torch.randint(0, 18, (n,), generator=g).float(),
        "reward": torch.randn(n, generator=g),
        "next_obs": torch.randint(0, 256, (n, *ATARI), dtype=torch.uint8, generator=g),
        "done": (torch.rand(n, generator=g) < 0.01).float()}, batch_size=[n])


# ---- C2: Rainbow PER + 3-step replay (BASELINE.json configs[1]) -----------------------------------------
class C2Reference:
    """The reference's ``PrioritizedReplayBuffer`` + ``MultiStepReplayBuffer`` pair at the config's
    capacity, driven as training/train_off_policy.py:363-376 drives it: per agent
    ``memory.sample(B, beta)`` -> ``n_step_memory.sample_from_indices(experiences["idxs"])`` ->
    ``memory.update_priorities(idxs, priorities)`` with ``idxs`` the (B,1) tensor of the batch and
    ``priorities`` an fp32 ndarray (algorithms/dqn_rainbow.py:487-488)."""

    ROW_BYTES = 2 * 84 * 84 * 4 + 12

    def __init__(self, cap=1_000_000, B=256, n_step=3, gamma=0.99, alpha=0.6, beta=0.4, num_envs=16,
                 fill_seconds=30.0, ram_fraction=0.6, seed=0):
        self.kind = kind()
        self.cap_config = cap
        if self.kind == "reference":
            # the reference's _init ZERO-FILLS max_size rows of every key (torch.zeros_like, :68): the whole
            # storage is resident, so the capacity this host can hold bounds max_size (stated in `sample`)
            fit = int(ram_fraction * mem_available_bytes() // (2 * self.ROW_BYTES))
            cap = max(4096, min(cap, fit))
        self.cap, self.B, self.beta = cap, B, beta
        rng = np.random.default_rng(seed)
        leaves = (np.abs(rng.standard_normal(cap)) + 1e-3) ** alpha
        self.prio = (np.abs(rng.standard_normal(B)) + 1e-6).astype(np.float32)
        g = torch.Generator().manual_seed(seed)
        t0 = time.perf_counter()
        if self.kind == "reference":
            ref = ref_loader.load_reference()
            self.per = ref.PrioritizedReplayBuffer(max_size=cap, alpha=alpha)
            self.nstep = ref.MultiStepReplayBuffer(max_size=cap, n_step=n_step, gamma=gamma)
            # real env steps through n-step -> PER (train_off_policy.py:341-345): the reference's own
            # _init allocates both storages (zeros_like of the expanded first row, replay_buffer.py:60-70)
            for _ in range(n_step + 2):
                one = self.nstep.add(atari_transition(ref.TensorDict, num_envs, g))
                if one is not None:
                    self.per.add(one)
            big = [buf.storage[k] for buf in (self.per, self.nstep) for k in ("obs", "next_obs")]
            self.filled = fill_rows(big, fill_seconds, float("inf"), seed)
            for buf in (self.per, self.nstep):
                buf._size, buf._cursor = cap, 0
                for k in ("action", "reward", "done"):
                    buf.storage[k][:] = torch.randn(buf.storage[k].shape, generator=g)
            self.per.tree_ptr = 0
            fill_tree_pair(self.per.sum_tree, self.per.min_tree, leaves)
            self.storage_rows = cap
        else:  # oracle port (no copy of the reference on this machine)
            self.per = xo.PerOracle(cap, alpha)
            self.per.size = cap
            cap2 = self.per.sum_tree.capacity
            self.per.sum_tree.tree[cap2:cap2 + cap] = leaves.tolist()
            self.per.min_tree.tree[cap2:cap2 + cap] = leaves.tolist()
            self.per.sum_tree.rebuild_from_leaves()
            self.per.min_tree.rebuild_from_leaves()
            rows = 4096
            self.storage_rows = self.filled = rows
            self.tables = [{"obs": rng.integers(0, 256, (rows, *ATARI), dtype=np.uint8),
                            "action": rng.standard_normal((rows, 1)).astype(np.float32),
                            "reward": rng.standard_normal((rows, 1)).astype(np.float32),
                            "next_obs": rng.integers(0, 256, (rows, *ATARI), dtype=np.uint8),
                            "done": rng.standard_normal((rows, 1)).astype(np.float32)} for _ in range(2)]
            self.gen = torch.Generator().manual_seed(seed)
        self.setup_s = time.perf_counter() - t0

    def agent_step(self):
        """One agent's learn-step buffer traffic; returns the number of sampled transitions."""
        if self.kind == "reference":
            experiences = self.per.sample(self.B, self.beta)
            n_step_experiences = self.nstep.sample_from_indices(experiences["idxs"])  # noqa: F841
            self.per.update_priorities(experiences["idxs"], self.prio)
        else:
            u = torch.rand(self.B, generator=self.gen).numpy()
            idx = self.per.sample_proportional(self.B, u)
            xo.gather(self.tables[0], idx % self.storage_rows)
            self.per.calculate_weights(idx, self.beta)
            xo.gather(self.tables[1], idx % self.storage_rows)
            self.per.update_priorities(idx, self.prio)
        return self.B

    def describe(self) -> str:
        if self.kind == "reference":
            return (f"unmodified reference classes ({ref_loader.REF_SOURCE} copy), 1 process, agents sequential on ONE "
                    f"shared memory; tree capacity 2^{int(np.log2(self.per.sum_tree.capacity))} filled by direct leaf "
                    f"assignment; max_size {self.cap} rows per buffer"
                    + ("" if self.cap == self.cap_config else f" (config: {self.cap_config}; bounded by host RAM — the "
                       f"reference zero-fills the whole storage, 2 x {self.cap * self.ROW_BYTES / 2**30:.0f} GiB here)")
                    + f", allocated by the reference's own _init, rows [0,{self.filled}) overwritten with random bytes "
                    f"(rest zeros); setup {self.setup_s:.0f} s untimed")
        return (f"ORACLE PORT (no reference copy on this machine): list trees at full capacity, numpy row storage "
                f"bounded to {self.storage_rows} rows (idx mod rows)")


def time_c2(seconds: float, max_agent_steps: int = 1 << 30, warmup_agent_steps: int = 2, **kw) -> dict:
    """sampled transitions / s of the C2 learn-step traffic on this host, one process."""
    w = C2Reference(**kw)
    torch.manual_seed(1234)
    for _ in range(warmup_agent_steps):
        w.agent_step()
    t0 = time.perf_counter()
    n = 0
    while n < max_agent_steps:
        w.agent_step()
        n += 1
        if time.perf_counter() - t0 > seconds:
            break
    wall = time.perf_counter() - t0
    return {"value": n * w.B / wall, "unit": "transitions/s", "cores": 1, "kind": w.kind,
            "sample": f"{n} agent-steps of {w.B} samples (sample + n-step sample_from_indices + update_priorities) "
                      f"in {wall:.1f} s; " + w.describe(),
            "agent_steps": n, "wall_s": round(wall, 2), **host_info()}


def _c2_worker(args):
    seconds, max_steps, seed, kw = args
    torch.set_num_threads(1)
    r = time_c2(seconds, max_steps, seed=seed, **kw)
    return r["agent_steps"], r["wall_s"], r["kind"]


def time_c2_population_parallel(seconds: float, workers: int, **kw) -> dict:
    """Upper bound the reference does NOT implement: ``workers`` independent processes, one agent and one
    PRIVATE memory each (the reference shares one memory and runs the agents one after the other).
    Reported beside the one-process headline, never as it."""
    import multiprocessing as mp

    kw = dict(kw)
    kw.setdefault("fill_seconds", 6.0)
    kw.setdefault("ram_fraction", 0.5 / workers)
    with mp.get_context("fork").Pool(workers) as pool:
        res = pool.map(_c2_worker, [(seconds, 1 << 30, s, kw) for s in range(workers)])
    steps, wall = sum(r[0] for r in res), max(r[1] for r in res)
    return {"value": steps * kw.get("B", 256) / wall, "unit": "transitions/s", "cores": workers, "kind": res[0][2],
            "sample": f"{workers} independent single-thread processes x 1 agent with a private memory each, "
                      f"{steps} agent-steps in {wall:.1f} s (not the reference's execution model: an upper bound)"}


# ---- C3: PPO RolloutBuffer GAE, 1024 envs x 128 steps x population 16 --------------------------------------
def time_c3(seconds: float, T=128, N=1024, P=16, gamma=0.99, lam=0.95) -> dict:
    """(t, env, agent) elements / s of RolloutBuffer.compute_returns_and_advantages (rollout_buffer.py:413-481,
    the method's own source, AST-extracted), called per agent as rollouts/on_policy.py:198-201 does."""
    rng = np.random.default_rng(3)
    data = [(rng.standard_normal((T, N)).astype(np.float32), rng.standard_normal((T, N)).astype(np.float32),
             rng.random((T, N)) < 0.02, rng.standard_normal(N).astype(np.float32), rng.random(N) < 0.02)
            for _ in range(P)]
    is_ref = ref_loader.available()
    fn = (lambda r, v, d, lv, ld: ref_loader.reference_gae(r, v, d, lv, ld, gamma, lam)) if is_ref else \
         (lambda r, v, d, lv, ld: xo.gae(r, v, d, lv, ld, gamma, lam))
    fn(*data[0])
    t0 = time.perf_counter()
    n = 0
    while True:
        fn(*data[n % P])
        n += 1
        if time.perf_counter() - t0 > seconds:
            break
    wall = time.perf_counter() - t0
    return {"value": n * T * N / wall, "unit": "steps/s", "cores": 1, "kind": "reference" if is_ref else "port",
            "sample": f"{n} per-agent calls of compute_returns_and_advantages ([{T},{N}]) in {wall:.1f} s, agents "
                      f"sequential, numpy single thread ({1e3 * wall / n:.1f} ms per agent)"}


# ---- C4: PER sum-tree at capacity 2^24, batch 4096 per agent -------------------------------------------------
def time_c4(seconds: float, cap=1 << 24, B=4096, alpha=0.6, beta=0.4) -> dict:
    """sampled transitions / s of the TREE operations of one agent's learn step on the unmodified
    PrioritizedReplayBuffer: _sample_proportional + _calculate_weights + update_priorities (no storage)."""
    rng = np.random.default_rng(4)
    leaves = (np.abs(rng.standard_normal(cap)) + 1e-3) ** alpha
    prio = (np.abs(rng.standard_normal(B)) + 1e-6).astype(np.float32)
    t0 = time.perf_counter()
    is_ref = ref_loader.available()
    if is_ref:
        per = ref_loader.load_reference().PrioritizedReplayBuffer(max_size=cap, alpha=alpha)
        per._size = cap
        fill_tree_pair(per.sum_tree, per.min_tree, leaves)

        def step():
            idx = per._sample_proportional(B)
            per._calculate_weights(idx, beta)
            per.update_priorities(idx, prio)
    else:
        per = xo.PerOracle(cap, alpha)
        per.size = cap
        s, m = xo.np_trees_from_leaves(leaves, cap)
        per.sum_tree.tree[:] = s.tolist()
        per.min_tree.tree[:] = m.tolist()
        g = torch.Generator().manual_seed(4)

        def step():
            idx = per.sample_proportional(B, torch.rand(B, generator=g).numpy())
            per.calculate_weights(idx, beta)
            per.update_priorities(idx, prio)
    setup = time.perf_counter() - t0
    torch.manual_seed(4)
    step()
    t0 = time.perf_counter()
    n = 0
    while True:
        step()
        n += 1
        if time.perf_counter() - t0 > seconds:
            break
    wall = time.perf_counter() - t0
    return {"value": n * B / wall, "unit": "transitions/s", "cores": 1, "kind": "reference" if is_ref else "port",
            "sample": f"{n} agent-steps of {B} (tree ops only: _sample_proportional + _calculate_weights + "
                      f"update_priorities) in {wall:.1f} s on 2^{int(np.log2(cap))}-leaf Python-list trees filled by direct leaf "
                      f"assignment (setup {setup:.0f} s untimed); 1 process"}


# ---- uniform replay: C5 (TD3, 10 M Humanoid rows, batch 1024) and C1 (DQN CartPole, 100 k rows, batch 64) ------
def _uniform_buffer(cap, row, num_envs, fill_seconds, ram_fraction, seed):
    """The reference's ReplayBuffer at ``cap`` rows; ``row(n, g)`` makes n transitions.  A few real adds let
    the reference's own _init allocate the storage; the rest is bulk-filled (untimed)."""
    g = torch.Generator().manual_seed(seed)
    ref = ref_loader.load_reference()
    probe = row(1, g)
    row_bytes = sum(max(v[0].numel(), 1) * v.element_size() for v in probe.values())
    cap = max(1024, min(cap, int(ram_fraction * mem_available_bytes() // row_bytes)))  # _init zero-fills it all
    buf = ref.ReplayBuffer(max_size=cap)
    for _ in range(3):
        buf.add(ref.TensorDict(row(num_envs, g), batch_size=[num_envs]))
    keys = list(buf.storage.keys())
    filled = fill_rows([buf.storage[k] for k in keys], fill_seconds, float("inf"), seed)
    buf._size, buf._cursor = cap, 0
    return buf, filled, cap


def time_uniform(seconds: float, cap: int, B: int, row, label: str, num_envs=16, fill_seconds=15.0,
                 ram_fraction=0.4, seed=5) -> dict:
    """sampled transitions / s of ReplayBuffer.sample(B) (replay_buffer.py:114-131: an O(size)
    torch.randperm per call + storage[indices]), one agent after the other in one process."""
    t0 = time.perf_counter()
    is_ref = ref_loader.available()
    if is_ref:
        cap_config = cap
        buf, filled, cap = _uniform_buffer(cap, row, num_envs, fill_seconds, ram_fraction, seed)

        def step():
            buf.sample(B)
    else:
        g = torch.Generator().manual_seed(seed)
        rows = min(cap, 65536)
        table = {k: v.numpy() for k, v in row(rows, g).items()}
        filled = rows

        def step():
            idx = torch.randperm(cap)[:B].numpy()
            xo.gather(table, idx % rows)
    setup = time.perf_counter() - t0
    torch.manual_seed(seed)
    step()
    t0 = time.perf_counter()
    n = 0
    while True:
        step()
        n += 1
        if time.perf_counter() - t0 > seconds:
            break
    wall = time.perf_counter() - t0
    return {"value": n * B / wall, "unit": "transitions/s", "cores": 1, "kind": "reference" if is_ref else "port",
            "sample": f"{label}: {n} calls of ReplayBuffer.sample({B}) in {wall:.1f} s ({1e3 * wall / n:.2f} ms per call, "
                      f"torch.randperm({cap}) inside every call); max_size {cap} rows"
                      + (f" (config: {cap_config}; bounded by host RAM)" if is_ref and cap != cap_config else "")
                      + f", [0,{filled}) overwritten with random bytes; "
                      f"1 process, torch threads {torch.get_num_threads()}; setup {setup:.0f} s untimed"}


def humanoid_row(n, g):
    return {"obs": torch.randn(n, 376, generator=g), "action": torch.rand(n, 17, generator=g) * 0.8 - 0.4,
            "reward": torch.randn(n, generator=g), "next_obs": torch.randn(n, 376, generator=g),
            "done": (torch.rand(n, generator=g) < 0.01).float()}


def cartpole_row(n, g):
    return {"obs": torch.randn(n, 4, generator=g), "action": torch.randint(0, 2, (n,), generator=g).float(),
            "reward": torch.ones(n), "next_obs": torch.randn(n, 4, generator=g),
            "done": (torch.rand(n, generator=g) < 0.02).float()}


def time_c5(seconds: float, cap=10_000_000, B=1024) -> dict:
    return time_uniform(seconds, cap, B, humanoid_row, "C5 TD3 uniform replay")


def time_c1(seconds: float, cap=100_000, B=64) -> dict:
    return time_uniform(seconds, cap, B, cartpole_row, "C1 DQN CartPole uniform replay", fill_seconds=5.0)


# ---- insert side: per env step, MultiStepReplayBuffer.add -> PrioritizedReplayBuffer.add ---------------------
def time_insert(seconds: float, cap=65536, num_envs=16, n_step=3, gamma=0.99, alpha=0.6) -> dict:
    """inserted rows / s of the per-step insert as training/train_off_policy.py:339-345 makes it: for each agent
    ``one_step = n_step_memory.add(transition)`` then ``memory.add(one_step)`` with ``num_envs`` Atari rows
    (the unmodified reference classes; every add walks both list trees once per new row)."""
    if not ref_loader.available():
        return {"value": None, "kind": "unavailable", "sample": "no copy of the reference on this machine"}
    ref = ref_loader.load_reference()
    per = ref.PrioritizedReplayBuffer(max_size=cap, alpha=alpha)
    nst = ref.MultiStepReplayBuffer(max_size=cap, n_step=n_step, gamma=gamma)
    g = torch.Generator().manual_seed(8)
    trs = [atari_transition(ref.TensorDict, num_envs, g) for _ in range(8)]

    def step(i):
        src = trs[i % 8]
        one = nst.add(ref.TensorDict({k: src[k] for k in src.keys()}, batch_size=[num_envs]))
        if one is not None:
            per.add(one)

    for i in range(n_step + 2):
        step(i)
    t0 = time.perf_counter()
    n = 0
    while True:
        step(n)
        n += 1
        if time.perf_counter() - t0 > seconds:
            break
    wall = time.perf_counter() - t0
    return {"value": n * num_envs / wall, "unit": "rows/s", "cores": 1, "kind": "reference",
            "sample": f"{n} agent-steps of {num_envs} rows (n_step_memory.add -> memory.add) in {wall:.1f} s "
                      f"({1e6 * wall / n:.0f} us per agent-step); max_size {cap}; 1 process"}
