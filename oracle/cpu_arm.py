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
