#!/usr/bin/env python
"""bench.py — the judged benchmark of the B200 experience-path engine.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c2|c3|c4|c5]

One "step" = one population step of the hot path on one batch of synthetic input.  The default
workload is BASELINE.json configs[1] (C2): Rainbow PER + 3-step replay, capacity 1M, batch 256,
population 8, 84x84x4 uint8 observations.  A step is, for every agent of the population:
PrioritizedReplayBuffer.sample(256, beta) [stratified sum-tree descent + IS weights + gather of
every key] + MultiStepReplayBuffer.sample_from_indices(idxs) [second gather] +
update_priorities(idxs, new_priorities) — SURVEY.md §8(d): 227 028 algorithmic bytes per sampled
transition.  The metric is sampled transitions per second.

  value : whole-job throughput, inputs (uniform draws, new priorities) already in HBM, the
          population batched into single launches, device-timed with CUDA events.
  e2e   : the same metric through the reference-facing Python API, one call per agent as
          train_off_policy.py makes them; uniforms are drawn on the host (torch mt19937) and
          priorities arrive as host arrays (dqn_rainbow.py:487), H2D inside the timed region,
          indices + IS weights read back D2H every step.
  roofline     : the dominant kernel (the row gather), algorithmic bytes / CUDA-event duration.
  cpu_baseline : the unmodified reference classes, timed on this host (N=1 only; oracle/cpu_arm.py).
  also         : device-timed numbers + rooflines for C3 (GAE) and C4 (PER tree 2^24), N=1 only.

`--impl reference` times the reference's own CPU implementation — the UNMODIFIED reference classes,
loaded by path (oracle/ref_loader.py) from /root/reference or from the byte-identical copy build()
staged under oracle/_ref — on the host, one process with the agents sequential on one shared memory
(the reference's execution model), same metric/config.

Multi-GPU (torchrun, one rank per GPU): weak scaling — every rank owns `population` agents; the
shared replay is replicated, so the (idx, priority) pairs of every rank are all-gathered (NCCL)
and applied on every replica in rank-major order; sampling and gathering are local.
"""

from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "per_sampled_transitions_per_s"
UNIT = "transitions/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c3", "c4", "c5"])
    ap.add_argument("--rows", type=int, default=0, help="override replay capacity (0 = config's)")
    ap.add_argument("--no-also", action="store_true", help="skip the secondary C3/C4 measurements")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    return ap.parse_args()


# =================================================================================================
# clocks: sampled DURING the timed region (pynvml in a thread; nvidia-smi fallback)
# =================================================================================================
class ClockSampler:
    BAD = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown")

    def __init__(self, index: int, period_s: float = 0.002):
        self.index, self.period = index, period_s
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._t = None
        self._nv = None
        try:
            import pynvml

            pynvml.nvmlInit()
            # honour CUDA_VISIBLE_DEVICES: map through the torch device's PCI bus id
            import torch

            bus = torch.cuda.get_device_properties(index).pci_bus_id
            dom = torch.cuda.get_device_properties(index).pci_domain_id
            dev = torch.cuda.get_device_properties(index).pci_device_id
            busid = f"{dom:08X}:{bus:02X}:{dev:02X}.0"
            try:
                self._h = pynvml.nvmlDeviceGetHandleByPciBusId(busid.encode())
            except Exception:
                self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self._nv = pynvml
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._nv = None

    def _poll_once(self):
        nv = self._nv
        mhz = nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)
        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)
        self.samples.append(mhz)
        for name, bit in (("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown),
                          ("hw_thermal_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown),
                          ("sw_thermal_slowdown", nv.nvmlClocksEventReasonSwThermalSlowdown),
                          ("sw_power_cap", nv.nvmlClocksEventReasonSwPowerCap)):
            if r & bit:
                self.reasons.add(name)

    def _run(self):
        while not self._stop.is_set():
            try:
                self._poll_once()
            except Exception:
                break
            self._stop.wait(self.period)

    def start(self):
        if self._nv is None:
            return
        self._stop.clear()
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self):
        if self._t is not None:
            self._stop.set()
            self._t.join()
            self._t = None

    def summary(self) -> dict:
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                    "samples": 0}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# =================================================================================================
# measured peaks
# =================================================================================================
def hbm_peak() -> tuple[float, str]:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


# =================================================================================================
# workload definitions (sizes from BASELINE.json / SURVEY.md §8d)
# =================================================================================================
ATARI = (84, 84, 4)
R_C2 = 2 * 84 * 84 * 4 + 4 + 4 + 4  # 56 460 B per stored transition row
C2 = dict(name="c2_rainbow_per_nstep3_cap1M_b256_pop8_obs84x84x4_u8", cap=1_000_000, B=256, P=8,
          n_step=3, gamma=0.99, alpha=0.6, beta=0.4, num_envs=16, L=20)
C3 = dict(name="c3_ppo_gae_1024envs_128steps_pop16", T=128, N=1024, P=16, gamma=0.99, lam=0.95)
C4 = dict(name="c4_per_tree_cap2p24_b4096_pop64", cap=1 << 24, B=4096, P=64, alpha=0.6, beta=0.4,
          L=24)
C5 = dict(name="c5_td3_uniform_cap10M_obs376_act17_b1024_pop32", cap=10_000_000, B=1024, P=32,
          obs=376, act=17)
R_C5 = 2 * 376 * 4 + 17 * 4 + 4 + 4  # 3 084 B


def per_bytes_per_sample(L: int) -> dict:
    """algorithmic bytes per sampled transition of the tree ops."""
    return {"descent": 8 * L + 4 + 8, "weights": 8 + 4,
            "update": 8 + 4 + 2 * (8 + L * 24)}


# =================================================================================================
# B200 arm
# =================================================================================================
def cuda_time_ms(fn, sync):
    import torch

    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync()
    a.record()
    fn()
    b.record()
    sync()
    return a.elapsed_time(b)


def graph_time_us(launches, reps: int):
    """Device time per launch of a list of closures, replayed from one CUDA graph (no host
    launch gaps).  Returns microseconds per closure call."""
    import torch

    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for fn in launches:  # warm-up outside capture
            fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for fn in launches:
            fn()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        g.replay()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / (reps * len(launches))


class C2Workload:
    """Rainbow PER + n-step replay (BASELINE.json configs[1])."""

    def __init__(self, args, rank, world, dev):
        import torch

        from agilerl_b200 import _native as nv
        from agilerl_b200 import xp
        from agilerl_b200.components import MultiStepReplayBuffer, PrioritizedReplayBuffer, TensorBatch

        self.torch, self.nv, self.xp, self.TensorBatch = torch, nv, xp, TensorBatch
        self.args, self.rank, self.world, self.dev = args, rank, world, dev
        c = dict(C2)
        self.c = c
        cap = args.rows or c["cap"]
        free, total = torch.cuda.mem_get_info(dev)
        need = 2 * cap * R_C2 + (6 << 30)
        if need > free:  # another tenant on the GPU: shrink and SAY so (config.rows)
            cap = int((free - (6 << 30)) // (2 * R_C2))
        self.cap = cap
        self.P, self.B = c["P"], c["B"]
        self.per = PrioritizedReplayBuffer(cap, alpha=c["alpha"], device=dev)
        self.nstep = MultiStepReplayBuffer(cap, n_step=c["n_step"], gamma=c["gamma"], device=dev)
        self.lib = nv.load()

    # -- synthetic transitions of the named shape ---------------------------------------------
    def transition(self, n, g):
        torch = self.torch
        return self.TensorBatch({
            "obs": torch.randint(0, 256, (n, *ATARI), dtype=torch.uint8, generator=g),
            "action": torch.randint(0, 18, (n,), generator=g).float(),
            "reward": torch.randn(n, generator=g),
            "next_obs": torch.randint(0, 256, (n, *ATARI), dtype=torch.uint8, generator=g),
            "done": (torch.rand(n, generator=g) < 0.01).float()}, batch_size=[n])

    def setup(self):
        torch, xp = self.torch, self.xp
        c, cap, dev = self.c, self.cap, self.dev
        g = torch.Generator().manual_seed(0)
        # real env steps through n-step -> PER (train_off_policy.py:341-345); allocates storage
        for _ in range(c["n_step"] + 2):
            one = self.nstep.add(self.transition(c["num_envs"], g))
            if one is not None:
                self.per.add(one)
        # bulk fill of the remaining rows straight into HBM (setup, untimed): random bytes
        # the shared memory is REPLICATED: storage and priorities are seeded identically on every rank;
        # only the agents' own uniform draws / new priorities (below) differ per rank
        gd = torch.Generator(device=dev).manual_seed(1)
        block = torch.randint(0, 256, (1 << 28,), dtype=torch.uint8, device=dev, generator=gd)
        for buf in (self.per, self.nstep):
            for path, t in buf._leaves.items():
                flat = t.view(-1).view(torch.uint8) if t.dtype != torch.uint8 else t.view(-1)
                if t.dtype == torch.uint8:
                    n = flat.numel()
                    off = 0
                    k = 0
                    while off < n:
                        m = min(block.numel() - 4096, n - off)
                        s = (k * 1031) % 4096  # shift the block so rows are not periodic copies
                        flat[off:off + m] = block[s:s + m]
                        off += m
                        k += 1
                else:
                    t.copy_(torch.randn(t.shape, device=dev, generator=gd).to(t.dtype))
            buf._size = cap
            buf._cursor = 0
        self.per.tree_ptr = 0
        del block
        # priorities of every stored row: |randn| + 1e-3 through the real update path
        pr = torch.randn(cap, device=dev, generator=gd).abs() + 1e-3
        self.per.update_priorities(torch.arange(cap, device=dev), pr)
        # resident inputs of the timed steps (rotated)
        n = self.P * self.B
        self.R = 4
        self.u = [torch.rand(n, generator=torch.Generator().manual_seed(10 + i + 1000 * self.rank)).to(dev)
                  for i in range(self.R)]
        gr = torch.Generator(device=dev).manual_seed(77 + self.rank)
        self.pr = [(torch.randn(n, device=dev, generator=gr).abs() + 1e-6) for _ in range(self.R)]
        self.idx = torch.empty(n, dtype=torch.int64, device=dev)
        self.w = torch.empty(n, dtype=torch.float32, device=dev)
        self.err = torch.zeros(1, dtype=torch.int32, device=dev)
        # two output sets per buffer; every set is written once per step (231 MB per step)
        self.outs = []
        for _ in range(2):
            o = {}
            for name, buf in (("per", self.per), ("nstep", self.nstep)):
                paths = list(buf._leaves.keys())
                out = [torch.empty((n, *buf._leaves[p].shape[1:]), dtype=buf._leaves[p].dtype, device=dev)
                       for p in paths]
                keys = xp.make_keys([(buf._leaves[p], out[i]) for i, p in enumerate(paths)])
                o[name] = (keys, len(paths), out, paths)
            self.outs.append(o)
        _, _, st, h = self.per._ensure_trees()
        self.h, self.state = h, self.per._state
        self.px = None
        if self.world > 1:
            self.all_idx = torch.empty(self.world * n, dtype=torch.int64, device=dev)
            self.all_pr = torch.empty(self.world * n, dtype=torch.float32, device=dev)
            from agilerl_b200 import peer

            if peer.available() and os.environ.get("XP_BENCH_NCCL") != "1":
                try:
                    # every rank's (indices, priorities) live in NVLink peer-mapped memory; the
                    # exchange is the library's own one-kernel all-gather (xp_peer_allgather)
                    self.px = peer.PeerExchange(12 * n, [(0, 8 * n, torch.int64), (8 * n, 4 * n, torch.float32)], dev)
                    self.idx_blk = [self.px.view(k, 0, n, torch.int64) for k in (0, 1)]
                    for k in (0, 1):  # the learner's new priorities are produced straight into the block
                        self.px.view(k, 8 * n, n, torch.float32).copy_(self.pr[k])
                except Exception as e:  # pragma: no cover - depends on the box
                    sys.stderr.write(f"bench.py: peer exchange unavailable ({e}); using NCCL all-gather\n")
                    self.px = None
        torch.cuda.synchronize()

    @property
    def units_per_step(self):
        return self.P * self.B


    def gather_bytes(self):
        """algorithmic bytes of ONE gather launch: (2R + 8) per sample."""
        return self.P * self.B * (2 * R_C2 + 8)

    def step(self, i, ev=None):
        import ctypes as C

        lib, c = self.lib, self.c
        stream = self.torch.cuda.current_stream().cuda_stream
        n = self.P * self.B
        u, pr, o = self.u[i % self.R], self.pr[i % self.R], self.outs[i % 2]
        chk = self.nv.check
        if self.px is not None:
            self.idx = self.idx_blk[self.px.current]  # sampled indices land in the peer-mapped block
        chk(lib.xp_per_sample(C.byref(self.h), u.data_ptr(), self.B, self.P, self.cap, c["beta"],
                              self.idx.data_ptr(), self.w.data_ptr(), stream))
        if ev is not None:
            ev[0].record()
        k = o["per"]
        chk(lib.xp_gather(k[0], k[1], self.idx.data_ptr(), n, self.cap, self.err.data_ptr(), stream))
        if ev is not None and len(ev) > 2:
            ev[2].record()
        k = o["nstep"]
        chk(lib.xp_gather(k[0], k[1], self.idx.data_ptr(), n, self.cap, self.err.data_ptr(), stream))
        if ev is not None:
            ev[1].record()
        if self.px is not None:
            all_idx, all_pr = self.px.allgather()
            chk(lib.xp_per_update(C.byref(self.h), self.state.data_ptr(), all_idx.data_ptr(),
                                  all_pr.data_ptr(), self.world * n, self.cap, c["alpha"],
                                  self.err.data_ptr(), stream))
        elif self.world > 1:
            import torch.distributed as dist

            # both all-gathers in ONE NCCL launch (ncclGroupStart/End)
            with dist._coalescing_manager(device=self.dev):
                dist.all_gather_into_tensor(self.all_idx, self.idx)
                dist.all_gather_into_tensor(self.all_pr, pr)
            chk(lib.xp_per_update(C.byref(self.h), self.state.data_ptr(), self.all_idx.data_ptr(),
                                  self.all_pr.data_ptr(), self.world * n, self.cap, c["alpha"],
                                  self.err.data_ptr(), stream))
        else:
            chk(lib.xp_per_update(C.byref(self.h), self.state.data_ptr(), self.idx.data_ptr(),
                                  pr.data_ptr(), n, self.cap, c["alpha"], self.err.data_ptr(), stream))
        if ev is not None and len(ev) > 3:
            ev[3].record()

    def capture_steps(self):
        """R consecutive steps (every rotating input set once; an even number of exchanges, so the two
        peer-mapped blocks alternate consistently) as ONE CUDA graph: at N > 1 no host work sits between the
        rendez-vous of consecutive steps (xp_peer_allgather_graph keeps the exchange epoch on the device)."""
        torch = self.torch
        side = torch.cuda.Stream(self.dev)
        side.wait_stream(torch.cuda.current_stream(self.dev))
        if self.px is not None:
            assert self.px.epoch % 2 == 0, "capture on an even exchange count"
        g = torch.cuda.CUDAGraph()
        n0 = self.lib.xp_launch_count()
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                for i in range(self.R):
                    self.step(i)
        self.launches_per_graph = int(self.lib.xp_launch_count() - n0)  # kernel nodes of one replay
        torch.cuda.current_stream(self.dev).wait_stream(side)
        if self.px is not None:
            self.px.epoch -= self.R  # nothing ran during capture
        return g

    def replay(self, g):
        g.replay()
        if self.px is not None:
            self.px.advance(self.R)

    def replica_checksum(self):
        """Two 64-bit checksums of everything the replicas must agree on after the timed steps: every
        (sum, min) node of the tree pair, max_priority, and the ring / tree cursors."""
        torch = self.torch
        _, _, st, _ = self.per._ensure_trees()
        bits = st.nodes.view(-1).view(torch.int64)[2:]  # node 0 is unused
        w = torch.arange(1, bits.numel() + 1, device=bits.device, dtype=torch.int64) * 2654435761 + 1
        host = torch.tensor([self.per._cursor, self.per.tree_ptr, self.per._size], dtype=torch.int64, device=bits.device)
        return torch.stack([bits.sum(), (bits * w).sum(), self.per._state.view(torch.int64)[0], host.sum()])

    def verify(self):
        """size-independent property at full size: the gathered rows equal torch indexing of the
        storage, indices are within [0, size), the tree root equals the sum of the leaves'
        pairwise reduction (checked in tests), error flag clear."""
        torch = self.torch
        self.step(0)
        torch.cuda.synchronize()
        assert int(self.err.item()) == 0, "device error flag set"
        idx = self.idx
        assert int(idx.min()) >= 0 and int(idx.max()) < self.cap
        o = self.outs[0]
        for name, buf in (("per", self.per), ("nstep", self.nstep)):
            _, _, out, paths = o[name]
            for t, p in zip(out, paths):
                sel = idx[:64]
                assert torch.equal(t[:64], buf._leaves[p][sel]), f"gather mismatch on {name}/{p}"

    # -- e2e through the public API ---------------------------------------------------------------
    def e2e_setup(self):
        torch = self.torch
        n = self.B
        g = torch.Generator().manual_seed(77)
        self.prio_host = [(torch.randn(n, generator=g).abs() + 1e-6).pin_memory().numpy()
                          for _ in range(self.P * 3)]
        self.res_idx_dev = torch.empty(self.P * n, 1, dtype=torch.int64, device=self.dev)
        self.res_w_dev = torch.empty(self.P * n, 1, dtype=torch.float32, device=self.dev)
        self.res_idx_host = torch.empty(self.P * n, 1, dtype=torch.int64).pin_memory()
        self.res_w_host = torch.empty(self.P * n, 1, dtype=torch.float32).pin_memory()
        self.res_ev = torch.cuda.Event()
        self.rep = None
        # set-up, outside the timed loop: the two buffers are filled in lock-step, so
        # memory.sample may gather both tables in one launch; the loop below keeps the reference's calls
        import agilerl_b200

        agilerl_b200.link_n_step(self.per, self.nstep)
        if self.world > 1:
            # the SAME algorithm as `value`: every agent's update is exchanged and applied on every replica
            from agilerl_b200.population import ReplicatedReplay

            self.rep = ReplicatedReplay(self.per)
        torch.manual_seed(1234 + self.rank)

    def e2e_step(self, i):
        torch = self.torch
        c = self.c
        idxs, ws = [], []
        for a in range(self.P):
            batch = self.per.sample(self.B, c["beta"])           # host draws + H2D u inside
            nb = self.nstep.sample_from_indices(batch["idxs"])    # noqa: F841 (n-step batch)
            idxs.append(batch["idxs"])
            ws.append(batch["weights"])
            prio = self.prio_host[(i * self.P + a) % len(self.prio_host)]
            if self.rep is not None:
                self.rep.update_priorities(batch["idxs"], prio)  # exchange (NVLink peer kernel) + replicated update
            else:
                self.per.update_priorities(batch["idxs"], prio)
        # D2H: the step's result (sampled indices + IS weights of all agents) into pinned memory
        n = self.P * self.B
        torch.cat(idxs, out=self.res_idx_dev)
        torch.cat(ws, out=self.res_w_dev)
        self.res_idx_host.copy_(self.res_idx_dev, non_blocking=True)
        self.res_w_host.copy_(self.res_w_dev, non_blocking=True)
        self.res_ev.record()
        self.res_ev.synchronize()
        return self.res_idx_host, self.res_w_host

    def e2e_bytes(self):
        n = self.P * self.B
        return n * 4 + n * 4, n * 8 + n * 4  # H2D: u + priorities; D2H: idxs + weights


def bench_gae(dev, sets=8, P=None, reps=8, rounds=5):
    """C3: one launch computes GAE for the whole population; cold-L2 by rotating `sets` buffer
    sets (> 126 MB in total).  One graph holds `rounds` passes over the sets (40 launches), so that the boundary
    between two graph replays — which is not the kernel's — is paid once per 40 launches, not once per 8."""
    import torch

    from agilerl_b200 import xp

    c = C3
    P = P or c["P"]
    T, N = c["T"], c["N"]
    gd = torch.Generator(device=dev).manual_seed(3)
    bufs = []
    for _ in range(sets):
        r = torch.randn(P, T, N, device=dev, generator=gd)
        v = torch.randn(P, T, N, device=dev, generator=gd)
        d = torch.rand(P, T, N, device=dev, generator=gd) < 0.02
        lv = torch.randn(P, N, device=dev, generator=gd).double()
        ld = (torch.rand(P, N, device=dev, generator=gd) < 0.02).double()
        bufs.append((r, v, d, lv, ld, torch.empty_like(r), torch.empty_like(r)))

    def mk(b):
        r, v, d, lv, ld, a, g = b
        return lambda: xp.gae(r, v, d, lv, ld, a, g, T, N, P, T * N, c["gamma"], c["lam"])

    us = graph_time_us([mk(b) for _ in range(rounds) for b in bufs], reps)
    steps = P * T * N
    return {"workload": c["name"] if P == c["P"] else f"{c['name']}_x{P // c['P']}",
            "metric": "gae_steps_per_s", "value": steps / (us * 1e-6), "unit": "steps/s",
            "us_per_launch": us, "launches_timed": reps * rounds * sets, "launches_per_graph": rounds * sets,
            "algorithmic_bytes": steps * 17,
            "l2": f"rotating {sets} buffer sets ({sets * steps * 17 / 1e6:.0f} MB)"}


def bench_gae_per_agent(dev, sets=8, reps=20):
    """C3 as the DROP-IN issues it: rollouts/on_policy.py:198-201 calls compute_returns_and_advantages per
    agent, i.e. 16 launches with P = 1 (32 CTAs each) — timed back to back over the agents of a slab, and one
    such launch alone (rotating 64 single-agent buffer sets, > L2 in total)."""
    import torch

    from agilerl_b200 import xp

    c = C3
    P, T, N = c["P"], c["T"], c["N"]
    gd = torch.Generator(device=dev).manual_seed(6)

    def mk_set(p_):
        r = torch.randn(p_, T, N, device=dev, generator=gd)
        v = torch.randn(p_, T, N, device=dev, generator=gd)
        d = torch.rand(p_, T, N, device=dev, generator=gd) < 0.02
        lv = torch.randn(p_, N, device=dev, generator=gd).double()
        ld = (torch.rand(p_, N, device=dev, generator=gd) < 0.02).double()
        return r, v, d, lv, ld, torch.empty_like(r), torch.empty_like(r)

    def one(b, p_):
        r, v, d, lv, ld, a, g = b
        return lambda: xp.gae(r[p_], v[p_], d[p_], lv[p_], ld[p_], a[p_], g[p_], T, N, 1, T * N, c["gamma"], c["lam"])

    slabs = [mk_set(P) for _ in range(sets)]
    us16 = graph_time_us([one(b, p_) for b in slabs for p_ in range(P)], reps) * P
    singles = [mk_set(1) for _ in range(64)]
    us1 = graph_time_us([one(b, 0) for b in singles], reps)
    steps = P * T * N
    return {"workload": c["name"] + "_per_agent_launches", "metric": "gae_steps_per_s",
            "value": steps / (us16 * 1e-6), "unit": "steps/s", "us_16_launches_of_P1": us16, "us_one_P1_launch": us1,
            "algorithmic_bytes": steps * 17, "roofline": roofline(steps * 17, us16),
            "roofline_one_P1_launch": roofline(T * N * 17, us1),
            "l2": f"rotating {sets} slabs of 16 agents / 64 single-agent sets (> L2 in total)"}


def bench_per_tree(dev, reps=10):
    """C4: PER tree cap 2^24, 64 agents x 4096 samples per launch; sample(+weights) and update."""
    import torch

    from agilerl_b200 import xp

    c = C4
    cap, B, P = c["cap"], c["B"], c["P"]
    nodes = torch.empty(2 * cap, 2, dtype=torch.float64, device=dev)
    aux = torch.zeros(2 * cap, dtype=torch.int32, device=dev)
    t = xp.make_tree(nodes[:, 0], nodes[:, 1], 2, cap, aux)
    xp.tree_init(t, torch.cuda.current_stream().cuda_stream)
    state = torch.ones(1, dtype=torch.float64, device=dev)
    gd = torch.Generator(device=dev).manual_seed(4)
    xp.per_add(t, state, 0, cap, cap, c["alpha"])
    pr0 = torch.randn(cap, device=dev, generator=gd).abs() + 1e-3
    xp.per_update(t, state, torch.arange(cap, device=dev), pr0, cap, c["alpha"])
    del pr0
    n = P * B
    R = 4
    u = [torch.rand(n, device=dev, generator=gd) for _ in range(R)]
    pr = [torch.randn(n, device=dev, generator=gd).abs() + 1e-6 for _ in range(R)]
    idx = [torch.empty(n, dtype=torch.int64, device=dev) for _ in range(R)]
    w = torch.empty(n, dtype=torch.float32, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)

    def sample(i):
        return lambda: xp.per_sample(t, u[i], B, P, cap, c["beta"], idx[i], w)

    def update(i):
        return lambda: xp.per_update(t, state, idx[i], pr[i], cap, c["alpha"], err)

    us_s = graph_time_us([sample(i) for i in range(R)], reps)
    us_u = graph_time_us([update(i) for i in range(R)], reps)
    seq = []
    for i in range(R):
        seq += [sample(i), update(i)]
    us_b = 2 * graph_time_us(seq, reps)
    assert int(err.item()) == 0
    # skewed batches: fresh transitions carry max_priority and are drawn together
    # (replay_buffer.py:306-309) — a share of the batch falls into the most recently written leaves
    skew = {}
    for share, hot in ((0.1, 4096), (0.3, 4096), (0.3, 65536)):
        sk = []
        for i in range(R):
            k = int(n * share)
            ix = torch.randint(0, cap, (n,), device=dev, generator=gd)
            ix[:k] = cap - hot + torch.randint(0, hot, (k,), device=dev, generator=gd)
            sk.append(ix[torch.randperm(n, device=dev, generator=gd)])
        us = graph_time_us([(lambda i=i: xp.per_update(t, state, sk[i], pr[i], cap, c["alpha"], err)) for i in range(R)], reps)
        skew[f"{int(share * 100)}pct_in_last_{hot}_leaves"] = round(us, 2)
    assert int(err.item()) == 0
    bps = per_bytes_per_sample(c["L"])
    tot = sum(bps.values())
    return {"workload": c["name"], "metric": METRIC, "value": n / (us_b * 1e-6), "unit": UNIT,
            "us_sample": us_s, "us_update": us_u, "us_sample_plus_update": us_b,
            "us_update_skewed": skew, "skew_worst_over_uniform": round(max(skew.values()) / us_u, 2),
            "algorithmic_bytes_per_sample": tot, "algorithmic_bytes": n * tot,
            "l2": "tree 805 MB > L2; top levels L2-resident by design"}


def bench_uniform(dev, reps=10, P=None):
    """C5: TD3 uniform replay, 10 M Humanoid-shape rows (30.8 GB), 32 agents x 1024 per launch:
    xp_randperm_prefix (torch.randperm(size)[:B] from the raw mt19937 words, replay_buffer.py:125)
    + xp_gather of every key (:126-131)."""
    import torch

    from agilerl_b200 import xp

    c = C5
    cap, B, P = c["cap"], c["B"], P or c["P"]
    n = P * B
    shapes = [("obs", (c["obs"],)), ("action", (c["act"],)), ("reward", (1,)), ("next_obs", (c["obs"],)),
              ("done", (1,))]
    gd = torch.Generator(device=dev).manual_seed(5)
    store = [torch.empty(cap, *sh, dtype=torch.float32, device=dev) for _, sh in shapes]
    for t in store:  # touch every page once; values are irrelevant to the timing
        t.view(-1)[::1024].fill_(1.0)
    R = 4
    outs = [[torch.empty(n, *sh, dtype=torch.float32, device=dev) for _, sh in shapes] for _ in range(R)]
    raw = [torch.randint(-2**31, 2**31 - 1, (n,), device=dev, generator=gd, dtype=torch.int64).to(torch.int32)
           for _ in range(R)]
    idx = [torch.empty(n, dtype=torch.int64, device=dev) for _ in range(R)]
    err = torch.zeros(1, dtype=torch.int32, device=dev)

    def draw(i):
        return lambda: xp.randperm_prefix(raw[i], B, P, cap, idx[i])

    def gather(i):
        return lambda: xp.gather(list(zip(store, outs[i])), idx[i], cap, err)

    for i in range(R):
        draw(i)()
    us_d = graph_time_us([draw(i) for i in range(R)], reps)
    us_g = graph_time_us([gather(i) for i in range(R)], reps)
    seq = []
    for i in range(R):
        seq += [draw(i), gather(i)]
    us_b = 2 * graph_time_us(seq, reps)
    assert int(err.item()) == 0
    for i in range(R):  # every agent's indices are distinct and in range (a randperm prefix)
        a = idx[i].view(P, B)
        assert int(a.min()) >= 0 and int(a.max()) < cap
        assert all(int(torch.unique(a[p]).numel()) == B for p in (0, P - 1))
    alg = n * (2 * R_C5 + 8)
    return {"workload": c["name"], "metric": "uniform_sampled_transitions_per_s", "value": n / (us_b * 1e-6),
            "unit": UNIT, "us_randperm_prefix": us_d, "us_gather": us_g, "us_sample": us_b, "row_bytes": R_C5,
            "algorithmic_bytes": alg, "gather_roofline": roofline(alg, us_g),
            "l2": "random rows of a 30.8 GB storage; 4 rotating output sets of 101 MB"}


def bench_insert(dev, reps=10):
    """Insert side (SURVEY.md 8d metric 3; train_off_policy.py:339-345): (a) the per-step path through the
    classes — for each of 8 agents MultiStepReplayBuffer.add(16 env rows) -> PrioritizedReplayBuffer.add —
    rows/s with device-resident transitions (host-driven: ~10 small launches per agent); (b) the ring-insert
    kernel on a bulk batch (offline ingestion, train_offline.py:190: ReplayBuffer.add of 16 384 Atari rows),
    roofline 2R bytes per row."""
    import torch

    from agilerl_b200 import xp
    from agilerl_b200.components import MultiStepReplayBuffer, PrioritizedReplayBuffer, TensorBatch

    c = C2
    cap, E, P = 65536, c["num_envs"], c["P"]
    per = PrioritizedReplayBuffer(cap, alpha=c["alpha"], device=dev)
    nst = MultiStepReplayBuffer(cap, n_step=c["n_step"], gamma=c["gamma"], device=dev)
    gd = torch.Generator(device=dev).manual_seed(8)

    def rows(n):
        return {"obs": torch.randint(0, 256, (n, *ATARI), dtype=torch.uint8, device=dev, generator=gd),
                "action": torch.randint(0, 18, (n,), device=dev, generator=gd).float(),
                "reward": torch.randn(n, device=dev, generator=gd),
                "next_obs": torch.randint(0, 256, (n, *ATARI), dtype=torch.uint8, device=dev, generator=gd),
                "done": (torch.rand(n, device=dev, generator=gd) < 0.01).float()}

    trs = [rows(E) for _ in range(8)]

    def step(i):
        for a in range(P):
            one = nst.add(TensorBatch(dict(trs[(i + a) % 8]), batch_size=[E]))
            if one is not None:
                per.add(one)

    for i in range(5):
        step(i)
    torch.cuda.synchronize()
    K = 50
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record()
    for i in range(K):
        step(i)
    b.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    api_ms = max(a.elapsed_time(b), wall * 1e3) / K
    # (b) bulk ring insert, kernel-timed, rotating two source batches and a moving cursor
    nb = 16384
    big = torch.empty(4 * nb, *ATARI, dtype=torch.uint8, device=dev)
    small = torch.empty(4 * nb, 1, dtype=torch.float32, device=dev)
    src = [rows(nb) for _ in range(2)]

    def ins(i):
        pairs = [(src[i % 2]["obs"], big), (src[i % 2]["reward"].view(nb, 1), small)]
        return lambda: xp.ring_insert(pairs, nb, (i * nb + 1234) % (4 * nb), 4 * nb)

    us = graph_time_us([ins(i) for i in range(4)], reps)
    alg = nb * 2 * (84 * 84 * 4 + 4)
    return {"workload": "c2_insert_nstep3_to_per_16envs_x_8agents", "metric": "inserted_rows_per_s",
            "value": P * E / (api_ms * 1e-3), "unit": "rows/s", "ms_per_population_step": api_ms,
            "api": "per agent: MultiStepReplayBuffer.add -> PrioritizedReplayBuffer.add (device-resident rows)",
            "bulk_ring_insert": {"rows": nb, "us_per_launch": us, "algorithmic_bytes": alg,
                                 "rows_per_s": nb / (us * 1e-6), "roofline": roofline(alg, us)},
            "l2": "bulk: 2 x 462 MB source batches into a 1.8 GB ring (> L2)"}


def bench_c1(dev):
    """C1 (BASELINE.json configs[0], the reference's CPU-runnable case): DQN CartPole uniform replay,
    capacity 100 000, batch 64, population 4 — through the public API, one ReplayBuffer.sample per agent
    (exact torch.randperm stream, replay_buffer.py:125), batch read back to the host every step."""
    import torch

    from agilerl_b200.components import ReplayBuffer, TensorBatch

    cap, B, P = 100_000, 64, 4
    buf = ReplayBuffer(cap, device=dev)
    g = torch.Generator().manual_seed(1)
    n = 10_000
    for _ in range(cap // n):
        buf.add(TensorBatch({"obs": torch.randn(n, 4, generator=g), "action": torch.randint(0, 2, (n,), generator=g).float(),
                             "reward": torch.ones(n), "next_obs": torch.randn(n, 4, generator=g),
                             "done": (torch.rand(n, generator=g) < 0.02).float()}, batch_size=[n]))
    torch.manual_seed(0)
    for _ in range(10):
        buf.sample(B)
    torch.cuda.synchronize()
    K = 300
    t0 = time.perf_counter()
    for _ in range(K):
        for _a in range(P):
            s_ = buf.sample(B)
        s_["reward"].cpu()
    wall = time.perf_counter() - t0
    return {"workload": "c1_dqn_cartpole_uniform_cap100k_b64_pop4", "metric": "uniform_sampled_transitions_per_s",
            "value": K * P * B / wall, "unit": UNIT, "ms_per_population_step": 1e3 * wall / K,
            "api": "ReplayBuffer.sample(64) per agent, host-timed incl. D2H of the last batch's rewards; "
                   "exact torch generator advance (size-1 words per call)",
            "note": "latency-bound: 4 x 64 rows of 44 B; the CPU-reference config, GPU number for completeness"}


def bench_per_tree_replicated(dev, rank, world, reps=10):
    """C4 strong scaling (BASELINE.json configs[3]: population 64 sharded over the GPUs): every rank samples for
    its 64/world agents from ITS replica of the 2^24-leaf tree, the (index, priority) pairs of all ranks are
    exchanged over NVLink peer memory (xp_peer_allgather_graph) and EVERY replica applies all 262 144 updates in
    rank-major order — the price of keeping the shared memory bit-identical.  One CUDA graph of 4 steps."""
    import torch
    import torch.distributed as dist

    from agilerl_b200 import peer, xp

    c = C4
    cap, B, P = c["cap"], c["B"], c["P"] // world
    n = P * B
    nodes = torch.empty(2 * cap, 2, dtype=torch.float64, device=dev)
    aux = torch.zeros(2 * cap, dtype=torch.int32, device=dev)
    t = xp.make_tree(nodes[:, 0], nodes[:, 1], 2, cap, aux)
    xp.tree_init(t, torch.cuda.current_stream().cuda_stream)
    state = torch.ones(1, dtype=torch.float64, device=dev)
    gd = torch.Generator(device=dev).manual_seed(4)  # the replicated tree: same seed on every rank
    xp.per_add(t, state, 0, cap, cap, c["alpha"])
    pr0 = torch.randn(cap, device=dev, generator=gd).abs() + 1e-3
    xp.per_update(t, state, torch.arange(cap, device=dev), pr0, cap, c["alpha"])
    del pr0
    gr = torch.Generator(device=dev).manual_seed(400 + rank)  # this rank's agents
    R = 4
    u = [torch.rand(n, device=dev, generator=gr) for _ in range(R)]
    px = peer.PeerExchange(12 * n, [(0, 8 * n, torch.int64), (8 * n, 4 * n, torch.float32)], dev)
    for k in (0, 1):
        px.view(k, 8 * n, n, torch.float32).copy_(torch.randn(n, device=dev, generator=gr).abs() + 1e-6)
    idx_blk = [px.view(k, 0, n, torch.int64) for k in (0, 1)]
    w = torch.empty(n, dtype=torch.float32, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)

    def step(i):
        def run():
            xp.per_sample(t, u[i], B, P, cap, c["beta"], idx_blk[i & 1], w)
            all_idx, all_pr = px.allgather()
            xp.per_update(t, state, all_idx, all_pr, cap, c["alpha"], err)
        return run

    torch.cuda.synchronize()
    dist.barrier()
    us = graph_time_us([step(i) for i in range(R)], reps)
    px.epoch = int(px._epoch_dev[0].item())  # the graph replays advanced the device epoch on their own
    px.check()
    assert int(err.item()) == 0
    tt = torch.tensor([us], device=dev, dtype=torch.float64)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    us = float(tt.item())
    bits = nodes.view(-1).view(torch.int64)[2:]
    ck = torch.stack([bits.sum(), (bits * (torch.arange(bits.numel(), device=dev) * 2654435761 + 1)).sum(),
                      state.view(torch.int64)[0]])
    allck = [torch.empty_like(ck) for _ in range(world)]
    dist.all_gather(allck, ck)
    total = c["P"] * B
    tot_bytes = total * sum(per_bytes_per_sample(c["L"]).values())
    return {"workload": c["name"] + f"_sharded_x{world}", "metric": METRIC, "value": total / (us * 1e-6), "unit": UNIT,
            "scaling": "strong", "agents_per_gpu": P, "us_sample_exchange_update": us,
            "updates_applied_per_replica": total, "replicas_identical": all(bool(torch.equal(allck[0], x)) for x in allck[1:]),
            "algorithmic_bytes_all_replicas": tot_bytes,
            "note": "per replica: 1/N of the descents, ALL of the updates (replicated tree) + the NVLink exchange"}


def bench_replicated_insert(dev, rank, world, steps=60):
    """Insert side at N > 1 (north_star: "each step's newly inserted transitions broadcast over NVLink"): every rank
    contributes the rows of its 8 agents x 16 envs per env step through population.ReplicatedReplay.add — ONE kernel
    per rank (xp_peer_ring_insert) signals, waits and pulls every rank's rows straight into the local ring, then the
    new leaves are set; all replicas insert world x 128 rows at the same cursor.  Host-inclusive wall time (it is a
    host-driven call per env step), max over ranks; replicas compared afterwards."""
    import torch
    import torch.distributed as dist

    from agilerl_b200.components import PrioritizedReplayBuffer, TensorBatch
    from agilerl_b200.population import ReplicatedReplay

    c = C2
    n = c["P"] * c["num_envs"]
    mem = PrioritizedReplayBuffer(32768, alpha=c["alpha"], device=dev)
    rep = ReplicatedReplay(mem)
    gd = torch.Generator(device=dev).manual_seed(90 + rank)
    rows = [{"obs": torch.randint(0, 256, (n, *ATARI), dtype=torch.uint8, device=dev, generator=gd),
             "action": torch.randint(0, 18, (n,), device=dev, generator=gd).float(),
             "reward": torch.randn(n, device=dev, generator=gd),
             "next_obs": torch.randint(0, 256, (n, *ATARI), dtype=torch.uint8, device=dev, generator=gd),
             "done": (torch.rand(n, device=dev, generator=gd) < 0.01).float()} for _ in range(4)]
    for i in range(5):
        rep.add(TensorBatch(dict(rows[i % 4]), batch_size=[n]))
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    for i in range(steps):
        rep.add(TensorBatch(dict(rows[i % 4]), batch_size=[n]))
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3 / steps
    tt = torch.tensor([ms], device=dev, dtype=torch.float64)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms = float(tt.item())
    if rep._ix is not None:
        rep._ix.check()
    ck = torch.stack([mem._leaves[("obs",)].view(-1)[::4099].long().sum(), mem._leaves[("reward",)].view(torch.int32).long().sum(),
                      mem._trees[2].nodes.view(-1).view(torch.int64)[2:].sum(),
                      torch.tensor(mem._cursor * 65537 + mem.tree_ptr, device=dev)])
    allck = [torch.empty_like(ck) for _ in range(world)]
    dist.all_gather(allck, ck)
    nbytes = world * n * R_C2
    return {"workload": f"c2_replicated_insert_{n}rows_per_rank_x{world}", "metric": "inserted_rows_per_s",
            "value": world * n / (ms * 1e-3), "unit": "rows/s (into EVERY replica)", "ms_per_env_step": ms,
            "bytes_pulled_per_replica_per_step": nbytes, "peer_kernel": bool(rep._ix_ok),
            "replicas_identical": all(bool(torch.equal(allck[0], x)) for x in allck[1:]),
            "api": "population.ReplicatedReplay.add (xp_peer_ring_insert + xp_per_add), host-inclusive"}


def also_multi_gpu(dev, rank, world):
    """The configs BASELINE.json names for 1/2/4/8 GPUs, per N: C3 and C5 shard by population with no
    exchange (each rank's agents own their rollout buffers / sample their replica); C4 is strong scaling with a
    replicated tree.  Every number is the max over ranks of a device-timed graph replay."""
    import torch
    import torch.distributed as dist

    def max_over_ranks(x):
        tt = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    out = []
    pl = max(1, C3["P"] // world)
    dist.barrier()
    g = bench_gae(dev, sets=8 * world, P=pl)
    us = max_over_ranks(g["us_per_launch"])
    steps = pl * world * C3["T"] * C3["N"]
    out.append({"workload": C3["name"] + f"_sharded_x{world}", "metric": "gae_steps_per_s", "value": steps / (us * 1e-6),
                "unit": "steps/s", "scaling": "strong", "agents_per_gpu": pl, "us_per_launch": us,
                "roofline_per_gpu": roofline(pl * C3["T"] * C3["N"] * 17, us), "exchange": "none"})
    torch.cuda.empty_cache()
    out.append(bench_per_tree_replicated(dev, rank, world))
    torch.cuda.empty_cache()
    out.append(bench_replicated_insert(dev, rank, world))
    torch.cuda.empty_cache()
    pl = max(1, C5["P"] // world)
    dist.barrier()
    c5 = bench_uniform(dev, P=pl)
    us = max_over_ranks(c5["us_sample"])
    out.append({"workload": C5["name"] + f"_sharded_x{world}", "metric": "uniform_sampled_transitions_per_s",
                "value": pl * world * C5["B"] / (us * 1e-6), "unit": UNIT, "scaling": "strong", "agents_per_gpu": pl,
                "us_sample": us, "roofline_per_gpu": roofline(pl * C5["B"] * (2 * R_C5 + 8), us), "exchange": "none"})
    return out


def also_single_gpu(dev, with_cpu=True):
    """C3 / C4 / C5 (+ C1, insert side) device-timed with rooflines, each beside the reference's CPU number."""
    import torch

    def cpu(entry, fn, seconds):
        if with_cpu:
            entry["cpu_baseline"] = cpu_leg(fn, seconds)
        return entry

    g = bench_gae(dev)
    g["roofline"] = roofline(g["algorithmic_bytes"], g["us_per_launch"], kernel="gae_", capture="ncu_gae.txt")
    cpu(g, "time_c3", 4.0)
    ga = bench_gae_per_agent(dev)
    g16 = bench_gae(dev, sets=2, P=C3["P"] * 16, reps=20)
    g16["roofline"] = roofline(g16["algorithmic_bytes"], g16["us_per_launch"])
    p = bench_per_tree(dev)
    p["roofline"] = roofline(p["algorithmic_bytes"], p["us_sample_plus_update"])
    p["roofline_sample"] = roofline(C4["P"] * C4["B"] * (8 * C4["L"] + 24), p["us_sample"],
                                    kernel="per_sample_kernel", capture="ncu_sample.txt")
    p["roofline_update"] = roofline(C4["P"] * C4["B"] * per_bytes_per_sample(C4["L"])["update"], p["us_update"],
                                    kernel="set_wbucket_kernel", capture="ncu_update.txt")
    cpu(p, "time_c4", 6.0)
    torch.cuda.empty_cache()
    c5 = bench_uniform(dev)
    c5["roofline"] = roofline(c5["algorithmic_bytes"], c5["us_sample"], kernel="move_kernel", capture="ncu_gather_c5.txt")
    cpu(c5, "time_c5", 6.0)
    torch.cuda.empty_cache()
    ins = cpu(bench_insert(dev), "time_insert", 3.0)
    torch.cuda.empty_cache()
    c1 = cpu(bench_c1(dev), "time_c1", 3.0)
    return [g, ga, g16, p, c5, ins, c1]


def ncu_traffic(kernel_substr: str, capture: str = "ncu_*.txt"):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of a kernel, from the newest
    committed ``profiles/r*_ncu_*.txt`` summary that contains it (tools/ncu_summary.py output of an
    ``ncu --set full`` capture); ``capture`` narrows the file name (the same kernel is captured on
    more than one workload); None when no capture is on disk."""
    import glob
    import re

    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    root = os.path.dirname(os.path.abspath(__file__))
    for path in sorted(glob.glob(os.path.join(root, "profiles", "r*_" + capture)), reverse=True):
        cur, rd, wr = None, None, None
        for ln in open(path, errors="replace"):
            if ln.startswith("Kernel Name"):
                if cur and rd is not None and wr is not None:
                    break
                cur, rd, wr = (kernel_substr in ln), None, None
            elif cur:
                m = re.match(r"dram__bytes_(read|write)\.sum\s+([0-9.]+)\s+(\w+)", ln)
                if m:
                    v = float(m.group(2)) * unit.get(m.group(3), 1.0)
                    if m.group(1) == "read":
                        rd = v
                    else:
                        wr = v
        if cur and rd is not None and wr is not None:
            return {"bytes": rd + wr, "source": os.path.relpath(path, root)}
    return None


NOMINAL_HBM_GBS = 8000.0  # north_star's "~8 TB/s peak"; SURVEY.md 8(d) asks for both denominators


def roofline(alg_bytes: float, us: float, traffic=None, kernel: str | None = None, capture: str | None = None) -> dict:
    """achieved = algorithmic bytes / event-timed duration; frac of the MEASURED copy bandwidth
    (MEASURED_PEAKS.json) and, beside it, of the nominal 8 TB/s.  ``kernel`` / ``capture``: read `traffic`
    (dram read + write bytes of one launch) from the committed ncu summary under profiles/."""
    peak, how = hbm_peak()
    ach = alg_bytes / (us * 1e-6) / 1e9
    src = None
    if traffic is None and kernel:
        t = ncu_traffic(kernel, capture or "ncu_*.txt")
        if t:
            traffic, src = t["bytes"], t["source"]
    return {"bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
            "frac": round(ach / peak, 4), "traffic": traffic, "peak_source": how,
            "frac_of_nominal_8tbs": round(ach / NOMINAL_HBM_GBS, 4), "traffic_source": src}


def cpu_leg(fn, seconds: float):
    """A bounded CPU-baseline leg (oracle/cpu_arm.py: the unmodified reference on this host); a failure
    there (say, too little host RAM) must not take the GPU numbers down with it."""
    try:
        from oracle import cpu_arm

        return getattr(cpu_arm, fn)(seconds)
    except Exception as e:  # pragma: no cover - depends on the host
        return {"value": None, "kind": "unavailable", "sample": f"{type(e).__name__}: {e}"}


def b200_arm(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the B200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from agilerl_b200 import _native as nv

    lib = nv.load()

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    if args.workload != "c2":
        raise SystemExit("only --workload c2 is a bench line; c3/c4 are reported under 'also'")
    wl = C2Workload(args, rank, world, dev)
    wl.setup()
    wl.verify()
    K, W = args.steps, max(args.warmup, 3)
    use_graph = world > 1 and os.environ.get("XP_BENCH_NOGRAPH") != "1"
    for i in range(W):
        wl.step(i)
    sync()
    clocks = ClockSampler(local)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if use_graph:
        # N > 1: R consecutive steps captured once and replayed — every kernel of the step is host-sync-free, so
        # the ranks meet only inside xp_peer_allgather.  One untimed replay after the host barrier aligns the
        # ranks ON THE DEVICE (its exchanges are rendez-vous) before event `a` is recorded.
        if wl.px is not None and wl.px.epoch % 2:
            wl.step(W)  # even exchange count before capture
        graph = wl.capture_steps()
        sync()
        wl.replay(graph)
        clocks.start()
        l0 = lib.xp_launch_count()
        a.record()
        for _ in range(K // wl.R):
            wl.replay(graph)
        for i in range(K % wl.R):
            wl.step(i)
        b.record()
        sync()
        # replays launch kernels without passing through the library's host entry points: count them from the capture
        l1 = lib.xp_launch_count() + (K // wl.R) * wl.launches_per_graph
        timed = []
    else:
        # the gather pair is event-timed on every 4th step of the timed region: an event record between two
        # kernels ends the programmatic-launch chain there, so timing every step would tax the step itself
        EV = 4
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) if i % EV == 0 else None
               for i in range(K)]
        clocks.start()
        l0 = lib.xp_launch_count()
        a.record()
        for i in range(K):
            wl.step(W + i, evs[i])
        b.record()
        sync()
        l1 = lib.xp_launch_count()
        timed = [e for e in evs if e is not None]
    ms = a.elapsed_time(b)
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = wl.units_per_step * world * K / (ms * 1e-3)
    # replicas must still be bit-identical after the timed steps (untimed check; fails the run otherwise)
    replicas_identical = None
    if world > 1:
        ck = wl.replica_checksum()
        allck = [torch.empty_like(ck) for _ in range(world)]
        dist.all_gather(allck, ck)
        replicas_identical = all(bool(torch.equal(allck[0], x)) for x in allck[1:])

    # per-kernel breakdown of the step (separate untimed pass; events between every launch)
    Kb = min(K, 50)
    bev = [[torch.cuda.Event(enable_timing=True) for _ in range(5)] for _ in range(Kb)]
    for i in range(Kb):
        bev[i][4].record()
        wl.step(W + i, bev[i])
    sync()
    breakdown = {n: round(sum(e[x].elapsed_time(e[y]) for e in bev) * 1e3 / Kb, 2)
                 for n, x, y in (("per_sample_us", 4, 0), ("gather_per_table_us", 0, 2),
                                 ("gather_nstep_table_us", 2, 1), ("per_update_us", 1, 3))}
    if timed:  # the two gather launches (PER table, n-step table; same indices) sit between one event pair
        gather_us = sum(x.elapsed_time(y) for x, y in timed) * 1e3 / len(timed) / 2
        gather_how = f"event-timed inside the timed region on every {EV}th step"
    else:      # graph replays hold no event nodes: the same launches event-timed in the pass right after
        gather_us = (breakdown["gather_per_table_us"] + breakdown["gather_nstep_table_us"]) / 2
        gather_how = f"event-timed in the {Kb}-step per-kernel pass that follows the (graph-replayed) timed region"

    # -- e2e through the public per-agent API ---------------------------------------------------
    e2e = None
    clocks.period = 0.025  # the e2e loop is host-bound Python: the sampling thread competes for the GIL, so poll 12x less often
    if not args.no_e2e:
        wl.e2e_setup()
        Ke = max(100, min(K, 200))  # its own step count (reported as e2e.steps): a 20-step K would time 7 ms of a host-bound loop
        for i in range(10):
            wl.e2e_step(i)
        sync()
        t0 = time.perf_counter()
        a.record()
        for i in range(Ke):
            wl.e2e_step(i)
        b.record()
        sync()
        wall = time.perf_counter() - t0
        ems = max(a.elapsed_time(b), wall * 1e3 * 0.0)  # device-side span of the same region
        if world > 1:
            t = torch.tensor([ems], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ems = float(t.item())
        bi, bo = wl.e2e_bytes()
        e2e = {"value": wl.units_per_step * world * Ke / (ems * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": bi, "d2h_bytes_per_step": bo, "steps": Ke,
               "ms_per_step": ems / Ke,
               "api": "per agent: PrioritizedReplayBuffer.sample + MultiStepReplayBuffer."
                      "sample_from_indices + update_priorities (host uniforms / host priorities); buffers linked once "
                      "at set-up with agilerl_b200.link_n_step" + ("; updates through population.ReplicatedReplay "
                      "(NVLink exchange + replicated update per agent)" if world > 1 else "")}
    clocks.stop()

    also = None
    cpu = None
    rows = wl.cap
    exchange = "peer" if wl.px is not None else "nccl"
    if wl.px is not None:
        wl.px.check()
    if world > 1 and not args.no_also:  # collective: every rank takes part
        del wl
        torch.cuda.empty_cache()
        also = also_multi_gpu(dev, rank, world)
    if rank == 0 and world == 1 and not args.no_also:
        del wl
        torch.cuda.empty_cache()
        also = also_single_gpu(dev, with_cpu=not args.no_cpu)
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = reference_c2(seconds=args.cpu_seconds, fill_seconds=20.0)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64 tree / u8 rows", "data": "synthetic",
            "config": {"workload": C2["name"], "rows": rows,
                       "batch": C2["B"], "population_per_gpu": C2["P"], "population": C2["P"] * world,
                       "row_bytes": R_C2, "tree_levels": C2["L"],
                       "l2": "inputs larger than L2: random rows of a 113 GB replay storage; "
                             "2 rotating output sets of 231 MB",
                       "parallelism": f"pop-shard x{world}, replicated replay"
                                      + ("" if world == 1 else (", exchange: xp_peer_allgather over NVLink peer memory"
                                                                if exchange == "peer" else ", exchange: NCCL all-gather"))},
            "e2e": e2e,
            "gpu_launches": int(l1 - l0),
            "clocks": clocks.summary(),
            "roofline": dict(roofline(C2["P"] * C2["B"] * (2 * R_C2 + 8), gather_us,
                                      (ncu_traffic("move_kernel<0>", "ncu_gather.txt") or {}).get("bytes")),
                             kernel="move_kernel<gather>", us_per_launch=round(gather_us, 2),
                             algorithmic_bytes_per_launch=C2["P"] * C2["B"] * (2 * R_C2 + 8),
                             launches_timed=2 * len(timed),
                             traffic_source=(ncu_traffic("move_kernel<0>", "ncu_gather.txt") or {}).get("source")),
            "step_breakdown": breakdown,
            "cpu_baseline": cpu,
            "also": also,
        }
        line["roofline"]["timed_how"] = gather_how
        if world > 1:
            line["replicas_identical"] = replicas_identical
            line["timed_region"] = ("CUDA-graph replays of 4 consecutive steps (sample, gather x2, NVLink exchange, "
                                    "replicated update); ranks aligned on the device by one untimed replay"
                                    if use_graph else "host-launched steps")
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
        if replicas_identical is False:
            raise SystemExit("bench.py: the replicas diverged during the timed steps")


# =================================================================================================
# reference arm / cpu_baseline: the UNMODIFIED reference classes on the host (oracle/cpu_arm.py)
# =================================================================================================
def reference_c2(seconds: float, max_agent_steps: int = 1 << 30, fill_seconds: float = 30.0) -> dict:
    """One process, agents sequential on one shared memory — the reference's own execution model
    (train_off_policy.py:46,249).  kind == "reference" wherever a copy of the reference is present
    (/root/reference here, the staged oracle/_ref on the GPU box), else the oracle port."""
    from oracle import cpu_arm

    c = C2
    return cpu_arm.time_c2(seconds, max_agent_steps, cap=c["cap"], B=c["B"], n_step=c["n_step"], gamma=c["gamma"],
                           alpha=c["alpha"], beta=c["beta"], num_envs=c["num_envs"], fill_seconds=fill_seconds)


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.workload != "c2":
        raise SystemExit("only --workload c2 is a bench line")
    from oracle import cpu_arm

    K, W = args.steps, args.warmup
    c = C2
    t0 = time.perf_counter()
    # a "step" = one population step: P agents x (sample + n-step sample_from_indices + update_priorities);
    # K + W of them, bounded to ~60 s of timed work so that the run ends within a few minutes whatever K is
    cpu = reference_c2(seconds=60.0, max_agent_steps=max((K + W) * c["P"], c["P"]))
    upper = None
    if not args.no_also:
        workers = max(1, min(c["P"], os.cpu_count() or 1))
        upper = cpu_arm.time_c2_population_parallel(12.0, workers, cap=c["cap"], B=c["B"])
    wall = time.perf_counter() - t0
    line = {"impl": "reference", "metric": METRIC, "value": cpu["value"], "unit": UNIT,
            "n_gpus": args.gpus, "steps": K, "warmup": W,
            "ms_per_step": 1e3 * c["B"] * c["P"] / cpu["value"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64 tree / u8 rows", "data": "synthetic",
            "config": {"workload": c["name"], "rows": c["cap"], "batch": c["B"],
                       "population_per_gpu": c["P"], "population": c["P"] * max(args.gpus, 1), "row_bytes": R_C2,
                       "tree_levels": c["L"],
                       "parallelism": "1 host process, agents sequential on one shared memory (the reference's model): its "
                                      "rate per sampled transition does not depend on the population size, so the same "
                                      "measurement stands beside every --gpus N"},
            "cpu_baseline": cpu,
            "population_parallel_upper_bound": upper,
            "e2e": {"value": cpu["value"], "unit": UNIT, "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": round(wall, 1), "agent_steps": cpu["agent_steps"]}
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    # stdout carries exactly ONE JSON line: anything native libraries print there (NCCL's version
    # banner, for one) is sent to stderr instead
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w", buffering=1)
    if args.impl == "reference":
        reference_arm(args)
    else:
        b200_arm(args)


if __name__ == "__main__":
    main()
