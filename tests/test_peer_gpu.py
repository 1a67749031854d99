"""xp_peer_allgather over NVLink peer memory: 2 / 4 / 8 ranks (one per GPU, as many as the box has) exchange their blocks with
the library's own kernel; the result must equal a rank-major all-gather for every epoch, including
block reuse (alternating blocks) and a replicated PER update applied identically on both ranks.
Skipped where the box has fewer GPUs than ranks."""

import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORLDS = [2, 4, 8]


def _spawn(worker, world, tmp_path, *extra):
    """One process per GPU (skips when the box has fewer than ``world`` GPUs)."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(worker, args=(world, port, str(tmp_path), *extra), nprocs=world, join=True)


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from agilerl_b200 import peer
    from agilerl_b200.components import PrioritizedReplayBuffer, TensorBatch

    assert peer.available()
    n = 300
    px = peer.PeerExchange(12 * n, [(0, 8 * n, torch.int64), (8 * n, 4 * n, torch.float32)], dev)
    mem = PrioritizedReplayBuffer(4096, alpha=0.6, device=dev)
    mem.add(TensorBatch({"obs": torch.randn(2048, 3), "reward": torch.zeros(2048)}, batch_size=[2048]))
    ok = True
    for step in range(7):
        k = px.current
        g = torch.Generator().manual_seed(100 * step + rank)
        idx = torch.randint(0, 2048, (n,), generator=g)
        pr = torch.rand(n, generator=g) + 0.01
        px.view(k, 0, n, torch.int64).copy_(idx)
        px.view(k, 8 * n, n, torch.float32).copy_(pr)
        all_idx, all_pr = px.allgather()
        exp_i, exp_p = [], []
        for r in range(world):
            gr = torch.Generator().manual_seed(100 * step + r)
            exp_i.append(torch.randint(0, 2048, (n,), generator=gr))
            exp_p.append(torch.rand(n, generator=gr) + 0.01)
        ok &= torch.equal(all_idx.cpu(), torch.cat(exp_i)) and torch.equal(all_pr.cpu(), torch.cat(exp_p))
        mem.update_priorities(all_idx, all_pr)
    px.check()
    mem.check_errors()
    torch.save({"ok": ok, "tree": mem._trees[2].nodes.cpu(), "maxp": mem.max_priority},
               os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", WORLDS)
def test_peer_allgather(tmp_path, world):
    _spawn(_worker, world, tmp_path)
    res = [torch.load(tmp_path / f"r{r}.pt", weights_only=False) for r in range(world)]
    assert all(r["ok"] for r in res)
    for r in res[1:]:  # replicas stay bit-identical
        assert torch.equal(res[0]["tree"], r["tree"]) and res[0]["maxp"] == r["maxp"]


def _worker_rr(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from agilerl_b200.components import PrioritizedReplayBuffer, TensorBatch
    from agilerl_b200.population import ReplicatedReplay

    mem = PrioritizedReplayBuffer(1024, alpha=0.6, device=dev)
    rep = ReplicatedReplay(mem)
    # shadow: ONE plain buffer that receives every rank's rows and (index, priority) pairs in rank-major order
    # through the ordinary single-process calls — what the replicas must equal (its row contents do not matter)
    exp = PrioritizedReplayBuffer(1024, alpha=0.6, device=dev)

    def everyone(t):
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return torch.cat(out)

    g = torch.Generator().manual_seed(7 + rank)
    for _ in range(3):
        rep.add(TensorBatch({"obs": torch.randn(8, 3, generator=g), "reward": torch.randn(8, generator=g)}, batch_size=[8]))
        exp.add(TensorBatch({"obs": torch.zeros(8 * world, 3), "reward": torch.zeros(8 * world)}, batch_size=[8 * world]))
    for step in range(5):
        idx = torch.randint(0, 24 * world, (16,), generator=g)
        pr = torch.rand(16, generator=g) + 0.05
        rep.update_priorities(idx, pr)
        exp.update_priorities(everyone(idx.to(dev)), everyone(pr.to(dev)).cpu().numpy())
    # the learn-step pattern: device indices straight from sample(), priorities as an fp32 host array
    # (dqn_rainbow.py:487) -> the one-call native path (FastPer.update_replicated; with B <= 1024 the exchange kernel
    # writes this rank's block itself, xp_peer_publish_allgather_graph)
    native = []
    for step in range(6):
        batch = rep.sample(16, 0.4)
        pr = (torch.rand(16, generator=g) + 0.05).numpy()
        keep = pr.copy()
        native.append(rep._native_update(batch["idxs"], pr) if step % 2 else None)
        if not step % 2:
            rep.update_priorities(batch["idxs"], pr)
        pr[:] = -7.0  # the host array is the caller's again once the call has returned
        exp.update_priorities(everyone(batch["idxs"].reshape(-1)), everyone(torch.from_numpy(keep).to(dev)).cpu().numpy())
    # a batch above XP_HOST_VEC_MAX: the native path stages it through the pinned ring and copy calls
    big_idx = torch.randint(0, 24 * world, (1500,), generator=g).to(dev)
    big_pr = (torch.rand(1500, generator=g) + 0.05).numpy()
    native.append(rep._native_update(big_idx, big_pr))
    exp.update_priorities(everyone(big_idx), everyone(torch.from_numpy(big_pr).to(dev)).cpu().numpy())
    mem.check_errors()
    rep._px.check()
    from agilerl_b200 import _hostaccel

    same = torch.equal(mem._trees[2].nodes, exp._trees[2].nodes) and mem.max_priority == exp.max_priority
    torch.save({"peer": rep._peer_ok, "tree": mem._trees[2].nodes.cpu(), "obs": mem._leaves[("obs",)][:24 * world].cpu(),
                "native": [x for x in native if x is not None], "accel": _hostaccel.load() is not None,
                "equals_one_buffer": bool(same)},
               os.path.join(out_dir, f"q{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", WORLDS)
def test_replicated_replay_uses_peer_exchange(tmp_path, world):
    _spawn(_worker_rr, world, tmp_path)
    q = [torch.load(tmp_path / f"q{r}.pt", weights_only=False) for r in range(world)]
    assert all(x["peer"] is True for x in q)
    assert all(all(x["native"]) for x in q if x["accel"]), "the one-call native replicated update was not taken"
    assert all(x["equals_one_buffer"] for x in q), "replicas differ from ONE buffer fed in rank-major order"
    for x in q[1:]:
        assert torch.equal(q[0]["tree"], x["tree"]) and torch.equal(q[0]["obs"], x["obs"])


def _rows(step, rank, n):
    g = torch.Generator().manual_seed(1000 * step + rank)
    return {"obs": torch.randint(0, 256, (n, 7), generator=g, dtype=torch.uint8),   # 7-byte rows: byte path
            "vec": torch.randn(n, 3, generator=g),                                  # 12-byte rows: 4-byte path
            "wide": torch.randn(n, 8, generator=g),                                 # 32-byte rows: 128-bit path
            "reward": torch.randn(n, generator=g)}


def _worker_insert(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from agilerl_b200.components import PrioritizedReplayBuffer, ReplayBuffer, TensorBatch
    from agilerl_b200.population import ReplicatedReplay

    ok, used = True, []
    for cls, cap in ((PrioritizedReplayBuffer, 20 * world), (ReplayBuffer, 26 * world)):
        n = 8
        mem, exp = cls(cap, device=dev), cls(cap, device=dev)
        rep = ReplicatedReplay(mem)
        for step in range(6):  # 6 x 8*world rows into 20*world / 26*world slots: wraps inside a rank's rows and between ranks
            rep.add(TensorBatch(_rows(step, rank, n), batch_size=[n]))
            cat = {k: torch.cat([_rows(step, r, n)[k] for r in range(world)]) for k in _rows(0, 0, 1)}
            exp.add(TensorBatch(cat, batch_size=[world * n]))
            for path in exp._leaves:
                ok &= torch.equal(mem._leaves[path], exp._leaves[path])
            ok &= (mem._cursor, mem._size, mem.counter) == (exp._cursor, exp._size, exp.counter)
        if cls is PrioritizedReplayBuffer:
            ok &= torch.equal(mem._trees[2].nodes, exp._trees[2].nodes) and mem.tree_ptr == exp.tree_ptr
        rep._ix.check()
        used.append(rep._ix_ok)
    torch.save({"ok": bool(ok), "used": used}, os.path.join(out_dir, f"i{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", WORLDS)
def test_replicated_add_is_one_peer_kernel(tmp_path, world):
    """ReplicatedReplay.add through xp_peer_ring_insert == ReplayBuffer.add of the rank-major
    concatenation (replay_buffer.py:100-107), wrap included, for every row width."""
    _spawn(_worker_insert, world, tmp_path)
    for r in range(world):
        res = torch.load(tmp_path / f"i{r}.pt", weights_only=False)
        assert res["used"] == [True, True], "peer insert path was not taken"
        assert res["ok"]


# ---- the exchange inside a CUDA graph (device-resident epoch) --------------------------------------------
def _worker_graph(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from agilerl_b200 import peer, xp
    from agilerl_b200.components import PrioritizedReplayBuffer, TensorBatch

    n, steps = 256, 9
    px = peer.PeerExchange(12 * n, [(0, 8 * n, torch.int64), (8 * n, 4 * n, torch.float32)], dev)
    mem = PrioritizedReplayBuffer(4096, alpha=0.6, device=dev)
    exp = PrioritizedReplayBuffer(4096, alpha=0.6, device=dev)
    for b in (mem, exp):
        b.add(TensorBatch({"obs": torch.zeros(4096, 2), "reward": torch.zeros(4096)}, batch_size=[4096]))
    _, _, st, h = mem._ensure_trees()

    def inputs(step, r):
        g = torch.Generator().manual_seed(31 * step + r)
        return torch.randint(0, 4096, (n,), generator=g), torch.rand(n, generator=g) + 0.01

    # the producers of a step write into a fixed staging pair; the graph copies it into the block the
    # device epoch selects... the blocks alternate per exchange, so the graph holds TWO exchanges (one per
    # block) and every replay advances the epoch by two.
    stage_i = [torch.empty(n, dtype=torch.int64, device=dev) for _ in range(2)]
    stage_p = [torch.empty(n, dtype=torch.float32, device=dev) for _ in range(2)]
    side = torch.cuda.Stream(dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for k in (0, 1):
                px.view(k, 0, n, torch.int64).copy_(stage_i[k])
                px.view(k, 8 * n, n, torch.float32).copy_(stage_p[k])
                all_idx, all_pr = px.allgather()
                xp.per_update(h, mem._state, all_idx, all_pr, mem.max_size, mem.alpha, st.err)
    torch.cuda.current_stream(dev).wait_stream(side)
    px.epoch = 0  # nothing ran during capture; the device epoch is still 0
    dist.barrier()
    for rep in range(steps):
        for k in (0, 1):
            i, p = inputs(2 * rep + k, rank)
            stage_i[k].copy_(i)
            stage_p[k].copy_(p)
        g.replay()
        px.advance(2)
        for k in (0, 1):  # the expected replica: every rank's block in rank-major order, through the plain API
            cat_i = torch.cat([inputs(2 * rep + k, r)[0] for r in range(world)])
            cat_p = torch.cat([inputs(2 * rep + k, r)[1] for r in range(world)])
            exp.update_priorities(cat_i.to(dev), cat_p.to(dev))
    torch.cuda.synchronize(dev)
    px.check()
    ok = torch.equal(mem._trees[2].nodes, exp._trees[2].nodes) and mem.max_priority == exp.max_priority
    ok &= int(px._epoch_dev[0]) == 2 * steps and int(px._epoch_dev[1]) == 0
    # an eager exchange after the replays continues the same epoch sequence
    k = px.current
    i, p = inputs(999, rank)
    px.view(k, 0, n, torch.int64).copy_(i)
    px.view(k, 8 * n, n, torch.float32).copy_(p)
    all_idx, _ = px.allgather()
    ok &= torch.equal(all_idx.cpu(), torch.cat([inputs(999, r)[0] for r in range(world)]))
    torch.save({"ok": bool(ok), "tree": mem._trees[2].nodes.cpu()}, os.path.join(out_dir, f"g{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", WORLDS)
def test_peer_exchange_inside_a_cuda_graph(tmp_path, world):
    """xp_peer_allgather_graph: exchange + replicated update captured ONCE and replayed — no host work
    between the rendez-vous of consecutive steps; replicas bit-identical to the eager rank-major update."""
    _spawn(_worker_graph, world, tmp_path)
    res = [torch.load(tmp_path / f"g{r}.pt", weights_only=False) for r in range(world)]
    assert all(r["ok"] for r in res)
    for r in res[1:]:
        assert torch.equal(res[0]["tree"], r["tree"])


# ---- a lagging rank ----------------------------------------------------------------------------------------
def _worker_lag(rank, world, port, out_dir, timeout_s):
    import time

    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from agilerl_b200 import _native as nv
    from agilerl_b200 import peer

    n = 64
    px = peer.PeerExchange(8 * n, [(0, 8 * n, torch.int64)], dev)
    if timeout_s:
        peer.PeerExchange.set_timeout(timeout_s)
    dist.barrier()
    raised, ok = False, True
    try:
        for step in range(3):
            if step == 1 and rank == world - 1:
                time.sleep(2.0)  # eval / checkpoint / GC on one rank
            px.view(px.current, 0, n, torch.int64).fill_(100 * step + rank)
            (out,) = px.allgather()
            torch.cuda.synchronize(dev)
            want = torch.arange(world).repeat_interleave(n) + 100 * step
            ok &= torch.equal(out.cpu(), want)
        px.check()
    except nv.NativeError:
        raised = True
    torch.save({"ok": bool(ok), "raised": raised}, os.path.join(out_dir, f"l{rank}_{int(timeout_s * 10)}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_lagging_rank_waits_by_default_and_is_reported_under_a_short_timeout(tmp_path):
    """a rank that lags 2 s must not silently leave stale data in the other replicas.  With the
    default wait (~30 s) the exchange simply completes; with a 0.5 s time-out the fast rank's kernel gives
    up, sets the flag, and the failure is RAISED (next call or check()) instead of being swallowed."""
    world = 2
    _spawn(_worker_lag, world, tmp_path, 0.0)
    for r in range(world):
        res = torch.load(tmp_path / f"l{r}_0.pt", weights_only=False)
        assert res["ok"] and not res["raised"]
    _spawn(_worker_lag, world, tmp_path, 0.5)
    res0 = torch.load(tmp_path / "l0_5.pt", weights_only=False)
    assert res0["raised"], "the fast rank must report the peer that did not arrive in time"
