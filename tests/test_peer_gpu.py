"""xp_peer_allgather over NVLink peer memory: two ranks (one per GPU) exchange their blocks with
the library's own kernel; the result must equal a rank-major all-gather for every epoch, including
block reuse (alternating blocks) and a replicated PER update applied identically on both ranks.
Needs two GPUs: skipped on a single-GPU box."""

import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


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


def test_peer_allgather_two_gpus(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r0 = torch.load(tmp_path / "r0.pt", weights_only=False)
    r1 = torch.load(tmp_path / "r1.pt", weights_only=False)
    assert r0["ok"] and r1["ok"]
    assert torch.equal(r0["tree"], r1["tree"]) and r0["maxp"] == r1["maxp"]  # replicas stay bit-identical


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
    g = torch.Generator().manual_seed(7 + rank)
    for _ in range(3):
        rep.add(TensorBatch({"obs": torch.randn(8, 3, generator=g), "reward": torch.randn(8, generator=g)}, batch_size=[8]))
    for step in range(5):
        idx = torch.randint(0, 48, (16,), generator=g)
        rep.update_priorities(idx, torch.rand(16, generator=g) + 0.05)
    mem.check_errors()
    torch.save({"peer": rep._peer_ok, "tree": mem._trees[2].nodes.cpu(), "obs": mem._leaves[("obs",)][:48].cpu()},
               os.path.join(out_dir, f"q{rank}.pt"))
    dist.destroy_process_group()


def test_replicated_replay_uses_peer_exchange(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker_rr, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    q0 = torch.load(tmp_path / "q0.pt", weights_only=False)
    q1 = torch.load(tmp_path / "q1.pt", weights_only=False)
    assert q0["peer"] is True and q1["peer"] is True
    assert torch.equal(q0["tree"], q1["tree"]) and torch.equal(q0["obs"], q1["obs"])


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
    for cls, cap in ((PrioritizedReplayBuffer, 40), (ReplayBuffer, 52)):
        n = 8
        mem, exp = cls(cap, device=dev), cls(cap, device=dev)
        rep = ReplicatedReplay(mem)
        for step in range(6):  # 6 x 16 rows into 40 / 52 slots: wraps inside a rank's rows and between ranks
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


def test_replicated_add_is_one_peer_kernel(tmp_path):
    """ReplicatedReplay.add through xp_peer_ring_insert == ReplayBuffer.add of the rank-major
    concatenation (replay_buffer.py:100-107), wrap included, for every row width."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker_insert, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        res = torch.load(tmp_path / f"i{r}.pt", weights_only=False)
        assert res["used"] == [True, True], "peer insert path was not taken"
        assert res["ok"]
