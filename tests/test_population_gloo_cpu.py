"""World-size-2 gloo test of the multi-GPU exchange logic (agilerl_b200/population.py): the
canonical rank-major ordering that keeps replicated replay memories identical."""

import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class FakeMemory:
    """Records what a replica would insert / update (stands in for the CUDA buffer on CPU)."""

    storage_device = torch.device("cpu")

    def __init__(self):
        self.rows, self.updates = [], []

    def add(self, batch):
        self.rows.append({k: v.clone() for k, v in batch.items()})

    def update_priorities(self, idx, prio):
        self.updates.append((idx.clone(), prio.clone()))


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from agilerl_b200.components import TensorBatch
    from agilerl_b200.population import ReplicatedReplay, ShardPlan, all_gather_rows

    plan = ShardPlan(population=4, world_size=world, rank=rank)
    mem = FakeMemory()
    rep = ReplicatedReplay(mem)
    n_envs = 3
    for step in range(2):
        rows = []
        for a in plan.local_agents:
            rows.append(torch.full((n_envs, 2), 100.0 * step + 10.0 * a) + torch.arange(n_envs).view(-1, 1))
        local = TensorBatch({"obs": torch.cat(rows), "done": torch.cat(rows)[:, 0] > 1e9}, batch_size=[len(rows) * n_envs])
        rep.add(local)
    idx = torch.tensor([10 * a + i for a in plan.local_agents for i in range(2)])
    rep.update_priorities(idx, idx.float() / 7)
    g = all_gather_rows(torch.tensor([[rank]]))
    torch.save({"rows": mem.rows, "updates": mem.updates, "ranks": g}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


def test_replicas_receive_identical_rows_in_canonical_order(tmp_path):
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r0 = torch.load(tmp_path / "r0.pt", weights_only=False)
    r1 = torch.load(tmp_path / "r1.pt", weights_only=False)
    assert r0["ranks"].view(-1).tolist() == [0, 1]
    for a, b in zip(r0["rows"], r1["rows"]):
        assert torch.equal(a["obs"], b["obs"]) and a["done"].dtype == torch.bool and torch.equal(a["done"], b["done"])
    # rank-major, agent-major: agents 0,1 (rank 0) then 2,3 (rank 1), each with its 3 env rows
    first = r0["rows"][0]["obs"][:, 0].tolist()
    assert first == [0.0, 1.0, 2.0, 10.0, 11.0, 12.0, 20.0, 21.0, 22.0, 30.0, 31.0, 32.0]
    (i0, p0), (i1, p1) = r0["updates"][0], r1["updates"][0]
    assert torch.equal(i0, i1) and torch.equal(p0, p1)
    assert i0.tolist() == [0, 1, 10, 11, 20, 21, 30, 31]
