"""Time xp_peer_allgather vs coalesced NCCL all-gathers for the C2 exchange (16 KB + 8 KB per rank)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
from agilerl_b200 import peer
n = 2048
px = peer.PeerExchange(12 * n, [(0, 8 * n, torch.int64), (8 * n, 4 * n, torch.float32)], dev)
a = torch.zeros(n, dtype=torch.int64, device=dev); b = torch.zeros(n, dtype=torch.float32, device=dev)
A = torch.zeros(world * n, dtype=torch.int64, device=dev); B = torch.zeros(world * n, dtype=torch.float32, device=dev)
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
def timeit(fn, iters=300):
    for _ in range(20): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    ev0.record()
    for _ in range(iters): fn()
    ev1.record(); torch.cuda.synchronize()
    return ev0.elapsed_time(ev1) * 1e3 / iters
def nccl():
    with dist._coalescing_manager(device=dev):
        dist.all_gather_into_tensor(A, a); dist.all_gather_into_tensor(B, b)
t_peer = timeit(px.allgather); t_nccl = timeit(nccl)
px.check()
if rank == 0: print(f"world {world}: xp_peer_allgather {t_peer:.2f} us   NCCL (2 coalesced all-gathers) {t_nccl:.2f} us")
# ---- insert path: C2 rows (16 envs x 4 agents per rank, 56 460 B per row) ------------------
from agilerl_b200.components import ReplayBuffer, TensorBatch
from agilerl_b200.population import ReplicatedReplay, all_gather_batch
rows = 64
batch = TensorBatch({"obs": torch.zeros(rows, 84, 84, 4, dtype=torch.uint8, device=dev),
                     "next_obs": torch.zeros(rows, 84, 84, 4, dtype=torch.uint8, device=dev),
                     "action": torch.zeros(rows, device=dev), "reward": torch.zeros(rows, device=dev),
                     "done": torch.zeros(rows, device=dev)}, batch_size=[rows])
m1, m2 = ReplayBuffer(65536, device=dev), ReplayBuffer(65536, device=dev)
rp = ReplicatedReplay(m1)
t_pi = timeit(lambda: rp.add(batch), 100)
assert rp._ix_ok, "peer insert not used"
rp._ix.check()
t_ni = timeit(lambda: m2.add(all_gather_batch(batch)), 100)
if rank == 0: print(f"world {world}: replicated add of {rows} C2 rows/rank ({rows * 56460 / 1e6:.1f} MB): peer ring insert "
                    f"{t_pi:.1f} us   NCCL all-gather per key + xp_ring_insert {t_ni:.1f} us (host-inclusive, stream-timed)")
dist.destroy_process_group()
