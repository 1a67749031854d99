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
dist.destroy_process_group()
