"""Probe torch symmetric memory on this box: allocate, rendezvous, peer pointers, barrier timing."""
import os, time
import torch, torch.distributed as dist

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
import torch.distributed._symmetric_memory as symm

t = symm.empty(4096, dtype=torch.int64, device=dev)
hdl = symm.rendezvous(t, dist.group.WORLD.group_name)
if rank == 0:
    print("handle attrs:", [a for a in dir(hdl) if not a.startswith("_")])
    print("buffer_ptrs", [hex(p) for p in hdl.buffer_ptrs], "world", hdl.world_size, "rank", hdl.rank)
t.fill_(rank + 1)
torch.cuda.synchronize(); dist.barrier()
hdl.barrier()
peer = hdl.get_buffer((rank + 1) % world, (4096,), torch.int64)
print(rank, "peer value", int(peer[0]))
torch.cuda.synchronize()
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for _ in range(20): hdl.barrier()
torch.cuda.synchronize(); dist.barrier()
ev0.record()
for _ in range(200): hdl.barrier()
ev1.record(); torch.cuda.synchronize()
if rank == 0: print("symm barrier us", ev0.elapsed_time(ev1) * 1e3 / 200)
a = torch.zeros(2048, dtype=torch.int64, device=dev); b = torch.zeros(world * 2048, dtype=torch.int64, device=dev)
for _ in range(20): dist.all_gather_into_tensor(b, a)
torch.cuda.synchronize(); dist.barrier()
ev0.record()
for _ in range(200): dist.all_gather_into_tensor(b, a)
ev1.record(); torch.cuda.synchronize()
if rank == 0: print("nccl all_gather 16KB us", ev0.elapsed_time(ev1) * 1e3 / 200)
dist.destroy_process_group()
