"""C5 gather (32 768 random rows of 1 504 B out of 10 M) against torch.index_select of the same rows and a
contiguous copy of the same byte count: what a random 1.5 KB-row gather can reach on this HBM."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, bench
from agilerl_b200 import xp
dev = torch.device("cuda", 0)
cap, n, W = 10_000_000, 32768, 376
store = torch.empty(cap, W, dtype=torch.float32, device=dev); store.view(-1)[::1024].fill_(1.0)
R = 4
idx = [torch.randint(0, cap, (n,), device=dev) for _ in range(R)]
outs = [torch.empty(n, W, dtype=torch.float32, device=dev) for _ in range(R)]
err = torch.zeros(1, dtype=torch.int32, device=dev)
us_xp = bench.graph_time_us([(lambda i=i: xp.gather([(store, outs[i])], idx[i], cap, err)) for i in range(R)], 20)
us_t = bench.graph_time_us([(lambda i=i: torch.index_select(store, 0, idx[i], out=outs[i])) for i in range(R)], 20)
src = [torch.empty(n, W, dtype=torch.float32, device=dev) for _ in range(40)]  # 40 x 49 MB > L2
us_c = bench.graph_time_us([(lambda i=i: outs[i % R].copy_(src[i])) for i in range(40)], 5)
b = 2 * n * W * 4
print(f"obs key only ({b/1e6:.0f} MB moved): xp_gather {us_xp:.1f} us = {b/us_xp/1e3:.0f} GB/s | torch.index_select {us_t:.1f} us = "
      f"{b/us_t/1e3:.0f} GB/s | contiguous copy_ {us_c:.1f} us = {b/us_c/1e3:.0f} GB/s")
# sorted indices: the same rows in ascending order (DRAM page locality upper bound for a gather)
sidx = [i.sort().values for i in idx]
us_s = bench.graph_time_us([(lambda i=i: xp.gather([(store, outs[i])], sidx[i], cap, err)) for i in range(R)], 20)
print(f"xp_gather with sorted indices {us_s:.1f} us = {b/us_s/1e3:.0f} GB/s")
