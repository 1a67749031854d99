"""Graph-timed xp_per_update at the batch sizes the replicated C2 exchange produces (cap 2^20)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, bench
from agilerl_b200 import xp
dev = torch.device("cuda", 0)
cap = 1 << 20
nodes = torch.empty(2 * cap, 2, dtype=torch.float64, device=dev)
aux = torch.zeros(2 * cap, dtype=torch.int32, device=dev)
t = xp.make_tree(nodes[:, 0], nodes[:, 1], 2, cap, aux)
xp.tree_init(t, torch.cuda.current_stream().cuda_stream)
state = torch.ones(1, dtype=torch.float64, device=dev)
xp.per_add(t, state, 0, cap, cap, 0.6)
err = torch.zeros(1, dtype=torch.int32, device=dev)
for B in (256, 2048, 4096, 8192, 16384, 32768):
    sets = [(torch.randint(0, cap, (B,), device=dev), torch.rand(B, device=dev) + 0.01) for _ in range(8)]
    fns = [(lambda i=i, p=p: xp.per_update(t, state, i, p, cap, 0.6, err)) for i, p in sets]
    print(B, "items:", round(bench.graph_time_us(fns, 10), 2), "us")
