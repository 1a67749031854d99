"""Schedule S vs L (XP_NO_S=1) on skewed large batches: a share of the 262 144 updates falls into the most
recently written leaves (fresh transitions carry the maximum priority and are drawn together)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from agilerl_b200 import xp
dev = torch.device("cuda", 0)
cap, n = 1 << 24, 262144
nodes = torch.empty(2 * cap, 2, dtype=torch.float64, device=dev)
aux = torch.zeros(2 * cap, dtype=torch.int32, device=dev)
t = xp.make_tree(nodes[:, 0], nodes[:, 1], 2, cap, aux)
xp.tree_init(t, torch.cuda.current_stream().cuda_stream)
state = torch.ones(1, dtype=torch.float64, device=dev)
xp.per_add(t, state, 0, cap, cap, 0.6)
err = torch.zeros(1, dtype=torch.int32, device=dev)
g = torch.Generator(device=dev).manual_seed(0)
for share, hot in ((0.0, 0), (0.1, 4096), (0.3, 4096), (0.3, 65536), (0.9, 1 << 20)):
    k = int(n * share)
    idx = torch.randint(0, cap, (n,), device=dev, generator=g)
    if k:
        idx[:k] = cap - hot + torch.randint(0, hot, (k,), device=dev, generator=g)
    idx = idx[torch.randperm(n, device=dev, generator=g)]
    pr = torch.rand(n, device=dev, generator=g) + 0.01
    for _ in range(2):
        xp.per_update(t, state, idx, pr, cap, 0.6, err)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        xp.per_update(t, state, idx, pr, cap, 0.6, err)
    b.record(); torch.cuda.synchronize()
    print(f"share {share:.1f} in the last {hot:>7} leaves: {a.elapsed_time(b) * 1e3 / 5:9.1f} us per update")
assert int(err.item()) == 0
