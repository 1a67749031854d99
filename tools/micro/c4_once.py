"""A few eager C4 sample + update steps, a bulk ring insert and C3 GAE launches (for ncu launch lists / captures)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from agilerl_b200 import xp
dev = torch.device("cuda", 0)
cap, B, P = 1 << 24, 4096, 64
nodes = torch.empty(2 * cap, 2, dtype=torch.float64, device=dev)
aux = torch.zeros(2 * cap, dtype=torch.int32, device=dev)
t = xp.make_tree(nodes[:, 0], nodes[:, 1], 2, cap, aux)
xp.tree_init(t, torch.cuda.current_stream().cuda_stream)
state = torch.ones(1, dtype=torch.float64, device=dev)
xp.per_add(t, state, 0, cap, cap, 0.6)
g = torch.Generator(device=dev).manual_seed(4)
xp.per_update(t, state, torch.arange(cap, device=dev), torch.randn(cap, device=dev, generator=g).abs() + 1e-3, cap, 0.6)
n = P * B
err = torch.zeros(1, dtype=torch.int32, device=dev)
idx = torch.empty(n, dtype=torch.int64, device=dev)
w = torch.empty(n, dtype=torch.float32, device=dev)
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4):
    u = torch.rand(n, device=dev, generator=g)
    pr = torch.randn(n, device=dev, generator=g).abs() + 1e-6
    xp.per_sample(t, u, B, P, cap, 0.4, idx, w)
    xp.per_update(t, state, idx, pr, cap, 0.6, err)
torch.cuda.synchronize()
assert int(err.item()) == 0
# bulk ring insert: 16 384 Atari rows (offline ingestion, train_offline.py:190)
nb = 16384
ring = torch.empty(2 * nb, 84, 84, 4, dtype=torch.uint8, device=dev)
rows = torch.randint(0, 256, (nb, 84, 84, 4), dtype=torch.uint8, device=dev, generator=g)
for i in range(3):
    xp.ring_insert([(rows, ring)], nb, (i * 5000) % (2 * nb), 2 * nb)
# C3 GAE, 16 agents in one launch, rotating sets
T, N, PP = 128, 1024, 16
sets = []
for _ in range(6):
    r = torch.randn(PP, T, N, device=dev, generator=g); v = torch.randn(PP, T, N, device=dev, generator=g)
    d = torch.rand(PP, T, N, device=dev, generator=g) < 0.02
    lv = torch.randn(PP, N, device=dev, generator=g).double(); ld = torch.zeros(PP, N, device=dev, dtype=torch.float64)
    sets.append((r, v, d, lv, ld, torch.empty_like(r), torch.empty_like(r)))
for r, v, d, lv, ld, a, gg in sets:
    xp.gae(r, v, d, lv, ld, a, gg, T, N, PP, T * N, 0.99, 0.95)
torch.cuda.synchronize()
print("ok")
