"""A few eager C3 GAE launches (for ncu captures of whichever kernel XP_GAE_KERNEL selects)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from agilerl_b200 import xp
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(4)
T, N, PP = 128, 1024, int(sys.argv[1]) if len(sys.argv) > 1 else 16
sets = []
for _ in range(4):
    r = torch.randn(PP, T, N, device=dev, generator=g); v = torch.randn(PP, T, N, device=dev, generator=g)
    d = torch.rand(PP, T, N, device=dev, generator=g) < 0.02
    lv = torch.randn(PP, N, device=dev, generator=g).double(); ld = torch.zeros(PP, N, device=dev, dtype=torch.float64)
    sets.append((r, v, d, lv, ld, torch.empty_like(r), torch.empty_like(r)))
for r, v, d, lv, ld, a, gg in sets:
    xp.gae(r, v, d, lv, ld, a, gg, T, N, PP, T * N, 0.99, 0.95)
torch.cuda.synchronize()
print("ok")
