import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from agilerl_b200 import xp
from oracle import xp_oracle as xo
dev = torch.device("cuda", 0)
for (P, T, N) in [(1, 128, 32), (1, 32, 32), (2, 64, 64), (1, 50, 32), (1, 50, 96), (3, 50, 96), (2, 257, 128)]:
    for use_gae in (True, False):
        rng = np.random.default_rng(P * 1000 + T)
        r = rng.standard_normal((P, T, N)).astype(np.float32)
        v = rng.standard_normal((P, T, N)).astype(np.float32)
        d = rng.random((P, T, N)) < 0.1
        lv = rng.standard_normal((P, N)); ld = (rng.random((P, N)) < 0.2).astype(np.float64)
        R, V, D = [torch.as_tensor(a).to(dev) for a in (r, v, d)]
        LV, LD = torch.as_tensor(lv).to(dev), torch.as_tensor(ld).to(dev)
        A = torch.full_like(R, float("nan")); G = torch.full_like(R, float("nan"))
        try:
            xp.gae(R, V, D, LV, LD, A, G, T, N, P, T * N, 0.99, 0.95, use_gae)
            torch.cuda.synchronize()
        except Exception as e:
            print((P, T, N), use_gae, "ERROR", repr(e)[:300]); sys.exit(1)
        ok = True
        for p in range(P):
            fn = xo.gae if use_gae else xo.mc_returns
            args = (r[p], v[p], d[p], lv[p], ld[p], 0.99) + ((0.95,) if use_gae else ())
            a, g = fn(*args)
            ea = np.array_equal(A[p].cpu().numpy(), a); eg = np.array_equal(G[p].cpu().numpy(), g)
            if not (ea and eg):
                ok = False
                bad = np.argwhere(A[p].cpu().numpy() != a)
                print((P, T, N), use_gae, "agent", p, "mismatch adv", len(bad), "first", bad[:3].tolist(), "nan", int(np.isnan(A[p].cpu().numpy()).sum()))
        print((P, T, N), use_gae, "ok" if ok else "MISMATCH")
