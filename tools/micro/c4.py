import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, bench
dev = torch.device("cuda", 0)
r = bench.bench_per_tree(dev)
print({k: (round(v, 2) if isinstance(v, float) else v) for k, v in r.items() if k.startswith("us_") or k == "value"})
if len(sys.argv) > 1:
    g = bench.bench_gae(dev)
    print("gae us", round(g["us_per_launch"], 2))
