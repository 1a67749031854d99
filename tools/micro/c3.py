"""Graph-timed C3 GAE launch (same code path as bench.py's `also` entry)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, bench
dev = torch.device("cuda", 0)
for _ in range(2):
    g = bench.bench_gae(dev)
    print("C3 gae us/launch", round(g["us_per_launch"], 2), "->", round(g["algorithmic_bytes"] / g["us_per_launch"] / 1e3), "GB/s")
g = bench.bench_gae(dev, sets=2, P=256, reps=20)
print("x16 gae us/launch", round(g["us_per_launch"], 2), "->", round(g["algorithmic_bytes"] / g["us_per_launch"] / 1e3), "GB/s")
