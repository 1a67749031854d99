"""Graph-timed C5 uniform sample (same code path as bench.py's `also` entry)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, bench
r = bench.bench_uniform(torch.device("cuda", 0))
print({k: round(v, 2) for k, v in r.items() if k.startswith("us_")}, "gather frac", r["gather_roofline"]["frac"])
