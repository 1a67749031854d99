"""Graph-timed per-agent (P = 1) GAE launches — what the drop-in issues (bench.py also[1])."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, bench
r = bench.bench_gae_per_agent(torch.device("cuda", 0))
print({k: round(v, 2) for k, v in r.items() if k.startswith("us_")})
