"""The per-rank C3 launches of bench.py --gpus 8 / 4 / 2 (sets = 8 * world, P = 16 / world) on ONE GPU: graph sizes and us per launch."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, bench
dev = torch.device("cuda", 0)
for sets, P in ((64, 2), (32, 4), (16, 8)):
    g = bench.bench_gae(dev, sets=sets, P=P)
    print("sets", sets, "P", P, "us/launch", round(g["us_per_launch"], 2), "launches per graph", g["launches_per_graph"])
