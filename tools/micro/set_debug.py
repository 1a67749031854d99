import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from agilerl_b200 import xp
dev = torch.device("cuda", 0)
cap, B = int(sys.argv[1]), int(sys.argv[2])
nodes = torch.empty(2 * cap, 2, dtype=torch.float64, device=dev)
aux = torch.ones(2 * cap, dtype=torch.int32, device=dev)
t = xp.make_tree(nodes[:, 0], nodes[:, 1], 2, cap, aux)
xp.tree_init(t, torch.cuda.current_stream().cuda_stream)
err = torch.zeros(1, dtype=torch.int32, device=dev)
rng = np.random.default_rng(0)
for rnd in range(3):
    idx = torch.as_tensor(rng.integers(0, cap, B)).to(dev)
    val = torch.as_tensor(rng.random(B) + 0.1).to(dev)
    xp.tree_set_batch(t, idx, val, err)
    torch.cuda.synchronize()
    leaves = nodes[cap:, 0]
    print(rnd, "root", float(nodes[1, 0]), "leafsum", float(leaves.sum()), "aux nonzero tickets", int(aux[cap:].abs().sum()))
