"""e2e leg of bench.py (per-agent API loop, full C2 buffer), 7 blocks of 200 steps: min / median ms per step.
Run twice, with and without XP_NO_HOST_VEC=1, to A/B the in-launch host vectors."""
import os, statistics, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.argv = ["bench.py"] + sys.argv[1:]
import bench, torch
args = bench.parse_args()
dev = torch.device("cuda", 0)
wl = bench.C2Workload(args, 0, 1, dev)
wl.setup(); wl.e2e_setup()
for i in range(20): wl.e2e_step(i)
torch.cuda.synchronize()
res = []
for blk in range(7):
    t0 = time.perf_counter()
    for i in range(200): wl.e2e_step(i)
    torch.cuda.synchronize()
    res.append((time.perf_counter() - t0) / 200 * 1e3)
print("host_vec" if os.environ.get("XP_NO_HOST_VEC", "0") != "1" else "ring    ",
      f"e2e ms/step min {min(res):.4f} median {statistics.median(res):.4f}  all {[round(r, 4) for r in res]}")
