"""Why does PrioritizedReplayBuffer.sample cost more right after a synchronize?  Variants, host time per call."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.argv = ["bench.py", "--rows", "65536"]
import bench, torch
args = bench.parse_args()
dev = torch.device("cuda", 0)
wl = bench.C2Workload(args, 0, 1, dev)
wl.setup(); wl.e2e_setup()
for i in range(5): wl.e2e_step(i)
torch.cuda.synchronize()
fp = wl.per._fast_per()

def timeit(name, fn, sync=True, n=200):
    tot = 0.0
    for _ in range(n):
        if sync: torch.cuda.synchronize()
        t = time.perf_counter(); r = fn(); tot += time.perf_counter() - t
        del r
    print(f"{name:46s} {tot / n * 1e6:8.1f} us")

timeit("per.sample (linked), sync before", lambda: wl.per.sample(wl.B, 0.4))
timeit("per.sample (linked), no sync", lambda: wl.per.sample(wl.B, 0.4), sync=False)
timeit("FastPer.sample native, sync before", lambda: fp.sample(wl.B, 0.4, wl.per._size, None))
wl.per.link_n_step(None)
timeit("per.sample (unlinked), sync before", lambda: wl.per.sample(wl.B, 0.4))
timeit("per.sample (unlinked), no sync", lambda: wl.per.sample(wl.B, 0.4), sync=False)
u = torch.rand(wl.B, device=dev)
timeit("per.sample(u=device uniforms), sync before", lambda: wl.per.sample(wl.B, 0.4, u=u))
import cProfile, pstats, io
pr = cProfile.Profile()
for _ in range(100):
    torch.cuda.synchronize(); pr.enable(); r = wl.per.sample(wl.B, 0.4); pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(8); print(s.getvalue()[:1800])
