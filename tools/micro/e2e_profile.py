"""Host-side profile of the per-agent API loop (bench.py e2e leg) on a small buffer."""
import cProfile, pstats, sys, os, time, io
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.argv = ["bench.py", "--rows", "65536"]
import bench, torch
args = bench.parse_args()
dev = torch.device("cuda", 0)
wl = bench.C2Workload(args, 0, 1, dev)
wl.setup(); wl.e2e_setup()
for i in range(5): wl.e2e_step(i)
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(50): wl.e2e_step(i)
torch.cuda.synchronize()
print("e2e ms/step", (time.perf_counter() - t0) / 50 * 1e3)
# per-call host time (no sync inside)
import collections
acc = collections.Counter()
for i in range(50):
    for a in range(wl.P):
        t = time.perf_counter(); b = wl.per.sample(wl.B, 0.4); acc["sample"] += time.perf_counter() - t
        t = time.perf_counter(); nb = wl.nstep.sample_from_indices(b["idxs"]); acc["nstep"] += time.perf_counter() - t
        t = time.perf_counter(); wl.per.update_priorities(b["idxs"], wl.prio_host[a]); acc["update"] += time.perf_counter() - t
    t = time.perf_counter(); torch.cuda.synchronize(); acc["sync"] += time.perf_counter() - t
for k, v in acc.items(): print(f"{k:8s} {v / 50 / (wl.P if k != 'sync' else 1) * 1e6:8.1f} us per call")
# the native calls alone (no Python wrappers)
fp = wl.per._fast_per()
idx = b["idxs"]
import numpy as np
prt = torch.from_numpy(wl.prio_host[0])
for name, fn in (("FastPer.sample", lambda: fp.sample(wl.B, 0.4, wl.per._size, None)), ("FastPer.update", lambda: fp.update(idx, prt))):
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(400): fn()
    dt = time.perf_counter() - t; torch.cuda.synchronize()
    print(f"{name:16s} {dt / 400 * 1e6:8.1f} us per call (native only)")
# pure host enqueue time: the GPU queue is drained before every call, only the call is timed
for name, fn in (("per.sample", lambda: wl.per.sample(wl.B, 0.4)), ("FastPer.sample", lambda: fp.sample(wl.B, 0.4, wl.per._size, None)),
                 ("FastPer.update", lambda: fp.update(idx, prt)), ("per.update_priorities", lambda: wl.per.update_priorities(idx, wl.prio_host[0]))):
    tot = 0.0
    for _ in range(200):
        torch.cuda.synchronize(); t = time.perf_counter(); r = fn(); tot += time.perf_counter() - t
    print(f"{name:22s} {tot / 200 * 1e6:8.1f} us host enqueue (queue drained)")
# GPU time of one agent's triple
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
torch.cuda.synchronize(); ev[0].record()
for _ in range(100):
    b = wl.per.sample(wl.B, 0.4); nb = wl.nstep.sample_from_indices(b["idxs"]); wl.per.update_priorities(b["idxs"], wl.prio_host[0])
ev[1].record(); torch.cuda.synchronize()
print(f"device span per agent triple (100 back to back): {ev[0].elapsed_time(ev[1]) * 10:.1f} us")
pr = cProfile.Profile(); pr.enable()
for i in range(50): wl.e2e_step(i)
torch.cuda.synchronize(); pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(28); print(s.getvalue()[:6000])
