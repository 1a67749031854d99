"""Host-side profile of the per-step insert path through the classes (MultiStepReplayBuffer.add -> PER.add)."""
import cProfile, io, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from agilerl_b200.components import MultiStepReplayBuffer, PrioritizedReplayBuffer, TensorBatch
dev = torch.device("cuda", 0)
cap, E = 65536, 16
per = PrioritizedReplayBuffer(cap, alpha=0.6, device=dev)
nst = MultiStepReplayBuffer(cap, n_step=3, gamma=0.99, device=dev)
g = torch.Generator(device=dev).manual_seed(0)
def rows(n):
    return {"obs": torch.randint(0, 256, (n, 84, 84, 4), dtype=torch.uint8, device=dev, generator=g),
            "action": torch.randint(0, 18, (n,), device=dev, generator=g).float(), "reward": torch.randn(n, device=dev, generator=g),
            "next_obs": torch.randint(0, 256, (n, 84, 84, 4), dtype=torch.uint8, device=dev, generator=g),
            "done": (torch.rand(n, device=dev, generator=g) < 0.01).float()}
trs = [rows(E) for _ in range(8)]
def step(i):
    one = nst.add(TensorBatch(dict(trs[i % 8]), batch_size=[E]))
    if one is not None:
        per.add(one)
for i in range(20): step(i)
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(400): step(i)
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f"host {1e6 * (t1 - t0) / 400:.1f} us per agent-step (16 rows); with final sync {1e6 * (t2 - t0) / 400:.1f}")
ta = tb = 0.0
for i in range(400):
    t = time.perf_counter(); one = nst.add(TensorBatch(dict(trs[i % 8]), batch_size=[E])); ta += time.perf_counter() - t
    t = time.perf_counter(); per.add(one); tb += time.perf_counter() - t
print(f"nstep.add {1e6 * ta / 400:.1f} us, per.add {1e6 * tb / 400:.1f} us")
pr = cProfile.Profile(); pr.enable()
for i in range(400): step(i)
pr.disable(); torch.cuda.synchronize()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(22); print(s.getvalue()[:5000])
