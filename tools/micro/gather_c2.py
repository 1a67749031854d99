"""C2-shaped gather micro-benchmark: 5 keys (2 x 28 224 B uint8 obs + 3 fp32 scalars), 2 048 rows
per launch from a table far larger than L2.  Graph-timed (launch-to-launch, rotating index sets
and output buffers), next to a plain device copy of the same byte count (the achievable floor for
a kernel of this size)."""
import sys

import torch

sys.path.insert(0, ".")
from agilerl_b200 import xp  # noqa: E402


def graph_us(fns, reps):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for f in fns:
            f()
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(reps):
                for f in fns:
                    f()
        g.replay()
        s.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e9
        for _ in range(5):
            e0.record(s)
            g.replay()
            e1.record(s)
            e1.synchronize()
            best = min(best, e0.elapsed_time(e1) * 1e3 / (reps * len(fns)))
    return best


def main():
    dev = torch.device("cuda:0")
    rows = int(sys.argv[1]) if len(sys.argv) > 1 else 300_000
    B = 2048
    obs = torch.empty(rows, 84, 84, 4, dtype=torch.uint8, device=dev)
    nobs = torch.empty_like(obs)
    obs.random_(0, 256)
    nobs.random_(0, 256)
    small = [torch.randn(rows, 1, device=dev) for _ in range(3)]
    srcs = [obs, nobs] + small
    sets = 6
    outs = [[torch.empty((B, *s.shape[1:]), dtype=s.dtype, device=dev) for s in srcs] for _ in range(sets)]
    idxs = [torch.randint(0, rows, (B,), device=dev) for _ in range(sets)]

    def mk(i):
        pairs = list(zip(srcs, outs[i]))
        return lambda: xp.gather(pairs, idxs[i], rows)

    us = graph_us([mk(i) for i in range(sets)], 10)
    nbytes = B * (2 * sum(s[0].numel() * s.element_size() for s in srcs) + 8)
    print(f"gather 5 keys B={B}: {us:.2f} us/launch  {nbytes / us / 1e3:.0f} GB/s")
    # check
    for s, o in zip(srcs, outs[0]):
        assert torch.equal(o, s[idxs[0]])

    # floor: plain copy of the same bytes (contiguous), rotating buffers
    n = nbytes // 2
    a = [torch.empty(n, dtype=torch.uint8, device=dev).random_(0, 256) for _ in range(sets)]
    b = [torch.empty(n, dtype=torch.uint8, device=dev) for _ in range(sets)]
    us_c = graph_us([(lambda i=i: b[i].copy_(a[i])) for i in range(sets)], 10)
    print(f"copy_ same bytes: {us_c:.2f} us/launch  {2 * n / us_c / 1e3:.0f} GB/s")
    # obs key alone
    def mk1(i):
        pairs = [(obs, outs[i][0])]
        return lambda: xp.gather(pairs, idxs[i], rows)
    us1 = graph_us([mk1(i) for i in range(sets)], 10)
    print(f"gather 1 key (28224 B rows): {us1:.2f} us/launch  {B * 2 * 28224 / us1 / 1e3:.0f} GB/s")


if __name__ == "__main__":
    main()
