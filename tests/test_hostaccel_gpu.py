"""The native host fast path (csrc/hostaccel.cpp) and the pure-Python host must be
indistinguishable: same draws from the torch CPU generator, same indices / weights / rows, same
trees after updates, same exceptions.  (The component parity tests run with whichever host path
is active; this file pins the two against each other.)"""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def C(cuda_device):
    import agilerl_b200.components as comps

    return comps


@pytest.fixture()
def accel():
    from agilerl_b200 import _hostaccel

    mod = _hostaccel.load()
    if mod is None:
        pytest.skip("host fast path not built")

    class Toggle:
        def off(self):
            _hostaccel._MOD = None

        def on(self):
            _hostaccel._MOD = mod

    t = Toggle()
    yield t
    t.on()


def _fill(C, cls, n=600, cap=1024, **kw):
    g = torch.Generator().manual_seed(5)
    b = cls(max_size=cap, device="cuda", **kw)
    for _ in range(n // 8):
        b.add(C.TensorBatch({
            "obs": torch.randint(0, 256, (8, 6, 6, 2), dtype=torch.uint8, generator=g),
            "action": torch.randint(0, 4, (8,), generator=g).float(),
            "reward": torch.randn(8, generator=g),
            "next_obs": torch.randint(0, 256, (8, 6, 6, 2), dtype=torch.uint8, generator=g),
            "done": (torch.rand(8, generator=g) < 0.1).float()}, batch_size=[8]))
    return b


def _same(a, b):
    assert list(a.keys()) == list(b.keys())
    for k in a.keys():
        assert a[k].dtype == b[k].dtype and a[k].shape == b[k].shape, k
        assert torch.equal(a[k], b[k]), k


def test_fast_path_is_active(C, accel):
    b = _fill(C, C.PrioritizedReplayBuffer, alpha=0.6)
    assert b._fast_table() is not None and b._fast_per() is not None


def test_per_sample_update_identical(C, accel):
    accel.off()
    slow = _fill(C, C.PrioritizedReplayBuffer, alpha=0.6)
    assert slow._fast_table() is None
    accel.on()
    fast = _fill(C, C.PrioritizedReplayBuffer, alpha=0.6)
    rng = np.random.default_rng(3)
    for it in range(6):
        prio = (np.abs(rng.standard_normal(64)) + 1e-6).astype(np.float32)
        torch.manual_seed(100 + it)
        accel.off()
        s = slow.sample(64, 0.4)
        slow.update_priorities(s["idxs"], prio)
        torch.manual_seed(100 + it)
        accel.on()
        f = fast.sample(64, 0.4)
        fast.update_priorities(f["idxs"], prio)
        _same(s, f)
        assert f["idxs"].shape == (64, 1) and f["weights"].shape == (64, 1)
    assert torch.equal(slow._trees[2].nodes, fast._trees[2].nodes)
    assert slow.max_priority == fast.max_priority
    # device priorities, (B,) indices
    pr = torch.rand(64, device="cuda") + 0.1
    idx = torch.randint(0, 600, (64,), device="cuda")
    accel.off()
    slow.update_priorities(idx, pr)
    accel.on()
    fast.update_priorities(idx, pr)
    assert torch.equal(slow._trees[2].nodes, fast._trees[2].nodes)


def test_gather_identical_and_fresh(C, accel):
    accel.on()
    b = _fill(C, C.MultiStepReplayBuffer, n_step=1)
    idx = torch.randint(0, len(b), (32, 1), device="cuda")
    f = b.sample_from_indices(idx)
    accel.off()
    b._ftab = None
    s = b.sample_from_indices(idx)
    accel.on()
    _same(s, f)
    assert f["obs"].shape == (32, 1, 6, 6, 2) or f["obs"].shape == (32, 6, 6, 2)
    g = b.sample_from_indices(idx)
    assert g["obs"].data_ptr() != f["obs"].data_ptr()  # fresh copies every call


def test_deferred_index_error_through_fast_path(C, accel):
    accel.on()
    b = _fill(C, C.PrioritizedReplayBuffer, alpha=0.6)
    bad = torch.tensor([5, 4096], device="cuda")
    b.update_priorities(bad, np.array([1.0, 2.0], dtype=np.float32))
    with pytest.raises(AssertionError):
        b.check_errors()


def test_link_n_step_serves_sample_from_indices_from_one_gather(C, cuda_device, accel):
    """agilerl_b200.link_n_step(memory, n_step_memory): PER.sample gathers both tables in one launch and the
    reference's next call, n_step_memory.sample_from_indices(batch["idxs"]), is served from it — the same
    rows as without the link; any other index tensor, a modified one, or an add in between fall back."""
    import agilerl_b200
    from agilerl_b200 import _native as nv

    lib = nv.load()

    def pair():
        per = C.PrioritizedReplayBuffer(512, alpha=0.6, device=cuda_device)
        nst = C.MultiStepReplayBuffer(512, n_step=3, gamma=0.9, device=cuda_device)
        g = torch.Generator().manual_seed(3)
        for _ in range(40):
            n = 8
            tr = C.TensorBatch({"obs": torch.randint(0, 255, (n, 2, 5), dtype=torch.uint8, generator=g),
                                "action": torch.randn(n, generator=g), "reward": torch.randn(n, generator=g),
                                "next_obs": torch.randint(0, 255, (n, 2, 5), dtype=torch.uint8, generator=g),
                                "done": (torch.rand(n, generator=g) < 0.1).float()}, batch_size=[n])
            one = nst.add(tr)
            if one is not None:
                per.add(one)
        return per, nst, tr

    per_a, nst_a, _ = pair()
    per_b, nst_b, tr = pair()
    agilerl_b200.link_n_step(per_b, nst_b)
    for rnd in range(3):
        torch.manual_seed(50 + rnd)
        ea = per_a.sample(32, 0.4)
        na = nst_a.sample_from_indices(ea["idxs"])
        torch.manual_seed(50 + rnd)
        c0 = lib.xp_launch_count()
        eb = per_b.sample(32, 0.4)
        c1 = lib.xp_launch_count()
        nb = nst_b.sample_from_indices(eb["idxs"])
        c2 = lib.xp_launch_count()
        assert c1 - c0 == 2 and c2 == c1          # descent + ONE gather for both tables; nothing for the n-step call
        assert torch.equal(ea["idxs"], eb["idxs"]) and torch.equal(ea["weights"], eb["weights"])
        assert nb.batch_size == na.batch_size == torch.Size([32, 1])
        for k in ("obs", "action", "reward", "next_obs", "done"):
            assert torch.equal(ea[k], eb[k]) and torch.equal(na[k], nb[k]), k
            assert nb[k].shape == na[k].shape
        nb["reward"].zero_()                       # fresh copies: mutating them leaves the storage alone
        assert float(nst_b.storage["reward"].abs().sum()) > 0
    # fall-backs: another index tensor / a modified one / an add since the prefetch
    eb = per_b.sample(16, 0.4)
    other = eb["idxs"].clone()
    c0 = lib.xp_launch_count()
    r1 = nst_b.sample_from_indices(other)
    assert lib.xp_launch_count() == c0 + 1 and torch.equal(r1["obs"], nst_b.storage["obs"][other.view(-1)].unsqueeze(1))
    eb = per_b.sample(16, 0.4)
    eb["idxs"].clamp_(max=3)
    r2 = nst_b.sample_from_indices(eb["idxs"])
    assert torch.equal(r2["reward"], nst_b.storage["reward"][eb["idxs"].view(-1)].unsqueeze(1))
    eb = per_b.sample(16, 0.4)
    one = nst_b.add(tr)
    per_b.add(one)
    r3 = nst_b.sample_from_indices(eb["idxs"])
    assert torch.equal(r3["reward"], nst_b.storage["reward"][eb["idxs"].view(-1)].unsqueeze(1))
    per_b.link_n_step(None)
    c0 = lib.xp_launch_count()
    per_b.sample(8, 0.4)
    assert lib.xp_launch_count() == c0 + 2
