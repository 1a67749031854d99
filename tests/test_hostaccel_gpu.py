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
