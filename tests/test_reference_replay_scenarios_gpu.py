"""The reference's own MultiStepReplayBuffer / PrioritizedReplayBuffer test scenarios
(/root/reference/tests/test_components/test_replay_buffer.py:432-1056), restated against
agilerl_b200.components with the same calls, in the same order, with the same assertions — the
buffers are constructed exactly as the reference tests construct them (default ``device="cpu"``:
batches come back on the CPU while the storage lives on the GPU).  ``TensorDict`` below is the
engine's ``TensorBatch``; every transition is built with ``unsqueeze(0)`` + ``batch_size = [1]`` as
there.  A last test walks every public method of every mirrored class once."""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def C(cuda_device):
    import agilerl_b200.components as comps

    return comps


def td1(C, d):
    data = C.TensorBatch(d)
    data = data.unsqueeze(0)
    data.batch_size = [1]
    return data


# ---- MultiStepReplayBuffer -------------------------------------------------------------------------
def nstep_buffer(C, **kw):
    b = C.MultiStepReplayBuffer(**kw)
    b.reward_key = "reward"
    b.done_key = "done"
    b.ns_key = "next_state"
    return b


def test_nstep_buffer_initialization(C):  # :433-444
    b = C.MultiStepReplayBuffer(1000, 3, 0.99)
    assert b.max_size == 1000 and b.n_step == 3 and b.gamma == 0.99 and len(b.n_step_buffer) == 0


def test_nstep_buffer_add(C):  # :448-492
    b = nstep_buffer(C, max_size=1000, n_step=3, gamma=0.99)
    b.add(td1(C, {"state": torch.tensor([1, 2, 3]), "action": torch.tensor([0]), "reward": torch.tensor([1.0]),
                  "next_state": torch.tensor([4, 5, 6]), "done": torch.tensor([False])}))
    assert len(b.n_step_buffer) == 1 and len(b) == 0
    for i in range(2):
        b.add(td1(C, {"state": torch.tensor([i + 4, i + 5, i + 6]), "action": torch.tensor([i + 1]),
                      "reward": torch.tensor([float(i + 2)]), "next_state": torch.tensor([i + 7, i + 8, i + 9]),
                      "done": torch.tensor([False])}))
    assert len(b) == 1 and len(b.n_step_buffer) == 3


def test_nstep_buffer_add_with_images_and_sample(C):  # :494-528 — calls the INHERITED sample()
    b = nstep_buffer(C, max_size=1000, n_step=3, gamma=0.99)
    for i in range(4):
        b.add(td1(C, {"state": torch.ones((3, 84, 84)) * i, "action": torch.tensor([i % 3]),
                      "reward": torch.tensor([float(i)]), "next_state": torch.ones((3, 84, 84)) * (i + 1),
                      "done": torch.tensor([False])}))
    assert len(b) >= 1
    samples = b.sample(1)
    assert samples["state"].shape[1:] == (3, 84, 84)
    assert samples["state"].device.type == "cpu"  # device="cpu" buffers hand out CPU batches


def test_nstep_reward_calculation_from_the_deque(C):  # :532-569 — fills n_step_buffer directly
    b = nstep_buffer(C, max_size=1000, n_step=3, gamma=0.9)
    b._storage = C.TensorBatch({}, batch_size=[0, 0])
    for i in range(3):
        b.n_step_buffer.append(td1(C, {"state": torch.tensor([i, i + 1, i + 2]), "action": torch.tensor([i]),
                                       "reward": torch.tensor([float(i + 1)]),
                                       "next_state": torch.tensor([i + 3, i + 4, i + 5]),
                                       "done": torch.tensor([False])}))
    n_step_data = b._get_n_step_info()
    expected_reward = 1.0 + 0.9 * 2.0 + 0.9 * 0.9 * 3.0
    assert torch.isclose(n_step_data["reward"], torch.tensor(expected_reward))
    # bit-level: fp32 product then fp32 sum, forward order (replay_buffer.py:244)
    want = np.float32(1.0) + np.float32(2.0) * np.float32(0.9) + np.float32(3.0) * np.float32(0.9**2)
    assert np.float32(n_step_data["reward"].item()) == want
    assert torch.all(n_step_data["next_state"] == b.n_step_buffer[2]["next_state"])
    assert torch.all(n_step_data["done"] == b.n_step_buffer[2]["done"])
    assert torch.all(n_step_data["state"] == b.n_step_buffer[0]["state"])


def test_nstep_early_termination_from_the_deque(C):  # :571-628
    b = nstep_buffer(C, max_size=1000, n_step=3, gamma=0.9)
    data = [td1(C, {"state": torch.tensor(s), "action": torch.tensor([a]), "reward": torch.tensor([r]),
                    "next_state": torch.tensor(ns), "done": torch.tensor([d])})
            for s, a, r, ns, d in (([1, 2, 3], 0, 1.0, [4, 5, 6], False), ([4, 5, 6], 1, 2.0, [7, 8, 9], True),
                                   ([7, 8, 9], 2, 3.0, [10, 11, 12], False))]
    for d in data:
        b.n_step_buffer.append(d)
    n_step_data = b._get_n_step_info()
    assert torch.isclose(n_step_data["reward"], torch.tensor(1.0 + 0.9 * 2.0))
    assert torch.all(n_step_data["next_state"] == data[1]["next_state"])
    assert torch.all(n_step_data["done"] == data[1]["done"])


@pytest.mark.parametrize("key", ["termination", "terminated"])
def test_multistep_done_key_detection(C, key):  # :630-666
    b = C.MultiStepReplayBuffer(max_size=100, n_step=3, gamma=0.99)
    b.reward_key = "reward"
    b.ns_key = "next_obs"
    for i in range(3):
        b.add(td1(C, {"state": torch.tensor([i]), "action": torch.tensor([i]),
                      "reward": torch.tensor([float(i + 1)]), "next_obs": torch.tensor([i + 1]),
                      key: torch.tensor([False])}))
    assert b.done_key == key


def test_nstep_sampling_with_indices(C):  # :670-704 — MultiStepReplayBuffer.sample(n, return_idx=True)
    b = nstep_buffer(C, max_size=1000, n_step=3, gamma=0.99)
    for i in range(10):
        b.add(td1(C, {"state": torch.tensor([i, i + 1, i + 2]), "action": torch.tensor([i % 3]),
                      "reward": torch.tensor([float(i)]), "next_state": torch.tensor([i + 3, i + 4, i + 5]),
                      "done": torch.tensor([i % 5 == 0])}))
    assert len(b) == 8
    samples = b.sample(min(3, len(b)), return_idx=True)
    assert "idxs" in samples
    assert len(samples["idxs"]) == min(3, len(b))
    assert all(0 <= idx < len(b) for idx in samples["idxs"])
    # and it IS torch.randperm(size)[:B] on the same seed, rows equal to the storage rows
    torch.manual_seed(11)
    want = torch.randperm(len(b))[:5]
    torch.manual_seed(11)
    s = b.sample(5, return_idx=True)
    assert torch.equal(s["idxs"].cpu(), want)
    assert torch.equal(s["state"].cpu(), b.storage["state"].cpu()[want])


def test_multistep_buffer_sample_from_indices(C):  # :708-730
    b = nstep_buffer(C, max_size=100, n_step=3, gamma=0.99)
    for i in range(6):
        b.add(td1(C, {"state": torch.tensor([i]), "action": torch.tensor([i % 2]),
                      "reward": torch.tensor([float(i)]), "next_state": torch.tensor([i + 1]),
                      "done": torch.tensor([i == 5])}))
    assert len(b) > 0
    samples = b.sample_from_indices(torch.tensor([0]))
    assert "state" in samples and samples.batch_size[0] == 1


# ---- PrioritizedReplayBuffer -----------------------------------------------------------------------
def per_row(C, i):
    return td1(C, {"state": torch.tensor([i, i + 1, i + 2]), "action": torch.tensor([i % 3]),
                   "reward": torch.tensor([float(i)])})


def test_prioritized_buffer_initialization(C):  # :735-747
    b = C.PrioritizedReplayBuffer(1000, 0.6)
    assert b.max_size == 1000 and b.alpha == 0.6 and b.max_priority == 1.0 and b.tree_ptr == 0
    assert isinstance(b.sum_tree, C.SumSegmentTree) and isinstance(b.min_tree, C.MinSegmentTree)


def test_prioritized_buffer_add_and_add_batch(C):  # :751-790
    b = C.PrioritizedReplayBuffer(max_size=100, alpha=0.6)
    b.add(td1(C, {"state": torch.tensor([1, 2, 3]), "action": torch.tensor([0]), "reward": torch.tensor([1.0])}))
    assert len(b) == 1 and b.tree_ptr == 1
    assert b.sum_tree[0] == b.max_priority**b.alpha and b.min_tree[0] == b.max_priority**b.alpha
    b = C.PrioritizedReplayBuffer(max_size=100, alpha=0.6)
    batch = C.TensorBatch({"state": torch.tensor([[1, 2, 3], [4, 5, 6]]), "action": torch.tensor([[0], [1]]),
                           "reward": torch.tensor([[1.0], [2.0]])})
    batch.batch_size = [2]
    b.add(batch)
    assert len(b) == 2 and b.tree_ptr == 2
    assert b.sum_tree[0] == b.max_priority**b.alpha and b.sum_tree[1] == b.max_priority**b.alpha


def test_prioritized_buffer_with_image_observations(C):  # :794-824
    b = C.PrioritizedReplayBuffer(max_size=100, alpha=0.6)
    for i in range(5):
        b.add(td1(C, {"state": torch.ones((3, 84, 84)) * i, "action": torch.tensor([i % 3]),
                      "reward": torch.tensor([float(i)])}))
    samples = b.sample(3, 0.4)
    assert isinstance(samples, C.TensorBatch) and "state" in samples
    assert samples["state"].shape == (3, 3, 84, 84) and "weights" in samples and "idxs" in samples


def test_prioritized_buffer_with_dict_observations(C):  # :826-864
    b = C.PrioritizedReplayBuffer(max_size=100, alpha=0.6)
    for i in range(5):
        obs = C.TensorBatch({"image": torch.ones((3, 84, 84)) * i, "vector": torch.tensor([i, i + 1, i + 2])})
        b.add(td1(C, {"state": obs, "action": torch.tensor([i % 3]), "reward": torch.tensor([float(i)])}))
    samples = b.sample(3, 0.4)
    assert isinstance(samples, C.TensorBatch) and isinstance(samples["state"], C.TensorBatch)
    assert samples["state"]["image"].shape == (3, 3, 84, 84) and samples["state"]["vector"].shape == (3, 3)
    assert "weights" in samples and "idxs" in samples


def test_prioritized_sampling(C):  # :866-897
    b = C.PrioritizedReplayBuffer(max_size=100, alpha=0.6)
    for i in range(10):
        b.add(per_row(C, i))
    samples = b.sample(5, 0.4)
    assert isinstance(samples, C.TensorBatch)
    for k in ("state", "action", "reward", "weights", "idxs"):
        assert k in samples
    assert samples.batch_size == torch.Size([5])
    assert samples["weights"].shape == (5, 1) and samples["idxs"].shape == (5, 1)


def test_prioritized_buffer_update_priorities(C):  # :901-936
    b = C.PrioritizedReplayBuffer(max_size=100, alpha=0.6)
    for i in range(5):
        b.add(per_row(C, i))
    indices = np.arange(3)
    new_priorities = torch.tensor([2.0, 3.0, 4.0])
    b.update_priorities(indices, new_priorities)
    for idx, priority in zip(indices, new_priorities, strict=False):
        assert b.sum_tree[idx.item()] == priority.item() ** b.alpha
        assert b.min_tree[idx.item()] == priority.item() ** b.alpha
    assert b.max_priority == 4.0


def test_prioritized_buffer_update_priorities_small_priority_clamped(C):  # :938-952
    b = C.PrioritizedReplayBuffer(max_size=10, alpha=0.6)
    for i in range(3):
        b.add(td1(C, {"state": torch.tensor([i]), "action": torch.tensor([0]), "reward": torch.tensor([1.0])}))
    b.update_priorities(torch.tensor([0, 1]), torch.tensor([1e-10, 1e-10]))
    assert b.sum_tree[0] >= 1e-5**b.alpha
    assert b.sum_tree[0] == 1e-5**b.alpha  # max(p, 1e-5) ** alpha exactly (:425,322)


def test_proportional_sampling_with_mocked_torch_rand(C):  # :956-1000 — replaces torch.rand itself
    b = C.PrioritizedReplayBuffer(max_size=100, alpha=0.6)
    for i in range(5):
        b.add(per_row(C, i))
        b._update_priority(i, float(i + 1))
    original_rand = torch.rand
    try:
        predetermined_values = [0.1, 0.3, 0.6, 0.8]
        counter = 0

        def mock_rand(size):
            nonlocal counter
            val = predetermined_values[counter % len(predetermined_values)]
            counter += 1
            return torch.tensor([val])

        torch.rand = mock_rand
        indices = b._sample_proportional(4)
        assert counter == 4  # one torch.rand(1) per stratum, as the reference's loop makes them
        assert all(0 <= idx < len(b) for idx in indices)
        assert indices.device.type == "cpu" and indices.dtype == torch.int64
    finally:
        torch.rand = original_rand
    # known answer from the reference's arithmetic on those draws (segment_tree.py:136-156)
    leaves = [float(i + 1) ** 0.6 for i in range(5)]
    total = 0.0
    # the tree's own pairwise order: ((l0+l1)+(l2+l3)) + ((l4+0)+(0+0)) ... capacity 128
    level = leaves + [0.0] * (128 - 5)
    tree = {128 + i: v for i, v in enumerate(level)}
    for node in range(127, 0, -1):
        tree[node] = tree[2 * node] + tree[2 * node + 1]
    total = tree[1]
    seg = total / 4
    want = []
    for i, u in enumerate((0.1, 0.3, 0.6, 0.8)):
        a_, b_ = seg * i, seg * (i + 1)
        ub = float(np.float32(u)) * (b_ - a_) + a_
        idx = 1
        while idx < 128:
            if tree[2 * idx] > ub:
                idx = 2 * idx
            else:
                ub -= tree[2 * idx]
                idx = 2 * idx + 1
        want.append(idx - 128)
    assert indices.tolist() == want


def test_weight_calculation(C):  # :1004-1039
    b = C.PrioritizedReplayBuffer(max_size=100, alpha=0.6)
    for i in range(5):
        b.add(per_row(C, i))
    priorities = [0.5, 1.0, 1.5, 2.0, 2.5]
    for i, priority in enumerate(priorities):
        b._update_priority(i, priority)
    beta = 0.4
    for i, priority in enumerate(priorities):
        weight = b._calculate_weights(torch.tensor([i]), beta)[0]
        p_sample = (priority**b.alpha) / b.sum_tree.sum()
        p_min = b.min_tree.min() / b.sum_tree.sum()
        expected_weight = ((p_sample * len(b)) ** -beta) / ((p_min * len(b)) ** -beta)
        assert torch.isclose(weight, torch.tensor(expected_weight), rtol=1e-5)  # the reference's own assertion
        assert weight.item() == np.float32(expected_weight)                     # and bit-exact


def test_prioritized_replay_buffer_update_priority_assert(C):  # :1043-1056
    b = C.PrioritizedReplayBuffer(max_size=10, alpha=0.6)
    b.add(td1(C, {"state": torch.tensor([1]), "action": torch.tensor([0]), "reward": torch.tensor([1.0])}))
    with pytest.raises(AssertionError):
        b._update_priority(100, 1.0)


# ---- uniform draws stay aligned over a RUN, not just per call ------------------------------------------
def test_two_uniform_samples_follow_the_reference_stream(C):
    """replay_buffer.py:125 consumes size-1 generator outputs per call; the second sample() of a seeded
    run matches the reference's second torch.randperm(size)[:B]."""
    b = C.ReplayBuffer(max_size=4096)
    b.add(C.TensorBatch({"x": torch.arange(3000, dtype=torch.float32).view(-1, 1)}, batch_size=[3000]))
    torch.manual_seed(21)
    want = [torch.randperm(3000)[:32] for _ in range(3)]
    tail = torch.rand(3)
    torch.manual_seed(21)
    got = [b.sample(32, return_idx=True)["idxs"].cpu() for _ in range(3)]
    for w, g in zip(want, got):
        assert torch.equal(w, g)
    assert torch.equal(tail, torch.rand(3))  # every other consumer of the global generator stays aligned too


# ---- one call of every public method of every mirrored class ------------------------------------------
def test_walk_every_public_method(C, cuda_device):
    from agilerl_b200.components import spaces

    def tr(n, k):
        return C.TensorBatch({"obs": torch.full((n, 4), float(k)), "action": torch.zeros(n), "reward": torch.ones(n),
                              "next_obs": torch.full((n, 4), float(k + 1)), "done": torch.zeros(n)}, batch_size=[n])

    for dev in ("cpu", cuda_device):
        rb = C.ReplayBuffer(64, device=dev)
        ns = C.MultiStepReplayBuffer(64, n_step=2, gamma=0.9, device=dev)
        per = C.PrioritizedReplayBuffer(64, alpha=0.5, device=dev)
        for k in range(6):
            rb.add(tr(4, k))
            one = ns.add(tr(4, k))
            if one is not None:
                per.add(one)
        for buf in (rb, ns, per):
            assert len(buf) == buf.size > 0 and not buf.is_full and buf.storage is not None
            assert buf.initialized and buf.counter == buf.size
        out_dev = torch.device(dev).type
        assert rb.sample(8)["obs"].device.type == out_dev
        assert rb.sample(8, return_idx=True)["idxs"].shape == (8,)
        assert ns.sample(8)["obs"].device.type == out_dev            # inherited ReplayBuffer.sample
        assert ns.sample(8, True)["idxs"].shape == (8,)
        assert ns.sample_from_indices(torch.tensor([[0], [1]]))["obs"].shape == (2, 1, 4)
        assert ns._get_n_step_info()["reward"].shape == (4, 1)
        s = per.sample(8, 0.4)
        assert s["weights"].shape == (8, 1) and s["idxs"].shape == (8, 1) and s["obs"].device.type == out_dev
        per.update_priorities(s["idxs"], np.full(8, 0.5, np.float32))
        per._update_priority(0, 2.0)
        assert per.max_priority == 2.0
        assert per._sample_proportional(4).shape == (4,)
        assert per._calculate_weights(torch.tensor([0, 1]), 0.4).shape == (2,)
        assert per.sum_tree.sum() > 0 and per.min_tree.min() > 0 and per.sum_tree.retrieve(0.0) >= 0
        per.check_errors()
        sm = C.Sampler(memory=per)
        assert sm.sample(4, 0.4)["idxs"].shape == (4, 1)
        assert C.Sampler(memory=ns).sample(torch.tensor([0, 1]))["obs"].shape[0] == 2
        assert C.Sampler(memory=rb).sample(4)["obs"].shape[0] == 4
        rb.size = 5
        assert rb._size == 5
        for buf in (rb, ns, per):
            buf.clear()
            assert len(buf) == 0 and buf.storage is None and not buf.initialized

    ro = C.RolloutBuffer(capacity=4, observation_space=spaces.Box(-1, 1, (3,)), action_space=spaces.Discrete(2),
                         num_envs=2, device=cuda_device)
    for t in range(4):
        ro.add(obs=np.zeros((2, 3), np.float32), action=np.zeros(2), reward=np.ones(2, np.float32),
               done=np.zeros(2, bool), value=np.zeros(2, np.float32), log_prob=np.zeros(2, np.float32),
               next_obs=np.zeros((2, 3), np.float32))
    assert ro.size() == 8 and ro.full and ro.pos == 4
    ro.compute_returns_and_advantages(np.zeros(2, np.float32), np.zeros(2, bool))
    assert ro.get()["returns"].shape[0] == 8 and ro.get(4)["returns"].shape[0] == 4
    assert ro.get_tensor_batch()["advantages"].shape[0] == 8 and ro.get_tensor_batch(3)["advantages"].shape[0] == 3
    ro.reset()
    assert ro.size() == 0 and not ro.full


def test_sample_from_indices_index_errors_and_negative_indices(C, cuda_device):
    """``storage[idxs]`` semantics (replay_buffer.py:196-204): host indices raise IndexError at once and negative
    ones wrap like Python's; device indices are checked by the kernel and reported by a later call."""
    b = C.MultiStepReplayBuffer(max_size=16, n_step=1, device=cuda_device)
    for i in range(16):
        b.add(C.TensorBatch({"obs": torch.full((1, 2), float(i)), "action": torch.zeros(1), "reward": torch.ones(1),
                             "next_obs": torch.zeros(1, 2), "done": torch.zeros(1)}, batch_size=[1]))
    got = b.sample_from_indices(torch.tensor([-1, 0, -16]))
    assert got["obs"][:, 0].tolist() == [15.0, 0.0, 0.0]
    with pytest.raises(IndexError):
        b.sample_from_indices(torch.tensor([3, 16]))
    with pytest.raises(IndexError):
        b.sample_from_indices(torch.tensor([-17]))
    out = b.sample_from_indices(torch.tensor([2, 99], device=cuda_device))  # device indices: no host sync here
    assert out["obs"][0, 0].item() == 2.0 and out["obs"][1].abs().sum().item() == 0.0
    with pytest.raises(IndexError):
        b.check_gather_errors()
    b.check_gather_errors()  # flag cleared


def test_uniform_sample_larger_than_the_prefix_kernel_table(C):
    """replay_buffer.py:125 with batch_size > 16 384 (beyond xp_randperm_prefix's shared-memory table): still the
    exact torch.randperm(size)[:batch_size], and batch_size > size returns size rows like the reference's slice."""
    b = C.ReplayBuffer(max_size=40_000)
    b.add(C.TensorBatch({"x": torch.arange(40_000, dtype=torch.float32).view(-1, 1)}, batch_size=[40_000]))
    torch.manual_seed(3)
    want = torch.randperm(40_000)[:20_000]
    torch.manual_seed(3)
    s = b.sample(20_000, return_idx=True)
    assert torch.equal(s["idxs"].cpu(), want) and torch.equal(s["x"].cpu().view(-1), want.float())
    torch.manual_seed(4)
    want = torch.randperm(40_000)[:50_000]
    torch.manual_seed(4)
    s = b.sample(50_000, return_idx=True)
    assert s["idxs"].shape == (40_000,) and torch.equal(s["idxs"].cpu(), want)
