"""Host-side logic that needs no GPU: containers, input formatting, dispatch, shape helpers, and
the rule that the product path refuses to run without the CUDA engine (no CPU fallback)."""

import numpy as np
import pytest
import torch

import agilerl_b200.components as C
from agilerl_b200 import _native as nv
from agilerl_b200.components import spaces
from agilerl_b200.population import ShardPlan

needs_no_gpu = pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")


def test_tensor_batch_mapping_and_indexing():
    b = C.TensorBatch({"obs": torch.arange(12.0).view(4, 3), "nested": {"a": torch.arange(4)}}, batch_size=[4])
    assert list(b.keys()) == ["obs", "nested"] and "obs" in b and b.get("zzz") is None and not b.is_empty()
    assert b[1]["obs"].tolist() == [3.0, 4.0, 5.0] and b[1].batch_size == torch.Size([])
    assert b[1:3].batch_size == torch.Size([2]) and b[torch.tensor([3, 0])]["nested"]["a"].tolist() == [3, 0]
    assert b[np.array([2])]["obs"].shape == (1, 3)
    c = b.clone()
    c["obs"].zero_()
    assert b["obs"].sum() > 0
    b[0:2] = C.TensorBatch({"obs": torch.ones(2, 3), "nested": {"a": torch.tensor([9, 9])}}, batch_size=[2])
    assert b["obs"][:2].sum() == 6 and b["nested"]["a"].tolist() == [9, 9, 2, 3]
    assert b.unsqueeze(0).batch_size == torch.Size([1, 4]) and b.view(2, 2)["obs"].shape == (2, 2, 3)
    assert b.to(torch.float64)["obs"].dtype == torch.float64
    assert C.TensorBatch({}, batch_size=[0]).is_empty()
    assert b.select("obs").exclude("obs").is_empty()


def test_transition_casts_like_reference():
    # agilerl/components/data.py:76-93 and tests/test_components/test_replay_data.py:82-126
    t = C.Transition(obs=np.zeros((2, 4), np.uint8), action=np.array([1, 2]), next_obs=np.ones((2, 4), np.uint8),
                     reward=np.array([0.5, 1.5], np.float64), done=np.array([True, False]), batch_size=[2])
    td = t.to_tensordict()
    assert td["obs"].dtype == torch.uint8 and td["action"].dtype == torch.float32
    assert td["reward"].dtype == torch.float32 and td["done"].dtype == torch.float32 and td["done"].tolist() == [1.0, 0.0]
    s = C.Transition(obs={"a": np.zeros(3)}, action=1, next_obs=(np.zeros(2), np.ones(1)), reward=1.0, done=False)
    assert s.reward.shape == (1,) and s.done.shape == (1,) and s.action.ndim == 0
    assert isinstance(s.obs, C.TensorBatch) and list(s.next_obs.keys()) == ["tuple_obs_0", "tuple_obs_1"]
    assert s.obs["a"].dtype == torch.float32


def test_class_identity_and_sampler_dispatch_without_gpu():
    assert issubclass(C.PrioritizedReplayBuffer, C.ReplayBuffer)
    assert issubclass(C.MultiStepReplayBuffer, C.ReplayBuffer)
    per, nst, std = C.PrioritizedReplayBuffer(64), C.MultiStepReplayBuffer(64), C.ReplayBuffer(64)
    assert C.Sampler(memory=per).sample.__func__ is C.Sampler.sample_per
    assert C.Sampler(memory=nst).sample.__func__ is C.Sampler.sample_n_step
    assert C.Sampler(memory=std).sample.__func__ is C.Sampler.sample_standard
    assert per.max_priority == 1.0 and per.tree_ptr == 0 and nst.n_step == 3 and nst.gamma == 0.99
    assert nst.reward_key == "reward" and nst.done_key is None and nst.ns_key == "next_obs"
    with pytest.raises(AssertionError):
        C.Sampler()

    class Sub(C.ReplayBuffer):  # the reference's tests subclass the buffers (test_train.py:432)
        pass

    assert isinstance(Sub(4), C.ReplayBuffer)


def test_space_helpers():
    assert spaces.get_obs_shape(spaces.Box(-1, 1, (84, 84, 4))) == (84, 84, 4)
    assert spaces.get_obs_shape(spaces.Discrete(5)) == (1,)
    assert spaces.get_obs_shape(spaces.MultiDiscrete([2, 3, 4])) == (3,)
    assert spaces.get_obs_shape(spaces.MultiBinary(6)) == (6,)
    assert spaces.get_obs_shape(spaces.Dict({"a": spaces.Box(0, 1, (2,)), "b": spaces.Discrete(2)})) == {"a": (2,), "b": (1,)}
    assert spaces.get_num_actions(spaces.Box(-1, 1, (17,))) == 17 and spaces.get_num_actions(spaces.Discrete(4)) == 1
    assert spaces.get_num_actions(spaces.MultiDiscrete([3, 3])) == 2 and spaces.get_num_actions(spaces.MultiBinary(5)) == 5


@needs_no_gpu
def test_product_path_fails_loudly_without_cuda():
    b = C.ReplayBuffer(8)
    with pytest.raises(nv.NativeError):
        b.add(C.TensorBatch({"x": torch.zeros(2, 2)}, batch_size=[2]))
    with pytest.raises(nv.NativeError):
        C.PrioritizedReplayBuffer(8).sum_tree
    with pytest.raises(nv.NativeError):
        C.RolloutBuffer(4, spaces.Box(-1, 1, (3,)), spaces.Discrete(2))
    with pytest.raises(nv.NativeError):
        C.SumSegmentTree(4)


def test_missing_library_is_an_error(monkeypatch, tmp_path):
    monkeypatch.setattr(nv, "_lib", None)
    monkeypatch.setenv("B200XP_LIB", str(tmp_path / "nope.so"))
    with pytest.raises(nv.NativeError):
        nv.load()


def test_product_never_imports_the_oracle():
    import os
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for dirpath, _, files in os.walk(os.path.join(root, "agilerl_b200")):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f


def test_shard_plan_partitions_population():
    for P, G in [(64, 8), (64, 1), (16, 4), (10, 4), (3, 8)]:
        seen = []
        for r in range(G):
            plan = ShardPlan(P, G, r)
            seen.extend(plan.local_agents)
            for a in plan.local_agents:
                assert plan.owner(a) == r
        assert seen == list(range(P))
    assert list(ShardPlan(64, 8, 3).local_agents) == list(range(24, 32))


# ---- reference tests/test_components/test_replay_data.py:80-170, restated (host-only formatters) -----
def test_transition_and_formatters_like_reference_data_tests():
    from agilerl_b200.components.data import Transition, to_tensordict, to_torch_tensor

    t = Transition(obs=(np.array([1.0, 2.0]), 3.0), action=1, reward=1.0, next_obs=(np.array([4.0, 5.0]), 6.0), done=False)
    assert "tuple_obs_0" in t.obs.keys() and "tuple_obs_1" in t.obs.keys()
    with pytest.raises(AssertionError, match="Expected all elements of the tuple"):
        Transition(obs=(np.array([1.0]), object()), action=0, reward=0.0, next_obs=np.array([2.0]), done=False)
    t = Transition(obs={"vec": np.array([1.0, 2.0])}, action=0, reward=1.0, next_obs={"vec": np.array([3.0, 4.0])}, done=False)
    assert "vec" in t.obs.keys() and "vec" in t.next_obs.keys()
    t = Transition(obs=np.array([1.0]), action=0, reward=1.0, next_obs=np.array([2.0]), done=False)
    assert t.reward.ndim >= 1 and t.done.ndim >= 1
    td = to_tensordict((np.array([1.0, 2.0]), 3.0))
    assert "tuple_obs_0" in td.keys() and td["tuple_obs_1"].dtype == torch.float32
    td = to_tensordict({"a": np.array([1.0]), "b": torch.tensor([2.0])})
    assert "a" in td.keys() and "b" in td.keys() and all(v.dtype == torch.float32 for v in td.values())
    with pytest.raises(AssertionError, match="Expected all values of the dict"):
        to_tensordict({"a": np.array([1.0]), "b": "invalid"})
    assert to_torch_tensor(np.array([1.0, 2.0])).dtype == torch.float32
    assert to_torch_tensor(42).item() == 42
    assert to_torch_tensor(True).dtype == torch.float32
    assert to_torch_tensor(torch.tensor([1.0], dtype=torch.float64)).dtype == torch.float32
    assert to_torch_tensor([1, 2, 3]).shape == (3,)
