"""Scenarios of the reference's own tests/test_components/test_rollout_buffer.py, restated against the
B200 RolloutBuffer (same constructor arguments, same calls, same assertions; storage on the GPU)."""

import numpy as np
import pytest
import torch

from agilerl_b200.components import RolloutBuffer, spaces

pytestmark = pytest.mark.gpu


def test_dict_observations_add_get(cuda_device):  # reference tests/test_components/test_rollout_buffer.pytest_rollout_buffer.py:53-75,278-324,842-884
    sp = spaces.Dict({"vector": spaces.Box(low=-1, high=1, shape=(2,), dtype=np.float32), "discrete": spaces.Discrete(3)})
    b = RolloutBuffer(capacity=5, observation_space=sp, action_space=spaces.Discrete(2), num_envs=2, device="cuda")
    assert b.buffer["observations"]["vector"].shape == (5, 2, 2) and b.buffer["observations"]["discrete"].shape == (5, 2, 1)
    obs = {"vector": np.array([[1.0, 2.0], [3.0, 4.0]]), "discrete": np.array([[0], [2]])}
    b.add(obs=obs, action=np.array([0, 1]), reward=np.array([1.0, -1.0]), done=np.array([False, True]),
          value=np.array([0.5, -0.5]), log_prob=np.array([0.1, 0.2]))
    assert b.pos == 1
    assert torch.allclose(b.buffer["observations"]["vector"][0].cpu(), torch.tensor(obs["vector"], dtype=torch.float32))
    assert torch.allclose(b.buffer["observations"]["discrete"][0].cpu(), torch.tensor(obs["discrete"], dtype=torch.float32))
    b.compute_returns_and_advantages(np.zeros(2), np.zeros(2))
    d = b.get()
    assert isinstance(d["observations"], dict) and d["observations"]["vector"].shape == (2, 2)
    tb = b.get_tensor_batch()
    assert tb["observations"]["vector"].shape == (2, 2)

def test_recurrent_hidden_states_through_bptt(cuda_device):  # reference tests/test_components/test_rollout_buffer.py:77-121,417-456,990-1147
    arch = {"h_actor": (1, 2, 4)}
    b = RolloutBuffer(capacity=5, observation_space=spaces.Box(low=-1, high=1, shape=(2,), dtype=np.float32),
                      action_space=spaces.Discrete(2), num_envs=2, device="cuda", recurrent=True, hidden_state_architecture=arch)
    assert b.buffer.get("hidden_states") is not None
    for t in range(5):
        b.add(obs=np.ones((2, 2)) * t, action=np.array([0, 1]), reward=np.array([1.0, -1.0]), done=np.array([False, t == 2]),
              value=np.array([0.5, -0.5]), log_prob=np.array([0.1, 0.2]), hidden_state={"h_actor": torch.randn(1, 2, 4)})
    assert "h_actor" in b.buffer["hidden_states"] and b.full
    b.compute_returns_and_advantages(np.zeros(2), np.zeros(2))
    b.prepare_sequence_tensors()
    mbs = list(b.get_minibatch_sequences(batch_size=2))
    assert len(mbs) >= 1
    padded, unpadded = mbs[0]  # rollout_buffer.py:922 yields (padded, unpadded)
    for k in ("observations", "actions", "pad_mask"):
        assert k in padded
    for k in ("log_probs", "advantages", "values", "returns"):
        assert k in unpadded
    assert "h_actor" in padded.get_non_tensor("initial_hidden_states")

def test_recurrent_without_architecture_raises(cuda_device):  # reference tests/test_components/test_rollout_buffer.py:123-139
    with pytest.raises(ValueError):
        RolloutBuffer(capacity=5, observation_space=spaces.Box(low=-1, high=1, shape=(2,), dtype=np.float32),
                      action_space=spaces.Discrete(2), num_envs=2, device="cuda", recurrent=True)

def test_action_space_shapes(cuda_device):  # reference tests/test_components/test_rollout_buffer.py:208-276,2031-2071
    for asp, act, shape in ((spaces.MultiDiscrete([3, 2]), np.array([[1, 0], [2, 1]]), (5, 2, 2)),
                            (spaces.Box(low=-1, high=1, shape=(3,), dtype=np.float32), np.ones((2, 3)), (5, 2, 3)),
                            (spaces.Discrete(4), np.array([1, 3]), (5, 2, 1))):
        b = RolloutBuffer(capacity=5, observation_space=spaces.Box(low=-1, high=1, shape=(2,), dtype=np.float32),
                          action_space=asp, num_envs=2, device="cuda")
        b.add(obs=np.ones((2, 2)), action=act, reward=np.ones(2), done=np.zeros(2, bool), value=np.ones(2), log_prob=np.ones(2))
        assert tuple(b.buffer["actions"].shape) == shape, (asp, b.buffer["actions"].shape)

def test_empty_oversize_batch_and_reset(cuda_device):  # reference tests/test_components/test_rollout_buffer.py:652-734,886-922,1770-1862
    b = RolloutBuffer(capacity=5, observation_space=spaces.Box(low=-1, high=1, shape=(2,), dtype=np.float32),
                      action_space=spaces.Discrete(2), num_envs=2, device="cuda")
    assert b.get() == {} and b.get_tensor_batch().is_empty()
    for t in range(3):
        b.add(obs=np.ones((2, 2)), action=np.array([0, 1]), reward=np.ones(2), done=np.zeros(2, bool), value=np.ones(2), log_prob=np.ones(2))
    import warnings
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        d = b.get(batch_size=100)
        assert d["rewards"].shape[0] == 6
    d = b.get(batch_size=4)
    assert d["rewards"].shape[0] == 4
    b.reset(); assert b.pos == 0 and not b.full and b.size() == 0

def test_setstate_reinitialises_invalid_buffer(cuda_device):  # reference tests/test_components/test_rollout_buffer.py:1908-2029
    import warnings

    def fresh():
        return RolloutBuffer(capacity=5, observation_space=spaces.Box(low=-1, high=1, shape=(2,), dtype=np.float32),
                             action_space=spaces.Discrete(2), num_envs=2, device="cuda")

    b = fresh()
    b.add(obs=np.ones((2, 2)), action=np.array([0, 1]), reward=np.ones(2), done=np.zeros(2, bool), value=np.ones(2),
          log_prob=np.ones(2))
    # a sound state round-trips with its contents
    nb = RolloutBuffer.__new__(RolloutBuffer)
    nb.__setstate__(b.__getstate__())
    assert nb.pos == 1 and torch.equal(nb.buffer["rewards"], b.buffer["rewards"])
    # missing / invalid buffer -> warning + fresh buffers (:1908-1953)
    for mutate in (lambda s: s.pop("buffer"), lambda s: s.__setitem__("buffer", "invalid"), lambda s: s.pop("_leaves")):
        st = b.__getstate__()
        mutate(st)
        nb = RolloutBuffer.__new__(RolloutBuffer)
        with pytest.warns(UserWarning, match="Re-initializing"):
            nb.__setstate__(st)
        assert nb.buffer["rewards"].shape == (5, 2) and float(nb.buffer["rewards"].abs().sum()) == 0.0
    # spaces missing as well -> second warning, early return, buffer left as found (:1955-1979)
    st = b.__getstate__()
    st["buffer"] = "invalid"
    del st["observation_space"]
    nb = RolloutBuffer.__new__(RolloutBuffer)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        nb.__setstate__(st)
    assert any("Observation" in str(m.message) for m in w) and nb.buffer == "invalid"
    # leading shape mismatch -> warning + re-initialisation at the expected shape (:1981-2029)
    st = b.__getstate__()
    st["num_envs"] = 3
    nb = RolloutBuffer.__new__(RolloutBuffer)
    with pytest.warns(UserWarning, match="expected"):
        nb.__setstate__(st)
    assert nb.buffer["rewards"].shape == (5, 3)


def test_single_env_and_next_obs(cuda_device):  # reference tests/test_components/test_rollout_buffer.py:326-359,2073-2110
    b = RolloutBuffer(capacity=4, observation_space=spaces.Box(low=-1, high=1, shape=(3,), dtype=np.float32),
                      action_space=spaces.Discrete(2), num_envs=1, device="cuda")
    b.add(obs=np.ones(3), action=np.array(1), reward=np.array(1.0), done=np.array(False), value=np.array(0.3), log_prob=np.array(-0.2), next_obs=np.ones(3) * 2)
    assert b.buffer["next_observations"][0].cpu().flatten().tolist() == [2.0, 2.0, 2.0]


# ---- replay buffers: reference tests/test_components/test_replay_buffer.py scenarios not covered elsewhere ----
def test_replay_dict_observations_size_setter_storage(cuda_device):  # :302-340, :398-430, :826-864
    from agilerl_b200.components import PrioritizedReplayBuffer, ReplayBuffer, TensorBatch

    for cls, kw in ((ReplayBuffer, {}), (PrioritizedReplayBuffer, {"alpha": 0.6})):
        buf = cls(max_size=100, device="cuda", **kw)
        for i in range(10):
            obs = TensorBatch({"image": torch.ones((1, 3, 84, 84)) * i, "vector": torch.tensor([[i, i + 1, i + 2]])},
                              batch_size=[1])
            buf.add(TensorBatch({"state": obs, "action": torch.tensor([[i % 3]]), "reward": torch.tensor([[float(i)]])},
                                batch_size=[1]))
        assert len(buf) == 10 and buf.storage is not None and buf.storage.batch_size[0] == 100
        s = buf.sample(5) if cls is ReplayBuffer else buf.sample(5, beta=0.4)
        assert tuple(s.batch_size) == (5,)
        assert s["state"]["image"].shape == (5, 3, 84, 84) and s["state"]["vector"].shape == (5, 3)
        assert s["action"].shape == (5, 1) and s["reward"].shape == (5, 1)
        # every sampled row is one of the stored transitions, image and vector from the SAME step
        img = s["state"]["image"][:, 0, 0, 0].cpu()
        assert torch.equal(img.to(s["state"]["vector"].dtype), s["state"]["vector"][:, 0].cpu())
        assert torch.equal(img, s["reward"][:, 0].cpu())
        if cls is PrioritizedReplayBuffer:
            assert s["weights"].shape == (5, 1) and s["idxs"].shape == (5, 1)
        buf.size = 5
        assert buf._size == 5 and len(buf) == 5


def test_arena_keeps_one_launch_through_clone_and_mutation(cuda_device):
    """The evolutionary loop deep-copies a winner's buffer (agent.clone(), algorithms/core/base.py:444-503) and
    re-creates buffers after a mutation (algorithms/ppo.py:336-339,360-378).  RolloutArena.adopt re-attaches such a
    buffer — the same object the agent holds — to the slab, so ONE launch still covers the population; a buffer of
    another capacity gets a launch of its own.  Results bit-equal to per-agent compute_returns_and_advantages."""
    import copy

    import agilerl_b200.components as C
    from agilerl_b200 import _native as nv
    from agilerl_b200.components import spaces

    lib = nv.load()
    P, T, N = 4, 12, 32
    osp, asp = spaces.Box(-1, 1, (5,)), spaces.Discrete(3)
    arena = C.RolloutArena(P, T, osp, asp, num_envs=N, device=cuda_device, gamma=0.97, gae_lambda=0.9)
    rng = np.random.default_rng(0)

    def fill(buf, steps):
        buf.reset()
        for _ in range(steps):
            buf.add(obs=rng.standard_normal((N, 5)).astype(np.float32), action=rng.integers(0, 3, N),
                    reward=rng.standard_normal(N).astype(np.float32), done=rng.random(N) < 0.1,
                    value=rng.standard_normal(N).astype(np.float32), log_prob=rng.standard_normal(N).astype(np.float32))

    for b in arena.buffers:
        fill(b, T)
    # tournament: agent 1 wins, its clone replaces agent 3; agent 2 is mutated to a shorter learn_step
    clone = copy.deepcopy(arena.buffers[1])
    assert clone._leaves[("rewards",)].data_ptr() != arena.slab[("rewards",)][1].data_ptr()  # standalone copy
    got = arena.adopt(3, clone)
    assert got is clone and arena.buffers[3] is clone
    assert clone._leaves[("rewards",)].data_ptr() == arena.slab[("rewards",)][3].data_ptr()  # re-bound to the slab
    assert torch.equal(clone.buffer["rewards"], arena.buffers[1].buffer["rewards"])           # contents came along
    mutated = C.RolloutBuffer(T // 2, osp, asp, num_envs=N, device=cuda_device, gamma=0.97, gae_lambda=0.9)
    arena.adopt(2, mutated)
    assert 2 in arena._foreign
    # next generation: everybody collects again through the objects the agents hold
    for b in arena.buffers:
        fill(b, b.capacity)
    lv = rng.standard_normal((P, N)).astype(np.float32)
    ld = rng.random((P, N)) < 0.2
    c0 = lib.xp_launch_count()
    arena.compute_returns_and_advantages(lv, ld)
    assert lib.xp_launch_count() - c0 == 2  # the slab in one launch + the buffer of another capacity
    for i, b in enumerate(arena.buffers):
        ref = C.RolloutBuffer(b.capacity, osp, asp, num_envs=N, device=cuda_device, gamma=0.97, gae_lambda=0.9)
        for k in ("rewards", "values", "dones"):
            ref.buffer[k][:] = b.buffer[k]
        ref.pos, ref.full = b.pos, b.full
        ref.compute_returns_and_advantages(lv[i], ld[i])
        assert torch.equal(ref.buffer["advantages"], b.buffer["advantages"]), i
        assert torch.equal(ref.buffer["returns"], b.buffer["returns"]), i
    # re-slab the whole population after the generation: the odd one out stays standalone
    arena2 = C.RolloutArena.from_buffers(arena.buffers)
    assert arena2._foreign == {2} and all(x is y for x, y in zip(arena2.buffers, arena.buffers))
    before = [b.buffer["advantages"].clone() for b in arena2.buffers]
    arena2.compute_returns_and_advantages(lv, ld)
    assert all(torch.equal(x, b.buffer["advantages"]) for x, b in zip(before, arena2.buffers))
