"""Parity of the public API (agilerl_b200.components) against the reference: ports of the
reference's own known-answer tests (tests/test_components/test_replay_buffer.py,
test_segment_tree.py, test_rollout_buffer.py, test_sampler.py — cited per test) and replays of
the golden scenarios generated from the unmodified reference (oracle/make_golden.py)."""

import copy
import pickle
import warnings

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import xp_oracle as xo

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def C(cuda_device):
    import agilerl_b200.components as comps

    return comps


def TB(C, d, n):
    return C.TensorBatch(d, batch_size=[n])


# ---- ReplayBuffer (reference test_replay_buffer.py:15-394) -----------------------------------------
def test_replay_buffer_init_state(C):
    b = C.ReplayBuffer(max_size=1000, device="cpu")
    assert b.max_size == 1000 and b.counter == 0 and b.initialized is False
    assert b._cursor == 0 and b.size == 0 and b.storage is None and len(b) == 0 and not b.is_full


def test_add_wraps_with_reference_layout(C):
    # :145-178 storage layout after the wrap split: [data2[1], data1[1], data2[0]]
    b = C.ReplayBuffer(max_size=3)
    b.add(TB(C, {"state": torch.tensor([[1, 2, 3], [4, 5, 6]]), "action": torch.tensor([[0], [1]]),
                 "reward": torch.tensor([[1.0], [2.0]])}, 2))
    assert len(b) == 2 and b._cursor == 2
    b.add(TB(C, {"state": torch.tensor([[7, 8, 9], [10, 11, 12]]), "action": torch.tensor([[2], [3]]),
                 "reward": torch.tensor([[3.0], [4.0]])}, 2))
    assert len(b) == 3 and b._cursor == 1 and b.counter == 4 and b.is_full
    st = b.storage[: b.size]
    assert torch.equal(st["state"].cpu(), torch.tensor([[10, 11, 12], [4, 5, 6], [7, 8, 9]]))
    assert torch.equal(st["reward"].cpu(), torch.tensor([[4.0], [2.0], [3.0]]))


def test_add_reshapes_1d_fields(C):
    # :180-215 (flat and nested)
    b = C.ReplayBuffer(max_size=10)
    b.add(TB(C, {"state": torch.tensor([[1, 2, 3], [4, 5, 6]]), "action": torch.tensor([0.0, 1.0]),
                 "reward": torch.tensor([1.0, 2.0])}, 2))
    s = b.sample(2)
    assert s["reward"].shape == (2, 1) and s["action"].shape == (2, 1)
    b2 = C.ReplayBuffer(max_size=10)
    b2.add(TB(C, {"state": {"scalar": torch.tensor([1.0, 2.0])}, "action": torch.tensor([[0], [1]]),
                  "reward": torch.tensor([[1.0], [2.0]])}, 2))
    assert b2.sample(2)["state"]["scalar"].shape == (2, 1)


@pytest.mark.parametrize("obs_shape,dtype", [((4,), torch.float32), ((3, 84, 84), torch.uint8)])
def test_sample_shapes_dtypes_and_device(C, cuda_device, obs_shape, dtype):
    # :239-366
    b = C.ReplayBuffer(max_size=64, device=cuda_device)
    for _ in range(5):
        n = 8
        b.add(TB(C, {"obs": torch.zeros(n, *obs_shape, dtype=dtype), "action": torch.zeros(n),
                     "reward": torch.ones(n), "next_obs": torch.ones(n, *obs_shape, dtype=dtype),
                     "done": torch.zeros(n)}, n))
    s = b.sample(16, return_idx=True)
    assert s.batch_size == torch.Size([16])
    assert s["obs"].shape == (16, *obs_shape) and s["obs"].dtype == dtype and s["obs"].is_cuda
    assert s["idxs"].shape == (16,) and s["idxs"].dtype == torch.int64
    assert len(torch.unique(s["idxs"])) == 16 and int(s["idxs"].max()) < b.size
    cpu_b = C.ReplayBuffer(max_size=8, device="cpu")
    cpu_b.add(TB(C, {"obs": torch.ones(2, 3)}, 2))
    assert cpu_b.sample(2)["obs"].device.type == "cpu"  # batches come back on `device`


def test_sample_is_torch_randperm_prefix_and_a_copy(C):
    # replay_buffer.py:125 — identical indices under the same seed; TD3 mutates samples in place
    b = C.ReplayBuffer(max_size=5000)
    data = torch.arange(5000, dtype=torch.float32).view(-1, 1)
    for i in range(0, 5000, 500):
        b.add(TB(C, {"x": data[i:i + 500]}, 500))
    for seed in (0, 3):
        torch.manual_seed(seed)
        want = torch.randperm(b.size)[:64]
        torch.manual_seed(seed)
        s = b.sample(64, return_idx=True)
        assert torch.equal(s["idxs"].cpu(), want)
        assert torch.equal(s["x"].cpu().view(-1), want.float())
    s["x"].zero_()
    assert float(b.storage["x"].abs().sum()) > 0
    # with exact_rng_advance the generator ends where the reference leaves it
    b.exact_rng_advance = True
    torch.manual_seed(9)
    b.sample(64)
    after = torch.rand(4)
    torch.manual_seed(9)
    torch.randperm(b.size)
    assert torch.equal(after, torch.rand(4))


def test_device_rng_mode_samples_without_replacement(C):
    b = C.ReplayBuffer(max_size=1000, rng="device", seed=5)
    b.add(TB(C, {"x": torch.arange(1000).view(-1, 1)}, 1000))
    a = b.sample(256, return_idx=True)["idxs"]
    c = b.sample(256, return_idx=True)["idxs"]
    assert len(torch.unique(a)) == 256 and not torch.equal(a, c)


def test_clear(C):
    b = C.ReplayBuffer(max_size=10)
    b.add(TB(C, {"x": torch.ones(4, 2)}, 4))
    b.clear()
    assert len(b) == 0 and b._cursor == 0 and b.storage is None and b.initialized is False
    b.add(TB(C, {"y": torch.ones(3, 5)}, 3))  # re-initialises with a new schema
    assert list(b.storage.keys()) == ["y"] and len(b) == 3


# ---- MultiStepReplayBuffer (reference :432-730) ------------------------------------------------------
def test_nstep_reward_kats(C):
    # :532-569  R = 1 + 0.9*2 + 0.81*3 ; :571-628 early termination
    b = C.MultiStepReplayBuffer(max_size=100, n_step=3, gamma=0.9)
    outs = []
    for r, d in ((1.0, 0.0), (2.0, 0.0), (3.0, 0.0)):
        outs.append(b.add(TB(C, {"obs": torch.full((1, 2), r), "action": torch.zeros(1, 1),
                                 "reward": torch.tensor([[r]]), "next_obs": torch.full((1, 2), r + 1),
                                 "done": torch.tensor([[d]])}, 1)))
    assert outs[0] is None and outs[1] is None and outs[2] is not None
    assert len(b) == 1 and len(b.n_step_buffer) == 3 and b.done_key == "done"
    st = b.storage[0]
    want = np.float32(1.0) + np.float32(2.0) * np.float32(0.9) + np.float32(3.0) * np.float32(0.9**2)
    assert st["reward"].item() == pytest.approx(1.0 + 0.9 * 2.0 + 0.81 * 3.0)
    assert np.float32(st["reward"].item()) == np.float32(want)
    assert torch.equal(st["next_obs"].cpu(), torch.full((2,), 4.0)) and st["done"].item() == 0.0
    assert torch.equal(outs[2]["obs"].cpu(), torch.full((1, 2), 1.0))  # oldest raw transition (:194)

    b = C.MultiStepReplayBuffer(max_size=100, n_step=3, gamma=0.9)
    for r, d in ((1.0, 0.0), (2.0, 1.0), (3.0, 0.0)):
        b.add(TB(C, {"obs": torch.full((1, 2), r), "action": torch.zeros(1, 1), "reward": torch.tensor([[r]]),
                     "next_obs": torch.full((1, 2), r + 1), "termination": torch.tensor([[d]])}, 1))
    st = b.storage[0]
    assert b.done_key == "termination"  # :630-666 key detection
    assert st["reward"].item() == pytest.approx(1.0 + 0.9 * 2.0) and st["termination"].item() == 1.0
    assert torch.equal(st["next_obs"].cpu(), torch.full((2,), 3.0))


def test_nstep_missing_keys_assert(C):
    b = C.MultiStepReplayBuffer(max_size=10, n_step=1)
    with pytest.raises(AssertionError):
        b.add(TB(C, {"obs": torch.zeros(1, 2), "reward": torch.zeros(1, 1), "next_obs": torch.zeros(1, 2)}, 1))


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_ring_and_nstep_golden_through_api(C, seed):
    g = load_golden(f"ring_nstep_s{seed}")
    max_size, N, n_step = int(g["max_size"]), int(g["N"]), int(g["n_step"])
    ring = C.ReplayBuffer(max_size=max_size)
    nst = C.MultiStepReplayBuffer(max_size=max_size, n_step=n_step, gamma=float(g["gamma"]))
    keys = ["obs", "action", "reward", "next_obs", "done"]
    for s in range(g["obs"].shape[0]):
        d = {k: torch.from_numpy(g[k][s]) for k in keys}
        ring.add(TB(C, dict(d), N))
        out = nst.add(TB(C, dict(d), N))
        assert (out is None) == bool(g["returned_none"][s])
    assert ring._cursor == int(g["ring_cursor"]) and ring.size == int(g["ring_size"])
    assert nst._cursor == int(g["nstep_cursor"]) and nst.size == int(g["nstep_size"])
    for k in keys:
        assert np.array_equal(ring.storage[k].cpu().numpy(), g[f"ring_{k}"]), k
        assert np.array_equal(nst.storage[k].cpu().numpy(), g[f"nstep_{k}"]), k
    idxs = torch.tensor([[0], [2], [1]])
    got = nst.sample_from_indices(idxs)
    assert got["obs"].shape[:2] == (3, 1)
    assert np.array_equal(got["reward"].cpu().numpy()[:, 0], g["nstep_reward"][[0, 2, 1]])


# ---- PrioritizedReplayBuffer (reference :734-1056) ---------------------------------------------------
def one(C, i):
    return TB(C, {"state": torch.tensor([[i, i + 1, i + 2]]), "action": torch.tensor([[i % 3]]),
                  "reward": torch.tensor([[float(i)]])}, 1)


def test_per_init_and_add(C):
    b = C.PrioritizedReplayBuffer(max_size=100, alpha=0.6)
    assert b.alpha == 0.6 and b.max_priority == 1.0 and b.tree_ptr == 0
    assert isinstance(b.sum_tree, C.SumSegmentTree) and isinstance(b.min_tree, C.MinSegmentTree)
    assert b.sum_tree.capacity == 128
    b.add(one(C, 0))
    assert len(b) == 1 and b.tree_ptr == 1
    assert b.sum_tree[0] == b.max_priority**b.alpha and b.min_tree[0] == b.max_priority**b.alpha
    b.add(TB(C, {"state": torch.tensor([[1, 2, 3], [4, 5, 6]]), "action": torch.tensor([[0], [1]]),
                 "reward": torch.tensor([[1.0], [2.0]])}, 2))
    assert len(b) == 3 and b.tree_ptr == 3 and b.sum_tree[2] == 1.0


def test_per_update_priorities_kats(C):
    # :901-952
    b = C.PrioritizedReplayBuffer(max_size=100, alpha=0.6)
    for i in range(5):
        b.add(one(C, i))
    b.update_priorities(np.arange(3), torch.tensor([2.0, 3.0, 4.0]))
    for idx, p in zip(range(3), (2.0, 3.0, 4.0)):
        assert b.sum_tree[idx] == p**b.alpha and b.min_tree[idx] == p**b.alpha
    assert b.max_priority == 4.0
    b.update_priorities(torch.tensor([[0], [1]]), torch.tensor([1e-10, 1e-10]))  # (B,1) indices
    assert b.sum_tree[0] == 1e-5**b.alpha and b.sum_tree[1] == 1e-5**b.alpha


def test_per_weight_formula_and_asserts(C):
    # :1004-1056
    b = C.PrioritizedReplayBuffer(max_size=100, alpha=0.6)
    for i in range(5):
        b.add(one(C, i))
    pri = [0.5, 1.0, 1.5, 2.0, 2.5]
    for i, p in enumerate(pri):
        b._update_priority(i, p)
    for i, p in enumerate(pri):
        w = b._calculate_weights(torch.tensor([i]), 0.4)[0]
        p_sample = (p**b.alpha) / b.sum_tree.sum()
        p_min = b.min_tree.min() / b.sum_tree.sum()
        want = ((p_sample * len(b)) ** -0.4) / ((p_min * len(b)) ** -0.4)
        assert w.item() == np.float32(want)  # bit-exact fp32, the reference asks rtol 1e-5
    with pytest.raises(AssertionError):
        b._update_priority(100, 1.0)
    with pytest.raises(AssertionError):
        b.update_priorities(torch.tensor([3, 100]), torch.tensor([1.0, 1.0]))
    idx = b._sample_proportional(4)
    assert idx.shape == (4,) and all(0 <= int(i) < len(b) for i in idx)


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_per_golden_scenario_through_api(C, seed):
    """Same adds / updates / torch seed as the reference run -> same trees, indices, weights."""
    g = load_golden(f"per_s{seed}")
    max_size = int(g["max_size"])
    b = C.PrioritizedReplayBuffer(max_size=max_size, alpha=float(g["alpha"]))
    n_added = 0
    for step, n in enumerate(g["add_sizes"], start=1):
        n = int(n)
        b.add(TB(C, {"obs": torch.full((n, 4), float(step)), "reward": torch.zeros(n)}, n))
        n_added += n
        if step == 3:
            b.update_priorities(torch.from_numpy(g["early_idx"]), g["early_prio"])
    b.update_priorities(torch.from_numpy(g["upd_idx"]), torch.from_numpy(g["upd_prio"]))
    assert b.tree_ptr == int(g["tree_ptr"]) and b.size == int(g["size"])
    assert np.array_equal(np.array(b.sum_tree.tree.tolist()), g["sum_tree_after_update"])
    assert np.array_equal(np.array(b.min_tree.tree.tolist()), g["min_tree_after_update"])
    assert b.max_priority == float(g["max_priority"])
    B = len(g["u"])
    torch.manual_seed(77 + seed)
    s = b.sample(B, beta=float(g["beta"]))
    assert s["idxs"].shape == (B, 1) and s["weights"].shape == (B, 1) and s["weights"].dtype == torch.float32
    assert np.array_equal(s["idxs"].cpu().numpy()[:, 0], g["sample_idx"])
    assert np.array_equal(s["weights"].cpu().numpy()[:, 0], g["weights"])
    assert torch.equal(s["obs"].cpu(), b.storage["obs"].cpu()[torch.from_numpy(g["sample_idx"])])


# ---- segment trees (reference test_segment_tree.py) ---------------------------------------------------
def test_segment_tree_api_kats(C):
    t = C.SumSegmentTree(4)
    t[2] = 1.0
    t[3] = 3.0
    assert t.sum() == 4.0 and t.sum(0, 2) == 0.0 and t.sum(2, -1) == 1.0 and t.sum(2, 4) == 4.0
    assert [t.retrieve(x) for x in (0.0, 0.5, 0.99, 1.01, 3.0, 4.0)] == [2, 2, 2, 3, 3, 3]
    assert t.capacity == 4 and len(t.tree) == 8 and t[3] == 3.0
    m = C.MinSegmentTree(4)
    m[0] = 1.0
    m[2] = 0.5
    m[3] = 3.0
    assert m.min() == 0.5 and m.min(0, 2) == 1.0 and m.min(3, 4) == 3.0 and m.tree[1] == 0.5
    with pytest.raises(AssertionError):
        C.SumSegmentTree(6)
    with pytest.raises(AssertionError):
        t.retrieve(-1.0)


# ---- Sampler (reference test_sampler.py:261-394) ---------------------------------------------------------
def test_sampler_dispatch(C):
    per = C.PrioritizedReplayBuffer(max_size=64)
    nst = C.MultiStepReplayBuffer(max_size=64, n_step=1)
    std = C.ReplayBuffer(max_size=64)
    for i in range(8):
        tr = C.Transition(obs=np.full((2, 3), i, np.float32), action=np.array([0, 1]),
                          next_obs=np.full((2, 3), i + 1, np.float32), reward=np.array([1.0, 2.0]),
                          done=np.array([0, 0]), batch_size=[2]).to_tensordict()
        one_step = nst.add(tr)
        per.add(one_step)
        std.add(tr)
    s_per, s_n, s_std = C.Sampler(memory=per), C.Sampler(memory=nst), C.Sampler(memory=std)
    assert s_per.per and not s_per.n_step and s_n.n_step and not s_std.per
    e = s_per.sample(4, 0.4)
    assert set(("weights", "idxs", "obs", "action", "reward", "next_obs", "done")) <= set(e.keys())
    assert e["action"].dtype == torch.float32 and e["reward"].shape == (4, 1)
    n_e = s_n.sample(e["idxs"])
    assert n_e["obs"].shape[0] == 4
    assert torch.equal(n_e["obs"].reshape(4, 3).cpu(), e["obs"].cpu())  # index-aligned buffers
    assert s_std.sample(4, True)["idxs"].shape == (4,)
    with pytest.raises(AssertionError):
        C.Sampler()


# ---- RolloutBuffer (reference test_rollout_buffer.py) -----------------------------------------------
def spaces_mod():
    from agilerl_b200.components import spaces

    return spaces


def make_rollout(C, capacity=8, num_envs=4, obs_dim=5, **kw):
    sp = spaces_mod()
    return C.RolloutBuffer(capacity=capacity, observation_space=sp.Box(-1, 1, (obs_dim,)),
                           action_space=sp.Discrete(3), num_envs=num_envs, device="cpu", **kw)


def fill(buf, steps, seed=0, obs_dim=5):
    rng = np.random.default_rng(seed)
    N = buf.num_envs
    rec = []
    for _ in range(steps):
        step = dict(obs=rng.standard_normal((N, obs_dim)).astype(np.float32), action=rng.integers(0, 3, N),
                    reward=rng.standard_normal(N).astype(np.float32), done=rng.random(N) < 0.1,
                    value=rng.standard_normal(N).astype(np.float32), log_prob=rng.standard_normal(N).astype(np.float32),
                    next_obs=rng.standard_normal((N, obs_dim)).astype(np.float32))
        buf.add(**step)
        rec.append(step)
    return rec


def test_rollout_shapes_and_dtypes(C):
    # :16-139
    sp = spaces_mod()
    b = make_rollout(C, capacity=16, num_envs=4)
    assert b.buffer["observations"].shape == (16, 4, 5) and b.buffer["actions"].shape == (16, 4, 1)
    assert b.buffer["rewards"].shape == (16, 4) and b.buffer["dones"].dtype == torch.bool
    assert b.buffer.get("hidden_states") is None and b.pos == 0 and not b.full and b.size() == 0
    img = C.RolloutBuffer(4, sp.Box(0, 255, (3, 8, 8), np.uint8), sp.Box(-1, 1, (2,)), num_envs=2)
    assert img.buffer["observations"].shape == (4, 2, 3, 8, 8) and img.buffer["observations"].dtype == torch.float32
    assert img.buffer["actions"].shape == (4, 2, 2)
    d = C.RolloutBuffer(4, sp.Dict({"a": sp.Box(-1, 1, (3,)), "b": sp.Discrete(4)}), sp.MultiDiscrete([2, 3]), num_envs=2)
    assert d.buffer["observations"]["a"].shape == (4, 2, 3) and d.buffer["observations"]["b"].shape == (4, 2, 1)
    assert d.buffer["actions"].shape == (4, 2, 2)
    r = C.RolloutBuffer(4, sp.Box(-1, 1, (3,)), sp.Discrete(2), num_envs=2, recurrent=True,
                        hidden_state_architecture={"h": (1, 2, 16)})
    assert r.buffer["hidden_states"]["h"].shape == (4, 2, 1, 16)
    with pytest.raises(ValueError):
        C.RolloutBuffer(4, sp.Box(-1, 1, (3,)), sp.Discrete(2), recurrent=True)


def test_rollout_add_overflow_and_wrap(C):
    # :458-512
    b = make_rollout(C, capacity=3, num_envs=2)
    rec = fill(b, 3)
    assert b.pos == 3 and b.full and b.size() == 6
    assert np.array_equal(b.buffer["observations"].cpu().numpy()[1], rec[1]["obs"])
    assert np.array_equal(b.buffer["dones"].cpu().numpy()[2], rec[2]["done"])
    assert np.array_equal(b.buffer["actions"].cpu().numpy()[0, :, 0], rec[0]["action"].astype(np.float32))
    with pytest.raises(ValueError):
        fill(b, 1)
    w = make_rollout(C, capacity=3, num_envs=2, wrap_at_capacity=True)
    rec = fill(w, 4)
    assert w.pos == 1 and w.full
    assert np.array_equal(w.buffer["rewards"].cpu().numpy()[0], rec[3]["reward"])
    b.reset()
    assert b.pos == 0 and not b.full and b.size() == 0


@pytest.mark.parametrize("seed", [0, 1, 2])
@pytest.mark.parametrize("use_gae", [True, False])
def test_rollout_returns_golden_through_api(C, seed, use_gae):
    g = load_golden(f"gae_s{seed}")
    T, N = g["rewards"].shape
    b = make_rollout(C, capacity=T + 3, num_envs=N, gamma=float(g["gamma"]),
                     gae_lambda=float(g["gae_lambda"]), use_gae=use_gae)
    for t in range(T):
        b.add(obs=np.zeros((N, 5), np.float32), action=np.zeros(N), reward=g["rewards"][t],
              done=g["dones"][t], value=g["values"][t], log_prob=np.zeros(N))
    # last_value may arrive (N,1) (reference test :553-555), tensors or arrays
    b.compute_returns_and_advantages(torch.from_numpy(g["last_value"]).view(N, 1), g["last_done"])
    adv = b.buffer["advantages"][:T].cpu().numpy()
    ret = b.buffer["returns"][:T].cpu().numpy()
    assert np.array_equal(adv, g["adv" if use_gae else "adv_mc"])
    assert np.array_equal(ret, g["ret" if use_gae else "ret_mc"])
    assert float(b.buffer["advantages"][T:].abs().sum()) == 0.0  # rows beyond pos untouched


def test_rollout_gae_properties_like_reference_tests(C):
    # :518-646 property checks (non-zero, finite)
    b = make_rollout(C, capacity=10, num_envs=3)
    fill(b, 10)
    b.compute_returns_and_advantages(np.zeros(3, np.float32), np.zeros(3, bool))
    a = b.buffer["advantages"].cpu().numpy()
    assert np.isfinite(a).all() and np.abs(a).sum() > 0


def test_rollout_get_tensor_batch(C, cuda_device):
    # :652-918
    b = make_rollout(C, capacity=8, num_envs=4)
    assert b.get_tensor_batch().is_empty() and b.get() == {}
    rec = fill(b, 6)
    tb = b.get_tensor_batch(device=cuda_device)
    assert tb.batch_size == torch.Size([24]) and tb["observations"].shape == (24, 5)
    # t-major flattening: flat = t * num_envs + env
    assert np.array_equal(tb["rewards"].cpu().numpy(), np.stack([r["reward"] for r in rec]).reshape(-1))
    assert tb["rewards"].data_ptr() == b.buffer["rewards"].data_ptr()  # resident: a view, no copy
    cpu = b.get_tensor_batch()
    assert cpu["rewards"].device.type == "cpu"
    torch.manual_seed(4)
    want = torch.randperm(24)[:10]
    torch.manual_seed(4)
    mb = b.get_tensor_batch(batch_size=10, device=cuda_device)
    assert mb["observations"].shape == (10, 5)
    assert torch.equal(mb["rewards"].cpu(), tb["rewards"].cpu()[want])
    with pytest.warns(UserWarning):
        allb = b.get_tensor_batch(batch_size=1000)
    assert allb["actions"].shape == (24, 1)
    np.random.seed(0)
    d = b.get(batch_size=7)
    assert d["observations"].shape == (7, 5) and isinstance(d["dones"], np.ndarray)
    assert b.get()["values"].shape == (24,)


def test_rollout_pickle_and_deepcopy(C):
    # :1864-2029 — clone() deep-copies the rollout buffer (core/base.py:499-500)
    b = make_rollout(C, capacity=6, num_envs=2)
    fill(b, 4)
    b.compute_returns_and_advantages(np.ones(2, np.float32), np.zeros(2, bool))
    for other in (pickle.loads(pickle.dumps(b)), copy.deepcopy(b)):
        assert other.pos == 4 and other.capacity == 6 and other.buffer["rewards"].is_cuda
        for k in ("observations", "rewards", "advantages", "returns", "dones"):
            assert torch.equal(other.buffer[k], b.buffer[k])
        other.buffer["rewards"].zero_()
        assert float(b.buffer["rewards"].abs().sum()) > 0


def test_rollout_arena_one_launch_equals_per_agent(C):
    sp = spaces_mod()
    P, T, N = 5, 12, 32
    arena = C.RolloutArena(P, T, sp.Box(-1, 1, (5,)), sp.Discrete(3), num_envs=N)
    solo = [make_rollout(C, capacity=T, num_envs=N) for _ in range(P)]
    for p in range(P):
        rec = fill(arena.buffers[p], T, seed=p)
        fill(solo[p], T, seed=p)
    rng = np.random.default_rng(99)
    lv = rng.standard_normal((P, N)).astype(np.float32)
    ld = rng.random((P, N)) < 0.2
    arena.compute_returns_and_advantages(lv, ld)
    for p in range(P):
        solo[p].compute_returns_and_advantages(lv[p], ld[p])
        assert torch.equal(arena.buffers[p].buffer["advantages"], solo[p].buffer["advantages"])
        assert torch.equal(arena.buffers[p].buffer["returns"], solo[p].buffer["returns"])
        r = solo[p].buffer
        a, g2 = xo.gae(r["rewards"].cpu().numpy(), r["values"].cpu().numpy(), r["dones"].cpu().numpy(),
                       lv[p], ld[p], 0.99, 0.95)
        assert np.array_equal(arena.buffers[p].buffer["advantages"].cpu().numpy(), a)


def test_per_update_device_indices_error_is_deferred_not_lost(C):
    """Device-resident indices are range-checked by the kernel; the AssertionError of
    replay_buffer.py:319 surfaces on the next call (or check_errors()) instead of blocking."""
    b = C.PrioritizedReplayBuffer(max_size=8, alpha=0.6, device="cuda:0")
    b.add(C.TensorBatch({"obs": torch.zeros(4, 2), "reward": torch.zeros(4)}, batch_size=[4]))
    bad = torch.tensor([1, 99], device="cuda:0")
    b.update_priorities(bad, torch.tensor([1.0, 2.0]))
    with pytest.raises(AssertionError):
        b.check_errors()
    b.update_priorities(torch.tensor([1, 2], device="cuda:0"), torch.tensor([1.0, 2.0]))
    b.check_errors()


# ---- truncated-BPTT sequences (reference test_rollout_buffer.py:924-1830) ---------------------------
def _recurrent_buffer(C, T, N, msl, kind="chunked", obs_dim=3, dones=None, seed=0, device="cuda"):
    from agilerl_b200.components import spaces
    from agilerl_b200.components.rollout_buffer import BPTTSequenceType

    g = torch.Generator().manual_seed(seed)
    b = C.RolloutBuffer(capacity=T, observation_space=spaces.Box(-1, 1, (obs_dim,), np.float32),
                        action_space=spaces.Discrete(2), num_envs=N, device=device, recurrent=True,
                        hidden_state_architecture={"h_actor": (1, N, 4), "h_critic": (2, N, 8)},
                        max_seq_len=msl, bptt_sequence_type=BPTTSequenceType(kind))
    data = torch.randn(T, N, obs_dim, generator=g)
    hid = {"h_actor": torch.randn(T, 1, N, 4, generator=g), "h_critic": torch.randn(T, 2, N, 8, generator=g)}
    for t in range(T):
        b.add(data[t].numpy(), np.arange(N) % 2, np.ones(N) * t, dones[t], np.full(N, 0.5), np.full(N, 0.1),
              hidden_state={k: v[t] for k, v in hid.items()})
    b.compute_returns_and_advantages(np.zeros(N), np.ones(N))
    return b, data.numpy(), hid


@pytest.mark.parametrize("seed", range(9))
def test_bptt_prepare_matches_reference_golden(C, seed):
    z = load_golden(f"bptt_s{seed}")
    msl = int(z["max_seq_len"])
    msl = None if msl < 0 else msl
    dones = z["dones"]
    T, N = dones.shape
    b, _, hid = _recurrent_buffer(C, T, N, msl, str(z["seq_type"]), obs_dim=3, dones=dones, seed=seed)
    # overwrite the observations with the golden data field so that `padded` is comparable
    b._leaves[("observations",)].copy_(torch.from_numpy(z["data"]))
    b.prepare_sequence_tensors()
    S, L = int(z["num_sequences"]), z["padded"].shape[1]
    assert b.num_sequences == S and b.max_sequence_length == int(z["max_sequence_length"])
    pd = b.padded_data
    assert pd.batch_size == torch.Size([S * L])
    assert np.array_equal(pd["observations"].cpu().numpy().reshape(S, L, 3), z["padded"])
    assert np.array_equal(pd["pad_mask"].cpu().numpy().reshape(S, L), z["pad_mask"])
    assert pd["pad_mask"].dtype == torch.bool
    assert np.array_equal(np.concatenate([s.numpy() for s in b.unpadded_slices]), z["unpadded"])
    assert [len(s) for s in b.unpadded_slices] == z["lengths"].tolist()
    # initial hidden states = hidden state stored at the first step of every sequence (:818-826)
    first = np.array([int(s[0]) for s in b.unpadded_slices])
    init = pd.get_non_tensor("initial_hidden_states")
    for k, v in hid.items():
        flat = v.permute(0, 2, 1, 3).reshape(T * N, v.shape[1], v.shape[3]).numpy()
        assert np.array_equal(init[k].cpu().numpy(), flat[first]), k
    assert "hidden_states" not in pd
    for k in ("log_probs", "advantages", "values", "returns"):
        assert torch.equal(b.unpadded_data[k], b._leaves[(k,)][:T].reshape(-1))


@pytest.mark.parametrize("kind,msl", [("chunked", 5), ("maximum", 7), ("fifty_percent_overlap", 6), ("chunked", None)])
def test_bptt_prepare_matches_oracle_large(C, kind, msl):
    T, N = 257, 130
    g = torch.Generator().manual_seed(11)
    dones = (torch.rand(T, N, generator=g) < 0.03).numpy()
    b, data, _ = _recurrent_buffer(C, T, N, msl, kind, dones=dones, seed=4)
    b.prepare_sequence_tensors()
    r = xo.bptt_prepare(dones, {"observations": data}, msl, kind)
    S = r["num_sequences"]
    L = r["pad_mask"].shape[1]
    assert b.num_sequences == S and b.max_sequence_length == r["max_sequence_length"]
    assert np.array_equal(b.padded_data["observations"].cpu().numpy().reshape(S, L, -1), r["padded"]["observations"])
    assert np.array_equal(b.padded_data["pad_mask"].cpu().numpy().reshape(S, L), r["pad_mask"])
    assert all(np.array_equal(a.numpy(), e) for a, e in zip(b.unpadded_slices, r["unpadded_slices"]))


def test_bptt_minibatches_cover_every_sequence_once(C):
    # :1065-1147, :1722-1775 — whole sequences, shuffled by torch.randperm on the CPU generator
    T, N, msl = 40, 6, 4
    g = torch.Generator().manual_seed(2)
    dones = (torch.rand(T, N, generator=g) < 0.1).numpy()
    b, data, hid = _recurrent_buffer(C, T, N, msl, "chunked", dones=dones, seed=1)
    b.prepare_sequence_tensors()
    S, L = b.num_sequences, b.max_sequence_length
    torch.manual_seed(77)
    order = torch.randperm(S)
    torch.manual_seed(77)
    seen, n_unp = [], 0
    init_all = b.padded_data.get_non_tensor("initial_hidden_states")
    for i, (padded, unpadded) in enumerate(b.get_minibatch_sequences(5)):
        sel = order[i * 5:(i + 1) * 5]
        k = len(sel)
        assert padded.batch_size == torch.Size([k * L])
        rows = (sel.view(-1, 1) * L + torch.arange(L)).view(-1).cuda()
        assert torch.equal(padded["observations"], b.padded_data["observations"][rows])
        assert torch.equal(padded["pad_mask"], b.padded_data["pad_mask"][rows])
        ih = padded.get_non_tensor("initial_hidden_states")
        assert torch.equal(ih["h_critic"], init_all["h_critic"][sel.cuda()])
        uidx = torch.cat([b.unpadded_slices[int(s)] for s in sel]).cuda()
        for key in ("log_probs", "advantages", "values", "returns"):
            assert torch.equal(unpadded[key], b.unpadded_data[key][uidx])
        assert int(padded["pad_mask"].sum()) == uidx.numel()
        seen += sel.tolist()
        n_unp += uidx.numel()
    assert sorted(seen) == list(range(S))
    with pytest.warns(UserWarning):
        assert len(list(b.get_minibatch_sequences(10 * S))) == 1


def test_bptt_errors_and_helpers(C):
    from agilerl_b200.components import spaces

    nb = C.RolloutBuffer(capacity=4, observation_space=spaces.Box(-1, 1, (2,), np.float32),
                         action_space=spaces.Discrete(2), num_envs=1, device="cuda")
    with pytest.raises(ValueError):
        nb.prepare_sequence_tensors()  # recurrent=False (:737-741)
    dones = np.zeros((6, 2), dtype=bool)
    b, _, _ = _recurrent_buffer(C, 6, 2, 3, dones=dones)
    with pytest.raises(ValueError):
        list(b.get_minibatch_sequences(2))  # before preparation (:1149-1168)
    b.reset()
    with pytest.raises(ValueError):
        b.prepare_sequence_tensors()  # empty buffer
    # _pad_sequences (:1170-1216)
    seqs = [torch.tensor([[1, 2], [3, 4]]), torch.tensor([[5, 6], [7, 8], [9, 10]]), torch.tensor([[11, 12]])]
    assert C.RolloutBuffer._pad_sequences(seqs).shape == (3, 3, 2)
    assert C.RolloutBuffer._pad_sequences(seqs, target_length=5).shape == (3, 5, 2)
    # _get_complete_sequences (:1512-1558)
    out, mx = b._get_complete_sequences(torch.arange(20).reshape(10, 2), [[3, 7, 9], [4, 9]])
    assert mx == 5 and [len(s) for s in out] == [3, 3, 2, 3, 3]
    # host-side data on request (:1466-1507)
    b2, _, _ = _recurrent_buffer(C, 6, 2, 3, dones=dones)
    b2.prepare_sequence_tensors(device="cpu")
    assert b2.padded_data["observations"].device.type == "cpu"
    p, u = next(iter(b2.get_minibatch_sequences(2)))
    assert p["observations"].device.type == "cpu" and u["returns"].device.type == "cpu"


# ---- PPO minibatch slicing (algorithms/ppo.py:838-848): buffer_td[minibatch_indices] --------------------
def test_tensor_batch_row_indexing_by_host_indices_is_one_gather(C):
    from agilerl_b200 import _native as nv
    from agilerl_b200.components import spaces

    T, N = 12, 5
    b = C.RolloutBuffer(capacity=T, observation_space=spaces.Dict({"a": spaces.Box(-1, 1, (3,), np.float32),
                                                                  "b": spaces.Box(-1, 1, (2, 2), np.float32)}),
                        action_space=spaces.Discrete(3), num_envs=N, device="cuda")
    g = torch.Generator().manual_seed(0)
    for t in range(T):
        b.add({"a": torch.randn(N, 3, generator=g).numpy(), "b": torch.randn(N, 2, 2, generator=g).numpy()},
              np.arange(N) % 3, np.ones(N), np.zeros(N), np.full(N, 0.5), np.full(N, 0.1))
    b.compute_returns_and_advantages(np.zeros(N), np.zeros(N))
    td = b.get_tensor_batch(device="cuda")
    adv = td.get("advantages")
    td["advantages"] = (adv - adv.mean()) / (adv.std() + 1e-8)   # :832-836, plain torch on device
    indices = np.arange(T * N)
    np.random.default_rng(1).shuffle(indices)
    lib = nv.load()
    for start in range(0, T * N, 16):
        mb_idx = indices[start:start + 16]
        l0 = lib.xp_launch_count()
        mb = td[mb_idx]
        assert lib.xp_launch_count() - l0 == 1  # every key in one launch
        ref = torch.from_numpy(mb_idx).cuda()
        assert mb.batch_size == torch.Size([len(mb_idx)])
        assert torch.equal(mb["observations"]["b"], td["observations"]["b"][ref])
        assert torch.equal(mb["advantages"], td["advantages"][ref])
        assert torch.equal(mb["dones"], td["dones"][ref]) and mb["dones"].dtype == torch.bool
    # negative / boolean / device indices keep torch's semantics
    assert torch.equal(td[np.array([-1, 0])]["returns"], td["returns"][torch.tensor([-1, 0]).cuda()])
    assert td[torch.tensor([1, 2]).cuda()].batch_size == torch.Size([2])
    with pytest.raises(IndexError):
        td[np.array([T * N])]
