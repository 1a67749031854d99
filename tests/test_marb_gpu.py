"""MultiAgentReplayBuffer on the engine vs the unmodified reference class (golden scenarios from
oracle/make_golden.py::gen_marb) and ports of the reference's own tests
(tests/test_components/test_multi_agent_replay_buffer.py, cited per test)."""

import random

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle.make_golden import flatten_marb, marb_scenario

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def M(cuda_device):
    from agilerl_b200.components import MultiAgentReplayBuffer

    return MultiAgentReplayBuffer


@pytest.mark.parametrize("seed", range(3))
def test_matches_reference_golden(M, seed):
    z = load_golden(f"marb_s{seed}")
    cap, fields, agents, saves, batches = marb_scenario(seed)
    buf = M(cap, fields, agents, device="cuda")
    random.seed(500 + seed)
    k = 0
    for i, (args, vect) in enumerate(saves):
        buf.save_to_memory(*args, is_vectorised=vect)
        if i in (len(saves) // 3, len(saves) - 1):
            for bs in batches:
                if bs <= len(buf):
                    s = flatten_marb(buf.sample(bs), fields, agents)
                    for name, t in s.items():
                        assert t.is_cuda and t.dtype == torch.float32, name
                        assert np.array_equal(t.cpu().numpy(), z[f"s{k}|{name}"], equal_nan=True), (k, name)
                    k += 1
    assert k == int(z["n_samples"]) and len(buf) == int(z["final_len"]) and buf.counter == int(z["counter"])
    # the generator was consumed exactly as random.sample(memory, k) consumes it
    state = random.getstate()
    random.seed(500 + seed)
    for i, (args, vect) in enumerate(saves):
        if i in (len(saves) // 3, len(saves) - 1):
            n = min((i + 1) * (4 if vect else 1), cap)
            for bs in batches:
                if bs <= n:
                    random.sample(range(n), bs)
    assert random.getstate() == state


def test_init_len_and_asserts(M):
    # :10-45
    b = M(100, ["state", "action", "reward"], ["agent1", "agent2"])
    assert len(b) == 0 and b.memory.maxlen == 100 and b.counter == 0 and b.device is None
    assert b.field_names == ["state", "action", "reward"] and b.agent_ids == ["agent1", "agent2"]
    for bad in ((0, ["s"], ["a"]), (10, [], ["a"]), (10, ["s"], [])):
        with pytest.raises(AssertionError):
            M(*bad)
    st = {"agent1": np.array([1, 2, 3]), "agent2": np.array([1, 2, 3])}
    ac = {"agent1": np.array([4, 5]), "agent2": np.array([4, 5])}
    rw = {"agent1": np.array([6]), "agent2": np.array([7])}
    for _ in range(3):
        b.save_to_memory(st, ac, rw)
    assert len(b) == 3 and b.counter == 3


def test_append_and_overwrite_oldest(M):
    # :48-101 deque(maxlen=2): the oldest experience falls out
    b = M(2, ["state", "action", "reward"], ["agent1", "agent2"])
    exps = []
    for i in range(3):
        e = ({"agent1": np.array([i, 2, 3]), "agent2": np.array([i, 2, 3])},
             {"agent1": np.array([4, i]), "agent2": np.array([4, i])},
             {"agent1": np.array([6 + i]), "agent2": np.array([7 + i])})
        exps.append(e)
        b.save_to_memory(*e)
    assert len(b.memory) == 2
    for pos, e in enumerate(exps[1:]):
        got = b.memory[pos]
        for f, want in zip(("state", "action", "reward"), e):
            for a in ("agent1", "agent2"):
                assert np.array_equal(getattr(got, f)[a], want[a])


def test_vectorised_rows_stay_aligned(M):
    # :230-251
    agents = ["agent1", "agent2"]
    b = M(100, ["environments", "state"], agents, "cuda")
    num_envs = 50
    environments = {a: np.array([[e] for e in range(num_envs)]) for a in agents}
    states = {a: np.array([[(e + 1) * (i + 1)] for e in range(num_envs)]) for i, a in enumerate(agents)}
    b.save_to_memory(states, environments, is_vectorised=True)
    assert len(b) == num_envs
    s_states, s_envs = b.sample(batch_size=25)
    for a in agents:
        for i, env in enumerate(s_envs[a]):
            assert states[a][int(env)] == s_states[a][i].cpu().numpy()


def test_sample_shapes_images_and_dicts(M):
    # :253-322
    b = M(100, ["state", "action", "reward"], ["agent1", "agent2"])
    state = {"agent1": {i: np.random.rand(3, 16, 16) for i in range(3)}, "agent2": np.random.rand(3, 16, 16)}
    action = {"agent1": np.array([4, 5]), "agent2": np.array([4, 5])}
    reward = {"agent1": np.array([6]), "agent2": np.array([7])}
    b.save_to_memory(state, action, reward)
    t = b.sample(1)
    assert len(t) == 3
    assert t[0]["agent2"].shape == (1, 3, 16, 16) and t[0]["agent1"][2].shape == (1, 3, 16, 16)
    assert t[1]["agent1"].shape == (1, 2) and t[2]["agent1"].shape == (1, 1)
    with pytest.raises(ValueError):
        b.sample(5)  # Sample larger than population


def test_process_transition_and_stack_helpers(M):
    # :326-636 (host-side helpers)
    b = M(100, ["state", "action", "reward", "next_state", "done"], ["agent1", "agent2"], device="cpu")
    e = b.experience({"agent1": np.array([1, 2, 3]), "agent2": np.array([1, 2, 3])},
                     {"agent1": np.array([4, 5]), "agent2": np.array([4, 5])},
                     {"agent1": np.array([6]), "agent2": np.array([7])},
                     {"agent1": np.array([2, 3, 4]), "agent2": np.array([2, 3, 4])},
                     {"agent1": np.array([1]), "agent2": np.array([0])})
    tr = b._process_transition([e, e, None], np_array=True)
    assert tr["state"]["agent1"].shape == (2, 3) and tr["done"]["agent1"].dtype == np.uint8
    tt = b._process_transition([e])
    assert tt["reward"]["agent2"].dtype == torch.float32 and tt["reward"]["agent2"].shape == (1, 1)
    assert M.stack_transitions([1.0, 2.0]).shape == (2, 1)
    assert M.stack_transitions([(np.zeros(2), 1), (np.ones(2), 2)])[1].shape == (2, 1)
