"""Pins the oracle (oracle/xp_oracle.py) to the reference: golden vectors generated from the
unmodified reference code (oracle/make_golden.py) and the reference's own known-answer tests
(tests/test_components/test_segment_tree.py, test_replay_buffer.py in the AgileRL tree)."""

import numpy as np
import pytest

from conftest import load_golden
from oracle import xp_oracle as xo


# ---- GAE / MC ---------------------------------------------------------------------------------
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_gae_matches_reference_bitwise(seed):
    g = load_golden(f"gae_s{seed}")
    adv, ret = xo.gae(g["rewards"], g["values"], g["dones"], g["last_value"], g["last_done"],
                      float(g["gamma"]), float(g["gae_lambda"]))
    assert np.array_equal(adv, g["adv"]) and np.array_equal(ret, g["ret"])
    adv, ret = xo.mc_returns(g["rewards"], g["values"], g["dones"], g["last_value"],
                             g["last_done"], float(g["gamma"]))
    assert np.array_equal(adv, g["adv_mc"]) and np.array_equal(ret, g["ret_mc"])


@pytest.mark.parametrize("seed", [0, 2])
def test_gae_scalar_restatement_is_the_same_rounding_sequence(seed):
    g = load_golden(f"gae_s{seed}")
    adv, ret = xo.gae_scalar(g["rewards"], g["values"], g["dones"], g["last_value"].astype(float),
                             g["last_done"].astype(float), float(g["gamma"]), float(g["gae_lambda"]))
    assert np.array_equal(adv, g["adv"]) and np.array_equal(ret, g["ret"])


@pytest.mark.parametrize("seed", [0, 1, 2, 3, 4])
def test_gae_f32_matches_reference_loop_bitwise(seed):
    """The all-fp32 loop of ppo.py:693-713 / ippo.py:705-724 (fixtures from the extracted loop)."""
    g = load_golden(f"gae32_s{seed}")
    adv, ret = xo.gae_f32(g["rewards"], g["values"], g["dones"], g["next_value"], g["next_done"],
                          float(g["gamma"]), float(g["gae_lambda"]))
    assert adv.dtype == np.float32
    assert np.array_equal(adv, g["adv"]) and np.array_equal(ret, g["ret"])


# ---- segment tree KATs (reference tests/test_components/test_segment_tree.py:39-129) -----------
def test_tree_set_and_sum_kat():
    t = xo.SegTree(4, "sum")
    t[2] = 1.0
    t[3] = 3.0
    assert np.isclose(t.operate(), 4.0)
    assert np.isclose(t.operate(0, 2), 0.0)
    assert np.isclose(t.operate(0, 3), 1.0)
    assert np.isclose(t.operate(2, 3), 1.0)
    assert np.isclose(t.operate(2, -1), 1.0)
    assert np.isclose(t.operate(2, 4), 4.0)


def test_tree_set_overlap_kat():
    t = xo.SegTree(4, "sum")
    t[2] = 1.0
    t[2] = 3.0
    assert np.isclose(t.operate(), 3.0)
    assert np.isclose(t.operate(2, 3), 3.0)
    assert np.isclose(t.operate(2, -1), 3.0)
    assert np.isclose(t.operate(2, 4), 3.0)
    assert np.isclose(t.operate(1, 2), 0.0)


def test_prefixsum_idx_kat():
    t = xo.SegTree(4, "sum")
    t[2] = 1.0
    t[3] = 3.0
    assert t.retrieve(0.0) == 2
    assert t.retrieve(0.5) == 2
    assert t.retrieve(0.99) == 2
    assert t.retrieve(1.01) == 3
    assert t.retrieve(3.00) == 3
    assert t.retrieve(4.00) == 3


def test_prefixsum_idx2_kat():
    t = xo.SegTree(4, "sum")
    t[0] = 0.5
    t[1] = 1.0
    t[2] = 1.0
    t[3] = 3.0
    assert t.retrieve(0.00) == 0
    assert t.retrieve(0.55) == 1
    assert t.retrieve(0.99) == 1
    assert t.retrieve(1.51) == 2
    assert t.retrieve(3.00) == 3
    assert t.retrieve(5.50) == 3


def test_max_interval_tree_kat():
    t = xo.SegTree(4, "min")
    t[0] = 1.0
    t[2] = 0.5
    t[3] = 3.0
    assert np.isclose(t.operate(), 0.5)
    assert np.isclose(t.operate(0, 2), 1.0)
    assert np.isclose(t.operate(0, 3), 0.5)
    assert np.isclose(t.operate(0, -1), 0.5)
    assert np.isclose(t.operate(2, 4), 0.5)
    assert np.isclose(t.operate(3, 4), 3.0)
    t[2] = 0.7
    assert np.isclose(t.operate(), 0.7)
    assert np.isclose(t.operate(0, 2), 1.0)
    assert np.isclose(t.operate(0, 3), 0.7)
    t[2] = 4.0
    assert np.isclose(t.operate(), 1.0)
    assert np.isclose(t.operate(0, 2), 1.0)
    assert np.isclose(t.operate(2, 3), 4.0)
    assert np.isclose(t.operate(2, -1), 4.0)
    assert np.isclose(t.operate(3, 4), 3.0)


def test_non_pow2_capacity_asserts():
    with pytest.raises(AssertionError):
        xo.SegTree(6, "sum")


# ---- PER scenario ------------------------------------------------------------------------------
def replay_per_scenario(per, g):
    """Drives any PER-like object (oracle or engine adapter) through the golden scenario."""
    n_added = 0
    for step, n in enumerate(g["add_sizes"], start=1):
        per.add(int(n))
        n_added += int(n)
        if step == 3:
            assert n_added == int(g["early_at"])
            per.update_priorities(g["early_idx"], g["early_prio"])
    per.update_priorities(g["upd_idx"], g["upd_prio"])


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_per_oracle_matches_reference(seed):
    g = load_golden(f"per_s{seed}")
    per = xo.PerOracle(int(g["max_size"]), float(g["alpha"]))
    replay_per_scenario(per, g)
    assert per.tree_ptr == int(g["tree_ptr"]) and per.size == int(g["size"])
    assert np.array_equal(np.array(per.sum_tree.tree), g["sum_tree_after_update"])
    assert np.array_equal(np.array(per.min_tree.tree), g["min_tree_after_update"])
    assert per.max_priority == float(g["max_priority"])
    idx = per.sample_proportional(len(g["u"]), g["u"])
    assert np.array_equal(idx, g["sample_idx"])
    w = per.calculate_weights(idx, float(g["beta"]))
    assert np.array_equal(w, g["weights"])


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_array_oracle_matches_reference_golden(seed):
    """The numpy restatements used at C4 size (np_trees_from_leaves / np_sample_proportional /
    np_calculate_weights) against the fixtures written by the unmodified reference."""
    g = load_golden(f"per_s{seed}")
    cap = len(g["sum_tree_after_update"]) // 2
    s_leaves, m_leaves = g["sum_tree_after_update"][cap:], g["min_tree_after_update"][cap:]
    filled = ~np.isinf(m_leaves)  # leaves the reference never set: 0.0 in the sum tree, +inf in the min tree
    n = int(filled.sum())
    assert filled[:n].all() and not filled[n:].any(), "filled leaves are a prefix in the golden scenarios"
    s, m = xo.np_trees_from_leaves(s_leaves[:n], cap)
    assert np.array_equal(s[1:], g["sum_tree_after_update"][1:])
    assert np.array_equal(m[1:], g["min_tree_after_update"][1:])
    idx = xo.np_sample_proportional(s, cap, len(g["u"]), g["u"])
    assert np.array_equal(idx, g["sample_idx"])
    w = xo.np_calculate_weights(s, m, cap, idx, int(g["size"]), float(g["beta"]))
    assert np.array_equal(w, g["weights"])


def test_array_oracle_matches_list_oracle_at_2p16():
    """Pin of the vectorised walk: every index and weight equal to SegTree.retrieve / PerOracle on a
    65 536-leaf tree, including the stratum edges u = 0 and u -> 1."""
    cap, B = 1 << 16, 4096
    rng = np.random.default_rng(3)
    leaves = (np.abs(rng.standard_normal(cap)) + 1e-3) ** 0.6
    per = xo.PerOracle(cap, 0.6)
    per.size = cap
    per.sum_tree.tree[cap:] = leaves.tolist()
    per.min_tree.tree[cap:] = leaves.tolist()
    per.sum_tree.rebuild_from_leaves()
    per.min_tree.rebuild_from_leaves()
    s, m = xo.np_trees_from_leaves(leaves, cap)
    assert np.array_equal(s[1:], np.array(per.sum_tree.tree)[1:]) and np.array_equal(m[1:], np.array(per.min_tree.tree)[1:])
    u = rng.random(B).astype(np.float32)
    u[::64] = 0.0
    u[1::64] = np.float32(1.0) - np.float32(2.0**-24)
    want = per.sample_proportional(B, u)
    got = xo.np_sample_proportional(s, cap, B, u)
    assert np.array_equal(got, want)
    assert np.array_equal(xo.np_calculate_weights(s, m, cap, got, cap, 0.4), per.calculate_weights(want, 0.4))
    ub = np.concatenate([rng.random(2000) * s[1], [0.0, s[1]]])
    assert np.array_equal(xo.np_retrieve(s, cap, ub), np.array([per.sum_tree.retrieve(float(x)) for x in ub]))


def test_bulk_rebuild_equals_sequential_sets():
    rng = np.random.default_rng(0)
    a, b = xo.SegTree(64, "sum"), xo.SegTree(64, "sum")
    vals = rng.random(64)
    for i in rng.permutation(64):
        a[int(i)] = float(vals[i])
    for i in range(64):
        b.tree[64 + i] = float(vals[i])
    b.rebuild_from_leaves()
    assert a.tree == b.tree


# ---- ring + n-step ------------------------------------------------------------------------------
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_ring_and_nstep_oracle_match_reference(seed):
    g = load_golden(f"ring_nstep_s{seed}")
    max_size, N, n_step = int(g["max_size"]), int(g["N"]), int(g["n_step"])
    keys = ["obs", "action", "reward", "next_obs", "done"]

    def step_dict(s):
        d = {k: g[k][s] for k in keys}
        return {k: (v.reshape(N, 1) if v.ndim == 1 else v) for k, v in d.items()}

    ring = {k: np.zeros_like(g[f"ring_{k}"]) for k in keys}
    nst = {k: np.zeros_like(g[f"nstep_{k}"]) for k in keys}
    cur = ncur = 0
    window = []
    for s in range(g["obs"].shape[0]):
        d = step_dict(s)
        cur = xo.ring_insert(ring, d, cur, max_size)
        window.append(d)
        window = window[-n_step:]
        assert (len(window) < n_step) == bool(g["returned_none"][s])
        if len(window) == n_step:
            ncur = xo.ring_insert(nst, xo.nstep_fold(window, float(g["gamma"])), ncur, max_size)
    assert cur == int(g["ring_cursor"]) and ncur == int(g["nstep_cursor"])
    for k in keys:
        assert np.array_equal(ring[k], g[f"ring_{k}"]), k
        assert np.array_equal(nst[k], g[f"nstep_{k}"]), k


def test_nstep_reward_kat():
    # reference tests/test_components/test_replay_buffer.py:532-569: 1 + 0.9*2 + 0.81*3
    steps = [{"reward": np.array([[r]], np.float32), "done": np.zeros((1, 1), np.float32),
              "next_obs": np.full((1, 2), r, np.float32)} for r in (1.0, 2.0, 3.0)]
    out = xo.nstep_fold(steps, 0.9)
    assert np.isclose(out["reward"][0, 0], 1.0 + 0.9 * 2.0 + 0.81 * 3.0)
    assert np.array_equal(out["next_obs"], steps[2]["next_obs"])
    # early termination (:571-628): done at the second transition stops the fold there
    steps[1]["done"][:] = 1.0
    out = xo.nstep_fold(steps, 0.9)
    assert np.isclose(out["reward"][0, 0], 1.0 + 0.9 * 2.0)
    assert np.array_equal(out["next_obs"], steps[1]["next_obs"]) and out["done"][0, 0] == 1.0


# ---- host RNG streams ---------------------------------------------------------------------------
def test_mt19937_stream_is_torch_rand():
    g = load_golden("torch_rng")
    for seed in (0, 1, 123456789):
        raw = xo.mt19937_raw(seed, 700)
        assert np.array_equal(xo.torch_rand_from_raw(raw), g[f"rand_{seed}"])


def test_randperm_prefix_is_torch_randperm():
    g = load_golden("torch_rng")
    for k, (seed, n, B) in enumerate(g["cases"]):
        raw = xo.mt19937_raw(int(seed), int(B))
        assert np.array_equal(xo.randperm_prefix(raw, int(B), int(n)), g[f"perm_{k}"]), (seed, n, B)


# ---- truncated-BPTT sequence preparation ---------------------------------------------------------
@pytest.mark.parametrize("seed", range(9))
def test_bptt_oracle_matches_reference_golden(seed):
    z = load_golden(f"bptt_s{seed}")
    msl = int(z["max_seq_len"])
    msl = None if msl < 0 else msl
    r = xo.bptt_prepare(z["dones"], {"x": z["data"]}, msl, str(z["seq_type"]))
    assert r["num_sequences"] == int(z["num_sequences"])
    assert r["max_sequence_length"] == int(z["max_sequence_length"])
    assert np.array_equal(np.concatenate(r["unpadded_slices"]), z["unpadded"])
    assert np.array_equal(np.array([len(s) for s in r["unpadded_slices"]]), z["lengths"])
    assert np.array_equal(r["pad_mask"], z["pad_mask"])
    assert np.array_equal(r["padded"]["x"], z["padded"])


def test_bptt_reference_kats():
    # tests/test_components/test_rollout_buffer.py:1218-1270: one 12-step episode, max_seq_len 4
    d = np.zeros((12, 1), dtype=bool)
    d[11] = True
    assert len(xo.bptt_sequences(d, 4, "chunked")[0]) == 12 // 4
    assert len(xo.bptt_sequences(d, 4, "maximum")[0]) == 12 - 4 + 1
    # chunking drops the remainder of an episode (algo_utils.py:243-248)
    d = np.zeros((10, 1), dtype=bool)
    assert [s[1:] for s in xo.bptt_sequences(d, 4, "chunked")[0]] == [(0, 4), (4, 4)]
    with pytest.raises(ValueError):
        xo.bptt_sequences(d, 1, "fifty_percent_overlap")


# ---- MultiAgentReplayBuffer ----------------------------------------------------------------------------
@pytest.mark.parametrize("seed", range(3))
def test_marb_oracle_matches_reference_golden(seed):
    import random

    from oracle.make_golden import flatten_marb, marb_scenario

    z = load_golden(f"marb_s{seed}")
    cap, fields, agents, saves, batches = marb_scenario(seed)
    buf = xo.MarbOracle(cap, fields, agents)
    random.seed(500 + seed)
    k = 0
    for i, (args, vect) in enumerate(saves):
        buf.save(*args, is_vectorised=vect)
        if i in (len(saves) // 3, len(saves) - 1):
            for bs in batches:
                if bs <= len(buf):
                    s = flatten_marb(buf.sample(bs), fields, agents)
                    for name, t in s.items():
                        assert t.dtype == np.float32
                        assert np.array_equal(t, z[f"s{k}|{name}"], equal_nan=True), (k, name)
                    k += 1
    assert k == int(z["n_samples"]) and len(buf) == int(z["final_len"]) and buf.counter == int(z["counter"])


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_llm_gae_oracle_matches_reference_methods(seed):
    """oracle gae_tokens / gae_turns vs PPO._compute_gae_returns_token / _compute_gae_returns
    (algorithms/ppo_llm.py:758-872) run unmodified: the token chain bit for bit; whatever sits behind
    a torch `.sum()` (whitening, mean pooling) within rtol 2e-5."""
    g = load_golden(f"llm_gae_s{seed}")
    gamma, lam = float(g["gamma"]), float(g["gae_lambda"])
    ret, adv = xo.gae_tokens(g["rewards"], g["values"], g["mask"], gamma, lam, False)
    assert np.array_equal(ret, g["tok_ret_w0"]) and np.array_equal(adv, g["tok_adv_w0"])
    ret, adv = xo.gae_tokens(g["rewards"], g["values"], g["mask"], gamma, lam, True)
    assert np.array_equal(ret, g["tok_ret_w1"])
    np.testing.assert_allclose(adv, g["tok_adv_w1"], rtol=2e-5, atol=1e-6)
    for red in ("final_value", "mean"):
        for w in (0, 1):
            ret, adv = xo.gae_turns(g["rewards"], g["values"], g["turn_ids"], gamma, lam, bool(w), red)
            np.testing.assert_allclose(ret, g[f"turn_ret_{red}_w{w}"], rtol=2e-5, atol=1e-6)
            np.testing.assert_allclose(adv, g[f"turn_adv_{red}_w{w}"], rtol=2e-5, atol=1e-6)


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_normalize_advantages_oracle_matches_reference_statement(seed):
    g = load_golden(f"advnorm_s{seed}")
    np.testing.assert_allclose(xo.normalize_advantages(g["advantages"]), g["normalized"], rtol=2e-5, atol=1e-6)
