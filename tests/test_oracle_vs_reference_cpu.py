"""Oracle against the UNMODIFIED reference, run live on fresh random inputs (beyond the committed
fixtures).  Only runs where the reference tree exists (the build container); on the GPU box, where
/root/reference is absent, every test here is skipped and the golden fixtures carry the pin."""

import numpy as np
import pytest
import torch

from oracle import ref_loader
from oracle import xp_oracle as xo

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")


@pytest.mark.parametrize("seed", range(6))
def test_gae_and_mc_live(seed):
    rng = np.random.default_rng(seed)
    T, N = int(rng.integers(1, 90)), int(rng.integers(1, 40))
    r = rng.standard_normal((T, N)).astype(np.float32)
    v = rng.standard_normal((T, N)).astype(np.float32)
    d = rng.random((T, N)) < 0.1
    lv = rng.standard_normal(N).astype(np.float32)
    ld = rng.random(N) < 0.3
    gamma, lam = float(rng.uniform(0.9, 1.0)), float(rng.uniform(0.8, 1.0))
    adv, ret = ref_loader.reference_gae(r, v, d, lv, ld, gamma, lam, use_gae=True)
    a, g = xo.gae(r, v, d, lv, ld, gamma, lam)
    assert np.array_equal(a, adv) and np.array_equal(g, ret)
    adv, ret = ref_loader.reference_gae(r, v, d, lv, ld, gamma, lam, use_gae=False)
    a, g = xo.mc_returns(r, v, d, lv, ld, gamma)
    assert np.array_equal(a, adv) and np.array_equal(g, ret)


@pytest.mark.parametrize("seed", range(6))
@pytest.mark.parametrize("site", ["ppo", "ippo"])
def test_gae_f32_live(seed, site):
    rng = np.random.default_rng(50 + seed)
    T, N = int(rng.integers(1, 90)), int(rng.integers(1, 40))
    r = rng.standard_normal((T, N)).astype(np.float32)
    v = rng.standard_normal((T, N)).astype(np.float32)
    d = rng.random((T, N)) < 0.1
    nv = rng.standard_normal(N).astype(np.float32)
    nd = (rng.random(N) < 0.3).astype(np.float32)
    gamma, lam = float(rng.uniform(0.9, 1.0)), float(rng.uniform(0.8, 1.0))
    adv, ret = ref_loader.reference_gae_f32(r, v, d, nv, nd, gamma, lam, site=site)
    a, g = xo.gae_f32(r, v, d, nv, nd, gamma, lam)
    assert np.array_equal(a, adv) and np.array_equal(g, ret)


@pytest.mark.parametrize("seed", range(4))
@pytest.mark.parametrize("seq_type", ["chunked", "maximum", "fifty_percent_overlap"])
def test_bptt_live(seed, seq_type):
    rng = np.random.default_rng(200 + seed)
    T, N = int(rng.integers(4, 40)), int(rng.integers(1, 6))
    msl = int(rng.integers(2, 9))
    d = rng.random((T, N)) < 0.15
    x = rng.standard_normal((T, N, 3)).astype(np.float32)
    slices, padded, mask, max_len, n_seq = ref_loader.reference_bptt(d, x, msl, seq_type)
    r = xo.bptt_prepare(d, {"x": x}, msl, seq_type)
    assert r["num_sequences"] == n_seq and r["max_sequence_length"] == max_len
    assert len(slices) == len(r["unpadded_slices"])
    for s, o in zip(slices, r["unpadded_slices"]):
        assert np.array_equal(s, o)
    assert np.array_equal(r["pad_mask"], mask) and np.array_equal(r["padded"]["x"], padded)


def test_array_oracle_vs_unmodified_reference_live():
    """The numpy restatement that serves as the C4-size oracle (tests/test_kernels_gpu.py) against the
    UNMODIFIED reference classes on a 2^14-leaf tree filled through the reference's own
    _update_priority calls: tree arrays, sampled indices (same torch seed) and IS weights."""
    ref = ref_loader.load_reference()
    cap, B = 1 << 14, 512
    rng = np.random.default_rng(17)
    buf = ref.PrioritizedReplayBuffer(max_size=cap, alpha=0.6)
    buf._size = cap
    pri = (np.abs(rng.standard_normal(cap)) + 1e-3).astype(np.float32)
    for i, p in enumerate(pri):
        buf._update_priority(i, float(p))
    s, m = xo.np_trees_from_leaves(np.array([float(p) ** 0.6 for p in pri]), cap)
    assert np.array_equal(s[1:], np.array(buf.sum_tree.tree)[1:]) and np.array_equal(m[1:], np.array(buf.min_tree.tree)[1:])
    torch.manual_seed(5)
    u = torch.rand(B).numpy()
    torch.manual_seed(5)
    idx = buf._sample_proportional(B)
    got = xo.np_sample_proportional(s, cap, B, u)
    assert np.array_equal(got, idx.numpy())
    assert np.array_equal(xo.np_calculate_weights(s, m, cap, got, cap, 0.4), buf._calculate_weights(idx, 0.4).numpy())


@pytest.mark.parametrize("seed", range(3))
def test_per_trees_sampling_weights_live(seed):
    ref = ref_loader.load_reference()
    g = torch.Generator().manual_seed(300 + seed)
    max_size = [64, 200, 512][seed]
    B = 32
    buf = ref.PrioritizedReplayBuffer(max_size=max_size, alpha=0.6)
    orc = xo.PerOracle(max_size, 0.6)
    added = 0
    while added < max_size + 20:
        n = int(torch.randint(1, 20, (1,), generator=g))
        buf.add(ref.TensorDict({"obs": torch.randn(n, 2, generator=g)}, batch_size=[n]))
        orc.add(n)
        added += n
        if added % 3 == 0:
            idx = torch.randint(0, buf.size, (8,), generator=g)
            pr = (torch.randn(8, generator=g).abs() * 2 + 1e-7).numpy()
            buf.update_priorities(idx, pr)
            orc.update_priorities(idx.numpy(), pr)
    assert np.array_equal(np.array(buf.sum_tree.tree), np.array(orc.sum_tree.tree))
    assert np.array_equal(np.array(buf.min_tree.tree), np.array(orc.min_tree.tree))
    assert buf.max_priority == orc.max_priority and buf.tree_ptr == orc.tree_ptr
    torch.manual_seed(9 + seed)
    u = torch.rand(B)
    torch.manual_seed(9 + seed)
    idx = buf._sample_proportional(B)
    w = buf._calculate_weights(idx, 0.5)
    oi = orc.sample_proportional(B, u.numpy())
    assert np.array_equal(idx.numpy().reshape(-1), oi)
    assert np.array_equal(w.numpy().reshape(-1), orc.calculate_weights(oi, 0.5))


@pytest.mark.parametrize("seed", [11, 12, 13])
def test_llm_gae_live(seed):
    """Fresh random cases through the reference's own token- / turn-level GAE methods."""
    rng = np.random.default_rng(seed)
    B, S, K = int(rng.integers(1, 40)), int(rng.integers(2, 300)), int(rng.integers(1, 6))
    tid = rng.integers(-1, K, (B, S))
    tid[rng.random((B, S)) < 0.4] = -1
    tid[0, 0] = K - 1
    mask = tid >= 0
    r = (rng.standard_normal((B, S)) * mask).astype(np.float32)
    v = (rng.standard_normal((B, S)) * mask).astype(np.float32)
    want = ref_loader.reference_llm_gae(r, v, mask, 0.99, 0.95, False)
    got = xo.gae_tokens(r, v, mask, 0.99, 0.95, False)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    for red in ("final_value", "mean"):
        want = ref_loader.reference_llm_gae(r, v, mask, 0.99, 0.95, True, turn_ids=tid, turn_value_reduction=red)
        got = xo.gae_turns(r, v, tid, 0.99, 0.95, True, red)
        np.testing.assert_allclose(got[0], want[0], rtol=2e-5, atol=1e-6)
        np.testing.assert_allclose(got[1], want[1], rtol=2e-5, atol=1e-6)
