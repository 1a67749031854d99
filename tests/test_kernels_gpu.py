"""Kernel-level parity (through the C-ABI) against the oracle and the committed golden vectors.

Bars: integer / index / tree work bit-exact; GAE and n-step bit-exact (stricter than the 1e-6
relative the north star asks); IS weights bit-exact in fp32 (device pow is the glibc algorithm)."""

import math

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import xp_oracle as xo

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def xp(cuda_device):
    from agilerl_b200 import xp as _xp

    return _xp


def dev(a, device, dtype=None):
    t = torch.as_tensor(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.to(device).contiguous()


# ---- GAE ----------------------------------------------------------------------------------------
def run_gae(xp, device, r, v, d, lv, ld, gamma, lam, use_gae=True):
    """r, v, d: [P,T,N] numpy; lv, ld: [P,N]."""
    P, T, N = r.shape
    R, V, D = dev(r, device), dev(v, device), dev(d, device, torch.bool)
    LV, LD = dev(lv, device, torch.float64), dev(ld, device, torch.float64)
    A, G = torch.full_like(R, float("nan")), torch.full_like(R, float("nan"))
    xp.gae(R, V, D, LV, LD, A, G, T, N, P, T * N, gamma, lam, use_gae)
    torch.cuda.synchronize()
    return A.cpu().numpy(), G.cpu().numpy()


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_gae_golden_bitwise(xp, cuda_device, seed):
    g = load_golden(f"gae_s{seed}")
    args = (g["rewards"][None], g["values"][None], g["dones"][None], g["last_value"][None],
            g["last_done"][None], float(g["gamma"]), float(g["gae_lambda"]))
    adv, ret = run_gae(xp, cuda_device, *args)
    assert np.array_equal(adv[0], g["adv"]) and np.array_equal(ret[0], g["ret"])
    adv, ret = run_gae(xp, cuda_device, *args, use_gae=False)
    assert np.array_equal(adv[0], g["adv_mc"]) and np.array_equal(ret[0], g["ret_mc"])


@pytest.mark.parametrize("kernel", ["", "tma", "spec", "tile"])
@pytest.mark.parametrize("shape", [(3, 50, 96), (2, 1, 32), (1, 7, 1), (4, 130, 64), (2, 33, 40), (1, 300, 33), (2, 257, 132),
                                   (2, 128, 64), (1, 144, 32), (1, 272, 48)])
@pytest.mark.parametrize("use_gae", [True, False])
def test_gae_population_vs_oracle(xp, cuda_device, shape, use_gae, kernel, monkeypatch):
    """Every kernel behind xp_gae on every shape (the knob is read per call; "" = the size-based default):
    gae_tma_kernel (one warp per strip; look-ahead stage recycling for T > 128), gae_spec_kernel (warp-specialised; aligned blocks for T % 16 == 0, the
    per-row path otherwise; stage recycling for T > 128), gae_tile_kernel (any alignment)."""
    monkeypatch.setenv("XP_GAE_KERNEL", kernel)
    P, T, N = shape
    rng = np.random.default_rng(P * 1000 + T)
    r = rng.standard_normal((P, T, N)).astype(np.float32)
    v = rng.standard_normal((P, T, N)).astype(np.float32)
    d = rng.random((P, T, N)) < 0.05
    lv = rng.standard_normal((P, N)).astype(np.float32)
    ld = rng.random((P, N)) < 0.1
    adv, ret = run_gae(xp, cuda_device, r, v, d, lv, ld, 0.99, 0.95, use_gae)
    for p in range(P):
        if use_gae:
            a, g = xo.gae(r[p], v[p], d[p], lv[p], ld[p], 0.99, 0.95)
        else:
            a, g = xo.mc_returns(r[p], v[p], d[p], lv[p], ld[p], 0.99)
        assert np.array_equal(adv[p], a) and np.array_equal(ret[p], g)


@pytest.mark.parametrize("shape", [(3, 128, 96), (2, 50, 64), (1, 300, 48), (16, 128, 1024)])
@pytest.mark.parametrize("use_gae", [True, False])
def test_gae_paired_tensor_maps(xp, cuda_device, shape, use_gae, monkeypatch):
    """gae_tma_kernel moves rewards+values (and advantages+returns) of a stage with ONE rank-4 bulk tensor copy when
    the second array lies a positive multiple of 16 bytes behind the first.  Inputs carved from one slab in both
    orders (values behind rewards: paired; rewards behind values: five rank-3 maps), and the paired layout again
    with XP_GAE_PAIR=0: all three bit-equal to the oracle (one agent) and to each other (all agents)."""
    monkeypatch.setenv("XP_GAE_KERNEL", "tma")
    P, T, N = shape
    rng = np.random.default_rng(P * 7 + T)
    r = rng.standard_normal((P, T, N)).astype(np.float32)
    v = rng.standard_normal((P, T, N)).astype(np.float32)
    d = rng.random((P, T, N)) < 0.05
    lv = rng.standard_normal((P, N)).astype(np.float32)
    ld = rng.random((P, N)) < 0.1
    n = P * T * N
    outs = []
    for order, pair in (("rv", "1"), ("vr", "1"), ("rv", "0")):
        monkeypatch.setenv("XP_GAE_PAIR", pair)
        slab_in = torch.empty(2 * n + 8, dtype=torch.float32, device=cuda_device)
        slab_out = torch.full((2 * n + 8,), float("nan"), dtype=torch.float32, device=cuda_device)
        a, b = slab_in[:n].view(P, T, N), slab_in[n + 4: 2 * n + 4].view(P, T, N)   # 16 bytes of padding between
        oa, ob = slab_out[:n].view(P, T, N), slab_out[n + 4: 2 * n + 4].view(P, T, N)
        R, V = (a, b) if order == "rv" else (b, a)
        A, G = (oa, ob) if order == "rv" else (ob, oa)
        R.copy_(torch.from_numpy(r)), V.copy_(torch.from_numpy(v))
        D = torch.from_numpy(d).to(cuda_device)
        LV, LD = dev(lv, cuda_device, torch.float64), dev(ld, cuda_device, torch.float64)
        xp.gae(R, V, D, LV, LD, A, G, T, N, P, T * N, 0.99, 0.95, use_gae)
        torch.cuda.synchronize()
        outs.append((A.cpu().numpy().copy(), G.cpu().numpy().copy()))
        assert torch.isnan(slab_out[n: n + 4]).all() and torch.isnan(slab_out[2 * n + 4:]).all()  # nothing spilled
    for a_, g_ in outs[1:]:
        assert np.array_equal(a_, outs[0][0]) and np.array_equal(g_, outs[0][1])
    p_ = P - 1
    if use_gae:
        ea, eg = xo.gae(r[p_], v[p_], d[p_], lv[p_], ld[p_], 0.99, 0.95)
    else:
        ea, eg = xo.mc_returns(r[p_], v[p_], d[p_], lv[p_], ld[p_], 0.99)
    assert np.array_equal(outs[0][0][p_], ea) and np.array_equal(outs[0][1][p_], eg)


def test_gae_empty_is_noop(xp, cuda_device):
    z = torch.empty(0, device=cuda_device)
    zd = torch.empty(0, device=cuda_device, dtype=torch.float64)
    xp.gae(z, z, z.bool(), zd, zd, z, z, 0, 4, 1, 0, 0.99, 0.95)


def test_gae_full_size_properties(xp, cuda_device):
    """C3 size (1024 envs x 128 steps x 16 agents): linearity-free properties that hold exactly:
    with gamma = 0 the advantage is r - v; with done everywhere episodes do not leak."""
    P, T, N = 16, 128, 1024
    g = torch.Generator(device="cpu").manual_seed(0)
    r = torch.randn(P, T, N, generator=g).numpy()
    v = torch.randn(P, T, N, generator=g).numpy()
    d = np.ones((P, T, N), dtype=bool)
    lv = torch.randn(P, N, generator=g).numpy()
    ld = np.ones((P, N), dtype=bool)
    adv, ret = run_gae(xp, cuda_device, r, v, d, lv, ld, 0.99, 0.95)
    assert np.array_equal(adv, (r.astype(np.float64) - v.astype(np.float64)).astype(np.float32))
    # and one agent against the oracle at the full size
    d = np.random.default_rng(1).random((P, T, N)) < 0.02
    ld = np.zeros((P, N), dtype=bool)
    adv, ret = run_gae(xp, cuda_device, r, v, d, lv, ld, 0.99, 0.95)
    a, g2 = xo.gae(r[5], v[5], d[5], lv[5], ld[5], 0.99, 0.95)
    assert np.array_equal(adv[5], a) and np.array_equal(ret[5], g2)


# ---- all-fp32 GAE of the legacy experiences path (ppo.py:686-713, ippo.py:703-724) ------------------
@pytest.mark.parametrize("seed", [0, 1, 2, 3, 4])
def test_gae_f32_golden_bitwise(cuda_device, seed):
    from agilerl_b200.returns import gae_from_experiences

    g = load_golden(f"gae32_s{seed}")
    adv, ret = gae_from_experiences(g["rewards"], g["values"], g["dones"], g["next_value"], g["next_done"],
                                    float(g["gamma"]), float(g["gae_lambda"]))
    assert adv.is_cuda and adv.dtype == torch.float32 and tuple(adv.shape) == g["adv"].shape
    assert np.array_equal(adv.cpu().numpy(), g["adv"]) and np.array_equal(ret.cpu().numpy(), g["ret"])


@pytest.mark.parametrize("shape", [(1, 1, 1), (3, 50, 96), (2, 9, 33), (16, 128, 1024), (1, 300, 7)])
def test_gae_f32_population_vs_oracle(cuda_device, shape):
    from agilerl_b200.returns import gae_from_experiences

    P, T, N = shape
    rng = np.random.default_rng(P * 77 + T)
    r = rng.standard_normal((P, T, N)).astype(np.float32)
    v = rng.standard_normal((P, T, N)).astype(np.float32)
    d = rng.random((P, T, N)) < 0.05
    nv = rng.standard_normal((P, N)).astype(np.float32)
    nd = (rng.random((P, N)) < 0.2).astype(np.float32)
    adv, ret = gae_from_experiences(r, v, d.astype(np.int64), nv, nd, 0.99, 0.95)
    adv, ret = adv.cpu().numpy(), ret.cpu().numpy()
    for p in range(0, P, max(1, P // 3)):
        a, g = xo.gae_f32(r[p], v[p], d[p], nv[p], nd[p], 0.99, 0.95)
        assert np.array_equal(adv[p], a) and np.array_equal(ret[p], g)


def test_gae_f32_single_env_and_errors(cuda_device):
    from agilerl_b200.returns import gae_from_experiences

    rng = np.random.default_rng(0)
    r, v = rng.standard_normal(40).astype(np.float32), rng.standard_normal(40).astype(np.float32)
    d = rng.random(40) < 0.1
    adv, ret = gae_from_experiences(r, v, d, np.float32(0.3), np.float32(0.0), 0.9, 0.8)
    a, g = xo.gae_f32(r[:, None], v[:, None], d[:, None], [0.3], [0.0], 0.9, 0.8)
    assert adv.shape == (40,) and np.array_equal(adv.cpu().numpy(), a[:, 0]) and np.array_equal(ret.cpu().numpy(), g[:, 0])
    with pytest.raises(ValueError):
        gae_from_experiences(r, v[:-1], d, 0.3, 0.0, 0.9, 0.8)
    with pytest.raises(ValueError):
        gae_from_experiences(r[:, None], v[:, None], d[:, None], [0.3, 0.1], [0.0], 0.9, 0.8)
    e = torch.empty(0, 4)
    adv, ret = gae_from_experiences(e, e, e, torch.zeros(4), torch.zeros(4), 0.9, 0.8)
    assert adv.shape == (0, 4)


# ---- trees --------------------------------------------------------------------------------------
class DevTree:
    """(sum,min) interleaved device tree driven through the C-ABI."""

    def __init__(self, xp, device, cap, interleaved=True):
        self.xp, self.device, self.cap = xp, device, cap
        if interleaved:
            self.nodes = torch.empty(2 * cap, 2, dtype=torch.float64, device=device)
            self.sum, self.min, self.stride = self.nodes[:, 0], self.nodes[:, 1], 2
        else:
            self.sum = torch.empty(2 * cap, dtype=torch.float64, device=device)
            self.min = torch.empty(2 * cap, dtype=torch.float64, device=device)
            self.stride = 1
        self.aux = torch.ones(2 * cap, dtype=torch.int32, device=device)  # init must zero it
        self.t = xp.make_tree(self.sum, self.min, self.stride, cap, self.aux)
        xp.tree_init(self.t, torch.cuda.current_stream(device).cuda_stream)
        self.err = torch.zeros(1, dtype=torch.int32, device=device)

    def set(self, idx, val):
        self.xp.tree_set_batch(self.t, dev(idx, self.device, torch.int64),
                               dev(val, self.device, torch.float64), self.err)

    def query(self, start=0, end=0):
        out = torch.zeros(2, dtype=torch.float64, device=self.device)
        self.xp.tree_query(self.t, start, end, out)
        return out.cpu().numpy()

    def arrays(self):
        torch.cuda.synchronize()
        return self.sum.cpu().numpy().copy(), self.min.cpu().numpy().copy()


def test_tree_kats(xp, cuda_device):
    # reference tests/test_components/test_segment_tree.py:39-129
    t = DevTree(xp, cuda_device, 4)
    t.set([2, 3], [1.0, 3.0])
    assert t.query()[0] == 4.0 and t.query(0, 2)[0] == 0.0 and t.query(0, 3)[0] == 1.0
    assert t.query(2, 3)[0] == 1.0 and t.query(2, -1)[0] == 1.0 and t.query(2, 4)[0] == 4.0
    r = xp.tree_retrieve(t.t, dev([0.0, 0.5, 0.99, 1.01, 3.0, 4.0], cuda_device, torch.float64))
    assert r.cpu().tolist() == [2, 2, 2, 3, 3, 3]
    t.set([2, 2], [1.0, 3.0])  # overwrite inside one batch: last writer wins
    assert t.query()[0] == 6.0 and t.query(2, 3)[0] == 3.0 and t.query(1, 2)[0] == 0.0
    m = DevTree(xp, cuda_device, 4)
    m.set([0, 2, 3], [1.0, 0.5, 3.0])
    assert m.query()[1] == 0.5 and m.query(0, 2)[1] == 1.0 and m.query(3, 4)[1] == 3.0
    m.set([2], [4.0])
    assert m.query()[1] == 1.0 and m.query(2, 3)[1] == 4.0 and m.query(2, -1)[1] == 4.0


def test_tree_non_pow2_capacity_is_assertion(xp, cuda_device):
    bad = xp.make_tree(torch.zeros(12, dtype=torch.float64, device=cuda_device), None, 1, 6, None)
    with pytest.raises(AssertionError):
        xp.tree_init(bad, 0)


# schedules of tree_set.cuh: M (B <= 16384, one launch), L (bucketed, large B), D (dense rebuild)
@pytest.mark.parametrize("cap,B,interleaved", [(1, 3, True), (2, 5, False), (64, 40, True),
                                               (4096, 256, True), (4096, 5000, False), (1 << 16, 20000, True),
                                               (1 << 14, 700, False), (1 << 16, 1500, True), (1 << 16, 2048, True),
                                               (1 << 18, 3000, True), (1 << 12, 1 << 12, True), (1 << 14, 20000, True),
                                               (1 << 18, 8192, False), (1 << 18, 30000, True), (1 << 20, 100000, True),
                                               (1 << 18, 16384, True), (1 << 20, 12000, True), (1 << 18, 16385, True)])
def test_tree_set_batch_equals_sequential_reference(xp, cuda_device, cap, B, interleaved):
    rng = np.random.default_rng(cap + B)
    t = DevTree(xp, cuda_device, cap, interleaved)
    s, m = xo.SegTree(cap, "sum"), xo.SegTree(cap, "min")
    for rnd in range(3):
        idx = rng.integers(0, cap, B)
        if B > 4:
            idx[B // 2:] = idx[: B - B // 2]  # forced duplicates
        val = rng.random(B) * 10.0 + 1e-3
        t.set(idx, val)
        for i, x in zip(idx, val):
            s[int(i)] = float(x)
            m[int(i)] = float(x)
        ds, dm = t.arrays()
        assert np.array_equal(ds, np.array(s.tree)) and np.array_equal(dm, np.array(m.tree))
        assert int(t.aux[cap:].abs().sum()) == 0  # last-writer tickets consumed
    # descents: bit-exact indices for random upper bounds, incl. both ends
    total = s.operate()
    ub = np.concatenate([rng.random(512) * total, [0.0, total]])
    got = xp.tree_retrieve(t.t, dev(ub, cuda_device, torch.float64)).cpu().numpy()
    assert np.array_equal(got, np.array([s.retrieve(float(x)) for x in ub]))
    # range queries in the reference's association order
    for _ in range(10):
        a, b = sorted(rng.integers(0, cap + 1, 2))
        if a == b:
            continue
        q = t.query(int(a), int(b))
        assert q[0] == s.operate(int(a), int(b)) and q[1] == m.operate(int(a), int(b))


@pytest.mark.parametrize("pattern", ["sorted", "one_bucket", "one_leaf", "two_leaves_far"])
@pytest.mark.parametrize("cap,B", [(1 << 10, 300), (1 << 16, 2000), (1 << 16, 9000), (1 << 18, 40000)])
def test_tree_set_batch_adversarial_patterns(xp, cuda_device, pattern, cap, B):
    """sorted input (the PER fast path), every update in one bucket subtree (oversize bucket:
    several passes), every update on one leaf (one long duplicate run), two far leaves."""
    rng = np.random.default_rng(cap + B)
    t = DevTree(xp, cuda_device, cap, True)
    s, m = xo.SegTree(cap, "sum"), xo.SegTree(cap, "min")
    base = rng.integers(0, cap, cap // 2)
    bv = rng.random(cap // 2) + 0.5
    t.set(base, bv)
    for i, x in zip(base, bv):
        s[int(i)] = float(x)
        m[int(i)] = float(x)
    if pattern == "sorted":
        idx = np.sort(rng.integers(0, cap, B))
    elif pattern == "one_bucket":
        idx = rng.integers(cap // 2, cap // 2 + 16, B)
    elif pattern == "one_leaf":
        idx = np.full(B, cap - 1)
    else:
        idx = np.where(rng.random(B) < 0.5, 0, cap - 1)
    val = rng.random(B) * 3.0 + 1e-3
    t.set(idx, val)
    for i, x in zip(idx, val):
        s[int(i)] = float(x)
        m[int(i)] = float(x)
    ds, dm = t.arrays()
    assert np.array_equal(ds, np.array(s.tree)) and np.array_equal(dm, np.array(m.tree))
    assert int(t.aux[cap:].abs().sum()) == 0


def test_tree_set_batch_sizes_may_change_between_calls(xp, cuda_device):
    """The bucket scratch of a large batch is laid out by the batch size; a later call with another
    size (other bucket count, one-launch path in between) must still find its counters zero."""
    cap = 1 << 18
    rng = np.random.default_rng(5)
    tree = DevTree(xp, cuda_device, cap)
    ref_s = np.zeros(2 * cap)
    ref_m = np.full(2 * cap, np.inf)
    for B in (20000, 200000, 3000, 70000, 20000, 262144):
        idx = rng.integers(0, cap, B)
        val = rng.random(B) + 0.01
        tree.set(idx, val)
        # sequential last-writer-wins leaves, then the bottom-up rebuild the invariant implies
        ref_s[cap + idx] = 0  # placeholder write so that fancy assignment below is last-wins
        for arr in (ref_s, ref_m):
            arr[cap + idx] = val  # numpy fancy assignment keeps the LAST value of duplicate indices
        for lvl in range(17, -1, -1):
            lo, hi = 1 << lvl, 2 << lvl
            ref_s[lo:hi] = ref_s[2 * lo:2 * hi:2] + ref_s[2 * lo + 1:2 * hi:2]
            ref_m[lo:hi] = np.minimum(ref_m[2 * lo:2 * hi:2], ref_m[2 * lo + 1:2 * hi:2])
        s_, m_ = tree.arrays()
        assert np.array_equal(s_[1:], ref_s[1:]) and np.array_equal(m_[1:], ref_m[1:]), B
        assert int(tree.err.item()) == 0


def _rebuild(leaves_s, leaves_m, cap):
    """Bottom-up rebuild in the tree's own pairwise order (what segment_tree.py:89-95 leaves behind)."""
    s = np.zeros(2 * cap)
    m = np.full(2 * cap, np.inf)
    s[cap:], m[cap:] = leaves_s, leaves_m
    lvl = cap >> 1
    while lvl >= 1:
        s[lvl:2 * lvl] = s[2 * lvl:4 * lvl:2] + s[2 * lvl + 1:4 * lvl:2]
        m[lvl:2 * lvl] = np.minimum(m[2 * lvl:4 * lvl:2], m[2 * lvl + 1:4 * lvl:2])
        lvl >>= 1
    return s, m


@pytest.mark.parametrize("cap,B,share,hot", [(1 << 20, 100000, 0.3, 4096), (1 << 20, 100000, 0.1, 64),
                                             (1 << 22, 262144, 0.3, 4096), (1 << 22, 262144, 0.9, 1 << 16),
                                             (1 << 20, 40000, 1.0, 8), (1 << 22, 262144, 0.5, 33 * 400)])
def test_tree_set_skewed_large_batches(xp, cuda_device, cap, B, share, hot):
    """Schedule S / L with HOT buckets: a share of a large batch falls into a few leaves (fresh
    transitions carry max_priority and are drawn together, replay_buffer.py:306-309).  Hot buckets are
    not climbed: their items leave last-writer tickets and the dense kernel rebuilds them.  Both entry
    points (values / priorities), twice in a row (scratch back to zero), every node compared."""
    rng = np.random.default_rng(cap + B + hot)
    t = DevTree(xp, cuda_device, cap)
    ls, lm = np.zeros(cap), np.full(cap, np.inf)
    state = torch.ones(1, dtype=torch.float64, device=cuda_device)
    for rnd in range(3):
        k = int(B * share)
        idx = rng.integers(0, cap, B)
        lo = cap - hot if rnd != 1 else cap // 2      # another hot range in the second round
        idx[:k] = lo + rng.integers(0, hot, k)
        idx = idx[rng.permutation(B)]
        if rnd < 2:
            val = rng.random(B) * 5.0 + 1e-3
            t.set(idx, val)
        else:                                          # the PER entry point: leaf = max(p, 1e-5) ** alpha
            pr = (rng.random(B) * 3.0 + 1e-3).astype(np.float32)
            xp.per_update(t.t, state, dev(idx, cuda_device, torch.int64), dev(pr, cuda_device, torch.float32),
                          cap, 0.6, t.err)
            val = np.array([max(float(p), 1e-5) ** 0.6 for p in pr])  # CPython pow, as the reference (:322,425)
        ls[idx] = val   # numpy fancy assignment: the LAST duplicate wins, as the sequential loop
        lm[idx] = val
        ws, wm = _rebuild(ls, lm, cap)
        ds, dm = t.arrays()
        assert np.array_equal(ds[1:], ws[1:]) and np.array_equal(dm[1:], wm[1:]), rnd
        assert int(t.err.item()) == 0
        assert int(t.aux[cap:].abs().sum()) == 0       # tickets consumed
        mb = min(32768, cap // 4)                      # tree_set.cuh max_buckets(): counters / hot flags, zero between calls
        assert int(t.aux[:16].abs().sum()) == 0 and int(t.aux[16:16 + mb].abs().sum()) == 0


def test_tree_set_out_of_range_flags(xp, cuda_device):
    t = DevTree(xp, cuda_device, 8)
    t.set([1, 8, -1], [1.0, 2.0, 3.0])
    torch.cuda.synchronize()
    assert int(t.err.item()) == 1
    assert t.query()[0] == 1.0


# ---- PER ----------------------------------------------------------------------------------------
class DevPer:
    def __init__(self, xp, device, max_size, alpha):
        cap = 1
        while cap < max_size:
            cap *= 2
        self.tree = DevTree(xp, device, cap)
        self.xp, self.device, self.max_size, self.alpha = xp, device, max_size, alpha
        self.state = torch.ones(1, dtype=torch.float64, device=device)  # max_priority = 1.0
        self.tree_ptr = 0
        self.size = 0

    def add(self, n):
        self.xp.per_add(self.tree.t, self.state, self.tree_ptr, n, self.max_size, self.alpha)
        self.tree_ptr = (self.tree_ptr + n) % self.max_size
        self.size = min(self.size + n, self.max_size)

    def update_priorities(self, idx, prio):
        self.xp.per_update(self.tree.t, self.state, dev(idx, self.device, torch.int64),
                           dev(prio, self.device, torch.float32), self.max_size, self.alpha,
                           self.tree.err)

    def sample(self, u, beta, P=1):
        B = len(u) // P
        idx = torch.empty(len(u), dtype=torch.int64, device=self.device)
        w = torch.empty(len(u), dtype=torch.float32, device=self.device)
        self.xp.per_sample(self.tree.t, dev(u, self.device, torch.float32), B, P, self.size, beta, idx, w)
        return idx.cpu().numpy(), w.cpu().numpy()


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_per_scenario_golden_bitwise(xp, cuda_device, seed):
    from test_oracle_golden import replay_per_scenario

    g = load_golden(f"per_s{seed}")
    per = DevPer(xp, cuda_device, int(g["max_size"]), float(g["alpha"]))
    replay_per_scenario(per, g)
    assert per.tree_ptr == int(g["tree_ptr"]) and per.size == int(g["size"])
    s, m = per.tree.arrays()
    assert np.array_equal(s, g["sum_tree_after_update"])
    assert np.array_equal(m, g["min_tree_after_update"])
    assert float(per.state.item()) == float(g["max_priority"])
    idx, w = per.sample(g["u"], float(g["beta"]))
    assert np.array_equal(idx, g["sample_idx"])
    assert np.array_equal(w, g["weights"])
    w2 = torch.empty(len(idx), dtype=torch.float32, device=cuda_device)
    xp.per_weights(per.tree.t, dev(idx, cuda_device), per.size, float(g["beta"]), w2)
    assert np.array_equal(w2.cpu().numpy(), g["weights"])
    assert int(per.tree.err.item()) == 0


def test_per_population_batch_vs_oracle(xp, cuda_device):
    """P agents x B strata in one launch == P sequential reference calls; cap 2^16."""
    max_size, alpha, B, P = 50000, 0.6, 512, 4
    rng = np.random.default_rng(3)
    per = DevPer(xp, cuda_device, max_size, alpha)
    ora = xo.PerOracle(max_size, alpha)
    per.add(max_size)
    # oracle: direct leaf fill + bottom-up pass (sequential adds at 50k leaves take too long)
    ora.size, ora.tree_ptr = max_size, 0
    for i in range(max_size):
        ora.sum_tree.tree[ora.sum_tree.capacity + i] = 1.0
        ora.min_tree.tree[ora.min_tree.capacity + i] = 1.0
    ora.sum_tree.rebuild_from_leaves()
    ora.min_tree.rebuild_from_leaves()
    for _ in range(2):
        idx = rng.integers(0, max_size, 4000)
        pr = (np.abs(rng.standard_normal(4000)) + 1e-6).astype(np.float32)
        per.update_priorities(idx, pr)
        ora.update_priorities(idx, pr)
    s, m = per.tree.arrays()
    assert np.array_equal(s, np.array(ora.sum_tree.tree)) and np.array_equal(m, np.array(ora.min_tree.tree))
    assert float(per.state.item()) == ora.max_priority
    u = rng.random(P * B).astype(np.float32)
    idx, w = per.sample(u, 0.4, P)
    for p in range(P):
        oi = ora.sample_proportional(B, u[p * B:(p + 1) * B])
        assert np.array_equal(idx[p * B:(p + 1) * B], oi)
        assert np.array_equal(w[p * B:(p + 1) * B], ora.calculate_weights(oi, 0.4))


@pytest.mark.parametrize("max_size,B,P", [(1000, 700, 25), (8000, 2048, 9), (50000, 1024, 17), (1 << 20, 4096, 5)])
def test_per_large_draw_team_descent_vs_oracle(xp, cuda_device, max_size, B, P):
    """Draws above 16 384 samples take the line-group / lane-team walk (descend_skip + descend_team3):
    same indices and weights as the reference's level-by-level retrieve (segment_tree.py:147-156),
    for tree depths 10, 13, 16, 20 (every remainder mod 3) and a ragged last warp."""
    alpha = 0.6
    rng = np.random.default_rng(max_size)
    per = DevPer(xp, cuda_device, max_size, alpha)
    ora = xo.PerOracle(max_size, alpha)
    per.add(max_size)
    ora.size, ora.tree_ptr = max_size, 0
    for i in range(max_size):
        ora.sum_tree.tree[ora.sum_tree.capacity + i] = 1.0
        ora.min_tree.tree[ora.min_tree.capacity + i] = 1.0
    ora.sum_tree.rebuild_from_leaves()
    ora.min_tree.rebuild_from_leaves()
    n_up = min(max_size, 6000)
    idx = rng.integers(0, max_size, n_up)
    pr = (np.abs(rng.standard_normal(n_up)) * 3 + 1e-6).astype(np.float32)
    per.update_priorities(idx, pr)
    ora.update_priorities(idx, pr)
    assert P * B > 16384
    u = rng.random(P * B).astype(np.float32)
    u[::97] = 0.0  # stratum lower edges
    got_i, got_w = per.sample(u, 0.4, P)
    for p in range(P):
        oi = ora.sample_proportional(B, u[p * B:(p + 1) * B])
        assert np.array_equal(got_i[p * B:(p + 1) * B], oi), p
        assert np.array_equal(got_w[p * B:(p + 1) * B], ora.calculate_weights(oi, 0.4)), p


def test_per_full_size_properties(xp, cuda_device):
    """C4 size (cap 2^24, 64 agents x 4096 strata): properties that hold exactly at any size.
    * after batched updates every internal node IS op(left, right) (segment_tree.py:89-95) and
      max_priority is the running maximum (:329);
    * the large-draw walk (line groups + lane teams) returns exactly the leaves of the plain
      level-by-level retrieve on the same upper bounds, and the fused IS weights equal xp_per_weights;
    * repeating an update with the same (idx, priority) leaves the tree unchanged (idempotence), and
      duplicate indices resolve to the LAST priority (sequential semantics, :423-428)."""
    cap, B, P, alpha, beta = 1 << 24, 4096, 64, 0.6, 0.4
    per = DevPer(xp, cuda_device, cap, alpha)
    per.add(cap)
    g = torch.Generator(device=cuda_device).manual_seed(0)
    n = P * B
    for _ in range(2):
        idx = torch.randint(0, cap, (n,), device=cuda_device, generator=g)
        pr = torch.randn(n, device=cuda_device, generator=g).abs() * 2 + 1e-6
        xp.per_update(per.tree.t, per.state, idx, pr, cap, alpha, per.tree.err)
    nodes = per.tree.nodes
    kids = nodes[2:].view(cap - 1, 2, 2)  # node i (1 <= i < cap): children 2i, 2i+1
    assert torch.equal(nodes[1:cap, 0], kids[:, 0, 0] + kids[:, 1, 0])
    assert torch.equal(nodes[1:cap, 1], torch.minimum(kids[:, 0, 1], kids[:, 1, 1]))
    assert float(per.state.item()) == max(1.0, float(pr.double().max().item()), float(per.state.item()))
    # last writer wins: the leaf of a duplicated index holds its last priority
    last = {}
    ci, cp = idx[-4096:].cpu().numpy(), pr[-4096:].cpu().numpy()
    for i, p_ in zip(ci, cp):
        last[int(i)] = float(p_)
    seen_later = set(idx[-4096:].cpu().numpy().tolist())
    some = list(last.items())[:512]
    leaves = nodes[cap + torch.tensor([k for k, _ in some], device=cuda_device), 0].cpu().numpy()
    want = np.array([math.pow(max(v, 1e-5), alpha) for _, v in some])
    assert np.array_equal(leaves, want) and seen_later
    # descent: team walk == plain walk, fused weights == xp_per_weights
    u = torch.rand(n, device=cuda_device, generator=g)
    got_i = torch.empty(n, dtype=torch.int64, device=cuda_device)
    got_w = torch.empty(n, dtype=torch.float32, device=cuda_device)
    xp.per_sample(per.tree.t, u, B, P, cap, beta, got_i, got_w)
    total = nodes[1, 0]
    seg = total / B
    i = torch.arange(B, device=cuda_device, dtype=torch.float64).repeat(P)
    a, b = seg * i, seg * (i + 1)
    ub = u.double() * (b - a) + a
    assert torch.equal(got_i, xp.tree_retrieve(per.tree.t, ub))
    w2 = torch.empty_like(got_w)
    xp.per_weights(per.tree.t, got_i, cap, beta, w2)
    assert torch.equal(got_w, w2)
    assert int(got_i.min()) >= 0 and int(got_i.max()) < cap
    # idempotence of the batched set
    before = nodes.clone()
    xp.per_update(per.tree.t, per.state, idx, pr, cap, alpha, per.tree.err)
    assert torch.equal(nodes, before) and int(per.tree.err.item()) == 0


@pytest.mark.parametrize("max_size,B", [(1000, 32), (1 << 12, 1), (50_000, 257), (1 << 20, 1024), (1 << 22, 512)])
def test_per_host_vector_variants_match_device_pointer_variants(xp, cuda_device, max_size, B):
    """xp_per_sample_hostu / xp_per_update_hostp (the fp32 vector travels inside the launch) against
    xp_per_sample (P = 1) / xp_per_update on twin trees: indices, weights, every tree node, max_priority
    and the error flag — bit-exact — through three rounds, duplicates and an out-of-range index included;
    and against the list oracle (replay_buffer.py:357-428) when the tree is small."""
    alpha, beta = 0.6, 0.4
    rng = np.random.default_rng(max_size + B)
    a, b = DevPer(xp, cuda_device, max_size, alpha), DevPer(xp, cuda_device, max_size, alpha)
    a.add(max_size), b.add(max_size)
    ora = None
    if max_size <= 50_000:                              # the list oracle finishes in seconds there
        ora = xo.PerOracle(max_size, alpha)
        ora.add(max_size)
    for rnd in range(3):
        idx = rng.integers(0, max_size, B)
        if B > 4:
            idx[B // 2:] = idx[: B - B // 2]          # duplicates: the last writer wins
        pr = (np.abs(rng.standard_normal(B)) * (3.0 if rnd else 0.5) + 1e-7).astype(np.float32)
        d_idx = dev(idx, cuda_device, torch.int64)
        xp.per_update(a.tree.t, a.state, d_idx, dev(pr, cuda_device, torch.float32), max_size, alpha, a.tree.err)
        keep = pr.copy()
        xp.per_update_hostp(b.tree.t, b.state, d_idx, pr, max_size, alpha, b.tree.err)
        pr[:] = -1.0                                    # the host array is the caller's again on return
        u = rng.random(B, dtype=np.float32)
        wi = torch.empty(B, dtype=torch.int64, device=cuda_device); ww = torch.empty(B, device=cuda_device)
        gi = torch.empty_like(wi); gw = torch.empty_like(ww)
        xp.per_sample(a.tree.t, dev(u, cuda_device, torch.float32), B, 1, max_size, beta, wi, ww)
        keep_u = u.copy()
        xp.per_sample_hostu(b.tree.t, u, B, max_size, beta, gi, gw)
        u[:] = 2.0
        torch.cuda.synchronize()
        assert torch.equal(gi, wi) and torch.equal(gw, ww), rnd
        assert torch.equal(a.tree.nodes, b.tree.nodes) and torch.equal(a.state, b.state), rnd
        gi2 = torch.empty_like(gi)
        xp.per_sample_hostu(b.tree.t, keep_u, B, max_size, beta, gi2, None)   # indices only
        assert torch.equal(gi2, wi)
        if ora is not None:
            ora.update_priorities(idx, keep)
            oi = ora.sample_proportional(B, keep_u)
            assert np.array_equal(gi.cpu().numpy(), oi), rnd
            assert np.array_equal(gw.cpu().numpy(), ora.calculate_weights(oi, beta)), rnd
            assert float(b.state.item()) == ora.max_priority
    # an index outside [0, max_size): flagged, row skipped — like xp_per_update
    bad = dev(np.array([0, max_size, 1]), cuda_device, torch.int64)
    three = np.array([0.5, 9.0, 0.25], dtype=np.float32)
    xp.per_update(a.tree.t, a.state, bad, dev(three, cuda_device, torch.float32), max_size, alpha, a.tree.err)
    xp.per_update_hostp(b.tree.t, b.state, bad, three, max_size, alpha, b.tree.err)
    torch.cuda.synchronize()
    assert int(a.tree.err.item()) == 1 and int(b.tree.err.item()) == 1
    assert torch.equal(a.tree.nodes, b.tree.nodes) and torch.equal(a.state, b.state)
    # limits
    big = torch.zeros(1025, dtype=torch.int64, device=cuda_device)
    with pytest.raises(ValueError):
        xp.per_update_hostp(b.tree.t, b.state, big, np.ones(1025, dtype=np.float32), max_size, alpha, b.tree.err)
    with pytest.raises(ValueError):
        xp.per_sample_hostu(b.tree.t, np.ones(1025, dtype=np.float32), 1025, max_size, beta, big, None)


def test_per_c4_size_indices_and_weights_vs_array_oracle(xp, cuda_device):
    """BASELINE.json config 4 (cap 2^24, 64 agents x 4 096 strata, L = 24) against an oracle OUTSIDE the
    library: oracle.xp_oracle.np_* (numpy restatement of segment_tree.py:136-156 / replay_buffer.py:357-409,
    pinned to the list oracle at 2^16 and to the unmodified reference at 2^14 on the CPU).  Every node of
    both device trees, every sampled index and every IS weight — bit-exact.  Leaves go in through the
    library's own update path (priority ** alpha on the device); the host computes them with CPython's pow."""
    cap, B, P, alpha, beta = 1 << 24, 4096, 64, 0.6, 0.4
    rng = np.random.default_rng(24)
    per = DevPer(xp, cuda_device, cap, alpha)
    per.add(cap)
    pri = (np.abs(rng.standard_normal(cap)) + 1e-3).astype(np.float32)
    idx_all = torch.arange(cap, device=cuda_device)
    xp.per_update(per.tree.t, per.state, idx_all, dev(pri, cuda_device, torch.float32), cap, alpha, per.tree.err)
    # a second, sparse, duplicate-laden update of the C4 batch size on top (schedule S)
    n = P * B
    up_i = rng.integers(0, cap, n)
    up_i[n // 2:] = up_i[: n - n // 2]
    up_p = (np.abs(rng.standard_normal(n)) + 1e-6).astype(np.float32)
    xp.per_update(per.tree.t, per.state, dev(up_i, cuda_device, torch.int64), dev(up_p, cuda_device, torch.float32),
                  cap, alpha, per.tree.err)
    pri[up_i] = up_p  # numpy keeps the LAST duplicate: sequential last-writer-wins (replay_buffer.py:423-428)
    leaves = np.fromiter((max(float(p), 1e-5) ** alpha for p in pri.tolist()), dtype=np.float64, count=cap)
    s, m = xo.np_trees_from_leaves(leaves, cap)
    ds, dm = per.tree.arrays()
    assert np.array_equal(ds[1:], s[1:]) and np.array_equal(dm[1:], m[1:])
    assert int(per.tree.err.item()) == 0
    u = rng.random(n).astype(np.float32)
    u[::4099] = 0.0
    got_i, got_w = per.sample(u, beta, P)
    for p in range(P):
        want = xo.np_sample_proportional(s, cap, B, u[p * B:(p + 1) * B])
        assert np.array_equal(got_i[p * B:(p + 1) * B], want), p
    for p in (0, 31, 63):  # weights: CPython pow per element on the host, three agents
        sl = slice(p * B, (p + 1) * B)
        assert np.array_equal(got_w[sl], xo.np_calculate_weights(s, m, cap, got_i[sl], cap, beta)), p


def test_pow_is_cpython_pow(xp, cuda_device):
    rng = np.random.default_rng(0)
    x = np.concatenate([np.maximum(np.abs(rng.standard_normal(200000)), 1e-5),
                        rng.random(100000) * 1e-3, np.exp(rng.uniform(-30, 30, 100000)),
                        [1.0, 1e-5, 0.5, 2.0, 1e-300, 1e300]])
    for y in (0.6, -0.4, 0.5, 1.0, 0.0, -1.0, 0.25):
        got = xp.pow_batch(dev(x, cuda_device, torch.float64), y).cpu().numpy()
        want = np.array([math.pow(float(a), y) for a in x])
        assert np.array_equal(got, want), y


# ---- ring / gather / n-step -----------------------------------------------------------------------
KEYS = ["obs", "action", "reward", "next_obs", "done"]


def step_tensors(g, s, N, device):
    out = {}
    for k in KEYS:
        v = g[k][s]
        out[k] = dev(v.reshape(N, 1) if v.ndim == 1 else v, device)
    return out


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_ring_and_nstep_golden(xp, cuda_device, seed):
    g = load_golden(f"ring_nstep_s{seed}")
    max_size, N, n_step, gamma = int(g["max_size"]), int(g["N"]), int(g["n_step"]), float(g["gamma"])
    ring = {k: torch.zeros_like(dev(g[f"ring_{k}"], cuda_device)) for k in KEYS}
    nst = {k: torch.zeros_like(dev(g[f"nstep_{k}"], cuda_device)) for k in KEYS}
    stage = {k: torch.zeros(n_step * N, *ring[k].shape[1:], dtype=ring[k].dtype, device=cuda_device) for k in KEYS}
    roles = [0, 0, 2, 1, 1]  # obs, action FIRST; reward fold; next_obs, done LAST
    gpow = [float(np.float32(gamma**i)) for i in range(n_step)]
    cur = ncur = count = 0
    for s in range(g["obs"].shape[0]):
        d = step_tensors(g, s, N, cuda_device)
        xp.ring_insert([(d[k], ring[k]) for k in KEYS], N, cur, max_size)
        cur = (cur + N) % max_size
        slot = count % n_step
        xp.ring_insert([(d[k], stage[k]) for k in KEYS], N, slot * N, n_step * N)
        count += 1
        if count >= n_step:
            xp.nstep_fold_insert([(stage[k], nst[k]) for k in KEYS], roles, 4, 4, n_step,
                                 count % n_step, gpow, N, ncur, max_size)
            ncur = (ncur + N) % max_size
    torch.cuda.synchronize()
    assert cur == int(g["ring_cursor"]) and ncur == int(g["nstep_cursor"])
    for k in KEYS:
        assert np.array_equal(ring[k].cpu().numpy(), g[f"ring_{k}"]), k
        assert np.array_equal(nst[k].cpu().numpy(), g[f"nstep_{k}"]), k


@pytest.mark.parametrize("row_shape,dtype", [((84, 84, 4), torch.uint8), ((376,), torch.float32),
                                            ((1,), torch.float32), ((3,), torch.uint8),
                                            ((5,), torch.int16), ((7,), torch.float64), ((1031,), torch.uint8)])
def test_gather_and_insert_all_row_widths(xp, cuda_device, row_shape, dtype):
    rows, B = 300, 77
    g = torch.Generator(device="cpu").manual_seed(0)
    src = torch.randint(0, 200, (rows, *row_shape), generator=g).to(dtype).to(cuda_device)
    idx = torch.randint(0, rows, (B,), generator=g).to(cuda_device)
    out = torch.empty(B, *row_shape, dtype=dtype, device=cuda_device)
    err = torch.zeros(1, dtype=torch.int32, device=cuda_device)
    xp.gather([(src, out)], idx, rows, err)
    assert torch.equal(out, src[idx]) and int(err.item()) == 0
    # wrap-around insert: cursor near the end
    store = torch.zeros(rows, *row_shape, dtype=dtype, device=cuda_device)
    xp.ring_insert([(out, store)], B, rows - 10, rows)
    want = torch.zeros_like(store)
    want[rows - 10:] = out[:10]
    want[: B - 10] = out[10:]
    assert torch.equal(store, want)
    # unaligned views (odd byte offsets) still move correctly
    if dtype == torch.uint8 and len(row_shape) == 1:
        flat = torch.randint(0, 255, (rows * row_shape[0] + 1,), generator=g).to(dtype).to(cuda_device)
        usrc = flat[1:].view(rows, *row_shape)
        xp.gather([(usrc, out)], idx, rows, err)
        assert torch.equal(out, usrc[idx])


def test_gather_out_of_range_flags_and_zeroes(xp, cuda_device):
    src = torch.arange(40, dtype=torch.float32, device=cuda_device).view(10, 4)
    idx = torch.tensor([0, 10, 3, -1], device=cuda_device)
    out = torch.full((4, 4), 7.0, device=cuda_device)
    err = torch.zeros(1, dtype=torch.int32, device=cuda_device)
    xp.gather([(src, out)], idx, 10, err)
    assert int(err.item()) == 1
    assert torch.equal(out[0], src[0]) and torch.equal(out[2], src[3])
    assert float(out[1].abs().sum()) == 0.0 and float(out[3].abs().sum()) == 0.0


# ---- uniform sampling ---------------------------------------------------------------------------
def test_randperm_prefix_golden(xp, cuda_device):
    g = load_golden("torch_rng")
    for k, (seed, n, B) in enumerate(g["cases"]):
        raw = xo.mt19937_raw(int(seed), int(B)).view(np.int32)
        got = xp.randperm_prefix(dev(raw, cuda_device), int(B), 1, int(n)).cpu().numpy()
        assert np.array_equal(got, g[f"perm_{k}"]), (seed, n, B)


def test_randperm_prefix_population_and_dense(xp, cuda_device):
    rng = np.random.default_rng(0)
    for n, B, P in [(4096, 4096, 3), (5000, 4096, 2), (10_000_000, 1024, 32), (16384, 16384, 1),
                    (65536, 4096, 2), (1000, 50, 3), (20000, 1024, 4), (800, 50, 2), (799, 50, 2)]:  # hash path from n >= 16 B
        raw = rng.integers(0, 2**32, P * B, dtype=np.uint64).astype(np.uint32)
        got = xp.randperm_prefix(dev(raw.view(np.int32), cuda_device), B, P, n).cpu().numpy()
        for p in (0, P - 1):
            want = xo.randperm_prefix(raw[p * B:(p + 1) * B], B, n)
            assert np.array_equal(got[p * B:(p + 1) * B], want)
            assert len(np.unique(got[p * B:(p + 1) * B])) == B  # without replacement


def test_philox_is_deterministic_and_counter_based(xp, cuda_device):
    a = xp.philox_u32(123, 0, 4096, cuda_device)
    b = xp.philox_u32(123, 0, 4096, cuda_device)
    c = xp.philox_u32(123, 256, 3072, cuda_device)  # 256 counters * 4 words = offset 1024
    assert torch.equal(a, b) and torch.equal(a[1024:], c)
    # Philox4x32-10 known answer (Random123 kat_vectors: ctr=0, key=0)
    z = xp.philox_u32(0, 0, 4, cuda_device).cpu().numpy().view(np.uint32)
    assert z.tolist() == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
