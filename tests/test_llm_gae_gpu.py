"""Token- and turn-level GAE of the LLM PPO path (algorithms/ppo_llm.py:758-872) on the CUDA engine,
against golden vectors written by the UNMODIFIED reference methods (oracle/make_golden.py llm_gae)
and against the oracle on ragged shapes.

Bars: the token-level scan without whitening is a chain of single fp32 ops -> bit-exact.  Anything
behind a `tensor.sum()` of the reference (whitening statistics, mean-pooled turn rewards / values)
has no fixed summation order in torch -> rtol 2e-5 / atol 1e-6 in fp32 (north_star: 1e-6 relative for
returns; whitening divides by a standard deviation, which amplifies the last-ulp noise of the mean)."""

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import xp_oracle as xo

pytestmark = pytest.mark.gpu
TOL = dict(rtol=2e-5, atol=1e-6)


@pytest.fixture(scope="module")
def returns(cuda_device):
    from agilerl_b200 import returns as r

    return r


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_token_gae_golden(returns, cuda_device, seed):
    g = load_golden(f"llm_gae_s{seed}")
    ret, adv = returns.gae_returns_token(g["rewards"], g["values"], g["mask"], float(g["gamma"]),
                                         float(g["gae_lambda"]), adv_whitening=False, device=cuda_device)
    assert np.array_equal(ret.cpu().numpy(), g["tok_ret_w0"])  # bit-exact
    assert np.array_equal(adv.cpu().numpy(), g["tok_adv_w0"])
    ret, adv = returns.gae_returns_token(g["rewards"], g["values"], g["mask"], float(g["gamma"]),
                                         float(g["gae_lambda"]), adv_whitening=True, device=cuda_device)
    assert np.array_equal(ret.cpu().numpy(), g["tok_ret_w1"])  # returns do not pass through the whitening
    np.testing.assert_allclose(adv.cpu().numpy(), g["tok_adv_w1"], **TOL)


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
@pytest.mark.parametrize("reduction", ["final_value", "mean"])
@pytest.mark.parametrize("whiten", [False, True])
def test_turn_gae_golden(returns, cuda_device, seed, reduction, whiten):
    g = load_golden(f"llm_gae_s{seed}")
    ret, adv = returns.gae_returns_turn(g["rewards"], g["values"], g["mask"], g["turn_ids"], float(g["gamma"]),
                                        float(g["gae_lambda"]), turn_value_reduction=reduction,
                                        adv_whitening=whiten, device=cuda_device)
    np.testing.assert_allclose(ret.cpu().numpy(), g[f"turn_ret_{reduction}_w{int(whiten)}"], **TOL)
    np.testing.assert_allclose(adv.cpu().numpy(), g[f"turn_adv_{reduction}_w{int(whiten)}"], **TOL)


@pytest.mark.parametrize("B,S", [(1, 1), (3, 7), (33, 63), (32, 64), (40, 65), (5, 1000), (70, 130)])
def test_token_gae_ragged_shapes_vs_oracle(returns, cuda_device, B, S):
    """Row counts off the 32-row CTA, sequence lengths off the 64-step tile and off the 16-byte
    vector path (S % 4 != 0), float masks, non-contiguous inputs."""
    rng = np.random.default_rng(B * 1000 + S)
    mask = rng.random((B, S)) < 0.6
    r = (rng.standard_normal((B, S)) * mask).astype(np.float32)
    v = (rng.standard_normal((B, S)) * mask).astype(np.float32)
    want_ret, want_adv = xo.gae_tokens(r, v, mask, 0.997, 0.9, False)
    rt = torch.from_numpy(np.ascontiguousarray(r.T)).to(cuda_device).T  # non-contiguous view
    ret, adv = returns.gae_returns_token(rt, torch.from_numpy(v), torch.from_numpy(mask.astype(np.float32)),
                                         0.997, 0.9, adv_whitening=False, device=cuda_device)
    assert np.array_equal(ret.cpu().numpy(), want_ret) and np.array_equal(adv.cpu().numpy(), want_adv)
    if mask.sum() >= 2:
        want_ret, want_adv = xo.gae_tokens(r, v, mask, 0.997, 0.9, True)
        ret, adv = returns.gae_returns_token(r, v, mask, 0.997, 0.9, adv_whitening=True, device=cuda_device)
        assert np.array_equal(ret.cpu().numpy(), want_ret)
        np.testing.assert_allclose(adv.cpu().numpy(), want_adv, **TOL)


@pytest.mark.parametrize("B,S,K", [(1, 5, 1), (9, 77, 4), (35, 300, 40), (4, 2100, 300)])
def test_turn_gae_vs_oracle(returns, cuda_device, B, S, K):
    """Interleaved / non-monotonic turn ids, many turns, rows without any turn."""
    rng = np.random.default_rng(B + S + K)
    tid = rng.integers(-1, K, (B, S))
    tid[rng.random((B, S)) < 0.3] = -1
    if B > 2:
        tid[1, :] = -1  # a sequence with no action token at all
    tid[0, 0] = K - 1
    mask = tid >= 0
    r = (rng.standard_normal((B, S)) * mask).astype(np.float32)
    v = (rng.standard_normal((B, S)) * mask).astype(np.float32)
    for red in ("final_value", "mean"):
        for w in (False, True):
            if w and B * K < 2:  # one turn in total: the reference's unbiased variance raises (llm_utils.py:236)
                with pytest.raises(ValueError):
                    xo.gae_turns(r, v, tid, 0.99, 0.95, w, red)
                with pytest.raises(ValueError):
                    returns.gae_returns_turn(r, v, mask, tid, 0.99, 0.95, turn_value_reduction=red,
                                             adv_whitening=w, device=cuda_device)
                continue
            want_ret, want_adv = xo.gae_turns(r, v, tid, 0.99, 0.95, w, red)
            ret, adv = returns.gae_returns_turn(r, v, mask, tid, 0.99, 0.95, turn_value_reduction=red,
                                                adv_whitening=w, device=cuda_device)
            np.testing.assert_allclose(ret.cpu().numpy(), want_ret, **TOL)
            np.testing.assert_allclose(adv.cpu().numpy(), want_adv, **TOL)


def test_llm_gae_errors(returns, cuda_device):
    r = np.zeros((2, 8), dtype=np.float32)
    m = np.zeros((2, 8), dtype=bool)
    m[0, 3] = True
    with pytest.raises(ValueError, match="at least 2 unmasked"):  # llm_utils.py:236-241
        returns.gae_returns_token(r, r, m, 0.99, 0.95, adv_whitening=True, device=cuda_device)
    ret, adv = returns.gae_returns_token(r, r, m, 0.99, 0.95, adv_whitening=False, device=cuda_device)
    assert ret.shape == (2, 8) and float(adv.abs().sum()) == 0.0
    with pytest.raises(ValueError):
        returns.gae_returns_turn(r, r, m, np.zeros((2, 8), dtype=np.int64), 0.99, 0.95, turn_value_reduction="sum",
                                 device=cuda_device)
    with pytest.raises(ValueError):
        returns.gae_returns_token(r, r[:, :4], m, 0.99, 0.95, device=cuda_device)


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_normalize_advantages_golden(returns, cuda_device, seed):
    """(adv - adv.mean()) / (adv.std() + 1e-8) of PPO's flat learn step (algorithms/ppo.py:829-833) against the
    reference's own statement run on the CPU; mean / std sit behind torch reductions -> rtol 2e-5."""
    g = load_golden(f"advnorm_s{seed}")
    out = returns.normalize_advantages(g["advantages"], device=cuda_device)
    assert out.shape == g["advantages"].shape and out.dtype == torch.float32
    np.testing.assert_allclose(out.cpu().numpy(), g["normalized"], **TOL)
    np.testing.assert_allclose(out.cpu().numpy(), xo.normalize_advantages(g["advantages"]), **TOL)
    # in place on a device-resident buffer, and deterministic run to run
    x = torch.from_numpy(g["advantages"]).to(cuda_device)
    y = returns.normalize_advantages(x, out=x)
    assert y.data_ptr() == x.data_ptr() and torch.equal(y, out)


def test_normalize_advantages_c3_size_and_single_value(returns, cuda_device):
    x = torch.randn(128 * 1024, device=cuda_device) * 2 + 1  # one C3 agent's rollout
    y = returns.normalize_advantages(x)
    assert abs(float(y.mean())) < 1e-5 and abs(float(y.std()) - 1.0) < 1e-5
    assert torch.isnan(returns.normalize_advantages(torch.ones(1, device=cuda_device))).all()  # torch: std of 1 value
