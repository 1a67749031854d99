"""xp_standardize_static — the no-allocation variant of xp_standardize (scratch in a file-scope
``static __device__`` array inside libb200xp.so; include/b200xp.h).  It has no counterpart in the
reference.  Called through the C-ABI (ctypes) and
checked against the existing entry point, which runs the same kernels in the same fixed summation order:
the outputs must be bit-identical (and, a fortiori, within the rtol 2e-5 / atol 1e-6 the standardize
tests use against the reference).  Two calls in a row: the array keeps the previous call's contents.

This file is collected LAST on purpose."""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib(cuda_device):
    from agilerl_b200 import _native as nv

    return nv.load()


def _stream(dev):
    return torch.cuda.current_stream(dev).cuda_stream


def _both(lib, x):
    """(xp_standardize with caller scratch, xp_standardize_static) on the same input."""
    from agilerl_b200 import _native as nv

    dev = x.device
    scratch = torch.zeros(520, dtype=torch.float64, device=dev)  # XP_WHITEN_SCRATCH doubles
    want = torch.empty_like(x)
    got = torch.full_like(x, float("nan"))
    with torch.cuda.device(dev):
        nv.check(lib.xp_standardize(x.data_ptr(), want.data_ptr(), x.numel(), scratch.data_ptr(), _stream(dev)),
                 "xp_standardize")
        # from here on the call under test: inputs and outputs are the caller's, nothing else is passed
        before = torch.cuda.memory_allocated(dev)
        nv.check(lib.xp_standardize_static(x.data_ptr(), got.data_ptr(), x.numel(), _stream(dev)),
                 "xp_standardize_static")
        assert torch.cuda.memory_allocated(dev) == before  # (torch's allocator was not involved either)
    torch.cuda.synchronize(dev)
    return want, got


def test_static_scratch_matches_caller_scratch_on_two_calls_in_a_row(lib, cuda_device):
    g = torch.Generator(device=cuda_device).manual_seed(11)
    # 1st call: many CTAs (256 partials); 2nd: a short, ragged input that uses 3 CTAs only — the stale
    # partials and statistics of the 1st call are still in the array; 3rd: large again
    for n, scale, shift in ((1 << 20, 3.0, -1.5), (2049, 0.25, 40.0), (1_234_567, 1.0, 0.0)):
        x = torch.randn(n, device=cuda_device, generator=g) * scale + shift
        want, got = _both(lib, x)
        assert torch.equal(got, want), n                       # same kernels, same order: same bits
        ref = (x - x.mean()) / (x.std() + 1e-8)                # algorithms/ppo.py:829-833 in torch
        np.testing.assert_allclose(got.cpu().numpy(), ref.cpu().numpy(), rtol=2e-5, atol=1e-6)


def test_static_scratch_in_place_and_back_to_back_on_one_stream(lib, cuda_device):
    """out may alias x; two calls queued without a synchronise in between (stream order serialises them)."""
    from agilerl_b200 import _native as nv

    g = torch.Generator(device=cuda_device).manual_seed(12)
    a = torch.randn(70_001, device=cuda_device, generator=g) * 2 + 1
    b = torch.rand(300_000, device=cuda_device, generator=g)
    wa, _ = _both(lib, a)
    wb, _ = _both(lib, b)
    with torch.cuda.device(cuda_device):
        nv.check(lib.xp_standardize_static(a.data_ptr(), a.data_ptr(), a.numel(), _stream(cuda_device)))
        nv.check(lib.xp_standardize_static(b.data_ptr(), b.data_ptr(), b.numel(), _stream(cuda_device)))
    torch.cuda.synchronize(cuda_device)
    assert torch.equal(a, wa) and torch.equal(b, wb)


def test_static_scratch_limits_and_empty(lib, cuda_device):
    from agilerl_b200 import _native as nv

    x = torch.ones(8, device=cuda_device)
    st = lib.xp_standardize_static(x.data_ptr(), x.data_ptr(), (1 << 26) + 1, None)
    assert st == nv.XP_ERR_INVALID
    msg = lib.xp_last_error().decode()
    assert "exceeds the maximum" in msg and str(1 << 26) in msg
    with pytest.raises(ValueError):
        nv.check(st)
    assert lib.xp_standardize_static(x.data_ptr(), x.data_ptr(), -1, None) == nv.XP_ERR_INVALID
    assert lib.xp_standardize_static(None, None, 0, None) == nv.XP_OK       # empty: no launch, no look-up
    assert lib.xp_standardize_static(None, x.data_ptr(), 8, None) == nv.XP_ERR_INVALID
    torch.cuda.synchronize(cuda_device)
    assert torch.equal(x.cpu(), torch.ones(8))                               # refused calls touched nothing
