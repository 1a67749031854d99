"""CPU checks of the C-ABI boundary: the library loads, exports every symbol include/b200xp.h
declares, argument validation maps to the reference's exceptions, and the host-only entry points
(mt19937 mirror, randperm prefix) reproduce torch's CPU stream.  No kernel is launched."""

import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT, load_golden
from agilerl_b200 import _native as nv
from oracle import xp_oracle as xo


@pytest.fixture(scope="module")
def lib():
    from agilerl_b200 import build

    build.build()
    return nv.load()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200xp.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(xp_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"libb200xp.so does not export {n}"
        assert n in nv._SIGNATURES, f"{n} has no ctypes signature in agilerl_b200/_native.py"
    assert set(nv._SIGNATURES) == set(names)
    assert lib.xp_version() == 1


def test_status_codes_map_to_reference_exceptions(lib):
    t = nv.XpTree()
    t.sum, t.min, t.stride, t.cap, t.aux = 8, 0, 1, 6, 0  # capacity 6: not a power of two
    st = lib.xp_tree_init(C.byref(t), None)
    assert st == nv.XP_ERR_INDEX and b"power of 2" in lib.xp_last_error()
    with pytest.raises(AssertionError):
        nv.check(st)  # segment_tree.py:20-23 raises AssertionError
    with pytest.raises(ValueError):
        nv.check(lib.xp_gae(0, 0, 0, 0, 0, 0, 0, -1, 4, 1, 0, 0.99, 0.95, None))
    with pytest.raises(ValueError):
        nv.check(lib.xp_randperm_prefix(8, 10, 1, 5, 8, None))  # B > n
    assert lib.xp_gather(None, 0, None, 0, 0, None, None) == nv.XP_OK  # empty batch is a no-op
    assert lib.xp_gae(0, 0, 0, 0, 0, 0, 0, 0, 4, 1, 0, 0.99, 0.95, None) == nv.XP_OK


def test_host_mt19937_is_torch_cpu_stream(lib):
    g = load_golden("torch_rng")
    for seed in (0, 1, 123456789):
        m = nv.XpMt19937()
        lib.xp_mt19937_seed(C.byref(m), seed & 0xFFFFFFFF)
        raw = np.empty(700, dtype=np.uint32)
        lib.xp_mt19937_draw(C.byref(m), raw.ctypes.data, 700)
        assert np.array_equal(raw, xo.mt19937_raw(seed, 700))
        assert np.array_equal(xo.torch_rand_from_raw(raw), g[f"rand_{seed}"])
    a, b = nv.XpMt19937(), nv.XpMt19937()
    lib.xp_mt19937_seed(C.byref(a), 5)
    lib.xp_mt19937_seed(C.byref(b), 5)
    lib.xp_mt19937_skip(C.byref(a), 1000)
    junk = np.empty(1000, dtype=np.uint32)
    lib.xp_mt19937_draw(C.byref(b), junk.ctypes.data, 1000)
    x, y = np.empty(5, np.uint32), np.empty(5, np.uint32)
    lib.xp_mt19937_draw(C.byref(a), x.ctypes.data, 5)
    lib.xp_mt19937_draw(C.byref(b), y.ctypes.data, 5)
    assert np.array_equal(x, y)


def test_host_randperm_prefix_golden(lib):
    g = load_golden("torch_rng")
    for k, (seed, n, B) in enumerate(g["cases"]):
        if B > 1100:
            continue  # the host restatement is O(B^2): small checks only
        raw = xo.mt19937_raw(int(seed), int(B))
        out = np.empty(int(B), dtype=np.int64)
        assert lib.xp_host_randperm_prefix(raw.ctypes.data, int(B), int(n), out.ctypes.data) == 0
        assert np.array_equal(out, g[f"perm_{k}"])


def test_torch_cpu_stream_mirror_and_writeback():
    from agilerl_b200.components.rng import TorchCpuStream

    for pre in (0, 5, 623, 624, 625, 2000):
        torch.manual_seed(11)
        if pre:
            torch.rand(pre)
        s = TorchCpuStream()
        got = s.rand_f32(300)
        s.commit()
        tail = torch.rand(7)
        torch.manual_seed(11)
        if pre:
            torch.rand(pre)
        assert torch.equal(got, torch.rand(300)) and torch.equal(tail, torch.rand(7))
    torch.manual_seed(2)
    s = TorchCpuStream()
    raw = s.draw_u32(64).numpy().view(np.uint32)
    s.skip(100000 - 1 - 64)
    s.commit()
    nxt = torch.rand(3)
    torch.manual_seed(2)
    want = torch.randperm(100000)[:64].numpy()
    assert np.array_equal(xo.randperm_prefix(raw, 64, 100000), want) and torch.equal(nxt, torch.rand(3))


def test_ctypes_signatures_match_the_header_argument_by_argument():
    """Binding drift check: for every prototype in include/b200xp.h the ctypes table has the same number
    of parameters and the same kind for each (pointer / 64-bit integer / 32-bit integer / double)."""
    text = open(os.path.join(ROOT, "include", "b200xp.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    protos = re.findall(r"\b[a-z_0-9]+(?:\s*\*)?\s+(xp_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", text, flags=re.S)
    assert len(protos) >= 30

    def kind_of_c(param: str) -> str:
        param = " ".join(param.split())
        if "*" in param:
            return "ptr"
        base = param.rsplit(" ", 1)[0] if " " in param else param
        if base in ("int64_t", "uint64_t"):
            return "i64"
        if base in ("int", "uint32_t", "xp_status_t"):
            return "i32"
        if base == "double":
            return "f64"
        raise AssertionError(f"unclassified parameter {param!r}")

    def kind_of_ctypes(t) -> str:
        if t is C.c_void_p or t is C.c_char_p or (isinstance(t, type) and issubclass(t, C._Pointer)):
            return "ptr"
        if t in (C.c_int64, C.c_uint64, C.c_longlong, C.c_ulonglong):
            return "i64"
        if t in (C.c_int, C.c_uint32, C.c_uint):
            return "i32"
        if t is C.c_double:
            return "f64"
        raise AssertionError(f"unclassified ctypes type {t}")

    for name, params in protos:
        params = params.strip()
        want = [] if params in ("", "void") else [kind_of_c(q) for q in params.split(",")]
        _, argtypes = nv._SIGNATURES[name]
        got = [kind_of_ctypes(t) for t in argtypes]
        assert got == want, f"{name}: header {want} vs _native.py {got}"
