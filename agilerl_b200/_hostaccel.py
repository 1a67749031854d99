"""Loader of the optional native host fast path (``csrc/hostaccel.cpp`` -> ``_lib/_hostaccel.so``).

The extension only shortens the Python host's own bookkeeping (output allocation, staging,
argument marshalling); every kernel still comes from libb200xp.so.  ``load()`` returns the bound
module, or ``None`` when it is not built, fails to import, or ``XP_NO_HOSTACCEL=1`` — the Python
classes then run their own equivalent code."""

from __future__ import annotations

import ctypes as C
import importlib.util
import os
import warnings

from . import _native as nv
from .build import LIB_DIR

_MOD = None
_TRIED = False


def load():
    global _MOD, _TRIED
    if _TRIED:
        return _MOD
    _TRIED = True
    if os.environ.get("XP_NO_HOSTACCEL") == "1":
        return None
    path = LIB_DIR / "_hostaccel.so"
    if not path.exists():
        return None
    try:
        import torch  # noqa: F401  (libtorch symbols must be loaded first)

        spec = importlib.util.spec_from_file_location("_hostaccel", str(path))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        lib = nv.load()

        def addr(fn) -> int:
            return C.cast(fn, C.c_void_p).value

        mod.bind(addr(lib.xp_gather), addr(lib.xp_per_sample), addr(lib.xp_per_update),
                 addr(lib.xp_copy_to_device), nv.check)
        mod.bind_peer(addr(lib.xp_peer_allgather_graph))
        if os.environ.get("XP_NO_HOST_VEC", "0") != "1":  # developer knob: stage the 1 KB vectors through the pinned ring
            mod.bind_host_vec(addr(lib.xp_per_sample_hostu), addr(lib.xp_per_update_hostp),
                              addr(lib.xp_peer_publish_allgather_graph))
        _MOD = mod
    except Exception as e:  # pragma: no cover - depends on the local toolchain
        warnings.warn(f"agilerl_b200: native host fast path unavailable ({e}); using the Python host")
        _MOD = None
    return _MOD
