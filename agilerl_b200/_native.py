"""ctypes binding of libb200xp.so (the C-ABI declared in include/b200xp.h).

There is no CPU fallback: if the library is missing or a call fails, this raises.  Tensors are
passed as raw device pointers (``tensor.data_ptr()``) and the current CUDA stream as a
``cudaStream_t``; PyTorch only provides memory and streams.
"""

from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_LIB_PATH = Path(__file__).resolve().parent / "_lib" / "libb200xp.so"

XP_OK = 0
XP_ERR_INVALID = -1
XP_ERR_INDEX = -2
XP_ERR_CUDA = -3
XP_ERR_STATE = -4
XP_MAX_KEYS = 24
XP_MAX_NSTEP = 32
XP_ROLE_FIRST, XP_ROLE_LAST, XP_ROLE_REWARD = 0, 1, 2

c_i64 = C.c_int64
c_vp = C.c_void_p


class XpKey(C.Structure):
    _fields_ = [("src", c_vp), ("dst", c_vp), ("row_bytes", c_i64)]


class XpTree(C.Structure):
    _fields_ = [("sum", c_vp), ("min", c_vp), ("stride", c_i64), ("cap", c_i64), ("aux", c_vp)]


class XpMt19937(C.Structure):
    _fields_ = [("mt", C.c_uint32 * 624), ("idx", C.c_int32)]


# name -> (restype, argtypes); must list every symbol include/b200xp.h declares.
_SIGNATURES = {
    "xp_version": (C.c_int, []),
    "xp_last_error": (C.c_char_p, []),
    "xp_launch_count": (C.c_uint64, []),
    "xp_device_info": (C.c_int, [C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "xp_copy_to_device": (C.c_int, [c_vp, c_vp, c_i64, c_vp]),
    "xp_copy_to_host": (C.c_int, [c_vp, c_vp, c_i64, c_vp]),
    "xp_bptt_plan": (C.c_int, [c_vp, c_i64, c_i64, c_i64, C.c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "xp_bptt_indices": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "xp_peer_allgather": (C.c_int, [c_vp, c_vp, C.c_int, C.c_int, C.c_uint32, c_vp, C.c_int, c_vp, c_vp]),
    "xp_peer_allgather_graph": (C.c_int, [c_vp, c_vp, c_vp, C.c_int, C.c_int, c_vp, c_vp, C.c_int, c_vp, c_vp]),
    "xp_peer_publish_allgather_graph": (C.c_int, [c_vp, c_vp, c_vp, C.c_int, C.c_int, c_vp, c_vp, C.c_int, c_vp, C.c_int,
                                                  c_vp, c_vp, c_vp]),
    "xp_peer_set_timeout_cycles": (None, [c_i64]),
    "xp_peer_ring_insert": (C.c_int, [c_vp, c_vp, C.c_int, C.c_int, C.c_uint32, c_vp, C.c_int, c_i64, c_i64,
                                      c_i64, c_vp, c_vp]),
    "xp_ring_insert": (C.c_int, [C.POINTER(XpKey), C.c_int, c_i64, c_i64, c_i64, c_vp]),
    "xp_gather": (C.c_int, [C.POINTER(XpKey), C.c_int, c_vp, c_i64, c_i64, c_vp, c_vp]),
    "xp_randperm_prefix": (C.c_int, [c_vp, c_i64, c_i64, c_i64, c_vp, c_vp]),
    "xp_philox_u32": (C.c_int, [C.c_uint64, C.c_uint64, c_vp, c_i64, c_vp]),
    "xp_tree_init": (C.c_int, [C.POINTER(XpTree), c_vp]),
    "xp_tree_set_batch": (C.c_int, [C.POINTER(XpTree), c_vp, c_vp, c_i64, c_vp, c_vp]),
    "xp_tree_query": (C.c_int, [C.POINTER(XpTree), c_i64, c_i64, c_vp, c_vp]),
    "xp_tree_retrieve": (C.c_int, [C.POINTER(XpTree), c_vp, c_i64, c_vp, c_vp]),
    "xp_per_add": (C.c_int, [C.POINTER(XpTree), c_vp, c_i64, c_i64, c_i64, C.c_double, c_vp]),
    "xp_per_sample": (C.c_int, [C.POINTER(XpTree), c_vp, c_i64, c_i64, c_i64, C.c_double, c_vp, c_vp, c_vp]),
    "xp_per_weights": (C.c_int, [C.POINTER(XpTree), c_vp, c_i64, c_i64, C.c_double, c_vp, c_vp]),
    "xp_per_update": (C.c_int, [C.POINTER(XpTree), c_vp, c_vp, c_vp, c_i64, c_i64, C.c_double, c_vp, c_vp]),
    "xp_per_sample_hostu": (C.c_int, [C.POINTER(XpTree), c_vp, c_i64, c_i64, C.c_double, c_vp, c_vp, c_vp]),
    "xp_per_update_hostp": (C.c_int, [C.POINTER(XpTree), c_vp, c_vp, c_vp, c_i64, c_i64, C.c_double, c_vp, c_vp]),
    "xp_pow_batch": (C.c_int, [c_vp, C.c_double, c_vp, c_i64, c_vp]),
    "xp_nstep_fold_insert": (C.c_int, [C.POINTER(XpKey), C.POINTER(C.c_int), C.c_int, C.c_int, C.c_int,
                                       C.c_int, C.c_int, C.POINTER(C.c_float), c_i64, c_i64, c_i64, c_vp]),
    "xp_gae": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_i64,
                         C.c_double, C.c_double, c_vp]),
    "xp_gae_tokens": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, C.c_double, C.c_double, C.c_int, c_vp]),
    "xp_gae_turns": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, C.c_int, C.c_double,
                               C.c_double, c_vp]),
    "xp_gae_turn_broadcast": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_vp]),
    "xp_standardize": (C.c_int, [c_vp, c_vp, c_i64, c_vp, c_vp]),
    "xp_standardize_static": (C.c_int, [c_vp, c_vp, c_i64, c_vp]),
    "xp_masked_whiten": (C.c_int, [c_vp, c_vp, c_vp, c_i64, C.c_int, c_vp, c_vp, c_vp]),
    "xp_gae_f32": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_i64,
                             C.c_double, C.c_double, c_vp]),
    "xp_mc_returns": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_i64,
                                C.c_double, c_vp]),
    "xp_mt19937_seed": (None, [C.POINTER(XpMt19937), C.c_uint32]),
    "xp_mt19937_draw": (None, [C.POINTER(XpMt19937), c_vp, c_i64]),
    "xp_mt19937_skip": (None, [C.POINTER(XpMt19937), c_i64]),
    "xp_host_randperm_prefix": (C.c_int, [c_vp, c_i64, c_i64, c_vp]),
}

_lib = None


class NativeError(RuntimeError):
    """libb200xp.so is missing or a call into it failed."""


def lib_path() -> Path:
    return Path(os.environ.get("B200XP_LIB", str(_LIB_PATH)))


def load():
    """Load libb200xp.so and bind every declared symbol. Raises NativeError if absent."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not path.exists():
        raise NativeError(
            f"{path} not found: build it with `python -m agilerl_b200.build` "
            "(there is no CPU fallback for the experience path)"
        )
    lib = C.CDLL(str(path))
    for name, (restype, argtypes) in _SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:  # pragma: no cover - build/ABI mismatch
            raise NativeError(f"{path} does not export {name}") from e
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def last_error() -> str:
    return load().xp_last_error().decode("utf-8", "replace")


def check(status: int, what: str = "") -> None:
    """Map an xp_status_t to the exception the reference raises for that condition."""
    if status == XP_OK:
        return
    msg = f"{what}: {last_error()}" if what else last_error()
    if status == XP_ERR_INDEX:
        raise AssertionError(msg)
    if status == XP_ERR_INVALID:
        raise ValueError(msg)
    raise NativeError(f"{msg} (status {status})")


def ptr(t) -> int:
    """Device (or host) address of a tensor, 0 for None."""
    return 0 if t is None else t.data_ptr()


def stream_ptr(device=None) -> int:
    import torch

    return torch.cuda.current_stream(device).cuda_stream


class _NoGuard:
    __slots__ = ()

    def __enter__(self):
        return None

    def __exit__(self, *a):
        return False


_NO_GUARD = _NoGuard()


def raw_stream(index: int) -> int:
    """cudaStream_t of torch's current stream on device ``index`` (no Python Stream object)."""
    import torch

    return torch._C._cuda_getCurrentRawStream(index)


def on_device(dev):
    """Device guard that costs nothing when ``dev`` already is the current device (the normal
    one-process-per-GPU case)."""
    import torch

    if torch._C._cuda_getDevice() == dev.index:
        return _NO_GUARD
    return torch.cuda.device(dev)


def require_cuda(device=None):
    """Return the torch CUDA device the engine stores on, or raise loudly."""
    import torch

    if not torch.cuda.is_available():
        raise NativeError(
            "agilerl_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback"
        )
    load()
    if device is None or str(device) == "cpu":
        return torch.device("cuda", torch.cuda.current_device())
    dev = torch.device(device)
    if dev.type != "cuda":
        raise NativeError(f"unsupported storage device {device!r}")
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    return dev
