"""Build recipe for libb200xp.so (sm_100a only, in-tree).

``python -m agilerl_b200.build`` compiles every ``csrc/*.cu`` with nvcc for
``compute_100a/sm_100a`` (cross-compiles without a GPU) and links them into
``agilerl_b200/_lib/libb200xp.so``.  The built library is git-ignored but travels to the GPU
box with the repo snapshot.
"""

from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
LIB_DIR = PKG / "_lib"
OBJ_DIR = LIB_DIR / "obj"
LIB = LIB_DIR / "libb200xp.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false",  # every fp op on this path is separately rounded, like the reference
    "-Xcompiler", "-fPIC,-O2,-Wall",
    "-I", str(ROOT / "include"), "-I", str(CSRC),
    "-DXP_BUILDING=1",
]


def _nvcc() -> str:
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found; libb200xp needs the CUDA 12.9 toolkit")
    return nvcc


def _digest(src: Path) -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    h.update(src.read_bytes())
    for hdr in sorted(list(CSRC.glob("*.cuh")) + list((ROOT / "include").glob("*.h"))):
        h.update(hdr.read_bytes())
    return h.hexdigest()


def _compile(src: Path, verbose: bool) -> Path:
    obj = OBJ_DIR / (src.stem + ".o")
    stamp = OBJ_DIR / (src.stem + ".sha")
    dg = _digest(src)
    if obj.exists() and stamp.exists() and stamp.read_text() == dg:
        return obj
    cmd = [_nvcc(), *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
    if verbose:
        sys.stderr.write(r.stderr)
    stamp.write_text(dg)
    return obj


def build(verbose: bool = False, force: bool = False) -> Path:
    """Compile and link libb200xp.so; returns its path. Incremental unless ``force``."""
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    if force:
        for f in OBJ_DIR.glob("*"):
            f.unlink()
    srcs = sorted(CSRC.glob("*.cu"))
    if not srcs:
        raise RuntimeError(f"no CUDA sources under {CSRC}")
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), srcs))
    newest = max(o.stat().st_mtime for o in objs)
    if LIB.exists() and LIB.stat().st_mtime >= newest and not force:
        return LIB
    cmd = [_nvcc(), "-shared", "-o", str(LIB), *map(str, objs),
           "-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static",
           "-Xlinker", "--exclude-libs,ALL"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


HOSTACCEL_SRC = CSRC / "hostaccel.cpp"
HOSTACCEL = LIB_DIR / "_hostaccel.so"


def build_hostaccel(force: bool = False) -> Path:
    """Compile the optional torch/pybind11 host fast path (g++, no CUDA code of its own)."""
    import sysconfig

    import torch
    from torch.utils import cpp_extension as ce

    LIB_DIR.mkdir(parents=True, exist_ok=True)
    h = hashlib.sha256()
    h.update(HOSTACCEL_SRC.read_bytes())
    h.update((ROOT / "include" / "b200xp.h").read_bytes())
    h.update(torch.__version__.encode())
    stamp = OBJ_DIR / "hostaccel.sha"
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    if HOSTACCEL.exists() and stamp.exists() and stamp.read_text() == h.hexdigest() and not force:
        return HOSTACCEL
    cuda_home = os.path.dirname(os.path.dirname(_nvcc()))
    inc = ce.include_paths() + [sysconfig.get_paths()["include"], os.path.join(cuda_home, "include"),
                                str(ROOT / "include")]
    lib = ce.library_paths()[0]
    cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-w", str(HOSTACCEL_SRC), "-o", str(HOSTACCEL),
           "-DTORCH_EXTENSION_NAME=_hostaccel", "-DTORCH_API_INCLUDE_EXTENSION_H",
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
           *[f"-I{i}" for i in inc], f"-L{lib}", "-ltorch", "-ltorch_cpu", "-ltorch_python", "-lc10",
           "-lc10_cuda", "-ltorch_cuda", f"-Wl,-rpath,{lib}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"g++ failed for hostaccel.cpp:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(h.hexdigest())
    return HOSTACCEL


if __name__ == "__main__":
    p = build(verbose="-v" in sys.argv, force="-f" in sys.argv)
    print(p)
    print(build_hostaccel(force="-f" in sys.argv))
