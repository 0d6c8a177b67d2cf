"""Loader/builder for the in-tree sm_100a kernel library ``libnxd_b200.so``.

The library is plain C-ABI (declarations in ``include/nxd_b200.h``): every entry
point takes raw device pointers, sizes and a ``cudaStream_t`` and returns 0 or a
non-zero error code (``nxd_last_error()`` gives the text).  Python binds it with
``ctypes`` and passes ``tensor.data_ptr()`` plus the current torch stream, so the
kernels run in-order with the rest of the eager program and are CUDA-graph
capturable.

There is no silent fallback on a GPU machine: if CUDA is available but the
library is missing, ``available()`` raises unless ``NXD_ALLOW_TORCH_FALLBACK=1``.
"""
from __future__ import annotations

import ctypes
import glob
import os
import subprocess
import sys
import threading
from concurrent.futures import ThreadPoolExecutor
from typing import List, Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_REPO = os.path.dirname(os.path.dirname(_HERE))
CSRC_DIR = os.path.join(_REPO, "csrc")
INCLUDE_DIR = os.path.join(_REPO, "include")
LIB_PATH = os.environ.get("NXD_B200_LIB") or os.path.join(_HERE, "libnxd_b200.so")   # override: kernel experiments
BUILD_DIR = os.environ.get("NXD_BUILD_DIR") or os.path.join(_REPO, "build", "obj")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr", "--use_fast_math",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
] + [f for f in os.environ.get("NXD_NVCC_EXTRA", "").split() if f]

_lib: Optional[ctypes.CDLL] = None
_lock = threading.Lock()
launch_count = 0  # number of our kernels' entry points invoked (bench "gpu_launches")


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def sources() -> List[str]:
    return sorted(glob.glob(os.path.join(CSRC_DIR, "*.cu")))


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = sources() + glob.glob(os.path.join(CSRC_DIR, "*.cuh")) + glob.glob(os.path.join(INCLUDE_DIR, "*.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every ``csrc/*.cu`` for sm_100a and link ``libnxd_b200.so`` in-tree.

    Works without a GPU (nvcc cross-compiles).  Per-file objects are cached under
    ``build/obj`` and rebuilt only when the source (or any header) is newer.
    """
    if not force and not _stale():
        return LIB_PATH
    os.makedirs(BUILD_DIR, exist_ok=True)
    hdrs = glob.glob(os.path.join(CSRC_DIR, "*.cuh")) + glob.glob(os.path.join(INCLUDE_DIR, "*.h"))
    hdr_time = max([os.path.getmtime(h) for h in hdrs], default=0.0)
    nvcc = _nvcc()
    logs = {}

    def compile_one(src: str) -> str:
        obj = os.path.join(BUILD_DIR, os.path.basename(src)[:-3] + ".o")
        if (not force and os.path.exists(obj) and os.path.getmtime(obj) > os.path.getmtime(src)
                and os.path.getmtime(obj) > hdr_time):
            return obj
        cmd = [nvcc] + NVCC_FLAGS + ["-I", INCLUDE_DIR, "-I", CSRC_DIR, "-c", src, "-o", obj]
        res = subprocess.run(cmd, capture_output=True, text=True)
        logs[src] = res.stderr
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{res.stdout}\n{res.stderr}")
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, sources()))
    cmd = [nvcc, "-shared", "-o", LIB_PATH] + objs + ["-lcudart"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    if logs:
        with open(os.path.join(BUILD_DIR, "ptxas_v.log"), "w") as f:
            for src, text in logs.items():
                f.write(f"==== {os.path.basename(src)}\n{text}\n")
    if verbose:
        for src, text in logs.items():
            sys.stderr.write(f"==== {src}\n{text}\n")
    return LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                build()
            handle = ctypes.CDLL(LIB_PATH)
            handle.nxd_last_error.restype = ctypes.c_char_p
            _lib = handle
    return _lib


def available() -> bool:
    """True when our CUDA kernels are usable (a GPU is present and the .so loads)."""
    if not torch.cuda.is_available():
        return False
    if os.environ.get("NXD_FORCE_TORCH", "0") == "1":
        return False
    try:
        lib()
        return True
    except Exception as e:
        if os.environ.get("NXD_ALLOW_TORCH_FALLBACK", "0") == "1":
            return False
        raise RuntimeError(
            f"CUDA is available but the sm_100a kernel library could not be loaded from {LIB_PATH}: {e}. "
            "Run `python -c 'import __graft_entry__ as g; g.build()'` (or set NXD_ALLOW_TORCH_FALLBACK=1)."
        ) from e


def kernel_launches() -> int:
    """Kernel launches made by the library in this process so far (counted at every ``<<<>>>`` in csrc/)."""
    fn = lib().nxd_kernel_launches
    fn.restype = ctypes.c_ulonglong
    return int(fn())


def stream_ptr() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def check(code: int, what: str) -> None:
    global launch_count
    launch_count += 1
    if code != 0:
        msg = lib().nxd_last_error()
        raise RuntimeError(f"{what} failed (code {code}): {msg.decode() if msg else ''}")


def ptr(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())
