"""ctypes binding of ``csrc/runtime/nxd_runtime.cpp`` (``libnxd_runtime.so``, host-only C++17, built in-tree with g++).

``SafetensorsFile``  mmap + multi-threaded copies of any set of tensors (or row ranges) into caller buffers -- the
                     checkpoint-loading half of the reference's ``ModelBuilder`` / runner (``trace/model_builder.py:302-321``
                     loads ``tp{r}_sharded_checkpoint.safetensors`` tensor by tensor through Python).
``NativeRouter``     (model key, batch, seq) -> compiled bucket lookup of ``NxDModel`` (``trace/model_builder.py:64-120``).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading
from typing import Dict, Iterable, List, Optional, Tuple

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_REPO = os.path.dirname(os.path.dirname(_HERE))
SRC = os.path.join(_REPO, "csrc", "runtime", "nxd_runtime.cpp")
LIB_PATH = os.path.join(_HERE, "libnxd_runtime.so")
_lib: Optional[C.CDLL] = None
_lock = threading.Lock()

_DTYPES = {"F64": torch.float64, "F32": torch.float32, "F16": torch.float16, "BF16": torch.bfloat16,
           "I64": torch.int64, "I32": torch.int32, "I16": torch.int16, "I8": torch.int8, "U8": torch.uint8,
           "BOOL": torch.bool, "F8_E4M3": getattr(torch, "float8_e4m3fn", torch.uint8),
           "F8_E5M2": getattr(torch, "float8_e5m2", torch.uint8)}


def build(force: bool = False) -> str:
    """Compile the runtime library in-tree (g++, ~2 s).  Returns its path."""
    if not force and os.path.exists(LIB_PATH) and os.path.getmtime(LIB_PATH) >= os.path.getmtime(SRC):
        return LIB_PATH
    tmp = LIB_PATH + f".tmp{os.getpid()}"
    cmd = [os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", "-o", tmp, SRC]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"building libnxd_runtime.so failed:\n{res.stderr}")
    os.replace(tmp, LIB_PATH)
    return LIB_PATH


def lib() -> C.CDLL:
    global _lib
    with _lock:
        if _lib is None:
            L = C.CDLL(build())
            L.nxd_rt_last_error.restype = C.c_char_p
            L.nxd_st_open.restype = C.c_void_p
            L.nxd_st_open.argtypes = [C.c_char_p]
            L.nxd_st_close.argtypes = [C.c_void_p]
            L.nxd_st_num_tensors.argtypes = [C.c_void_p]
            L.nxd_st_tensor_info.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_int, C.c_char_p, C.c_int,
                                             C.POINTER(C.c_int64), C.POINTER(C.c_int), C.POINTER(C.c_uint64)]
            L.nxd_st_find.argtypes = [C.c_void_p, C.c_char_p]
            L.nxd_st_read_many.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_void_p), C.POINTER(C.c_uint64),
                                           C.c_int, C.c_int]
            L.nxd_st_read_rows.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_uint64, C.c_int]
            L.nxd_router_create.restype = C.c_void_p
            L.nxd_router_destroy.argtypes = [C.c_void_p]
            L.nxd_router_add.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.c_int]
            L.nxd_router_route_exact.argtypes = [C.c_void_p, C.c_int, C.c_int]
            L.nxd_router_route_padded.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_int),
                                                  C.POINTER(C.c_int)]
            _lib = L
    return _lib


def available() -> bool:
    try:
        lib()
        return True
    except Exception:  # noqa: BLE001 - no compiler on the box
        return False


def _err() -> str:
    return (lib().nxd_rt_last_error() or b"").decode()


class SafetensorsFile:
    """One mmapped ``.safetensors`` file.  ``read(names)`` returns CPU tensors filled by ``threads`` copy threads
    (pinned when ``pin_memory``), ``read_rows`` only rows ``[r0, r1)`` of a tensor."""

    def __init__(self, path: str):
        self.path = path
        self._h = lib().nxd_st_open(path.encode())
        if not self._h:
            raise OSError(_err())
        self.info: Dict[str, Tuple[int, torch.dtype, Tuple[int, ...], int]] = {}
        name, dt = C.create_string_buffer(1024), C.create_string_buffer(32)
        shape, nd, nb = (C.c_int64 * 8)(), C.c_int(), C.c_uint64()
        for i in range(lib().nxd_st_num_tensors(self._h)):
            lib().nxd_st_tensor_info(self._h, i, name, 1024, dt, 32, shape, C.byref(nd), C.byref(nb))
            self.info[name.value.decode()] = (i, _DTYPES[dt.value.decode()], tuple(shape[d] for d in range(nd.value)),
                                              int(nb.value))

    def keys(self) -> List[str]:
        return list(self.info)

    def read(self, names: Optional[Iterable[str]] = None, threads: int = 8, pin_memory: bool = False
             ) -> Dict[str, torch.Tensor]:
        names = list(self.info) if names is None else list(names)
        pin = pin_memory and torch.cuda.is_available()
        out, idx, ptrs, caps = {}, [], [], []
        for n in names:
            i, dt, shape, nbytes = self.info[n]
            t = torch.empty(shape, dtype=dt, pin_memory=pin)
            out[n] = t
            idx.append(i)
            ptrs.append(t.data_ptr())
            caps.append(nbytes)
        k = len(names)
        rc = lib().nxd_st_read_many(self._h, (C.c_int * k)(*idx), (C.c_void_p * k)(*ptrs), (C.c_uint64 * k)(*caps), k,
                                    threads)
        if rc != 0:
            raise RuntimeError(_err())
        return out

    def read_rows(self, name: str, r0: int, r1: int, threads: int = 4, pin_memory: bool = False) -> torch.Tensor:
        i, dt, shape, _ = self.info[name]
        pin = pin_memory and torch.cuda.is_available()
        t = torch.empty((r1 - r0,) + tuple(shape[1:]), dtype=dt, pin_memory=pin)
        rc = lib().nxd_st_read_rows(self._h, i, r0, r1, t.data_ptr(), t.numel() * t.element_size(), threads)
        if rc != 0:
            raise RuntimeError(_err())
        return t

    def close(self) -> None:
        if self._h:
            lib().nxd_st_close(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


def load_safetensors(path: str, device=None, threads: int = 8) -> Dict[str, torch.Tensor]:
    """Whole file -> dict of tensors; with a CUDA ``device`` the host copies land in pinned memory and move to the GPU
    with non-blocking copies."""
    cuda = device is not None and torch.device(device).type == "cuda"
    with SafetensorsFile(path) as f:
        host = f.read(threads=threads, pin_memory=cuda)
    if device is None or torch.device(device).type == "cpu":
        return host
    out = {k: v.to(device, non_blocking=True) for k, v in host.items()}
    if cuda:
        torch.cuda.current_stream().synchronize()
    return out


class NativeRouter:
    """Bucket lookup in C++: ``add(key, batch, seq, graph_id)``, ``route_exact(batch, seq)``, ``route_padded(key, ...)``."""

    def __init__(self):
        self._h = lib().nxd_router_create()

    def add(self, key: str, batch: int, seq: int, graph: int) -> None:
        if lib().nxd_router_add(self._h, key.encode(), batch, seq, graph) != 0:
            raise ValueError(_err())

    def route_exact(self, batch: int, seq: int) -> int:
        return int(lib().nxd_router_route_exact(self._h, batch, seq))

    def route_padded(self, key: str, batch: int, seq: int) -> Tuple[int, int, int]:
        ob, os_ = C.c_int(), C.c_int()
        g = int(lib().nxd_router_route_padded(self._h, key.encode(), batch, seq, C.byref(ob), C.byref(os_)))
        return (g, ob.value, os_.value) if g >= 0 else (-1, 0, 0)

    def __del__(self):
        try:
            if self._h:
                lib().nxd_router_destroy(self._h)
                self._h = None
        except Exception:  # noqa: BLE001
            pass
