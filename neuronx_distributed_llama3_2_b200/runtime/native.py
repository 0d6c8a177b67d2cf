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
            L.nxd_kv_create.restype = C.c_void_p
            L.nxd_kv_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
            L.nxd_kv_destroy.argtypes = [C.c_void_p]
            L.nxd_kv_num_free.argtypes = [C.c_void_p]
            L.nxd_kv_line_tokens.restype = C.c_int64
            L.nxd_kv_line_tokens.argtypes = [C.c_void_p, C.c_int]
            L.nxd_kv_pages_needed.argtypes = [C.c_void_p, C.c_int, C.c_int64]
            L.nxd_kv_reserve.argtypes = [C.c_void_p, C.c_int, C.c_int64]
            L.nxd_kv_free_line.argtypes = [C.c_void_p, C.c_int]
            L.nxd_kv_truncate.argtypes = [C.c_void_p, C.c_int, C.c_int64]
            L.nxd_kv_fork.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
            L.nxd_kv_page_shared.argtypes = [C.c_void_p, C.c_int, C.c_int64]
            L.nxd_kv_unshare.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.POINTER(C.c_int), C.POINTER(C.c_int)]
            L.nxd_kv_block_row.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int32), C.c_int]
            L.nxd_sched_create.restype = C.c_void_p
            L.nxd_sched_create.argtypes = [C.c_void_p, C.c_int]
            L.nxd_sched_destroy.argtypes = [C.c_void_p]
            L.nxd_sched_submit.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int]
            L.nxd_sched_admit.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int), C.c_int]
            L.nxd_sched_step.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.c_int, C.POINTER(C.c_int64),
                                         C.POINTER(C.c_int), C.POINTER(C.c_int64), C.POINTER(C.c_int), C.c_int]
            L.nxd_sched_num_waiting.argtypes = [C.c_void_p]
            L.nxd_sched_num_running.argtypes = [C.c_void_p]
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


class PagedKVPool:
    """Host-side page allocator of a paged KV cache (``nxd_kv_*``): cache lines own ordered page lists, pages are
    reference counted (``fork`` shares full pages copy-on-write), the pool may be smaller than ``lines * max_pages``."""

    def __init__(self, num_pages: int, page_size: int, max_lines: int, max_pages_per_line: int):
        self.num_pages, self.page_size, self.max_lines, self.max_pages = num_pages, page_size, max_lines, max_pages_per_line
        self._h = lib().nxd_kv_create(num_pages, page_size, max_lines, max_pages_per_line)
        if not self._h:
            raise ValueError(_err())

    @property
    def num_free(self) -> int:
        return lib().nxd_kv_num_free(self._h)

    def line_tokens(self, line: int) -> int:
        return int(lib().nxd_kv_line_tokens(self._h, line))

    def pages_needed(self, line: int, n_tokens: int) -> int:
        return lib().nxd_kv_pages_needed(self._h, line, n_tokens)

    def reserve(self, line: int, n_tokens: int) -> int:
        """Room for ``n_tokens`` tokens in total on ``line``; pages newly allocated.  ``MemoryError`` when the pool is dry."""
        rc = lib().nxd_kv_reserve(self._h, line, n_tokens)
        if rc < 0:
            raise MemoryError(_err())
        return rc

    def free_line(self, line: int) -> int:
        return lib().nxd_kv_free_line(self._h, line)

    def truncate(self, line: int, n_tokens: int) -> int:
        return lib().nxd_kv_truncate(self._h, line, n_tokens)

    def fork(self, src: int, dst: int) -> Optional[Tuple[int, int]]:
        """``dst`` shares ``src``'s full pages; returns ``(src_page, dst_page)`` to copy on the device when the last page
        was partially filled, else ``None``."""
        a, b = C.c_int(-1), C.c_int(-1)
        if lib().nxd_kv_fork(self._h, src, dst, C.byref(a), C.byref(b)) != 0:
            raise MemoryError(_err())
        return (a.value, b.value) if a.value >= 0 else None

    def page_shared(self, line: int, pos: int) -> bool:
        return lib().nxd_kv_page_shared(self._h, line, pos) == 1

    def unshare(self, line: int, pos: int) -> Optional[Tuple[int, int]]:
        a, b = C.c_int(-1), C.c_int(-1)
        if lib().nxd_kv_unshare(self._h, line, pos, C.byref(a), C.byref(b)) < 0:
            raise MemoryError(_err())
        return (a.value, b.value) if a.value >= 0 else None

    def block_row(self, line: int, fill: int = 0) -> Tuple[List[int], int]:
        buf = (C.c_int32 * self.max_pages)()
        n = lib().nxd_kv_block_row(self._h, line, buf, fill)
        return list(buf), n

    def close(self) -> None:
        if self._h:
            lib().nxd_kv_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


class NativeScheduler:
    """Continuous-batching admission / retirement / preemption over a ``PagedKVPool`` (``nxd_sched_*``).  FIFO admission
    while a cache line is free and the pool holds the prompt plus one page of head room per running sequence; when a
    running sequence cannot get its next page the youngest are preempted back to the front of the queue."""

    def __init__(self, pool: PagedKVPool, max_running: int):
        self.pool = pool
        self._h = lib().nxd_sched_create(pool._h, max_running)
        if not self._h:
            raise ValueError(_err())
        self._cap = max(max_running, 1)

    def submit(self, req_id: int, prompt_len: int, max_new_tokens: int) -> None:
        if lib().nxd_sched_submit(self._h, req_id, prompt_len, max_new_tokens) != 0:
            raise ValueError(_err())

    def admit(self) -> List[Tuple[int, int]]:
        ids, lines = (C.c_int64 * self._cap)(), (C.c_int * self._cap)()
        n = lib().nxd_sched_admit(self._h, ids, lines, self._cap)
        return [(int(ids[i]), int(lines[i])) for i in range(n)]

    def step(self, finished: Iterable[int] = ()) -> Tuple[List[int], List[int], int]:
        """After one model call for every running request.  Returns ``(retired ids, preempted ids, still running)``."""
        fin = list(finished)
        arr = (C.c_int64 * max(len(fin), 1))(*fin)
        done, pre = (C.c_int64 * self._cap)(), (C.c_int64 * self._cap)()
        nd, np_ = C.c_int(0), C.c_int(0)
        running = lib().nxd_sched_step(self._h, arr, len(fin), done, C.byref(nd), pre, C.byref(np_), self._cap)
        return [int(done[i]) for i in range(nd.value)], [int(pre[i]) for i in range(np_.value)], running

    @property
    def num_waiting(self) -> int:
        return lib().nxd_sched_num_waiting(self._h)

    @property
    def num_running(self) -> int:
        return lib().nxd_sched_num_running(self._h)

    def close(self) -> None:
        if self._h:
            lib().nxd_sched_destroy(self._h)
            self._h = None
