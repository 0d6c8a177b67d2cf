"""Process-wide backend selection and fallback accounting for the hot ops.

Three switches decide which implementation an op takes on a CUDA tensor:

* ``gemm``        ``tcgen05`` (our kernel, default) | ``cublas`` (``torch.matmul``)
* ``attn``        ``nxd`` (our tcgen05 flash attention, default) | ``sdpa`` (PyTorch SDPA / cuDNN)
* ``fused_comm``  ``True`` (AG->GEMM / GEMM->RS / GEMM->AR kernels over peer memory, default) | ``False`` (NCCL + GEMM)

The defaults come from ``NXD_GEMM_BACKEND`` / ``NXD_ATTN_BACKEND`` / ``NXD_FUSED_COMM`` once, at import; after that
they are *arguments*: ``set_backends(...)`` or the ``use_backends(...)`` context manager change them for the calls
that follow, so one process can time the library arm (``cublas`` + ``sdpa`` + NCCL: the comparison arm of
BASELINE.md section 3) right after the native arm on the same weights and batches (``bench.py``).

Every time an op that *wanted* its native kernel drops to a library path on a CUDA tensor (odd shape, unsupported
head dim, ragged rows, ...) it calls ``note_fallback``: the event is counted per (op, reason), logged once, and the
counters are printed in the bench line (``fallbacks`` must be empty for the headline configuration).  Choosing the
library arm explicitly is not a fallback and is not counted.
"""
from __future__ import annotations

import contextlib
import os
import threading
from collections import Counter
from typing import Dict, Optional

_lock = threading.Lock()


class _State:
    gemm: str = os.environ.get("NXD_GEMM_BACKEND", "tcgen05")
    attn: str = os.environ.get("NXD_ATTN_BACKEND", "nxd")
    fused_comm: bool = os.environ.get("NXD_FUSED_COMM", "1") == "1"
    fused_moe: bool = os.environ.get("NXD_FUSED_MOE", "1") == "1"


_state = _State()
_fallbacks: Counter = Counter()
_logged = set()


def gemm_backend() -> str:
    return _state.gemm


def attn_backend() -> str:
    return _state.attn


def fused_comm_enabled() -> bool:
    return _state.fused_comm


def fused_moe_enabled() -> bool:
    return _state.fused_moe and _state.fused_comm


def set_backends(gemm: Optional[str] = None, attn: Optional[str] = None, fused_comm: Optional[bool] = None,
                 fused_moe: Optional[bool] = None) -> Dict[str, object]:
    """Change the process-wide selection; returns the previous values (pass them back to restore)."""
    prev = {"gemm": _state.gemm, "attn": _state.attn, "fused_comm": _state.fused_comm, "fused_moe": _state.fused_moe}
    if gemm is not None:
        assert gemm in ("tcgen05", "cublas"), gemm
        _state.gemm = gemm
    if attn is not None:
        assert attn in ("nxd", "sdpa", "ref"), attn
        _state.attn = attn
    if fused_comm is not None:
        _state.fused_comm = bool(fused_comm)
    if fused_moe is not None:
        _state.fused_moe = bool(fused_moe)
    return prev


LIBRARY_ARM = {"gemm": "cublas", "attn": "sdpa", "fused_comm": False}
NATIVE_ARM = {"gemm": "tcgen05", "attn": "nxd", "fused_comm": True}


@contextlib.contextmanager
def use_backends(**kw):
    prev = set_backends(**kw)
    try:
        yield
    finally:
        set_backends(**prev)


def describe() -> Dict[str, object]:
    return {"gemm": _state.gemm, "attn": _state.attn, "fused_comm": _state.fused_comm}


def note_fallback(op: str, reason: str) -> None:
    """An op on a CUDA tensor left its native kernel for a library path although the native backend is selected."""
    key = f"{op}: {reason}"
    with _lock:
        _fallbacks[key] += 1
        first = key not in _logged
        _logged.add(key)
    if first:
        from ..utils.logger import get_logger

        get_logger().warning("native kernel not used -- %s (counted in ops.backend.fallback_counts())", key)


def fallback_counts() -> Dict[str, int]:
    with _lock:
        return dict(_fallbacks)


def reset_fallback_counts() -> None:
    with _lock:
        _fallbacks.clear()
