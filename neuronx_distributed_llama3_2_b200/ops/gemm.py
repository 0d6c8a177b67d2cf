"""ctypes binding of the tcgen05 GEMM (``csrc/gemm_tcgen05.cu``)."""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from . import _lib

_LAYOUT = {"nt": 0, "nn": 1, "tn": 2}
_c = ctypes


def gemm(a: torch.Tensor, b: torch.Tensor, layout: str, out: Optional[torch.Tensor] = None,
         out_dtype: Optional[torch.dtype] = None, accumulate: bool = False, block_n: int = 0) -> torch.Tensor:
    """``D = A_op @ B_op^T`` on the 5th-gen tensor cores.  ``a``/``b`` are 2-D bf16 with unit
    inner stride (row pitch may exceed the row length, e.g. a column slice of a fused tensor)."""
    assert a.dim() == 2 and b.dim() == 2 and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    if a.stride(1) != 1:
        a = a.contiguous()
    if b.stride(1) != 1:
        b = b.contiguous()
    if layout == "nt":
        m, k = a.shape
        n = b.shape[0]
        assert b.shape[1] == k
    elif layout == "nn":
        m, k = a.shape
        n = b.shape[1]
        assert b.shape[0] == k
    else:
        k, m = a.shape
        n = b.shape[1]
        assert b.shape[0] == k
    if out is None:
        dt = out_dtype or torch.bfloat16
        assert not accumulate
        out = torch.empty((m, n), dtype=dt, device=a.device)
    assert out.shape == (m, n) and out.stride(1) == 1
    if out.dtype == torch.bfloat16:
        assert not accumulate, "accumulate needs an fp32 output"
        mode = 0
    elif out.dtype == torch.float32:
        mode = 2 if accumulate else 1
    else:
        raise TypeError(f"unsupported GEMM output dtype {out.dtype}")
    rc = _lib.lib().nxd_gemm_bf16(
        _c.c_int(_LAYOUT[layout]), _lib.ptr(a), _lib.ptr(b), _lib.ptr(out), _c.c_int(m), _c.c_int(n), _c.c_int(k),
        _c.c_longlong(a.stride(0)), _c.c_longlong(b.stride(0)), _c.c_longlong(out.stride(0)), _c.c_int(mode),
        _c.c_int(block_n), _lib.stream_ptr())
    _lib.check(rc, "nxd_gemm_bf16")
    return out


def gemm_grouped(a: torch.Tensor, b: torch.Tensor, layout: str, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``E`` GEMMs of one shape in ONE launch (``nxd_gemm_grouped_bf16``: 3-D tensor maps, the persistent tile loop runs
    over every expert's tiles; rows past an expert's extent are out of bounds of the map, so the row count needs no
    alignment).  ``a``/``b`` are 3-D bf16 with a unit inner stride:
    ``nn``  ``a [E, M, K] @ b [E, K, N]``     ``nt``  ``a [E, M, K] @ b [E, N, K]^T``     ``tn``  ``a [E, K, M]^T @ b [E, K, N]``."""
    assert a.dim() == 3 and b.dim() == 3 and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    if a.stride(2) != 1:
        a = a.contiguous()
    if b.stride(2) != 1:
        b = b.contiguous()
    E = a.shape[0]
    if layout == "nt":
        m, k, n = a.shape[1], a.shape[2], b.shape[1]
    elif layout == "nn":
        m, k, n = a.shape[1], a.shape[2], b.shape[2]
    else:
        k, m, n = a.shape[1], a.shape[2], b.shape[2]
    if out is None:
        out = torch.empty((E, m, n), dtype=torch.bfloat16, device=a.device)
    assert out.shape == (E, m, n) and out.stride(2) == 1 and out.dtype == torch.bfloat16
    rc = _lib.lib().nxd_gemm_grouped_bf16(
        _c.c_int(_LAYOUT[layout]), _lib.ptr(a), _lib.ptr(b), _lib.ptr(out), _c.c_int(E), _c.c_int(m), _c.c_int(n),
        _c.c_int(k), _c.c_longlong(a.stride(1)), _c.c_longlong(b.stride(1)), _c.c_longlong(out.stride(1)),
        _c.c_longlong(a.stride(0)), _c.c_longlong(b.stride(0)), _c.c_longlong(out.stride(0)), _c.c_int(0),
        _lib.stream_ptr())
    _lib.check(rc, "nxd_gemm_grouped_bf16")
    return out
