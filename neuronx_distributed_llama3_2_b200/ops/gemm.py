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


SMALLM_MAX_ROWS = 128


def gemm_smallm(x: torch.Tensor, w: torch.Tensor, residual: Optional[torch.Tensor] = None, splits: int = 0
                ) -> Optional[torch.Tensor]:
    """``x[B, K] @ w[N, K]^T (+ residual)`` for ``B <= 128`` rows on the tensor cores with the WEIGHT rows as the UMMA M
    dimension (``csrc/gemm_smallm.cu``): a pure stream of the weight matrix, split-K over a thread-block cluster for
    narrow layers.  The batched token-generation GEMM (reference: the parallel linear layers of
    ``parallel_layers/layers.py`` at ``[batch, 1, hidden]``).  ``None`` when the shape does not qualify."""
    B, K = x.shape
    N = w.shape[0]
    if (not (x.is_cuda and w.is_cuda) or B == 0 or B > SMALLM_MAX_ROWS or x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16 or K % 8
            or x.stride(1) != 1 or w.stride(1) != 1 or x.stride(0) % 8 or w.stride(0) % 8
            or x.data_ptr() % 16 or w.data_ptr() % 16):
        return None
    out = torch.empty((B, N), dtype=torch.bfloat16, device=x.device)
    ldr = 0
    if residual is not None:
        residual = residual.reshape(B, N)
        if residual.stride(1) != 1:
            residual = residual.contiguous()
        ldr = residual.stride(0)
    rc = _lib.lib().nxd_gemm_smallm_bf16(
        _lib.ptr(x), _lib.ptr(w), _lib.ptr(residual), _lib.ptr(out), _c.c_int(B), _c.c_int(N), _c.c_int(K),
        _c.c_longlong(x.stride(0)), _c.c_longlong(w.stride(0)), _c.c_longlong(out.stride(0)), _c.c_longlong(ldr),
        _c.c_int(splits), _lib.stream_ptr())
    _lib.check(rc, "nxd_gemm_smallm_bf16")
    return out
