"""GEMM dispatch: the three operand layouts a Linear needs in fwd + bwd.

All matrices are row-major 2-D views:

* ``gemm_nt(a[M,K], b[N,K]) -> [M,N]``    forward     ``Y = X W^T``
* ``gemm_nn(a[M,K], b[K,N]) -> [M,N]``    dgrad       ``dX = dY W``
* ``gemm_tn(a[K,M], b[K,N]) -> [M,N]``    wgrad       ``dW = dY^T X``

On a B200 with the kernel library loaded, bf16 inputs go to the hand-written
tcgen05/TMA kernel in ``csrc/gemm_tcgen05.cu`` (all three layouts are handled by
operand-major bits in the UMMA descriptors, no transposed copies).  Everything
else (CPU tests, fp32, odd shapes) takes the plain ``torch.matmul`` path.

``ops.backend.set_backends(gemm=...)`` (default from ``NXD_GEMM_BACKEND``) = ``tcgen05`` | ``cublas`` selects the
library GEMM instead; the latter is the NCCL+cuBLAS *baseline* arm.  A CUDA call that wanted the kernel but did not
qualify is counted by ``ops.backend.note_fallback``.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib, backend


def _backend() -> str:
    return backend.gemm_backend()


def _use_kernel(a: torch.Tensor, b: torch.Tensor, m: int, n: int, k: int, layout: str = "nt") -> bool:
    if not (a.is_cuda and b.is_cuda):
        return False
    if _backend() != "tcgen05":
        return False
    if not _lib.available():
        return False
    if a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16:
        backend.note_fallback("gemm", f"dtype {a.dtype}/{b.dtype} (kernel is bf16)")
        return False
    # TMA needs 16-byte aligned row pitches (K and N multiples of 8 elements); M is free (OOB rows are clipped).
    if (k % 8) or (n % 8) or (layout == "tn" and m % 8):
        backend.note_fallback("gemm", f"{layout} M={m} N={n} K={k}: a leading dimension is not a multiple of 8")
        return False
    return True


def gemm_nt(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None,
            out_dtype: Optional[torch.dtype] = None, accumulate: bool = False) -> torch.Tensor:
    """``a[M,K] @ b[N,K]^T``.  ``accumulate`` adds into ``out`` (fp32 grad accumulation)."""
    m, k = a.shape
    n = b.shape[0]
    if (m <= 128 and out is None and not accumulate and out_dtype in (None, torch.bfloat16) and a.is_cuda
            and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and _backend() == "tcgen05" and _lib.available()):
        # token generation: weight-streaming kernels instead of a 256-row tensor-core tile that is mostly padding --
        # FMA GEMV up to 8 rows, above that the tensor-core kernel with the weight rows as the UMMA M dimension
        if m <= 8:
            from .fused_comm import skinny_gemm

            res = skinny_gemm(a, b)
        else:
            from .gemm import gemm_smallm

            res = gemm_smallm(a, b)
        if res is not None:
            return res
    if _use_kernel(a, b, m, n, k):
        from .gemm import gemm

        return gemm(a, b, "nt", out=out, out_dtype=out_dtype, accumulate=accumulate)
    res = torch.matmul(a, b.t())
    return _finish(res, out, out_dtype, accumulate)


def gemm_nn(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None,
            out_dtype: Optional[torch.dtype] = None, accumulate: bool = False) -> torch.Tensor:
    """``a[M,K] @ b[K,N]``."""
    m, k = a.shape
    n = b.shape[1]
    if _use_kernel(a, b, m, n, k, "nn"):
        from .gemm import gemm

        return gemm(a, b, "nn", out=out, out_dtype=out_dtype, accumulate=accumulate)
    res = torch.matmul(a, b)
    return _finish(res, out, out_dtype, accumulate)


def gemm_tn(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None,
            out_dtype: Optional[torch.dtype] = None, accumulate: bool = False) -> torch.Tensor:
    """``a[K,M]^T @ b[K,N]``."""
    k, m = a.shape
    n = b.shape[1]
    if _use_kernel(a, b, m, n, k, "tn"):
        from .gemm import gemm

        return gemm(a, b, "tn", out=out, out_dtype=out_dtype, accumulate=accumulate)
    if out is not None and not accumulate and out.dtype == a.dtype and out.is_contiguous():
        return torch.mm(a.t(), b, out=out)
    res = torch.matmul(a.t(), b)
    return _finish(res, out, out_dtype, accumulate)


def _finish(res, out, out_dtype, accumulate):
    if out is not None:
        if accumulate:
            out.add_(res.to(out.dtype))
        else:
            out.copy_(res)
        return out
    if out_dtype is not None and res.dtype != out_dtype:
        res = res.to(out_dtype)
    return res
