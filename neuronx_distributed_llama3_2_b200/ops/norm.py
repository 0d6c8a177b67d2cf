"""RMSNorm (optionally fused with the residual add) -- autograd wrapper."""
from __future__ import annotations

import ctypes as _c
from typing import Tuple

import torch

from . import _lib


def _use_kernel(x: torch.Tensor) -> bool:
    return x.is_cuda and x.dtype == torch.bfloat16 and x.shape[-1] % 8 == 0 and _lib.available()


def rms_norm_ref(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    xf = x.float()
    y = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    return (y * w.float()).to(x.dtype)


class _RMSNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, eps, residual):
        H = x.shape[-1]
        x2 = x.reshape(-1, H)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        res2 = res_out = None
        if residual is not None:
            res2 = residual.reshape(-1, H)
            if not res2.is_contiguous():
                res2 = res2.contiguous()
            res_out = torch.empty_like(x2)
        wc = w if w.dtype == torch.bfloat16 else w.to(torch.bfloat16)
        rc = _lib.lib().nxd_rmsnorm_fwd(_lib.ptr(x2), _lib.ptr(res2), _lib.ptr(wc), _lib.ptr(y), _lib.ptr(res_out),
                                        _lib.ptr(rstd), _c.c_int(rows), _c.c_int(H), _c.c_float(eps),
                                        _lib.stream_ptr())
        _lib.check(rc, "nxd_rmsnorm_fwd")
        normed_in = res_out if residual is not None else x2
        ctx.save_for_backward(normed_in, wc, rstd)
        ctx.has_residual = residual is not None
        ctx.w_dtype = w.dtype
        ctx.shape = x.shape
        if residual is not None:
            return y.reshape(x.shape), res_out.reshape(x.shape)
        return y.reshape(x.shape), None

    @staticmethod
    def backward(ctx, dy, dres_out):
        xin, wc, rstd = ctx.saved_tensors
        rows, H = xin.shape
        dy2 = dy.reshape(rows, H)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dres2 = None
        if dres_out is not None:
            dres2 = dres_out.reshape(rows, H)
            if not dres2.is_contiguous():
                dres2 = dres2.contiguous()
        dx = torch.empty_like(xin)
        P = _lib.lib().nxd_rmsnorm_bwd_num_partials(_c.c_int(rows))
        scratch = torch.empty(P * H, dtype=torch.float32, device=xin.device)
        dw = torch.empty(H, dtype=torch.float32, device=xin.device)
        rc = _lib.lib().nxd_rmsnorm_bwd(_lib.ptr(dy2), _lib.ptr(xin), _lib.ptr(wc), _lib.ptr(rstd), _lib.ptr(dres2),
                                        _lib.ptr(dx), _lib.ptr(None), _lib.ptr(dw), _c.c_int(0), _lib.ptr(scratch),
                                        _c.c_int(rows), _c.c_int(H), _lib.stream_ptr())
        _lib.check(rc, "nxd_rmsnorm_bwd")
        dx = dx.reshape(ctx.shape)
        # d(x + r) flows equally to x and to the residual input
        return dx, dw.to(ctx.w_dtype), None, (dx if ctx.has_residual else None)


def rms_norm(x: torch.Tensor, w: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
    if _use_kernel(x):
        return _RMSNormFn.apply(x, w, eps, None)[0]
    return rms_norm_ref(x, w, eps)


def fused_add_rms_norm(x: torch.Tensor, residual: torch.Tensor, w: torch.Tensor,
                       eps: float = 1e-6) -> Tuple[torch.Tensor, torch.Tensor]:
    """``s = x + residual; return rmsnorm(s), s`` in one pass over HBM."""
    if _use_kernel(x):
        return _RMSNormFn.apply(x, w, eps, residual)
    s = x + residual
    return rms_norm_ref(s, w, eps), s
