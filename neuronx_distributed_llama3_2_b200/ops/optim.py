"""Flat-buffer optimizer kernels: sum of squares, clip coefficient, fused AdamW."""
from __future__ import annotations

import ctypes as _c
from typing import Optional

import torch

from . import _lib

_scratch = {}


def _sumsq_scratch(device) -> torch.Tensor:
    key = (device.type, device.index)
    if key not in _scratch:
        n = _lib.lib().nxd_sumsq_num_partials()
        _scratch[key] = torch.empty(n, dtype=torch.float32, device=device)
    return _scratch[key]


def _dtype_code(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return 0
    if t.dtype == torch.bfloat16:
        return 1
    raise TypeError(f"unsupported dtype {t.dtype}")


def _use_kernel(t: torch.Tensor) -> bool:
    return t.is_cuda and t.dtype in (torch.float32, torch.bfloat16) and t.is_contiguous() and _lib.available()


def sumsq_(x: torch.Tensor, out: torch.Tensor, accumulate: bool = False) -> torch.Tensor:
    """``out[0] (+)= sum(x**2)`` (fp32 scalar tensor on the same device), deterministic two-stage."""
    if _use_kernel(x):
        rc = _lib.lib().nxd_sumsq(_lib.ptr(x), _c.c_longlong(x.numel()), _c.c_int(_dtype_code(x)),
                                  _lib.ptr(_sumsq_scratch(x.device)), _lib.ptr(out), _c.c_int(1 if accumulate else 0),
                                  _lib.stream_ptr())
        _lib.check(rc, "nxd_sumsq")
        return out
    s = x.float().pow(2).sum()
    if accumulate:
        out.add_(s)
    else:
        out.copy_(s.reshape(out.shape))
    return out


def clip_coef_(total_sumsq: torch.Tensor, max_norm: float, coef: torch.Tensor,
               norm_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``coef = min(1, max_norm / (sqrt(total) + 1e-6))`` computed on device (no host sync)."""
    if total_sumsq.is_cuda and _lib.available():
        rc = _lib.lib().nxd_clip_coef(_lib.ptr(total_sumsq), _c.c_float(max_norm), _lib.ptr(coef),
                                      _lib.ptr(norm_out), _lib.stream_ptr())
        _lib.check(rc, "nxd_clip_coef")
        return coef
    norm = total_sumsq.sqrt()
    if norm_out is not None:
        norm_out.copy_(norm.reshape(norm_out.shape))
    c = torch.clamp(max_norm / (norm + 1e-6), max=1.0) if max_norm > 0 else torch.ones_like(norm)
    coef.copy_(c.reshape(coef.shape))
    return coef


def adamw_step_(master: torch.Tensor, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor, grad: torch.Tensor,
                param_out: Optional[torch.Tensor], lr: float, beta1: float, beta2: float, eps: float,
                weight_decay: float, step: int, grad_scale: float = 1.0,
                clip_coef: Optional[torch.Tensor] = None) -> None:
    """One fused AdamW update on flat buffers.  ``master/exp_avg/exp_avg_sq`` fp32, ``grad`` fp32
    or bf16, ``param_out`` (bf16, optional) receives the rounded new parameter."""
    if _use_kernel(master) and _use_kernel(grad):
        rc = _lib.lib().nxd_adamw_step(
            _lib.ptr(master), _lib.ptr(exp_avg), _lib.ptr(exp_avg_sq), _lib.ptr(grad), _c.c_int(_dtype_code(grad)),
            _lib.ptr(param_out), _c.c_longlong(master.numel()), _c.c_float(lr), _c.c_float(beta1), _c.c_float(beta2),
            _c.c_float(eps), _c.c_float(weight_decay), _c.c_int(step), _c.c_float(grad_scale), _lib.ptr(clip_coef),
            _lib.stream_ptr())
        _lib.check(rc, "nxd_adamw_step")
        return
    g = grad.float() * grad_scale
    if clip_coef is not None:
        g = g * clip_coef
    exp_avg.mul_(beta1).add_(g, alpha=1 - beta1)
    exp_avg_sq.mul_(beta2).addcmul_(g, g, value=1 - beta2)
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    master.mul_(1 - lr * weight_decay)
    master.addcdiv_(exp_avg / bc1, (exp_avg_sq / bc2).sqrt().add_(eps), value=-lr)
    if param_out is not None:
        param_out.copy_(master.to(param_out.dtype))


def scale_(x: torch.Tensor, scale: float = 1.0, dev_scale: Optional[torch.Tensor] = None) -> torch.Tensor:
    if _use_kernel(x):
        rc = _lib.lib().nxd_scale_inplace(_lib.ptr(x), _c.c_longlong(x.numel()), _c.c_int(_dtype_code(x)),
                                          _c.c_float(scale), _lib.ptr(dev_scale), _lib.stream_ptr())
        _lib.check(rc, "nxd_scale_inplace")
        return x
    x.mul_(scale)
    if dev_scale is not None:
        x.mul_(dev_scale.to(x.dtype))
    return x
