"""Row statistics + backward for vocab-parallel cross entropy."""
from __future__ import annotations

import ctypes as _c

import torch

from . import _lib


def _use_kernel(logits: torch.Tensor) -> bool:
    return (logits.is_cuda and logits.dtype == torch.bfloat16 and logits.shape[-1] % 8 == 0
            and logits.is_contiguous() and _lib.available())


def ce_local_stats(logits2d: torch.Tensor, target: torch.Tensor, vocab_start: int) -> torch.Tensor:
    """[T, 4] fp32: (row max, sum exp(z - max), z[target] if local else 0, sum z)."""
    T, V = logits2d.shape
    if _use_kernel(logits2d):
        stats = torch.empty((T, 4), dtype=torch.float32, device=logits2d.device)
        tgt = target.to(torch.int64).contiguous()
        rc = _lib.lib().nxd_ce_stats(_lib.ptr(logits2d), _lib.ptr(tgt), _lib.ptr(stats), _c.c_longlong(T),
                                     _c.c_int(V), _c.c_int(vocab_start), _lib.stream_ptr())
        _lib.check(rc, "nxd_ce_stats")
        return stats
    z = logits2d.float()
    m = z.max(dim=-1).values
    s = torch.exp(z - m[:, None]).sum(-1)
    local = target - vocab_start
    ok = (local >= 0) & (local < V)
    tl = torch.where(ok, z.gather(1, local.clamp(0, V - 1)[:, None]).squeeze(1), torch.zeros_like(m))
    return torch.stack([m, s, tl, z.sum(-1)], dim=1)


def ce_backward(logits2d: torch.Tensor, target: torch.Tensor, lse: torch.Tensor, grad_out: torch.Tensor,
                vocab_start: int, smoothing: float, vocab_global: int) -> torch.Tensor:
    """``g * (softmax - (1-a) onehot - a/V)``; rows with target < 0 get zero grad."""
    T, V = logits2d.shape
    if _use_kernel(logits2d):
        grad = torch.empty_like(logits2d)
        tgt = target.to(torch.int64).contiguous()
        rc = _lib.lib().nxd_ce_bwd(_lib.ptr(logits2d), _lib.ptr(tgt), _lib.ptr(lse.contiguous()),
                                   _lib.ptr(grad_out), _lib.ptr(grad), _c.c_longlong(T), _c.c_int(V),
                                   _c.c_int(vocab_start), _c.c_float(smoothing), _c.c_int(vocab_global),
                                   _lib.stream_ptr())
        _lib.check(rc, "nxd_ce_bwd")
        return grad
    z = logits2d.float()
    p = torch.exp(z - lse[:, None]) - smoothing / vocab_global
    local = target - vocab_start
    ok = (local >= 0) & (local < V)
    onehot = torch.zeros_like(p)
    onehot.scatter_(1, local.clamp(0, V - 1)[:, None], ok[:, None].float())
    p = p - (1.0 - smoothing) * onehot
    g = torch.where(target < 0, torch.zeros_like(grad_out), grad_out)
    return (p * g[:, None]).to(logits2d.dtype)
