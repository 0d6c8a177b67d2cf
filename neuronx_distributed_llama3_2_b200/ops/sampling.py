"""On-device sampling kernels."""
from __future__ import annotations

import ctypes as _c
from typing import Optional

import torch

from . import _lib


def _code(t):
    return 0 if t.dtype == torch.float32 else 1


def _use_kernel(logits: torch.Tensor) -> bool:
    return (logits.is_cuda and logits.dtype in (torch.float32, torch.bfloat16) and logits.stride(-1) == 1
            and _lib.available())


def argmax(logits: torch.Tensor) -> torch.Tensor:
    """Greedy token per row of ``[B, V]`` (ties -> lowest index)."""
    if _use_kernel(logits):
        B, V = logits.shape
        out = torch.empty(B, dtype=torch.int64, device=logits.device)
        rc = _lib.lib().nxd_argmax(_lib.ptr(logits), _c.c_int(_code(logits)), _c.c_int(B), _c.c_int(V),
                                   _c.c_longlong(logits.stride(0)), _lib.ptr(out), _lib.stream_ptr())
        _lib.check(rc, "nxd_argmax")
        return out
    return torch.argmax(logits, dim=-1)


def topk_sample(logits: torch.Tensor, top_k: int, top_p: float = 1.0, temperature: float = 1.0,
                uniform: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Top-k (optionally top-p) multinomial draw per row by inverse CDF against ``uniform[B]``."""
    B, V = logits.shape
    if uniform is None:
        uniform = torch.rand(B, device=logits.device, dtype=torch.float32)
    top_k = min(top_k, V)
    if _use_kernel(logits) and top_k <= 1024:
        out = torch.empty(B, dtype=torch.int64, device=logits.device)
        rc = _lib.lib().nxd_topk_sample(_lib.ptr(logits), _c.c_int(_code(logits)), _c.c_int(B), _c.c_int(V),
                                        _c.c_longlong(logits.stride(0)), _c.c_int(top_k), _c.c_float(top_p),
                                        _c.c_float(temperature), _lib.ptr(uniform.float().contiguous()),
                                        _lib.ptr(out), _lib.stream_ptr())
        _lib.check(rc, "nxd_topk_sample")
        return out
    return topk_sample_ref(logits, top_k, top_p, temperature, uniform)


def topk_sample_ref(logits, top_k, top_p, temperature, uniform):
    vals, idx = torch.sort(logits.float(), dim=-1, descending=True, stable=True)
    vals, idx = vals[:, :top_k], idx[:, :top_k]
    probs = torch.softmax(vals / max(temperature, 1e-6), dim=-1)
    if 0.0 < top_p < 1.0:
        c = probs.cumsum(-1)
        keep = (c - probs) < top_p
        probs = probs * keep
    mass = probs.sum(-1, keepdim=True)
    cdf = probs.cumsum(-1)
    u = uniform.float().reshape(-1, 1) * mass
    pick = (cdf <= u).sum(-1).clamp(max=top_k - 1)
    # never pick a zero-probability (filtered) slot
    last_valid = (probs > 0).sum(-1) - 1
    pick = torch.minimum(pick, last_valid)
    return idx.gather(1, pick[:, None]).squeeze(1)
