"""SwiGLU on the fused gate|up tensor."""
from __future__ import annotations

import ctypes as _c

import torch
import torch.nn.functional as F

from . import _lib


def swiglu_ref(gate_up: torch.Tensor) -> torch.Tensor:
    g, u = gate_up.chunk(2, dim=-1)
    return F.silu(g) * u


class _SwiGLUFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gate_up):
        I = gate_up.shape[-1] // 2
        gu = gate_up.reshape(-1, 2 * I)
        if not gu.is_contiguous():
            gu = gu.contiguous()
        out = torch.empty((gu.shape[0], I), dtype=gu.dtype, device=gu.device)
        rc = _lib.lib().nxd_swiglu_fwd(_lib.ptr(gu), _lib.ptr(out), _c.c_longlong(gu.shape[0]), _c.c_int(I),
                                       _lib.stream_ptr())
        _lib.check(rc, "nxd_swiglu_fwd")
        ctx.save_for_backward(gu)
        ctx.shape = gate_up.shape
        return out.reshape(gate_up.shape[:-1] + (I,))

    @staticmethod
    def backward(ctx, dy):
        (gu,) = ctx.saved_tensors
        I = gu.shape[1] // 2
        dy2 = dy.reshape(-1, I)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dgu = torch.empty_like(gu)
        rc = _lib.lib().nxd_swiglu_bwd(_lib.ptr(gu), _lib.ptr(dy2), _lib.ptr(dgu), _c.c_longlong(gu.shape[0]),
                                       _c.c_int(I), _lib.stream_ptr())
        _lib.check(rc, "nxd_swiglu_bwd")
        return dgu.reshape(ctx.shape)


def swiglu(gate_up: torch.Tensor) -> torch.Tensor:
    """``silu(gate) * up`` where ``gate_up[..., :I]`` is gate and ``[..., I:]`` is up."""
    I = gate_up.shape[-1] // 2
    if gate_up.is_cuda and gate_up.dtype == torch.bfloat16 and I % 8 == 0 and _lib.available():
        return _SwiGLUFn.apply(gate_up)
    return swiglu_ref(gate_up)
