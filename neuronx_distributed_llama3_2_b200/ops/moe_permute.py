"""Token dispatch / combine of the capacity-factor MoE path as two kernels with autograd (``csrc/moe_permute.cu``;
reference ``src/neuronx_distributed/modules/moe/expert_mlps.py:183-265``: gather into the expert-aligned ``[E, C, H]``
buffer by ``token_assignments``, affinity-weighted scatter-combine back to ``[T, H]``)."""
from __future__ import annotations

import ctypes as _c

import torch

from . import _lib


def usable(x: torch.Tensor) -> bool:
    return x.is_cuda and x.dtype == torch.bfloat16 and x.shape[-1] % 8 == 0 and _lib.available()


def _gather(x, slot_token, scale, slots, H):
    out = torch.empty((slots, H), dtype=torch.bfloat16, device=x.device)
    rc = _lib.lib().nxd_moe_gather(_lib.ptr(x), _lib.ptr(slot_token), _lib.ptr(scale), _lib.ptr(out),
                                   _c.c_longlong(slots), _c.c_int(H), _lib.stream_ptr())
    _lib.check(rc, "nxd_moe_gather")
    return out


def _combine(y, slot_of, w, T, k, H):
    out = torch.empty((T, H), dtype=torch.bfloat16, device=y.device)
    rc = _lib.lib().nxd_moe_combine(_lib.ptr(y), _lib.ptr(slot_of), _lib.ptr(w), _lib.ptr(out), _c.c_longlong(T),
                                    _c.c_int(k), _c.c_int(H), _lib.stream_ptr())
    _lib.check(rc, "nxd_moe_combine")
    return out


class _Dispatch(torch.autograd.Function):
    """``x [T, H]`` -> expert-aligned ``[slots, H]``: slot ``r`` holds token ``slot_token[r]`` (``-1`` = empty, zeros).
    Backward: every token sums the gradient rows of the slots it occupies (``slot_of [T, k]``, ``kept [T, k]``)."""

    @staticmethod
    def forward(ctx, x, slot_token, slot_of, kept_f):
        x = x.contiguous()
        ctx.save_for_backward(slot_of, kept_f)
        ctx.T = x.shape[0]
        return _gather(x, slot_token, None, slot_token.numel(), x.shape[1])

    @staticmethod
    def backward(ctx, dy):
        slot_of, kept_f = ctx.saved_tensors
        dy = dy.contiguous()
        return _combine(dy, slot_of, kept_f, ctx.T, slot_of.shape[1], dy.shape[1]), None, None, None


class _Combine(torch.autograd.Function):
    """``out[t] = sum_j w[t, j] * y[slot_of[t, j]]`` with ``w`` the (masked, possibly normalised) affinities in fp32.
    Backward: ``dy[r] = w_of_slot[r] * dout[slot_token[r]]`` (a scaled gather) and ``dw[t, j] = <dout[t], y[slot]>``."""

    @staticmethod
    def forward(ctx, y, w, slot_of, slot_token, kept_u8):
        y = y.contiguous()
        w = w.float().contiguous()
        ctx.save_for_backward(y, w, slot_of, slot_token, kept_u8)
        return _combine(y, slot_of, w, w.shape[0], w.shape[1], y.shape[1])

    @staticmethod
    def backward(ctx, dout):
        y, w, slot_of, slot_token, kept_u8 = ctx.saved_tensors
        dout = dout.contiguous()
        T, k = w.shape
        H = y.shape[1]
        # affinity of the (token, choice) pair that fills each slot
        slot_scale = torch.zeros(slot_token.numel(), dtype=torch.float32, device=y.device)
        keep = kept_u8.bool()
        slot_scale[slot_of[keep].long()] = w[keep]
        dy = _gather(dout, slot_token, slot_scale, slot_token.numel(), H)
        dw = torch.empty((T, k), dtype=torch.float32, device=y.device)
        rc = _lib.lib().nxd_moe_combine_dw(_lib.ptr(y), _lib.ptr(dout), _lib.ptr(slot_of), _lib.ptr(kept_u8), _lib.ptr(dw),
                                           _c.c_longlong(T), _c.c_int(k), _c.c_int(H), _lib.stream_ptr())
        _lib.check(rc, "nxd_moe_combine_dw")
        return dy, dw, None, None, None


def dispatch(x: torch.Tensor, slot_token: torch.Tensor, slot_of: torch.Tensor, kept: torch.Tensor) -> torch.Tensor:
    """``slot_token [E*C] int32`` (token per slot or -1), ``slot_of [T, k] int32`` (slot per (token, choice)),
    ``kept [T, k] bool``."""
    return _Dispatch.apply(x, slot_token, slot_of, kept.float().contiguous())


def combine(y: torch.Tensor, w: torch.Tensor, slot_of: torch.Tensor, slot_token: torch.Tensor,
            kept: torch.Tensor) -> torch.Tensor:
    return _Combine.apply(y, w, slot_of, slot_token, kept.to(torch.uint8).contiguous())
