"""ctypes binding + autograd wrapper of the hand-written flash attention (``csrc/flash_attn.cu``)."""
from __future__ import annotations

import ctypes as _c
import math
import os
from typing import Optional

import torch

from . import _lib


def _dense_like(t: torch.Tensor) -> torch.Tensor:
    """Uninitialised dense tensor with ``t``'s shape and dim ordering (so a ``[S,B,H,D]``-stored
    ``[B,H,S,D]`` view yields an output that reshapes to ``[S,B,H*D]`` without a copy)."""
    order = sorted(range(t.dim()), key=lambda d: (t.stride(d), -d), reverse=True)
    shape = [t.shape[d] for d in order]
    buf = torch.empty(shape, dtype=t.dtype, device=t.device)
    inv = [order.index(d) for d in range(t.dim())]
    return buf.permute(inv)


def _fix(t: torch.Tensor) -> torch.Tensor:
    if t.stride(-1) != 1 or any(s % 8 for s in t.stride()[:-1]):
        return t.contiguous()
    return t


def _strides(*ts) -> "_c.Array":
    vals = []
    for t in ts:
        vals += [t.stride(0), t.stride(1), t.stride(2)]
    return (_c.c_longlong * len(vals))(*vals)


class _FlashAttnFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, scale, causal, dropout_p=0.0, seed=0):
        q, k, v = _fix(q), _fix(k), _fix(v)
        B, Hq, Sq, D = q.shape
        Hkv, Sk = k.shape[1], k.shape[2]
        o = _dense_like(q)
        lse = torch.empty((B, Hq, Sq), dtype=torch.float32, device=q.device)
        # "tc": tcgen05/TMEM/TMA kernel (csrc/flash_attn_tc.cu); "mma": mma.sync kernel (csrc/flash_attn.cu)
        impl = os.environ.get("NXD_ATTN_FWD", "tc")
        if dropout_p > 0.0:
            # dropout lives in the tcgen05 kernels only (mask = hash of (seed, batch, head, query, key))
            rc = _lib.lib().nxd_flash_attn_fwd_tc_dropout(
                _lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(o), _lib.ptr(lse), _strides(q, k, v, o), _c.c_int(B),
                _c.c_int(Hq), _c.c_int(Hkv), _c.c_int(Sq), _c.c_int(Sk), _c.c_int(D), _c.c_float(scale),
                _c.c_int(1 if causal else 0), _c.c_float(dropout_p), _c.c_ulonglong(seed), _lib.stream_ptr())
            _lib.check(rc, "nxd_flash_attn_fwd_tc_dropout")
        else:
            fn = _lib.lib().nxd_flash_attn_fwd_tc if impl == "tc" else _lib.lib().nxd_flash_attn_fwd
            rc = fn(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(o), _lib.ptr(lse), _strides(q, k, v, o), _c.c_int(B),
                    _c.c_int(Hq), _c.c_int(Hkv), _c.c_int(Sq), _c.c_int(Sk), _c.c_int(D), _c.c_float(scale),
                    _c.c_int(1 if causal else 0), _lib.stream_ptr())
            _lib.check(rc, f"nxd_flash_attn_fwd[{impl}]")
        ctx.save_for_backward(q, k, v, o, lse)
        ctx.scale, ctx.causal, ctx.dropout_p, ctx.seed = scale, causal, float(dropout_p), int(seed)
        return o

    @staticmethod
    def backward(ctx, do):
        q, k, v, o, lse = ctx.saved_tensors
        do = _fix(do)
        B, Hq, Sq, D = q.shape
        Hkv, Sk = k.shape[1], k.shape[2]
        dq, dk, dv = _dense_like(q), _dense_like(k), _dense_like(v)
        if D in (64, 128) and (os.environ.get("NXD_ATTN_BWD", "tc") == "tc" or ctx.dropout_p > 0.0):
            # tcgen05 backward (csrc/flash_attn_bwd_tc.cu): dQ is accumulated in fp32 through TMA reduce-add
            sq_pad = (Sq + 63) // 64 * 64
            lse_pad = torch.empty(B * Hq * sq_pad, dtype=torch.float32, device=q.device)
            delta_pad = torch.empty_like(lse_pad)
            dq_acc = torch.zeros((B, Hq, D, sq_pad), dtype=torch.float32, device=q.device)   # d-major accumulator
            # few (kv block, kv head) pairs per GPU (tensor-parallel shards): several CTAs share a K/V block and
            # reduce-add partial dK / dV into fp32 accumulators
            nsplit = int(_lib.lib().nxd_flash_attn_bwd_tc_num_splits(_c.c_int(B), _c.c_int(Hq), _c.c_int(Hkv),
                                                                     _c.c_int(Sq), _c.c_int(Sk)))
            dkv_acc = (torch.zeros((2, B, Hkv, Sk, D), dtype=torch.float32, device=q.device)
                       if nsplit > 1 and os.environ.get("NXD_ATTN_BWD_SPLIT", "1") == "1" else None)
            args = (_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(o), _lib.ptr(do), _lib.ptr(lse), _lib.ptr(lse_pad),
                    _lib.ptr(delta_pad), _lib.ptr(dq_acc), _lib.ptr(dkv_acc), _lib.ptr(dq), _lib.ptr(dk), _lib.ptr(dv),
                    _strides(q, k, v, o, do, dq, dk, dv), _c.c_int(B), _c.c_int(Hq), _c.c_int(Hkv), _c.c_int(Sq),
                    _c.c_int(Sk), _c.c_int(D), _c.c_float(ctx.scale), _c.c_int(1 if ctx.causal else 0))
            if ctx.dropout_p > 0.0:
                rc = _lib.lib().nxd_flash_attn_bwd_tc_dropout(*args, _c.c_float(ctx.dropout_p), _c.c_ulonglong(ctx.seed),
                                                              _lib.stream_ptr())
            else:
                rc = _lib.lib().nxd_flash_attn_bwd_tc(*args, _lib.stream_ptr())
            _lib.check(rc, "nxd_flash_attn_bwd_tc")
            return dq, dk, dv, None, None, None, None
        delta = torch.empty_like(lse)
        rc = _lib.lib().nxd_flash_attn_bwd(
            _lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(o), _lib.ptr(do), _lib.ptr(lse), _lib.ptr(delta),
            _lib.ptr(dq), _lib.ptr(dk), _lib.ptr(dv), _strides(q, k, v, o, do, dq, dk, dv), _c.c_int(B), _c.c_int(Hq),
            _c.c_int(Hkv), _c.c_int(Sq), _c.c_int(Sk), _c.c_int(D), _c.c_float(ctx.scale),
            _c.c_int(1 if ctx.causal else 0), _lib.stream_ptr())
        _lib.check(rc, "nxd_flash_attn_bwd")
        return dq, dk, dv, None, None, None, None


def flash_attn_nxd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, softmax_scale: Optional[float] = None,
                   causal: bool = True, dropout_p: float = 0.0, seed: int = 0) -> torch.Tensor:
    scale = softmax_scale if softmax_scale is not None else 1.0 / math.sqrt(q.shape[-1])
    return _FlashAttnFn.apply(q, k, v, float(scale), bool(causal), float(dropout_p), int(seed))


def dropout_keep_mask(seed: int, B: int, H: int, Sq: int, Sk: int, dropout_p: float, device=None) -> torch.Tensor:
    """The keep mask ``[B, H, Sq, Sk]`` the kernels generate for ``(seed, dropout_p)`` (``csrc/attn_tc_common.cuh``
    ``drop_keep``), restated with integer tensor arithmetic: the oracle of the dropout tests."""
    M = 0xFFFFFFFF
    seed32 = (seed ^ (seed >> 32)) & M
    thr = min(M, int(dropout_p * 4294967296.0)) if dropout_p > 0 else 0
    i = torch.arange(Sq, dtype=torch.int64, device=device).view(1, 1, Sq, 1)
    j = torch.arange(Sk, dtype=torch.int64, device=device).view(1, 1, 1, Sk)
    bh = torch.arange(B * H, dtype=torch.int64, device=device).view(B, H, 1, 1)
    hs = seed32 ^ ((bh * 0x9E3779B1) & M)
    x = ((i * 0x85EBCA77) & M) ^ ((j * 0xC2B2AE3D) & M) ^ hs
    x = x ^ (x >> 16)
    x = (x * 0x7FEB352D) & M
    x = x ^ (x >> 15)
    x = (x * 0x846CA68B) & M
    x = x ^ (x >> 16)
    return x >= thr
