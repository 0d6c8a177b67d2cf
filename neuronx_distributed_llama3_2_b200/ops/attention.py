"""Attention ops.

* ``flash_attn_func`` -- training / prefill causal attention with native GQA.  API parity with the
  reference's ``nki_flash_attn_func(q, k, v, dropout_p, softmax_scale, causal, ...)``
  (``src/neuronx_distributed/kernels/flash_attn.py:151-191``; tensors ``[B, H, S, D]``), without its
  ``S % 2048`` restriction and without ``repeat_kv`` (the reference expands KV heads first,
  ``examples/training/llama/modeling_llama_nxd.py:409-410``).
* ``decode_attention`` / ``kv_cache_update`` -- token generation over a contiguous or paged cache.

Backends for ``flash_attn_func`` (``ops.backend.set_backends(attn=...)``, default from ``NXD_ATTN_BACKEND``):
  ``nxd``   hand-written sm_100a tcgen05 kernel (``csrc/flash_attn_tcgen05.cu``)  [default on GPU]
  ``sdpa``  ``torch.nn.functional.scaled_dot_product_attention`` -- library call, used only as the
            baseline arm and as the numerics oracle
  ``ref``   plain fp32 PyTorch math (CPU tests)
"""
from __future__ import annotations

import ctypes as _c
import math
import os
from typing import Optional

import torch
import torch.nn.functional as F

from . import _lib, backend


def attention_ref(q, k, v, causal=True, softmax_scale=None, dropout_p=0.0, seed=0):
    """fp32 reference.  q [B,Hq,S,D], k/v [B,Hkv,Sk,D].  Dropout uses the kernels' keep mask (same seed -> same mask)."""
    B, Hq, S, D = q.shape
    Hkv = k.shape[1]
    scale = softmax_scale if softmax_scale is not None else 1.0 / math.sqrt(D)
    g = Hq // Hkv
    kf = k.float().repeat_interleave(g, dim=1)
    vf = v.float().repeat_interleave(g, dim=1)
    s = torch.matmul(q.float(), kf.transpose(-1, -2)) * scale
    if causal:
        Sk = k.shape[2]
        mask = torch.ones(S, Sk, dtype=torch.bool, device=q.device).tril(diagonal=Sk - S)
        s = s.masked_fill(~mask, float("-inf"))
    p = torch.softmax(s, dim=-1)
    if dropout_p > 0.0:
        from .flash_attn import dropout_keep_mask

        keep = dropout_keep_mask(seed, B, Hq, S, k.shape[2], dropout_p, device=q.device)
        p = p * keep / (1.0 - dropout_p)
    return torch.matmul(p, vf).to(q.dtype)


def _backend(q: torch.Tensor) -> str:
    if not q.is_cuda:
        return "ref"
    be = backend.attn_backend()
    if be == "nxd":
        if not _lib.available():
            return "sdpa"
        if q.dtype != torch.bfloat16 or q.shape[-1] not in (64, 128):
            backend.note_fallback("flash_attn", f"dtype {q.dtype}, head dim {q.shape[-1]} (kernel: bf16, D in 64/128)")
            return "sdpa"
    return be


def flash_attn_func(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, dropout_p: float = 0.0,
                    softmax_scale: Optional[float] = None, causal: bool = True, mixed_precision: bool = True,
                    seed: Optional[int] = None) -> torch.Tensor:
    """q ``[B, Hq, S, D]``, k/v ``[B, Hkv, S, D]`` (any strides with unit stride on D) -> ``[B, Hq, S, D]``.

    ``dropout_p`` / ``seed`` as in the reference (``kernels/flash_attn.py:122-127,151-160``): the softmax probabilities
    are dropped with probability ``dropout_p`` and rescaled by ``1 / (1 - dropout_p)`` before the PV product; the mask is
    a deterministic function of ``seed`` (drawn from the torch generator when ``None``), so backward regenerates it."""
    be = _backend(q)
    if dropout_p > 0.0 and seed is None:
        seed = int(torch.randint(0, 2 ** 62, (1,)).item())
    if be == "nxd":
        from .flash_attn import flash_attn_nxd

        return flash_attn_nxd(q, k, v, softmax_scale, causal, dropout_p, seed or 0)
    if be == "sdpa":
        return F.scaled_dot_product_attention(q, k, v, dropout_p=dropout_p, is_causal=causal, scale=softmax_scale,
                                              enable_gqa=(q.shape[1] != k.shape[1]))
    return attention_ref(q, k, v, causal, softmax_scale, dropout_p, seed or 0)


nki_flash_attn_func = flash_attn_func  # reference-compatible alias


# ---------------------------------------------------------------------------------------------
# decode path
# ---------------------------------------------------------------------------------------------

_decode_scratch = {}


def kv_cache_update(cache_k: torch.Tensor, cache_v: torch.Tensor, new_k: torch.Tensor, new_v: torch.Tensor,
                    positions: torch.Tensor, seq_ids: Optional[torch.Tensor] = None,
                    block_table: Optional[torch.Tensor] = None, page_size: int = 0) -> None:
    """Write ``new_k/new_v [B, Hkv, T, D]`` at rows ``positions[b] + t`` of cache line ``seq_ids[b]``
    (default ``b``).  Cache: contiguous ``[Bmax, Hkv, Smax, D]`` or paged ``[pages, Hkv, page, D]``."""
    B, Hkv, T, D = new_k.shape
    if cache_k.is_cuda and cache_k.dtype == torch.bfloat16 and _lib.available():
        pos = positions.to(torch.int32).contiguous()
        sid = seq_ids.to(torch.int32).contiguous() if seq_ids is not None else None
        bt = block_table.to(torch.int32).contiguous() if block_table is not None else None
        # keep the contiguous copies alive until the launch: two temporaries of equal size would otherwise share one
        # allocator block (the first is freed before the second is made) and K would be read from V's bytes
        nk, nv = new_k.contiguous(), new_v.contiguous()
        rc = _lib.lib().nxd_kv_cache_update(
            _lib.ptr(nk), _lib.ptr(nv), _lib.ptr(cache_k), _lib.ptr(cache_v),
            _lib.ptr(pos), _lib.ptr(sid), _lib.ptr(bt), _c.c_longlong(cache_k.stride(0)),
            _c.c_longlong(cache_k.stride(1)), _c.c_int(page_size), _c.c_int(bt.shape[1] if bt is not None else 0),
            _c.c_int(B), _c.c_int(Hkv), _c.c_int(T), _c.c_int(D), _lib.stream_ptr())
        _lib.check(rc, "nxd_kv_cache_update")
        return
    for b in range(B):
        line = int(seq_ids[b]) if seq_ids is not None else b
        p0 = int(positions[b])
        for t in range(T):
            pos = p0 + t
            if block_table is not None:
                page = int(block_table[line, pos // page_size])
                cache_k[page, :, pos % page_size] = new_k[b, :, t]
                cache_v[page, :, pos % page_size] = new_v[b, :, t]
            else:
                cache_k[line, :, pos] = new_k[b, :, t]
                cache_v[line, :, pos] = new_v[b, :, t]


def decode_attention(q: torch.Tensor, cache_k: torch.Tensor, cache_v: torch.Tensor, lengths: torch.Tensor,
                     softmax_scale: Optional[float] = None, block_table: Optional[torch.Tensor] = None,
                     page_size: int = 0, seq_ids: Optional[torch.Tensor] = None) -> torch.Tensor:
    """One new token per sequence: q ``[B, Hq, D]`` attends to cache rows ``[0, lengths[b])``.  ``seq_ids [B]``
    (continuous batching): batch row ``b`` reads cache line ``seq_ids[b]`` (contiguous cache) or block-table row
    ``seq_ids[b]`` (paged cache) IN PLACE -- no gathered copy of the lines."""
    B, Hq, D = q.shape
    Hkv = cache_k.shape[1]
    scale = softmax_scale if softmax_scale is not None else 1.0 / math.sqrt(D)
    if q.is_cuda and q.dtype == torch.bfloat16 and D in (64, 128) and _lib.available():
        lens = lengths.to(torch.int32).contiguous()
        max_len = cache_k.shape[2] if block_table is None else block_table.shape[1] * page_size
        n_splits = _lib.lib().nxd_decode_attn_num_splits(_c.c_int(B), _c.c_int(Hkv), _c.c_int(max_len))
        need = B * Hq * n_splits * (D + 2)
        key = (q.device.index, need)
        scratch = _decode_scratch.get(key)
        if scratch is None:
            scratch = torch.empty(need, dtype=torch.float32, device=q.device)
            _decode_scratch.clear()
            _decode_scratch[key] = scratch
        out = torch.empty_like(q)
        bt = block_table.to(torch.int32).contiguous() if block_table is not None else None
        qc = q.contiguous()
        sid = seq_ids.to(torch.int32).contiguous() if seq_ids is not None else None
        rc = _lib.lib().nxd_decode_attn_lines(
            _lib.ptr(qc), _lib.ptr(cache_k), _lib.ptr(cache_v), _lib.ptr(lens), _lib.ptr(bt), _lib.ptr(sid),
            _c.c_longlong(cache_k.stride(0)), _c.c_longlong(cache_k.stride(1)), _c.c_int(page_size),
            _c.c_int(bt.shape[1] if bt is not None else 0), _lib.ptr(out), _lib.ptr(scratch), _c.c_int(B),
            _c.c_int(Hq), _c.c_int(Hkv), _c.c_int(D), _c.c_int(n_splits), _c.c_float(scale), _lib.stream_ptr())
        _lib.check(rc, "nxd_decode_attn_lines")
        return out
    if seq_ids is not None:      # reference path: gather the lines (or the table rows), then the plain oracle
        if block_table is not None:
            block_table = block_table[seq_ids.long()]
        else:
            cache_k, cache_v = cache_k.index_select(0, seq_ids.long()), cache_v.index_select(0, seq_ids.long())
    return decode_attention_ref(q, cache_k, cache_v, lengths, scale, block_table, page_size)


def decode_attention_ref(q, cache_k, cache_v, lengths, scale, block_table=None, page_size=0):
    B, Hq, D = q.shape
    Hkv = cache_k.shape[1]
    g = Hq // Hkv
    outs = []
    for b in range(B):
        L = int(lengths[b])
        if block_table is not None:
            pages = block_table[b, : (L + page_size - 1) // page_size].long()
            kk = cache_k[pages].permute(1, 0, 2, 3).reshape(Hkv, -1, D)[:, :L]
            vv = cache_v[pages].permute(1, 0, 2, 3).reshape(Hkv, -1, D)[:, :L]
        else:
            kk, vv = cache_k[b, :, :L], cache_v[b, :, :L]
        kk = kk.float().repeat_interleave(g, dim=0)
        vv = vv.float().repeat_interleave(g, dim=0)
        s = torch.einsum("hd,hld->hl", q[b].float(), kk) * scale
        p = torch.softmax(s, dim=-1)
        outs.append(torch.einsum("hl,hld->hd", p, vv))
    return torch.stack(outs).to(q.dtype)


def __getattr__(name):   # ``NKIAttnFunc``: the autograd function behind ``flash_attn_func`` (reference kernels/flash_attn.py:56)
    if name == "NKIAttnFunc":
        from .flash_attn import _FlashAttnFn

        return _FlashAttnFn
    raise AttributeError(name)
