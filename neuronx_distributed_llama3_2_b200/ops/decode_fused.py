"""Bindings of the fused token-generation projections (``csrc/decode_fused.cu``): RMSNorm + fused QKV GEMV + RoPE + KV-cache
write in one launch, RMSNorm + gate/up GEMV + SwiGLU in one launch.  With the two row-parallel GEMVs that take the
residual in their epilogue (``fused_comm.skinny_gemm``) and the decode attention (split + merge) a decoder layer of the
token-generation graph is six launches (reference graph: ``examples/inference/modules/model_base.py:459-518``).

Batches of 9 .. 128 sequences take the tensor-core counterparts (``csrc/gemm_smallm.cu``: weight rows as the UMMA M
dimension, cluster split-K): fused QKV GEMM + RoPE + cache write, gate/up GEMM + SwiGLU (<= 64 rows), and the plain GEMM
with the residual in its epilogue; the two RMSNorms stay separate launches there (eight launches per layer)."""
from __future__ import annotations

import ctypes as _c
from typing import Optional

import torch

from . import _lib

MAX_ROWS = 8
BATCHED_MAX_ROWS = 128          # gemm_smallm kernels
BATCHED_SWIGLU_MAX_ROWS = 64


def qualifies(x: torch.Tensor) -> bool:
    return (x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2 and 1 <= x.shape[0] <= MAX_ROWS
            and x.shape[1] % 8 == 0 and x.stride(1) == 1 and x.stride(0) % 8 == 0 and _lib.available())


def qkv_rope_cache(x: torch.Tensor, norm_w: Optional[torch.Tensor], eps: float, w_qkv: torch.Tensor, cos: torch.Tensor,
                   sin: torch.Tensor, positions: torch.Tensor, cache_k: torch.Tensor, cache_v: torch.Tensor, n_q: int,
                   n_kv: int, D: int, seq_ids: Optional[torch.Tensor] = None, block_table: Optional[torch.Tensor] = None,
                   page_size: int = 0) -> torch.Tensor:
    """``x [B, H]`` -> ``q [B, n_q, D]`` (rotated); the rotated K row and the V row of every sequence are written into
    the cache at ``positions[b]``.  ``positions`` / ``seq_ids`` / ``block_table`` must be int32."""
    B, K = x.shape
    assert w_qkv.shape == ((n_q + 2 * n_kv) * D, K) and w_qkv.stride(1) == 1
    assert positions.dtype == torch.int32 and cos.dtype == torch.float32 and cos.shape[1] == D // 2
    q = torch.empty((B, n_q, D), dtype=torch.bfloat16, device=x.device)
    rc = _lib.lib().nxd_decode_qkv_rope_cache(
        _lib.ptr(x), _c.c_longlong(x.stride(0)), _lib.ptr(norm_w), _c.c_float(eps), _lib.ptr(w_qkv),
        _c.c_longlong(w_qkv.stride(0)), _lib.ptr(cos), _lib.ptr(sin), _lib.ptr(positions), _lib.ptr(q), _lib.ptr(cache_k),
        _lib.ptr(cache_v), _lib.ptr(seq_ids), _lib.ptr(block_table), _c.c_longlong(cache_k.stride(0)),
        _c.c_longlong(cache_k.stride(1)), _c.c_int(page_size), _c.c_int(block_table.shape[1] if block_table is not None else 0),
        _c.c_int(B), _c.c_int(K), _c.c_int(n_q), _c.c_int(n_kv), _c.c_int(D), _lib.stream_ptr())
    _lib.check(rc, "nxd_decode_qkv_rope_cache")
    return q


def gate_up_swiglu(x: torch.Tensor, norm_w: Optional[torch.Tensor], eps: float, w_gate_up: torch.Tensor) -> torch.Tensor:
    """``silu(W_gate n(x)) * (W_up n(x))`` for ``x [B, H]``; ``w_gate_up [2 I, H]`` holds the gate rows first."""
    B, K = x.shape
    I = w_gate_up.shape[0] // 2
    assert w_gate_up.shape[1] == K and w_gate_up.stride(1) == 1
    out = torch.empty((B, I), dtype=torch.bfloat16, device=x.device)
    rc = _lib.lib().nxd_decode_gate_up_swiglu(
        _lib.ptr(x), _c.c_longlong(x.stride(0)), _lib.ptr(norm_w), _c.c_float(eps), _lib.ptr(w_gate_up),
        _c.c_longlong(w_gate_up.stride(0)), _lib.ptr(out), _c.c_longlong(out.stride(0)), _c.c_int(B), _c.c_int(K),
        _c.c_int(I), _lib.stream_ptr())
    _lib.check(rc, "nxd_decode_gate_up_swiglu")
    return out


def qualifies_batched(x: torch.Tensor) -> bool:
    return (x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2 and MAX_ROWS < x.shape[0] <= BATCHED_MAX_ROWS
            and x.shape[1] % 8 == 0 and x.stride(1) == 1 and x.stride(0) % 8 == 0 and _lib.available())


def qkv_rope_cache_batched(xn: torch.Tensor, w_qkv: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor,
                           positions: torch.Tensor, cache_k: torch.Tensor, cache_v: torch.Tensor, n_q: int, n_kv: int,
                           D: int, seq_ids: Optional[torch.Tensor] = None, block_table: Optional[torch.Tensor] = None,
                           page_size: int = 0, splits: int = 0) -> torch.Tensor:
    """``xn [B, H]`` (already normalised, ``B <= 128``) -> ``q [B, n_q, D]`` rotated; rotated K rows and V rows go into
    the cache at ``positions[b]`` -- ONE tensor-core launch (``nxd_smallm_qkv_rope_cache``)."""
    B, K = xn.shape
    assert w_qkv.shape == ((n_q + 2 * n_kv) * D, K) and w_qkv.stride(1) == 1
    assert positions.dtype == torch.int32 and cos.dtype == torch.float32 and cos.shape[1] == D // 2
    q = torch.empty((B, n_q, D), dtype=torch.bfloat16, device=xn.device)
    rc = _lib.lib().nxd_smallm_qkv_rope_cache(
        _lib.ptr(xn), _c.c_longlong(xn.stride(0)), _lib.ptr(w_qkv), _c.c_longlong(w_qkv.stride(0)), _lib.ptr(cos),
        _lib.ptr(sin), _lib.ptr(positions), _lib.ptr(q), _lib.ptr(cache_k), _lib.ptr(cache_v), _lib.ptr(seq_ids),
        _lib.ptr(block_table), _c.c_longlong(cache_k.stride(0)), _c.c_longlong(cache_k.stride(1)), _c.c_int(page_size),
        _c.c_int(block_table.shape[1] if block_table is not None else 0), _c.c_int(B), _c.c_int(K), _c.c_int(n_q),
        _c.c_int(n_kv), _c.c_int(D), _c.c_int(splits), _lib.stream_ptr())
    _lib.check(rc, "nxd_smallm_qkv_rope_cache")
    return q


def gate_up_swiglu_batched(xn: torch.Tensor, w_gate_up: torch.Tensor, splits: int = 0) -> torch.Tensor:
    """``silu(xn W_gate^T) * (xn W_up^T)`` for ``xn [B, H]``, ``B <= 64``; ``w_gate_up [2 I, H]`` gate rows first: one
    launch with two accumulators per CTA (``nxd_smallm_gate_up_swiglu``)."""
    B, K = xn.shape
    I = w_gate_up.shape[0] // 2
    assert B <= BATCHED_SWIGLU_MAX_ROWS and w_gate_up.shape[1] == K and w_gate_up.stride(1) == 1 and I % 8 == 0
    out = torch.empty((B, I), dtype=torch.bfloat16, device=xn.device)
    rc = _lib.lib().nxd_smallm_gate_up_swiglu(
        _lib.ptr(xn), _c.c_longlong(xn.stride(0)), _lib.ptr(w_gate_up), _c.c_longlong(w_gate_up.stride(0)), _lib.ptr(out),
        _c.c_longlong(out.stride(0)), _c.c_int(B), _c.c_int(K), _c.c_int(I), _c.c_int(splits), _lib.stream_ptr())
    _lib.check(rc, "nxd_smallm_gate_up_swiglu")
    return out
