"""Rotary position embedding, incl. Llama-3 frequency scaling.

Reference: HF ``apply_rotary_pos_emb`` used at
``examples/training/llama/modeling_llama_nxd.py:398-399``; Llama-3 scaling
``examples/inference/llama3/neuron_modeling_llama.py:234-298``.
"""
from __future__ import annotations

import ctypes as _c
import math
from typing import Optional, Tuple

import torch

from . import _lib


def llama3_scale_inv_freq(inv_freq: torch.Tensor, factor: float = 32.0, low_freq_factor: float = 1.0,
                          high_freq_factor: float = 4.0, original_max_position: int = 8192) -> torch.Tensor:
    """Llama-3.x "llama3" rope scaling: long wavelengths are slowed by ``factor``, short ones
    kept, the band between interpolated smoothly."""
    low_wavelen = original_max_position / low_freq_factor
    high_wavelen = original_max_position / high_freq_factor
    wavelen = 2 * math.pi / inv_freq
    scaled = torch.where(wavelen > low_wavelen, inv_freq / factor, inv_freq)
    smooth = (original_max_position / wavelen - low_freq_factor) / (high_freq_factor - low_freq_factor)
    smoothed = (1 - smooth) * inv_freq / factor + smooth * inv_freq
    mid = (wavelen <= low_wavelen) & (wavelen >= high_wavelen)
    return torch.where(mid, smoothed, scaled)


def build_rope_cache(max_pos: int, head_dim: int, theta: float = 500000.0, rope_scaling: Optional[dict] = None,
                     device=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """fp32 ``cos``/``sin`` tables of shape ``[max_pos, head_dim/2]``."""
    inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float32, device=device) / head_dim))
    if rope_scaling and rope_scaling.get("rope_type", rope_scaling.get("type")) == "llama3":
        inv_freq = llama3_scale_inv_freq(
            inv_freq, rope_scaling.get("factor", 32.0), rope_scaling.get("low_freq_factor", 1.0),
            rope_scaling.get("high_freq_factor", 4.0), rope_scaling.get("original_max_position_embeddings", 8192))
    t = torch.arange(max_pos, dtype=torch.float32, device=device)
    freqs = torch.outer(t, inv_freq)
    return freqs.cos().contiguous(), freqs.sin().contiguous()


def _positions(T: int, pos_div: int, pos_mod: int, positions, device):
    if positions is not None:
        return positions.reshape(-1).long()
    return (torch.arange(T, device=device) // pos_div) % pos_mod


def rope_ref(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, n_rot_heads: int, pos_div: int, pos_mod: int,
             positions=None, inverse: bool = False) -> torch.Tensor:
    """x: [T, heads, D]; rotate the first ``n_rot_heads`` heads (rotate-half convention)."""
    T, Hh, D = x.shape
    pos = _positions(T, pos_div, pos_mod, positions, x.device)
    c = cos[pos].unsqueeze(1).float()
    s = sin[pos].unsqueeze(1).float()
    if inverse:
        s = -s
    xr = x[:, :n_rot_heads].float()
    x1, x2 = xr[..., : D // 2], xr[..., D // 2:]
    out = torch.cat([x1 * c - x2 * s, x2 * c + x1 * s], dim=-1).to(x.dtype)
    return torch.cat([out, x[:, n_rot_heads:]], dim=1)


class _RopeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cos, sin, n_rot_heads, pos_div, pos_mod, positions):
        ctx.args = (n_rot_heads, pos_div, pos_mod)
        ctx.save_for_backward(cos, sin, positions if positions is not None else torch.empty(0))
        ctx.has_pos = positions is not None
        out = x.contiguous().clone()
        _launch(out, cos, sin, n_rot_heads, pos_div, pos_mod, positions, False)
        return out

    @staticmethod
    def backward(ctx, dy):
        cos, sin, positions = ctx.saved_tensors
        n_rot_heads, pos_div, pos_mod = ctx.args
        dx = dy.contiguous().clone()
        _launch(dx, cos, sin, n_rot_heads, pos_div, pos_mod, positions if ctx.has_pos else None, True)
        return dx, None, None, None, None, None, None


def _launch(x, cos, sin, n_rot_heads, pos_div, pos_mod, positions, inverse):
    T, Hh, D = x.shape
    pos_t = None
    if positions is not None:
        pos_t = positions.reshape(-1).to(torch.int32).contiguous()
    rc = _lib.lib().nxd_rope_inplace(_lib.ptr(x), _lib.ptr(cos), _lib.ptr(sin), _lib.ptr(pos_t), _c.c_longlong(T),
                                     _c.c_int(Hh), _c.c_int(n_rot_heads), _c.c_int(D), _c.c_int(pos_div),
                                     _c.c_int(pos_mod), _c.c_int(1 if inverse else 0), _lib.stream_ptr())
    _lib.check(rc, "nxd_rope_inplace")


def apply_rope(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, n_rot_heads: int, pos_div: int = 1,
               pos_mod: int = 1 << 30, positions: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Rotate heads ``[0, n_rot_heads)`` of ``x [T, heads, D]``.  Token ``t`` sits at position
    ``positions[t]`` or ``(t // pos_div) % pos_mod`` ([S,B] layout: div=B, mod=S; [B,S]: div=1, mod=S)."""
    if x.is_cuda and x.dtype == torch.bfloat16 and x.shape[-1] % 16 == 0 and _lib.available():
        return _RopeFn.apply(x, cos, sin, n_rot_heads, pos_div, pos_mod, positions)
    return rope_ref(x, cos, sin, n_rot_heads, pos_div, pos_mod, positions)


def rope_inplace_(x: torch.Tensor, cos, sin, n_rot_heads: int, pos_div: int = 1, pos_mod: int = 1 << 30,
                  positions: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Inference variant: rotate in place (no autograd)."""
    if x.is_cuda and x.dtype == torch.bfloat16 and _lib.available():
        assert x.is_contiguous()
        _launch(x, cos, sin, n_rot_heads, pos_div, pos_mod, positions, False)
        return x
    x.copy_(rope_ref(x, cos, sin, n_rot_heads, pos_div, pos_mod, positions))
    return x
