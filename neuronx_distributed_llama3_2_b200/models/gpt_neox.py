"""GPT-NeoX / CodeGen (GPT-J style) decoder family on the tensor-parallel layers (reference
``examples/training/tp_dp_gpt_neox_hf_pretrain/tp_dp_gpt_neox_20b_hf_pretrain/modeling_gpt_neox_nxd.py`` and
``examples/training/codegen25``).

Both families share the block ``x + attn(ln1(x)) + mlp(ln2(x))`` ("parallel residual"; GPT-J / CodeGen use ONE
LayerNorm for both branches), LayerNorm with bias, GELU MLP, rotary embedding on the first ``rotary_ndims`` of every
head, and an untied output embedding.  They differ in the rotary pairing (NeoX: halves, GPT-J: interleaved) and in
biases -- all configuration here.

Layout follows the rest of the framework: sequence-first ``[S, B, H]`` activations, sequence parallelism keeps the
residual stream sharded on ``S``, the fused QKV weight is head-major (``[heads, 3, D]`` rows, the NeoX checkpoint
layout) so a contiguous split over TP ranks is a split over heads.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import attention as attn_ops
from ..parallel import mappings
from ..parallel import parallel_state as ps
from ..parallel.layer_norm import LayerNorm
from ..parallel.layers import ColumnParallelLinear, ParallelEmbedding, RowParallelLinear
from .llama import CausalLMLoss, _init_fn


@dataclass
class GPTNeoXConfig:
    vocab_size: int = 50432
    hidden_size: int = 6144
    num_hidden_layers: int = 44
    num_attention_heads: int = 64
    intermediate_size: int = 24576
    rotary_pct: float = 0.25
    rotary_emb_base: float = 10000.0
    rotary_interleaved: bool = False          # False: NeoX (rotate halves); True: GPT-J / CodeGen (rotate even/odd pairs)
    use_parallel_residual: bool = True
    shared_layernorm: bool = False            # GPT-J / CodeGen: one LayerNorm feeds attention and MLP
    qkv_bias: bool = True
    layer_norm_eps: float = 1e-5
    max_position_embeddings: int = 2048
    initializer_range: float = 0.02
    hidden_act: str = "gelu"
    sequence_parallel_enabled: bool = False
    activation_checkpointing: bool = False
    torch_dtype: torch.dtype = torch.bfloat16
    extra: dict = field(default_factory=dict)

    @property
    def head_dim(self) -> int:
        return self.hidden_size // self.num_attention_heads

    @property
    def rotary_ndims(self) -> int:
        return int(self.head_dim * self.rotary_pct) // 2 * 2


def gpt_neox_20b(**kw) -> GPTNeoXConfig:
    return GPTNeoXConfig(**kw)


def codegen25_7b(**kw) -> GPTNeoXConfig:
    base = dict(vocab_size=51200, hidden_size=4096, num_hidden_layers=32, num_attention_heads=32, intermediate_size=16384,
                rotary_pct=0.5, rotary_interleaved=True, shared_layernorm=True, qkv_bias=False, hidden_act="gelu_new")
    base.update(kw)
    return GPTNeoXConfig(**base)


def tiny_gpt_neox(**kw) -> GPTNeoXConfig:
    base = dict(vocab_size=512, hidden_size=128, num_hidden_layers=2, num_attention_heads=4, intermediate_size=256,
                max_position_embeddings=256, torch_dtype=torch.float32)
    base.update(kw)
    return GPTNeoXConfig(**base)


def _rotary_tables(S: int, ndims: int, base: float, device) -> tuple:
    inv = 1.0 / (base ** (torch.arange(0, ndims, 2, dtype=torch.float32, device=device) / ndims))
    ang = torch.outer(torch.arange(S, dtype=torch.float32, device=device), inv)      # [S, ndims / 2]
    return ang.cos(), ang.sin()


def apply_partial_rotary(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, ndims: int, interleaved: bool) -> torch.Tensor:
    """``x [S, B, heads, D]``: rotate the first ``ndims`` features of every head, pass the rest through."""
    if ndims == 0:
        return x
    rot, rest = x[..., :ndims].float(), x[..., ndims:]
    c, s = cos[:, None, None, :], sin[:, None, None, :]
    if interleaved:
        a, b = rot[..., 0::2], rot[..., 1::2]
        out = torch.stack((a * c - b * s, b * c + a * s), dim=-1).flatten(-2)
    else:
        a, b = rot[..., : ndims // 2], rot[..., ndims // 2:]
        out = torch.cat((a * c - b * s, b * c + a * s), dim=-1)
    return torch.cat((out.to(x.dtype), rest), dim=-1)


class GPTNeoXAttention(nn.Module):
    def __init__(self, config: GPTNeoXConfig, device=None):
        super().__init__()
        self.config = config
        tp = ps.get_tensor_model_parallel_size()
        assert config.num_attention_heads % tp == 0, "attention heads must divide by tp"
        self.heads_local = config.num_attention_heads // tp
        sp, dt, init = config.sequence_parallel_enabled, config.torch_dtype, _init_fn(config.initializer_range)
        self.query_key_value = ColumnParallelLinear(config.hidden_size, 3 * config.hidden_size, bias=config.qkv_bias,
                                                    gather_output=False, init_method=init, sequence_parallel_enabled=sp,
                                                    dtype=dt, device=device)
        self.dense = RowParallelLinear(config.hidden_size, config.hidden_size, bias=config.qkv_bias,
                                       input_is_parallel=True, init_method=init, sequence_parallel_enabled=sp, dtype=dt,
                                       device=device)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        cfg, h, D = self.config, self.heads_local, self.config.head_dim
        qkv = self.query_key_value(x)                                    # [S, B, h * 3 * D], head-major
        S, B = qkv.shape[0], qkv.shape[1]
        qkv = qkv.reshape(S, B, h, 3, D)
        cos, sin = _rotary_tables(S, cfg.rotary_ndims, cfg.rotary_emb_base, qkv.device)
        q = apply_partial_rotary(qkv[:, :, :, 0], cos, sin, cfg.rotary_ndims, cfg.rotary_interleaved)
        k = apply_partial_rotary(qkv[:, :, :, 1], cos, sin, cfg.rotary_ndims, cfg.rotary_interleaved)
        v = qkv[:, :, :, 2]
        o = attn_ops.flash_attn_func(q.permute(1, 2, 0, 3), k.permute(1, 2, 0, 3), v.permute(1, 2, 0, 3), causal=True)
        return self.dense(o.permute(2, 0, 1, 3).reshape(S, B, h * D))


class GPTNeoXMLP(nn.Module):
    def __init__(self, config: GPTNeoXConfig, device=None):
        super().__init__()
        sp, dt, init = config.sequence_parallel_enabled, config.torch_dtype, _init_fn(config.initializer_range)
        self.dense_h_to_4h = ColumnParallelLinear(config.hidden_size, config.intermediate_size, bias=True,
                                                  gather_output=False, init_method=init, sequence_parallel_enabled=sp,
                                                  dtype=dt, device=device)
        self.dense_4h_to_h = RowParallelLinear(config.intermediate_size, config.hidden_size, bias=True,
                                               input_is_parallel=True, init_method=init, sequence_parallel_enabled=sp,
                                               dtype=dt, device=device)
        self.approx = "tanh" if config.hidden_act in ("gelu_new", "gelu_fast", "gelu_pytorch_tanh") else "none"

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.dense_4h_to_h(F.gelu(self.dense_h_to_4h(x), approximate=self.approx))


class GPTNeoXLayer(nn.Module):
    def __init__(self, config: GPTNeoXConfig, device=None):
        super().__init__()
        self.config = config
        sp, dt = config.sequence_parallel_enabled, config.torch_dtype
        self.input_layernorm = LayerNorm(config.hidden_size, eps=config.layer_norm_eps, sequence_parallel_enabled=sp,
                                         dtype=dt, device=device)
        if not config.shared_layernorm:
            self.post_attention_layernorm = LayerNorm(config.hidden_size, eps=config.layer_norm_eps,
                                                      sequence_parallel_enabled=sp, dtype=dt, device=device)
        self.attention = GPTNeoXAttention(config, device=device)
        self.mlp = GPTNeoXMLP(config, device=device)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        ln1 = self.input_layernorm(x)
        attn = self.attention(ln1)
        if self.config.use_parallel_residual:
            ln2 = ln1 if self.config.shared_layernorm else self.post_attention_layernorm(x)
            return x + attn + self.mlp(ln2)
        x = x + attn
        return x + self.mlp(self.post_attention_layernorm(x))


class GPTNeoXForCausalLM(nn.Module):
    """``forward(input_ids [B, S], labels=None)``: vocab-parallel logits ``[S, B, V/tp]`` or the mean next-token loss."""

    def __init__(self, config: GPTNeoXConfig, device=None):
        super().__init__()
        self.config = config
        dt, init = config.torch_dtype, _init_fn(config.initializer_range)
        self.embed_in = ParallelEmbedding(config.vocab_size, config.hidden_size, init_method=init, dtype=dt,
                                          device=device, sequence_parallel_enabled=True)   # always [S(/tp), B, H]
        self.layers = nn.ModuleList([GPTNeoXLayer(config, device=device) for _ in range(config.num_hidden_layers)])
        self.final_layer_norm = LayerNorm(config.hidden_size, eps=config.layer_norm_eps,
                                          sequence_parallel_enabled=config.sequence_parallel_enabled, dtype=dt,
                                          device=device)
        self.embed_out = ColumnParallelLinear(config.hidden_size, config.vocab_size, bias=False, gather_output=False,
                                              init_method=init,
                                              sequence_parallel_enabled=config.sequence_parallel_enabled, dtype=dt,
                                              device=device)
        self.loss_fn = CausalLMLoss()

    def forward(self, input_ids: torch.Tensor, labels: Optional[torch.Tensor] = None):
        from torch.utils.checkpoint import checkpoint

        cfg = self.config
        hidden = self.embed_in(input_ids)
        if not cfg.sequence_parallel_enabled and ps.get_tensor_model_parallel_size() > 1:
            hidden = mappings.gather_from_sequence_parallel_region(hidden, to_model_parallel=False)
        for layer in self.layers:
            if cfg.activation_checkpointing and self.training and torch.is_grad_enabled():
                hidden = checkpoint(layer, hidden, use_reentrant=False)
            else:
                hidden = layer(hidden)
        logits = self.embed_out(self.final_layer_norm(hidden))
        if labels is None:
            return logits
        return self.loss_fn(logits, labels)


__all__ = ["GPTNeoXConfig", "GPTNeoXForCausalLM", "GPTNeoXLayer", "apply_partial_rotary", "codegen25_7b", "gpt_neox_20b",
           "tiny_gpt_neox"]
