"""DBRX-style fine-grained MoE decoder (reference ``examples/inference/dbrx/neuron_modeling_dbrx.py``):
the Mixtral block with LayerNorm (no bias) instead of RMSNorm, the fused QKV output clamped to ``[-clip_qkv, clip_qkv]``,
16 experts with top-4 routing and affinities normalised over the selected experts.

The parameter layout is the one of ``MixtralForCausalLM`` (fused GQA QKV, ``MoE(RouterTopK, ExpertMLPs)``), so the
expert-parallel / tensor-parallel sharding and the checkpoint tooling apply unchanged; HF DBRX names
(``transformer.blocks.N.norm_attn_norm.attn.Wqkv`` ...) map onto them in ``convert_dbrx_hf_state_dict``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from .mixtral import MixtralConfig, MixtralDecoderLayer, MixtralForCausalLM


@dataclass
class DbrxConfig(MixtralConfig):
    clip_qkv: Optional[float] = 8.0
    layer_norm_eps: float = 1e-5
    model_type: str = "dbrx"


def dbrx_instruct(**kw) -> DbrxConfig:
    base = dict(vocab_size=100352, hidden_size=6144, intermediate_size=10752, num_hidden_layers=40,
                num_attention_heads=48, num_key_value_heads=8, head_dim=128, rope_theta=5e5, rope_scaling=None,
                max_position_embeddings=32768, tie_word_embeddings=False, num_local_experts=16, num_experts_per_tok=4,
                clip_qkv=8.0)
    base.update(kw)
    return DbrxConfig(**base)


def tiny_dbrx(**kw) -> DbrxConfig:
    base = dict(vocab_size=512, hidden_size=128, intermediate_size=192, num_hidden_layers=2, num_attention_heads=4,
                num_key_value_heads=2, head_dim=32, max_position_embeddings=512, torch_dtype=torch.float32,
                num_local_experts=4, num_experts_per_tok=2, capacity_factor=2.0, tie_word_embeddings=False,
                rope_scaling=None, clip_qkv=0.5)
    base.update(kw)
    return DbrxConfig(**base)


class _BiaslessLayerNorm(nn.Module):
    """LayerNorm with a weight only; exposes ``variance_epsilon`` like ``RMSNorm`` so both fit the same call sites."""

    def __init__(self, hidden_size: int, eps: float, sequence_parallel_enabled: bool, dtype, device=None):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden_size, dtype=dtype, device=device))
        self.variance_epsilon = eps
        setattr(self.weight, "sequence_parallel_enabled", sequence_parallel_enabled)

    def forward(self, hidden: torch.Tensor, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        if residual is not None:
            hidden = hidden + residual
        return F.layer_norm(hidden.float(), hidden.shape[-1:], self.weight.float(), None, self.variance_epsilon).to(hidden.dtype)


def _clip_hook(clip: float):
    """Forward hook that clamps the (fused or split) QKV projection output: DBRX ``clip_qkv``.  A hook, not a wrapper
    module, so the parameter names stay those of the Mixtral block."""
    def hook(_module, _inputs, out):
        if isinstance(out, tuple):
            return tuple(o.clamp(-clip, clip) for o in out)
        return out.clamp(-clip, clip)
    return hook


class DbrxDecoderLayer(MixtralDecoderLayer):
    def __init__(self, config: DbrxConfig, layer_idx: int, device=None):
        super().__init__(config, layer_idx, device=device)
        sp, dt = config.sequence_parallel_enabled, config.torch_dtype
        self.input_layernorm = _BiaslessLayerNorm(config.hidden_size, config.layer_norm_eps, sp, dt, device)
        self.post_attention_layernorm = _BiaslessLayerNorm(config.hidden_size, config.layer_norm_eps, sp, dt, device)
        if config.clip_qkv:
            self.self_attn.qkv_proj.register_forward_hook(_clip_hook(float(config.clip_qkv)))

    def forward(self, hidden, residual, cos, sin):
        if residual is not None:
            hidden = hidden + residual
        residual = hidden
        a = self.self_attn(self.input_layernorm(hidden), cos, sin)
        residual = residual + a
        h = self.post_attention_layernorm(residual)
        if self.is_moe:
            out, router_logits = self.mlp(h)
        else:
            out, router_logits = self.mlp(h), None
        return out, residual, router_logits


class DbrxForCausalLM(MixtralForCausalLM):
    """``forward(input_ids, labels)`` as ``MixtralForCausalLM``; blocks and final norm are the DBRX variants."""

    def __init__(self, config: DbrxConfig, device=None):
        super().__init__(config, device=device)
        self.layers = nn.ModuleList([DbrxDecoderLayer(config, i, device=device) for i in range(config.num_hidden_layers)])
        self.norm = _BiaslessLayerNorm(config.hidden_size, config.layer_norm_eps, config.sequence_parallel_enabled,
                                       config.torch_dtype, device)


def convert_dbrx_hf_state_dict(hf: Dict[str, torch.Tensor], config: DbrxConfig) -> Dict[str, torch.Tensor]:
    """HF ``DbrxForCausalLM`` names -> this model's names (un-sharded).  Experts: HF stores ``w1 / v1 / w2`` of all
    experts stacked as ``[E * I, H]``; here gate|up are fused per expert as ``[E, H, 2I]`` and down as ``[E, I, H]``."""
    E, I, H = config.num_local_experts, config.intermediate_size, config.hidden_size
    out: Dict[str, torch.Tensor] = {}
    for k, v in hf.items():
        if k == "transformer.wte.weight":
            out["embed_tokens.weight"] = v
        elif k == "transformer.norm_f.weight":
            out["norm.weight"] = v
        elif k == "lm_head.weight":
            out["lm_head.weight"] = v
    for i in range(config.num_hidden_layers):
        src, dst = f"transformer.blocks.{i}.", f"layers.{i}."
        out[dst + "self_attn.qkv_proj.weight_qkv"] = hf[src + "norm_attn_norm.attn.Wqkv.weight"]
        out[dst + "self_attn.o_proj.weight"] = hf[src + "norm_attn_norm.attn.out_proj.weight"]
        out[dst + "input_layernorm.weight"] = hf[src + "norm_attn_norm.norm_1.weight"]
        out[dst + "post_attention_layernorm.weight"] = hf[src + "norm_attn_norm.norm_2.weight"]
        out[dst + "mlp.router.linear_router.weight"] = hf[src + "ffn.router.layer.weight"]
        w1 = hf[src + "ffn.experts.mlp.w1"].reshape(E, I, H)
        v1 = hf[src + "ffn.experts.mlp.v1"].reshape(E, I, H)
        w2 = hf[src + "ffn.experts.mlp.w2"].reshape(E, I, H)
        out[dst + "mlp.expert_mlps.mlp_op.gate_up_proj.weight"] = torch.cat([w1, v1], dim=1).transpose(1, 2).contiguous()
        out[dst + "mlp.expert_mlps.mlp_op.down_proj.weight"] = w2.contiguous()
    return out


__all__ = ["DbrxConfig", "DbrxDecoderLayer", "DbrxForCausalLM", "convert_dbrx_hf_state_dict", "dbrx_instruct", "tiny_dbrx"]
