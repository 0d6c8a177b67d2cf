"""Mixtral-style sparse-MoE decoder for training with TP + SP + expert parallelism (reference
``examples/training/mixtral/modeling_mixtral_moe_nxd.py:350-381``: the dense MLP of a Llama block is replaced by
``MoE(RouterTopK, ExpertMLPs)`` every ``moe_frequency`` layers, router logits of all layers feed
``load_balancing_loss_func``)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import torch
import torch.nn as nn
from torch.utils.checkpoint import checkpoint

from ..modules.moe import ExpertMLPs, MoE, RouterTopK, load_balancing_loss_func
from ..ops import norm as norm_ops
from ..parallel.layer_norm import RMSNorm
from ..parallel.layers import ColumnParallelLinear, ParallelEmbedding
from .llama import CausalLMLoss, LlamaAttention, LlamaConfig, LlamaMLP, RotaryTables, _init_fn


@dataclass
class MixtralConfig(LlamaConfig):
    num_local_experts: int = 8
    num_experts_per_tok: int = 2
    capacity_factor: Optional[float] = 2.0
    moe_frequency: int = 1
    router_aux_loss_coef: float = 0.02
    output_router_logits: bool = True
    model_type: str = "mixtral"


def mixtral_8x7b(**kw) -> MixtralConfig:
    base = dict(vocab_size=32000, hidden_size=4096, intermediate_size=14336, num_hidden_layers=32,
                num_attention_heads=32, num_key_value_heads=8, head_dim=128, rope_theta=1e6, rope_scaling=None,
                max_position_embeddings=32768, tie_word_embeddings=False)
    base.update(kw)
    return MixtralConfig(**base)


def tiny_mixtral(**kw) -> MixtralConfig:
    base = dict(vocab_size=512, hidden_size=128, intermediate_size=256, num_hidden_layers=2, num_attention_heads=4,
                num_key_value_heads=2, head_dim=32, max_position_embeddings=512, torch_dtype=torch.float32,
                num_local_experts=4, num_experts_per_tok=2, capacity_factor=2.0, tie_word_embeddings=False,
                rope_scaling=None)
    base.update(kw)
    return MixtralConfig(**base)


class MixtralDecoderLayer(nn.Module):
    def __init__(self, config: MixtralConfig, layer_idx: int, device=None):
        super().__init__()
        sp, dt = config.sequence_parallel_enabled, config.torch_dtype
        self.self_attn = LlamaAttention(config, device=device)
        self.is_moe = (layer_idx % config.moe_frequency) == 0
        if self.is_moe:
            init = _init_fn(config.initializer_range)
            router = RouterTopK(config.num_local_experts, config.num_experts_per_tok, config.hidden_size, dtype=dt,
                                device=device if device is not None else torch.device("cpu"))
            experts = ExpertMLPs(config.num_local_experts, config.num_experts_per_tok, config.hidden_size,
                                 config.intermediate_size, "silu", True, config.capacity_factor,
                                 normalize_top_k_affinities=True, init_method=init, output_layer_init_method=init,
                                 dtype=dt, device=device if device is not None else torch.device("cpu"))
            self.mlp = MoE(router, experts, return_router_logits=True, sequence_parallel_enabled=sp)
        else:
            self.mlp = LlamaMLP(config, device=device)
        self.input_layernorm = RMSNorm(config.hidden_size, config.rms_norm_eps, sequence_parallel_enabled=sp, dtype=dt,
                                       device=device)
        self.post_attention_layernorm = RMSNorm(config.hidden_size, config.rms_norm_eps, sequence_parallel_enabled=sp,
                                                dtype=dt, device=device)

    def forward(self, hidden, residual, cos, sin):
        eps = self.input_layernorm.variance_epsilon
        if residual is None:
            residual = hidden
            h = norm_ops.rms_norm(hidden, self.input_layernorm.weight, eps)
        else:
            h, residual = norm_ops.fused_add_rms_norm(hidden, residual, self.input_layernorm.weight, eps)
        a = self.self_attn(h, cos, sin)
        h, residual = norm_ops.fused_add_rms_norm(a, residual, self.post_attention_layernorm.weight, eps)
        if self.is_moe:
            out, router_logits = self.mlp(h)
        else:
            out, router_logits = self.mlp(h), None
        return out, residual, router_logits


class MixtralForCausalLM(nn.Module):
    """``forward(input_ids, labels)`` -> ``lm_loss + router_aux_loss_coef * load_balancing_loss``."""

    _fx_wrap_functions = [load_balancing_loss_func]       # kept opaque by the pipeline tracer (pipeline/trace.py)

    def __init__(self, config: MixtralConfig, device=None):
        super().__init__()
        self.config = config
        dt = config.torch_dtype
        self.embed_tokens = ParallelEmbedding(config.vocab_size, config.hidden_size,
                                              init_method=_init_fn(config.initializer_range), dtype=dt, device=device,
                                              sequence_parallel_enabled=True)
        self.layers = nn.ModuleList([MixtralDecoderLayer(config, i, device=device)
                                     for i in range(config.num_hidden_layers)])
        self.norm = RMSNorm(config.hidden_size, config.rms_norm_eps,
                            sequence_parallel_enabled=config.sequence_parallel_enabled, dtype=dt, device=device)
        self.rotary = RotaryTables(config)
        self.lm_head = ColumnParallelLinear(config.hidden_size, config.vocab_size, bias=False, gather_output=False,
                                            init_method=_init_fn(config.initializer_range),
                                            sequence_parallel_enabled=config.sequence_parallel_enabled, dtype=dt,
                                            device=device)
        self.loss_fn = CausalLMLoss()

    def forward(self, input_ids: torch.Tensor, labels: Optional[torch.Tensor] = None):
        from ..parallel import mappings
        from ..parallel import parallel_state as ps

        cfg = self.config
        hidden = self.embed_tokens(input_ids)
        if not cfg.sequence_parallel_enabled and ps.get_tensor_model_parallel_size() > 1:
            hidden = mappings.gather_from_sequence_parallel_region(hidden, to_model_parallel=False)
        cos, sin = self.rotary(input_ids)
        residual = None
        all_router_logits: List[torch.Tensor] = []
        for layer in self.layers:
            if cfg.activation_checkpointing and self.training and torch.is_grad_enabled():
                hidden, residual, rl = checkpoint(layer, hidden, residual, cos, sin, use_reentrant=False)
            else:
                hidden, residual, rl = layer(hidden, residual, cos, sin)
            if rl is not None:
                all_router_logits.append(rl)
        hidden = self.norm(hidden, residual)
        logits = self.lm_head(hidden)
        if labels is None:
            return logits
        loss = self.loss_fn(logits, labels)
        if cfg.output_router_logits and all_router_logits:
            # the list goes in as is: under pipeline tracing the call is one opaque node (``_fx_wrap_functions``) whose
            # inputs are the per-layer router logits carried from their stages to the last one
            aux = load_balancing_loss_func(all_router_logits, cfg.num_local_experts, cfg.num_experts_per_tok)
            loss = loss + cfg.router_aux_loss_coef * aux
        return loss


def convert_mixtral_hf_state_dict(hf, config: MixtralConfig):
    """HF ``MixtralForCausalLM`` names -> this model's names, un-sharded (reference
    ``examples/training/mixtral/convert_checkpoints.py``): q/k/v fused into ``qkv_proj.weight_qkv``; the per-expert
    ``w1`` (gate) / ``w3`` (up) ``[I, H]`` become one ``[E, H, 2I]`` slab, ``w2`` ``[H, I]`` becomes ``[E, I, H]``."""
    E = config.num_local_experts
    out = {"embed_tokens.weight": hf["model.embed_tokens.weight"], "norm.weight": hf["model.norm.weight"]}
    if "lm_head.weight" in hf:
        out["lm_head.weight"] = hf["lm_head.weight"]
    for i in range(config.num_hidden_layers):
        src, dst = f"model.layers.{i}.", f"layers.{i}."
        out[dst + "self_attn.qkv_proj.weight_qkv"] = torch.cat([hf[src + f"self_attn.{n}_proj.weight"] for n in "qkv"], 0)
        out[dst + "self_attn.o_proj.weight"] = hf[src + "self_attn.o_proj.weight"]
        out[dst + "input_layernorm.weight"] = hf[src + "input_layernorm.weight"]
        out[dst + "post_attention_layernorm.weight"] = hf[src + "post_attention_layernorm.weight"]
        moe = src + "block_sparse_moe."
        out[dst + "mlp.router.linear_router.weight"] = hf[moe + "gate.weight"]
        gate = torch.stack([hf[moe + f"experts.{e}.w1.weight"] for e in range(E)])       # [E, I, H]
        up = torch.stack([hf[moe + f"experts.{e}.w3.weight"] for e in range(E)])
        down = torch.stack([hf[moe + f"experts.{e}.w2.weight"] for e in range(E)])       # [E, H, I]
        out[dst + "mlp.expert_mlps.mlp_op.gate_up_proj.weight"] = torch.cat([gate, up], dim=1).transpose(1, 2).contiguous()
        out[dst + "mlp.expert_mlps.mlp_op.down_proj.weight"] = down.transpose(1, 2).contiguous()
    return out


def convert_mixtral_state_dict_to_hf(sd, config: MixtralConfig):
    """Inverse of ``convert_mixtral_hf_state_dict``."""
    E, I = config.num_local_experts, config.intermediate_size
    qs, kvs = config.num_attention_heads * config.head_dim, config.num_key_value_heads * config.head_dim
    out = {"model.embed_tokens.weight": sd["embed_tokens.weight"], "model.norm.weight": sd["norm.weight"]}
    if "lm_head.weight" in sd:
        out["lm_head.weight"] = sd["lm_head.weight"]
    for i in range(config.num_hidden_layers):
        src, dst = f"layers.{i}.", f"model.layers.{i}."
        q, k, v = torch.split(sd[src + "self_attn.qkv_proj.weight_qkv"], [qs, kvs, kvs], dim=0)
        out[dst + "self_attn.q_proj.weight"], out[dst + "self_attn.k_proj.weight"] = q, k
        out[dst + "self_attn.v_proj.weight"] = v
        out[dst + "self_attn.o_proj.weight"] = sd[src + "self_attn.o_proj.weight"]
        out[dst + "input_layernorm.weight"] = sd[src + "input_layernorm.weight"]
        out[dst + "post_attention_layernorm.weight"] = sd[src + "post_attention_layernorm.weight"]
        moe = dst + "block_sparse_moe."
        out[moe + "gate.weight"] = sd[src + "mlp.router.linear_router.weight"]
        gu = sd[src + "mlp.expert_mlps.mlp_op.gate_up_proj.weight"].transpose(1, 2)          # [E, 2I, H]
        down = sd[src + "mlp.expert_mlps.mlp_op.down_proj.weight"].transpose(1, 2)           # [E, H, I]
        for e in range(E):
            out[moe + f"experts.{e}.w1.weight"] = gu[e, :I].contiguous()
            out[moe + f"experts.{e}.w3.weight"] = gu[e, I:].contiguous()
            out[moe + f"experts.{e}.w2.weight"] = down[e].contiguous()
    return out
