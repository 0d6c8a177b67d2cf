"""Llama-3.x training model on the tensor/sequence-parallel layers.

Reference: ``examples/training/llama/modeling_llama_nxd.py:134-775`` (``LlamaRMSNorm``, ``LlamaMLP`` with the
fused stride-2 gate_up projection, ``LlamaAttention`` on ``GQAQKVColumnParallelLinear`` + flash attention,
``LlamaModel`` entering/leaving the sequence-parallel region, ``LlamaForCausalLM`` with vocab-parallel
cross entropy) plus the Llama-3.2 specifics of this fork: tied ``lm_head``/``embed_tokens``
(``examples/inference/modules/model_base.py:620``) and "llama3" RoPE scaling
(``examples/inference/llama3/neuron_modeling_llama.py:234-298``).

B200 design notes
  * the residual stream is always sequence-first ``[S(/tp), B, H]``; the embedding emits the SP shard
    directly with a reduce-scatter, the block never transposes;
  * Q, K, V come out of ONE fused projection as ``[S, B, heads, D]`` and are consumed *in place* by the
    flash-attention kernel through strides (no permute/contiguous, no ``repeat_kv``), RoPE rotates q and
    k heads of that buffer in one launch;
  * residual add + RMSNorm are one kernel (``fused_add_rms_norm``); the layer returns ``(mlp_out, residual)``
    and the next layer's input norm performs the pending add;
  * logits stay bf16 ``[S, B, V/tp]``; the CE kernels read them once forward and once backward (the
    reference upcasts them to fp64).
"""
from __future__ import annotations

from dataclasses import asdict, dataclass, field
from typing import Any, Dict, Optional, Tuple

import torch
import torch.nn as nn
from torch.utils.checkpoint import checkpoint

from ..modules.qkv_linear import GQAQKVColumnParallelLinear
from ..ops import activation as act_ops
from ..ops import attention as attn_ops
from ..ops import norm as norm_ops
from ..ops import rope as rope_ops
from ..parallel import mappings
from ..parallel import parallel_state as ps
from ..parallel.layer_norm import RMSNorm
from ..parallel.layers import ColumnParallelLinear, ParallelEmbedding, RowParallelLinear
from ..parallel.loss_functions import parallel_cross_entropy


@dataclass
class LlamaConfig:
    vocab_size: int = 128256
    hidden_size: int = 3072
    intermediate_size: int = 8192
    num_hidden_layers: int = 28
    num_attention_heads: int = 24
    num_key_value_heads: int = 8
    head_dim: Optional[int] = None
    rms_norm_eps: float = 1e-5
    rope_theta: float = 500000.0
    rope_scaling: Optional[Dict[str, Any]] = field(default_factory=lambda: {
        "rope_type": "llama3", "factor": 32.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
        "original_max_position_embeddings": 8192})
    max_position_embeddings: int = 131072
    tie_word_embeddings: bool = True
    initializer_range: float = 0.02
    # parallel / runtime switches (ad-hoc HF-config attributes in the reference,
    # examples/training/llama/tp_zero1_llama_hf_pretrain/tp_zero1_llama_hf_pretrain.py:134-147)
    sequence_parallel_enabled: bool = True
    kv_shared_group_size: int = 1
    fuse_qkv: bool = True
    use_flash_attention: bool = True
    selective_checkpoint_enabled: bool = False
    activation_checkpointing: bool = False
    torch_dtype: torch.dtype = torch.bfloat16
    model_type: str = "llama"

    def __post_init__(self):
        if self.head_dim is None:
            self.head_dim = self.hidden_size // self.num_attention_heads

    def to_dict(self) -> Dict[str, Any]:
        d = asdict(self)
        d["torch_dtype"] = str(self.torch_dtype).replace("torch.", "")
        return d

    @property
    def num_params(self) -> int:
        h, i, v, L = self.hidden_size, self.intermediate_size, self.vocab_size, self.num_hidden_layers
        attn = h * self.head_dim * (2 * self.num_attention_heads + 2 * self.num_key_value_heads)
        per_layer = attn + 3 * h * i + 2 * h
        return L * per_layer + v * h * (1 if self.tie_word_embeddings else 2) + h


def llama3_2_1b(**kw) -> LlamaConfig:
    """HF ``meta-llama/Llama-3.2-1B`` architecture."""
    base = dict(hidden_size=2048, intermediate_size=8192, num_hidden_layers=16, num_attention_heads=32,
                num_key_value_heads=8, head_dim=64)
    base.update(kw)
    return LlamaConfig(**base)


def llama3_2_3b(**kw) -> LlamaConfig:
    """HF ``meta-llama/Llama-3.2-3B`` architecture."""
    base = dict(hidden_size=3072, intermediate_size=8192, num_hidden_layers=28, num_attention_heads=24,
                num_key_value_heads=8, head_dim=128)
    base.update(kw)
    return LlamaConfig(**base)


def llama3_1_8b(**kw) -> LlamaConfig:
    base = dict(hidden_size=4096, intermediate_size=14336, num_hidden_layers=32, num_attention_heads=32,
                num_key_value_heads=8, head_dim=128, tie_word_embeddings=False,
                rope_scaling={"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
                              "original_max_position_embeddings": 8192})
    base.update(kw)
    return LlamaConfig(**base)


def tiny_llama(**kw) -> LlamaConfig:
    """BASELINE.json config #1: 2 layers, hidden 256 (CPU/gloo plumbing tests)."""
    base = dict(vocab_size=512, hidden_size=256, intermediate_size=512, num_hidden_layers=2, num_attention_heads=8,
                num_key_value_heads=2, head_dim=32, max_position_embeddings=512, torch_dtype=torch.float32)
    base.update(kw)
    return LlamaConfig(**base)


_LLAMA31_ROPE = {"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
                 "original_max_position_embeddings": 8192}


def _preset(doc: str, **base):
    def make(**kw) -> LlamaConfig:
        cfg = dict(base)
        cfg.update(kw)
        return LlamaConfig(**cfg)

    make.__doc__ = doc
    return make


# the other sizes the reference ships configs for (``examples/training/llama/tp_zero1_llama_hf_pretrain/*_config_*`` and
# ``tp_pp_llama_hf_pretrain/*_config_*``)
llama2_7b = _preset("Llama-2 7B (reference 7B_config_llama2)", vocab_size=32000, hidden_size=4096, intermediate_size=11008,
                    num_hidden_layers=32, num_attention_heads=32, num_key_value_heads=32, head_dim=128, rope_theta=10000.0,
                    rope_scaling=None, max_position_embeddings=4096, tie_word_embeddings=False)
llama2_13b = _preset("Llama-2 13B (reference 13B_config_llama2)", vocab_size=32000, hidden_size=5120,
                     intermediate_size=13824, num_hidden_layers=40, num_attention_heads=40, num_key_value_heads=40,
                     head_dim=128, rope_theta=10000.0, rope_scaling=None, max_position_embeddings=4096,
                     tie_word_embeddings=False)
llama2_70b = _preset("Llama-2 70B (reference 70B_config_llama2)", vocab_size=32000, hidden_size=8192,
                     intermediate_size=28672, num_hidden_layers=80, num_attention_heads=64, num_key_value_heads=8,
                     head_dim=128, rope_theta=10000.0, rope_scaling=None, max_position_embeddings=4096,
                     tie_word_embeddings=False)
llama3_8b = _preset("Llama-3 8B (reference 8B_config_llama3)", hidden_size=4096, intermediate_size=14336,
                    num_hidden_layers=32, num_attention_heads=32, num_key_value_heads=8, head_dim=128, rope_scaling=None,
                    max_position_embeddings=8192, tie_word_embeddings=False)
llama3_70b = _preset("Llama-3 70B (reference 70B_config_llama3)", hidden_size=8192, intermediate_size=28672,
                     num_hidden_layers=80, num_attention_heads=64, num_key_value_heads=8, head_dim=128, rope_scaling=None,
                     max_position_embeddings=8192, tie_word_embeddings=False)
llama3_1_70b = _preset("Llama-3.1 70B (reference 70B_config_llama3.1)", hidden_size=8192, intermediate_size=28672,
                       num_hidden_layers=80, num_attention_heads=64, num_key_value_heads=8, head_dim=128,
                       rope_scaling=dict(_LLAMA31_ROPE), tie_word_embeddings=False)

PRESETS = {"llama3.2-1b": llama3_2_1b, "llama3.2-3b": llama3_2_3b, "llama3.1-8b": llama3_1_8b, "tiny": tiny_llama,
           "llama2-7b": llama2_7b, "llama2-13b": llama2_13b, "llama2-70b": llama2_70b, "llama3-8b": llama3_8b,
           "llama3-70b": llama3_70b, "llama3.1-70b": llama3_1_70b}


def config_from_json(config_cls, path_or_dict, **overrides):
    """An HF-style ``config.json`` (path, directory or dict) -> ``config_cls`` (any config dataclass of this package);
    unknown keys are dropped, ``overrides`` win."""
    import json
    import os
    from dataclasses import fields

    d = path_or_dict
    if not isinstance(d, dict):
        p = os.path.join(d, "config.json") if os.path.isdir(d) else d
        with open(p) as f:
            d = json.load(f)
    known = {f.name for f in fields(config_cls)}
    kw = {k: v for k, v in d.items() if k in known and k != "torch_dtype"}
    if isinstance(d.get("torch_dtype"), str):
        kw["torch_dtype"] = getattr(torch, d["torch_dtype"])
    kw.update(overrides)
    return config_cls(**kw)


def config_from_hf_json(path_or_dict, **overrides) -> LlamaConfig:
    """``config_from_json`` for the Llama family.  The reference's scripts pass ``--model_path <dir with config.json>``."""
    return config_from_json(LlamaConfig, path_or_dict, **overrides)


def _init_fn(std: float):
    def f(w):
        return nn.init.normal_(w, mean=0.0, std=std)

    return f


class LlamaMLP(nn.Module):
    """``down(silu(gate) * up)`` with gate|up fused in one column-parallel weight sharded with
    ``stride=2`` (reference ``modeling_llama_nxd.py:153-211``)."""

    def __init__(self, config: LlamaConfig, device=None):
        super().__init__()
        sp, dt = config.sequence_parallel_enabled, config.torch_dtype
        init = _init_fn(config.initializer_range)
        self.gate_up_proj = ColumnParallelLinear(config.hidden_size, 2 * config.intermediate_size, bias=False,
                                                 gather_output=False, stride=2, init_method=init,
                                                 sequence_parallel_enabled=sp, dtype=dt, device=device)
        self.down_proj = RowParallelLinear(config.intermediate_size, config.hidden_size, bias=False,
                                           input_is_parallel=True, init_method=init,
                                           sequence_parallel_enabled=sp, dtype=dt, device=device)
        self.selective = config.selective_checkpoint_enabled

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        gate_up = self.gate_up_proj(x)
        if self.selective and self.training and torch.is_grad_enabled():
            inter = checkpoint(act_ops.swiglu, gate_up, use_reentrant=False)
        else:
            inter = act_ops.swiglu(gate_up)
        return self.down_proj(inter)


class LlamaAttention(nn.Module):
    """GQA attention: fused QKV projection -> RoPE -> causal flash attention -> o_proj
    (reference ``modeling_llama_nxd.py:238-440``)."""

    def __init__(self, config: LlamaConfig, device=None):
        super().__init__()
        self.config = config
        tp = ps.get_tensor_model_parallel_size()
        self.head_dim = config.head_dim
        self.num_heads = config.num_attention_heads
        self.num_kv_heads = config.num_key_value_heads
        kvm = config.kv_shared_group_size
        if tp > self.num_kv_heads * kvm:
            kvm = tp // self.num_kv_heads  # replicate KV heads so every rank owns one
        self.kv_size_multiplier = kvm
        assert self.num_heads % tp == 0, "attention heads must divide by tp (use pad_model otherwise)"
        self.q_heads_local = self.num_heads // tp
        self.kv_heads_local = max(1, self.num_kv_heads * kvm // tp)
        sp, dt = config.sequence_parallel_enabled, config.torch_dtype
        init = _init_fn(config.initializer_range)
        self.qkv_proj = GQAQKVColumnParallelLinear(
            config.hidden_size, [self.num_heads * self.head_dim, self.num_kv_heads * self.head_dim], bias=False,
            gather_output=False, kv_size_multiplier=kvm, fuse_qkv=config.fuse_qkv, init_method=init,
            sequence_parallel_enabled=sp, dtype=dt, device=device)
        self.o_proj = RowParallelLinear(self.num_heads * self.head_dim, config.hidden_size, bias=False,
                                        input_is_parallel=True, init_method=init, sequence_parallel_enabled=sp,
                                        dtype=dt, device=device)

    def preshard_hook(self, model_state_dict: dict, prefix: str) -> None:
        """Called on an UN-SHARDED state dict right before it is cut for a TP rank (``get_sharded_checkpoint`` /
        ``parallel.checkpointing``; reference ``GQAQKVColumnParallelLinear.preshard_hook``, ``modules/qkv_linear.py``).
        With replicated KV heads rank ``r`` holds KV head ``r % n_kv``, so the query heads -- and the matching columns of
        the output projection -- are re-ordered for a contiguous split to hand every rank the queries of ITS KV head.
        (Checkpoints written by the offline converter are permuted and sharded already and never pass through here.)"""
        base = prefix[: -len("weight")]
        for name, child in self.named_children():      # the walk stops at a module with a hook: pass it on ourselves
            if hasattr(child, "preshard_hook"):
                child.preshard_hook(model_state_dict, base + name + ".weight")
        m = self.kv_size_multiplier
        if m <= 1:
            return
        from ..parallel import parallel_state as ps
        from ..scripts.checkpoint_converter import q_head_permutation

        D = self.head_dim
        perm = q_head_permutation(self.num_heads, self.num_kv_heads, ps.get_tensor_model_parallel_size(), m)
        rows = torch.tensor([h * D + d for h in perm for d in range(D)])
        qs = self.num_heads * D
        for key in (base + "qkv_proj.weight_qkv", base + "qkv_proj.weight_q", base + "qkv_proj.bias_qkv",
                    base + "qkv_proj.bias_q"):
            t = model_state_dict.get(key)
            if t is not None:
                model_state_dict[key] = torch.cat([t[:qs][rows.to(t.device)], t[qs:]], dim=0)
        o = model_state_dict.get(base + "o_proj.weight")
        if o is not None:
            model_state_dict[base + "o_proj.weight"] = o[:, rows.to(o.device)]

    def forward(self, x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
        nq, nkv, D = self.q_heads_local, self.kv_heads_local, self.head_dim
        if self.config.fuse_qkv:
            qkv = self.qkv_proj(x, return_fused=True)            # [S, B, (nq + 2 nkv) * D]
        else:
            qkv = torch.cat(self.qkv_proj(x), dim=-1)
        S, B = qkv.shape[0], qkv.shape[1]
        qkv = qkv.reshape(S * B, nq + 2 * nkv, D)
        # rotate q and k heads (the first nq + nkv heads) in one pass; token t = s * B + b sits at position s
        qkv = rope_ops.apply_rope(qkv, cos, sin, nq + nkv, pos_div=B, pos_mod=S)
        qkv = qkv.reshape(S, B, nq + 2 * nkv, D)
        q = qkv[:, :, :nq].permute(1, 2, 0, 3)                   # [B, nq, S, D] strided views, no copies
        k = qkv[:, :, nq:nq + nkv].permute(1, 2, 0, 3)
        v = qkv[:, :, nq + nkv:].permute(1, 2, 0, 3)
        o = attn_ops.flash_attn_func(q, k, v, causal=True)       # [B, nq, S, D] stored as [S, B, nq, D]
        o = o.permute(2, 0, 1, 3).reshape(S, B, nq * D)
        return self.o_proj(o)


class LlamaDecoderLayer(nn.Module):
    def __init__(self, config: LlamaConfig, device=None):
        super().__init__()
        sp, dt = config.sequence_parallel_enabled, config.torch_dtype
        self.self_attn = LlamaAttention(config, device=device)
        self.mlp = LlamaMLP(config, device=device)
        self.input_layernorm = RMSNorm(config.hidden_size, config.rms_norm_eps, sequence_parallel_enabled=sp,
                                       dtype=dt, device=device)
        self.post_attention_layernorm = RMSNorm(config.hidden_size, config.rms_norm_eps,
                                                sequence_parallel_enabled=sp, dtype=dt, device=device)

    def forward(self, hidden: torch.Tensor, residual: Optional[torch.Tensor], cos: torch.Tensor,
                sin: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """``hidden`` is the previous block's un-added output, ``residual`` the stream it must be added to."""
        eps = self.input_layernorm.variance_epsilon
        if residual is None:
            residual = hidden
            h = norm_ops.rms_norm(hidden, self.input_layernorm.weight, eps)
        else:
            h, residual = norm_ops.fused_add_rms_norm(hidden, residual, self.input_layernorm.weight, eps)
        a = self.self_attn(h, cos, sin)
        h, residual = norm_ops.fused_add_rms_norm(a, residual, self.post_attention_layernorm.weight, eps)
        return self.mlp(h), residual


class RotaryTables(nn.Module):
    """cos/sin tables ``[S, D/2]`` for the sequence length of the batch (cached per (S, device)).  A pipeline-trace
    leaf: the FX graph only sees ``rotary(input_ids) -> (cos, sin)``."""

    _nxd_fx_leaf = True

    def __init__(self, config: LlamaConfig):
        super().__init__()
        self.head_dim, self.theta, self.scaling = config.head_dim, config.rope_theta, config.rope_scaling
        self._cache: Dict[Tuple[int, str], Tuple[torch.Tensor, torch.Tensor]] = {}

    def forward(self, input_ids: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        key = (input_ids.shape[1], str(input_ids.device))
        if key not in self._cache:
            self._cache = {key: rope_ops.build_rope_cache(key[0], self.head_dim, self.theta, self.scaling,
                                                          device=input_ids.device)}
        return self._cache[key]


class CausalLMLoss(nn.Module):
    """Mean next-token loss from vocab-parallel logits ``[S, B, V/tp]`` and labels ``[B, S]`` (shifted here,
    reference ``modeling_llama_nxd.py:733-747``).  A pipeline-trace leaf."""

    _nxd_fx_leaf = True

    def forward(self, logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        B, S = labels.shape
        shifted = torch.full((S, B), -100, dtype=torch.long, device=labels.device)
        shifted[:-1] = labels[:, 1:].t()
        per_token = parallel_cross_entropy(logits, shifted)        # [S, B] fp32
        valid = shifted >= 0
        return (per_token * valid).sum() / valid.sum().clamp(min=1)


class LlamaModel(nn.Module):
    def __init__(self, config: LlamaConfig, device=None):
        super().__init__()
        self.config = config
        dt = config.torch_dtype
        self.embed_tokens = ParallelEmbedding(config.vocab_size, config.hidden_size,
                                              init_method=_init_fn(config.initializer_range), dtype=dt, device=device,
                                              sequence_parallel_enabled=True)  # always emits [S(/tp), B, H]
        self.layers = nn.ModuleList([LlamaDecoderLayer(config, device=device)
                                     for _ in range(config.num_hidden_layers)])
        self.norm = RMSNorm(config.hidden_size, config.rms_norm_eps,
                            sequence_parallel_enabled=config.sequence_parallel_enabled, dtype=dt, device=device)
        self.rotary = RotaryTables(config)

    def forward(self, input_ids: torch.Tensor) -> torch.Tensor:
        """``input_ids [B, S]`` -> final normed hidden states ``[S(/tp), B, H]``."""
        cfg = self.config
        hidden = self.embed_tokens(input_ids)  # [S/tp, B, H] (reduce-scatter) -- SP region entered here
        if not cfg.sequence_parallel_enabled and ps.get_tensor_model_parallel_size() > 1:
            hidden = mappings.gather_from_sequence_parallel_region(hidden, to_model_parallel=False)
        cos_sin = self.rotary(input_ids)
        cos, sin = cos_sin[0], cos_sin[1]
        residual = None
        for layer in self.layers:
            if cfg.activation_checkpointing and self.training and torch.is_grad_enabled():
                hidden, residual = checkpoint(layer, hidden, residual, cos, sin, use_reentrant=False)
            else:
                hidden, residual = layer(hidden, residual, cos, sin)
        return self.norm(hidden, residual)   # fused (hidden + residual) -> RMSNorm


class LlamaForCausalLM(nn.Module):
    """``forward(input_ids, labels)`` returns the mean next-token loss (labels are shifted inside,
    reference ``modeling_llama_nxd.py:733-747``); without labels returns vocab-parallel logits."""

    def __init__(self, config: LlamaConfig, device=None):
        super().__init__()
        self.config = config
        self.model = LlamaModel(config, device=device)
        sp = config.sequence_parallel_enabled
        self.lm_head = ColumnParallelLinear(config.hidden_size, config.vocab_size, bias=False, gather_output=False,
                                            init_method=_init_fn(config.initializer_range),
                                            sequence_parallel_enabled=sp, dtype=config.torch_dtype,
                                            device="meta" if config.tie_word_embeddings else device)
        self.loss_fn = CausalLMLoss()
        if config.tie_word_embeddings:
            self.tie_weights()

    def tie_weights(self) -> None:
        self.lm_head.weight = self.model.embed_tokens.weight

    def get_input_embeddings(self):
        return self.model.embed_tokens

    def get_output_embeddings(self):
        return self.lm_head

    def forward(self, input_ids: torch.Tensor, labels: Optional[torch.Tensor] = None,
                attention_mask: Optional[torch.Tensor] = None):
        hidden = self.model(input_ids)                 # [S/tp, B, H]
        logits = self.lm_head(hidden)                  # [S, B, V/tp]  (AG->GEMM under SP)
        if labels is None:
            return logits
        return self.loss_fn(logits, labels)


def flops_per_token(cfg: LlamaConfig, seq_len: int) -> float:
    """Model FLOPs per trained token (fwd + bwd = 3x fwd; causal attention counted at half)."""
    h, L = cfg.hidden_size, cfg.num_hidden_layers
    qkv = h * cfg.head_dim * (cfg.num_attention_heads + 2 * cfg.num_key_value_heads)
    dense = L * (qkv + cfg.num_attention_heads * cfg.head_dim * h + 3 * h * cfg.intermediate_size) + cfg.vocab_size * h
    attn = L * 2 * seq_len * cfg.num_attention_heads * cfg.head_dim * 0.5
    return 3.0 * 2.0 * (dense + attn)
