"""BERT encoder for masked-LM + next-sentence pre-training on the tensor-parallel layers (reference
``examples/training/tp_dp_bert_hf_pretrain/tp_dp_bert_large_hf_pretrain_hdf5.py``, which swaps the HF BERT linears for
Column / RowParallelLinear and the word embedding for ParallelEmbedding).

Post-LayerNorm blocks, bidirectional attention (flash attention without the causal mask; an explicit padding mask
takes the masked-softmax path), GELU MLP, MLM decoder tied to the vocab-parallel word embedding, pooler + 2-way
next-sentence head.  Activations are sequence-first ``[S, B, H]``; with sequence parallelism the residual stream is
sharded on ``S`` and the replicated parameters that only see a sequence shard (LayerNorms, MLM transform) carry the
``sequence_parallel_enabled`` tag so that their gradients are all-reduced over TP.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import attention as attn_ops
from ..parallel import mappings
from ..parallel import parallel_state as ps
from ..parallel.layer_norm import LayerNorm
from ..parallel.layers import ColumnParallelLinear, ParallelEmbedding, RowParallelLinear
from ..parallel.loss_functions import parallel_cross_entropy
from .llama import _init_fn


@dataclass
class BertConfig:
    vocab_size: int = 30528                 # 30522 padded to a multiple of 64 (TP shards, tensor-core tiles)
    hidden_size: int = 1024
    num_hidden_layers: int = 24
    num_attention_heads: int = 16
    intermediate_size: int = 4096
    max_position_embeddings: int = 512
    type_vocab_size: int = 2
    layer_norm_eps: float = 1e-12
    initializer_range: float = 0.02
    sequence_parallel_enabled: bool = False
    activation_checkpointing: bool = False
    torch_dtype: torch.dtype = torch.bfloat16

    @property
    def head_dim(self) -> int:
        return self.hidden_size // self.num_attention_heads


def bert_large(**kw) -> BertConfig:
    return BertConfig(**kw)


def tiny_bert(**kw) -> BertConfig:
    base = dict(vocab_size=512, hidden_size=128, num_hidden_layers=2, num_attention_heads=4, intermediate_size=256,
                max_position_embeddings=128, torch_dtype=torch.float32)
    base.update(kw)
    return BertConfig(**base)


def _tag_sp(module: nn.Module, enabled: bool) -> None:
    for p in module.parameters():
        setattr(p, "sequence_parallel_enabled", enabled)


class BertEmbeddings(nn.Module):
    def __init__(self, config: BertConfig, device=None):
        super().__init__()
        dt, sp = config.torch_dtype, config.sequence_parallel_enabled
        self.sp = sp
        self.word_embeddings = ParallelEmbedding(config.vocab_size, config.hidden_size,
                                                 init_method=_init_fn(config.initializer_range), dtype=dt, device=device,
                                                 sequence_parallel_enabled=True)       # always [S(/tp), B, H]
        self.position_embeddings = nn.Embedding(config.max_position_embeddings, config.hidden_size, dtype=dt, device=device)
        self.token_type_embeddings = nn.Embedding(config.type_vocab_size, config.hidden_size, dtype=dt, device=device)
        self.LayerNorm = LayerNorm(config.hidden_size, eps=config.layer_norm_eps, sequence_parallel_enabled=sp, dtype=dt,
                                   device=device)
        # no sequence-parallel tag on these two: the scatter below all-gathers their gradient in its backward, so every
        # rank already holds the complete gradient
        for emb in (self.position_embeddings, self.token_type_embeddings):
            nn.init.normal_(emb.weight, std=config.initializer_range)

    def forward(self, input_ids: torch.Tensor, token_type_ids: Optional[torch.Tensor]) -> torch.Tensor:
        B, S = input_ids.shape
        tp = ps.get_tensor_model_parallel_size()
        words = self.word_embeddings(input_ids)                              # [S/tp, B, H] (or [S, B, H] at tp = 1)
        if token_type_ids is None:
            token_type_ids = torch.zeros_like(input_ids)
        pos = torch.arange(S, device=input_ids.device)
        extra = self.position_embeddings(pos)[:, None, :] + self.token_type_embeddings(token_type_ids.t())   # [S, B, H]
        if tp > 1 and self.sp:
            extra = mappings.scatter_to_sequence_parallel_region(extra)
        elif tp > 1:
            words = mappings.gather_from_sequence_parallel_region(words, to_model_parallel=False)
        return self.LayerNorm(words + extra)


class BertSelfAttention(nn.Module):
    def __init__(self, config: BertConfig, device=None):
        super().__init__()
        self.config = config
        tp = ps.get_tensor_model_parallel_size()
        assert config.num_attention_heads % tp == 0
        self.heads_local = config.num_attention_heads // tp
        sp, dt, init = config.sequence_parallel_enabled, config.torch_dtype, _init_fn(config.initializer_range)
        # one fused projection, head-major rows [heads, (q, k, v), D]: a TP split is a split over heads
        self.query_key_value = ColumnParallelLinear(config.hidden_size, 3 * config.hidden_size, bias=True,
                                                    gather_output=False, init_method=init, sequence_parallel_enabled=sp,
                                                    dtype=dt, device=device)
        self.dense = RowParallelLinear(config.hidden_size, config.hidden_size, bias=True, input_is_parallel=True,
                                       init_method=init, sequence_parallel_enabled=sp, dtype=dt, device=device)

    def forward(self, x: torch.Tensor, key_padding_mask: Optional[torch.Tensor]) -> torch.Tensor:
        h, D = self.heads_local, self.config.head_dim
        qkv = self.query_key_value(x)
        S, B = qkv.shape[0], qkv.shape[1]
        qkv = qkv.reshape(S, B, h, 3, D)
        q, k, v = (qkv[:, :, :, i].permute(1, 2, 0, 3) for i in range(3))   # [B, h, S, D] views
        if key_padding_mask is None:
            o = attn_ops.flash_attn_func(q, k, v, causal=False)
        else:   # padded batch: masked softmax (keys of padding tokens get -inf)
            scores = torch.matmul(q.float(), k.float().transpose(-1, -2)) / math.sqrt(D)
            scores = scores.masked_fill(~key_padding_mask[:, None, None, :].bool(), float("-inf"))
            o = torch.matmul(torch.softmax(scores, dim=-1), v.float()).to(q.dtype)
        return self.dense(o.permute(2, 0, 1, 3).reshape(S, B, h * D))


class BertLayer(nn.Module):
    def __init__(self, config: BertConfig, device=None):
        super().__init__()
        sp, dt, init = config.sequence_parallel_enabled, config.torch_dtype, _init_fn(config.initializer_range)
        self.attention = BertSelfAttention(config, device=device)
        self.attention_layernorm = LayerNorm(config.hidden_size, eps=config.layer_norm_eps, sequence_parallel_enabled=sp,
                                             dtype=dt, device=device)
        self.intermediate = ColumnParallelLinear(config.hidden_size, config.intermediate_size, bias=True,
                                                 gather_output=False, init_method=init, sequence_parallel_enabled=sp,
                                                 dtype=dt, device=device)
        self.output = RowParallelLinear(config.intermediate_size, config.hidden_size, bias=True, input_is_parallel=True,
                                        init_method=init, sequence_parallel_enabled=sp, dtype=dt, device=device)
        self.output_layernorm = LayerNorm(config.hidden_size, eps=config.layer_norm_eps, sequence_parallel_enabled=sp,
                                          dtype=dt, device=device)

    def forward(self, x: torch.Tensor, key_padding_mask: Optional[torch.Tensor]) -> torch.Tensor:
        x = self.attention_layernorm(x + self.attention(x, key_padding_mask))
        return self.output_layernorm(x + self.output(F.gelu(self.intermediate(x))))


class BertForPreTraining(nn.Module):
    """``forward(input_ids [B, S], token_type_ids, attention_mask, labels [B, S] (-100 = not masked),
    next_sentence_label [B])`` -> ``mlm_loss + nsp_loss``; without labels ``(vocab-parallel logits [S, B, V/tp],
    nsp logits [B, 2])``."""

    def __init__(self, config: BertConfig, device=None):
        super().__init__()
        self.config = config
        sp, dt, init = config.sequence_parallel_enabled, config.torch_dtype, _init_fn(config.initializer_range)
        self.embeddings = BertEmbeddings(config, device=device)
        self.layers = nn.ModuleList([BertLayer(config, device=device) for _ in range(config.num_hidden_layers)])
        # MLM head: transform on the (possibly sequence-sharded) hidden states, decoder tied to the word embedding
        self.transform = nn.Linear(config.hidden_size, config.hidden_size, dtype=dt, device=device)
        self.transform_layernorm = LayerNorm(config.hidden_size, eps=config.layer_norm_eps, sequence_parallel_enabled=sp,
                                             dtype=dt, device=device)
        self.decoder = ColumnParallelLinear(config.hidden_size, config.vocab_size, bias=False, gather_output=False,
                                            init_method=init, sequence_parallel_enabled=sp, dtype=dt, device=device)
        self.decoder.weight = self.embeddings.word_embeddings.weight
        self.decoder_bias = nn.Parameter(torch.zeros(self.decoder.weight.shape[0], dtype=dt, device=device))
        setattr(self.decoder_bias, "tensor_model_parallel", True)
        setattr(self.decoder_bias, "partition_dim", 0)
        setattr(self.decoder_bias, "partition_stride", 1)
        self.pooler = nn.Linear(config.hidden_size, config.hidden_size, dtype=dt, device=device)
        self.seq_relationship = nn.Linear(config.hidden_size, 2, dtype=dt, device=device)
        for lin in (self.transform, self.pooler, self.seq_relationship):
            nn.init.normal_(lin.weight, std=config.initializer_range)
            nn.init.zeros_(lin.bias)
        _tag_sp(self.transform, sp)

    def _cls_token(self, hidden: torch.Tensor) -> torch.Tensor:
        """Hidden state of position 0 on every rank (with sequence parallelism only TP rank 0 holds it)."""
        if not (self.config.sequence_parallel_enabled and ps.get_tensor_model_parallel_size() > 1):
            return hidden[0]
        mine = hidden[0] if ps.get_tensor_model_parallel_rank() == 0 else torch.zeros_like(hidden[0])
        return mappings.reduce_from_tensor_model_parallel_region(mine)

    def forward(self, input_ids: torch.Tensor, token_type_ids: Optional[torch.Tensor] = None,
                attention_mask: Optional[torch.Tensor] = None, labels: Optional[torch.Tensor] = None,
                next_sentence_label: Optional[torch.Tensor] = None):
        from torch.utils.checkpoint import checkpoint

        cfg = self.config
        if (attention_mask is not None and not isinstance(attention_mask, torch.fx.Proxy)
                and bool(attention_mask.all())):       # nothing is padded: take the flash-attention path
            attention_mask = None
        hidden = self.embeddings(input_ids, token_type_ids)
        for layer in self.layers:
            if cfg.activation_checkpointing and self.training and torch.is_grad_enabled():
                hidden = checkpoint(layer, hidden, attention_mask, use_reentrant=False)
            else:
                hidden = layer(hidden, attention_mask)
        mlm = self.transform_layernorm(F.gelu(self.transform(hidden)))
        logits = self.decoder(mlm) + self.decoder_bias                    # [S, B, V/tp]
        nsp_logits = self.seq_relationship(torch.tanh(self.pooler(self._cls_token(hidden))))
        if labels is None:
            return logits, nsp_logits
        per_token = parallel_cross_entropy(logits, labels.t().contiguous())          # [S, B]
        valid = labels.t() >= 0
        loss = (per_token * valid).sum() / valid.sum().clamp(min=1)
        if next_sentence_label is not None:
            loss = loss + F.cross_entropy(nsp_logits.float(), next_sentence_label)
        return loss


def convert_bert_hf_state_dict(hf, config: BertConfig):
    """HF ``BertForPreTraining`` names -> this model's names, un-sharded.  The separate query / key / value projections
    are interleaved per head into the fused ``[heads, (q, k, v), D]`` rows, so that a contiguous TP split is a split
    over heads; the vocabulary is zero-padded up to ``config.vocab_size`` (padding rows never win: bias ``-inf``-like)."""
    nh, D, H = config.num_attention_heads, config.head_dim, config.hidden_size

    def pad_vocab(t, fill=0.0):
        extra = config.vocab_size - t.shape[0]
        if extra <= 0:
            return t
        return torch.cat([t, torch.full((extra,) + tuple(t.shape[1:]), fill, dtype=t.dtype)], dim=0)

    out = {"embeddings.word_embeddings.weight": pad_vocab(hf["bert.embeddings.word_embeddings.weight"]),
           "embeddings.position_embeddings.weight": hf["bert.embeddings.position_embeddings.weight"],
           "embeddings.token_type_embeddings.weight": hf["bert.embeddings.token_type_embeddings.weight"],
           "embeddings.LayerNorm.weight": hf["bert.embeddings.LayerNorm.weight"],
           "embeddings.LayerNorm.bias": hf["bert.embeddings.LayerNorm.bias"],
           "transform.weight": hf["cls.predictions.transform.dense.weight"],
           "transform.bias": hf["cls.predictions.transform.dense.bias"],
           "transform_layernorm.weight": hf["cls.predictions.transform.LayerNorm.weight"],
           "transform_layernorm.bias": hf["cls.predictions.transform.LayerNorm.bias"],
           "decoder_bias": pad_vocab(hf["cls.predictions.bias"], fill=-1e4),
           "pooler.weight": hf["bert.pooler.dense.weight"], "pooler.bias": hf["bert.pooler.dense.bias"],
           "seq_relationship.weight": hf["cls.seq_relationship.weight"],
           "seq_relationship.bias": hf["cls.seq_relationship.bias"]}
    out["decoder.weight"] = out["embeddings.word_embeddings.weight"]
    for i in range(config.num_hidden_layers):
        src, dst = f"bert.encoder.layer.{i}.", f"layers.{i}."
        ws = [hf[src + f"attention.self.{n}.weight"].reshape(nh, D, H) for n in ("query", "key", "value")]
        bs = [hf[src + f"attention.self.{n}.bias"].reshape(nh, D) for n in ("query", "key", "value")]
        out[dst + "attention.query_key_value.weight"] = torch.stack(ws, dim=1).reshape(3 * H, H)
        out[dst + "attention.query_key_value.bias"] = torch.stack(bs, dim=1).reshape(3 * H)
        for a, b in (("attention.output.dense", "attention.dense"), ("attention.output.LayerNorm", "attention_layernorm"),
                     ("intermediate.dense", "intermediate"), ("output.dense", "output"),
                     ("output.LayerNorm", "output_layernorm")):
            out[dst + b + ".weight"] = hf[src + a + ".weight"]
            out[dst + b + ".bias"] = hf[src + a + ".bias"]
    return out


__all__ = ["BertConfig", "BertForPreTraining", "BertLayer", "bert_large", "convert_bert_hf_state_dict", "tiny_bert"]
