"""Llama-3.x inference model with KV cache: context encoding + token generation.

Reference: ``examples/inference/modules/model_base.py:52-518`` (``NeuronBaseModel``: KV cache as aliased graph state,
context-encoding / token-generation branches, last-token gather, on-device sampling),
``examples/inference/modules/attention/attention_base.py:27-200`` (prefill flash attention vs decode softmax over the
cached + the new key), ``examples/inference/llama3/neuron_modeling_llama.py`` (Llama-3 RoPE scaling, tied lm_head).

The module shares its parameter names with the training model (``models/llama.py``), so trainer checkpoints load
unchanged.  Differences from the training forward:
  * batch-first ``[B, T, H]`` activations, no sequence parallelism; RowParallel layers all-reduce (fused GEMM->AR
    when the symmetric-memory context is up);
  * context encoding (T > 1): causal flash attention over the prompt, K/V written into cache rows ``[0, T)`` of cache
    line ``seq_ids[b]`` by the in-place KV kernel;
  * token generation (T == 1): K/V of the new token written at ``position_ids[b]``, then the split-K decode attention
    kernel over ``[0, position_ids[b]]`` -- the reference computes scores against the whole padded cache with a mask and
    re-emits the cache tensor every step;
  * logits only for the last real token of every sequence; sampling happens on device.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn as nn

from ..models.llama import LlamaConfig, LlamaForCausalLM
from ..models.dbrx import DbrxConfig, DbrxForCausalLM
from ..models.mixtral import MixtralConfig, MixtralForCausalLM
from ..ops import activation as act_ops
from ..ops import attention as attn_ops
from ..ops import comm_linear
from ..ops import norm as norm_ops
from ..ops import rope as rope_ops
from ..parallel import mappings
from ..parallel import parallel_state as ps
from .config import InferenceConfig
from .kv_cache import KVCache
from .sampling import Sampler


class KVCacheInferenceMixin:
    """Context-encoding / token-generation forward over a KV cache, shared by the model families.  The host class is a
    training model (same parameter names, so trainer checkpoints load unchanged); it tells the mixin where its pieces
    are through ``_decoder_layers / _embedding / _final_norm`` and how a layer's feed-forward runs (``_mlp``)."""

    def _init_inference(self, inference_config: Optional[InferenceConfig]) -> None:
        self.inference_config = inference_config or InferenceConfig()
        self.kv_cache: Optional[KVCache] = None
        self.sampler = Sampler(self.inference_config)
        self._rope_tables = None
        self.num_medusa_heads = 0
        if self.inference_config.num_medusa_heads > 0:
            from .medusa import add_medusa_heads

            first = next(self.parameters())
            add_medusa_heads(self, self.inference_config.num_medusa_heads, device=first.device)
        self.eval()

    # -- where the host model keeps its modules (Llama layout by default) ------------------------------------------
    def _decoder_layers(self):
        return self.model.layers

    def _embedding(self):
        return self.model.embed_tokens

    def _final_norm(self):
        return self.model.norm

    def _mlp(self, layer, h: torch.Tensor) -> torch.Tensor:
        return layer.mlp.down_proj(act_ops.swiglu(layer.mlp.gate_up_proj(h)))

    # -- state -------------------------------------------------------------------------------------------------
    def init_kv_cache(self, device=None, dtype=None) -> KVCache:
        ic, cfg = self.inference_config, self.config
        attn0 = self._decoder_layers()[0].self_attn
        device = device or next(self.parameters()).device
        self.kv_cache = KVCache(cfg.num_hidden_layers, ic.max_batch_size, attn0.kv_heads_local, ic.seq_len, cfg.head_dim,
                                dtype=dtype or cfg.torch_dtype, device=device, paged=ic.paged_kv, page_size=ic.page_size,
                                num_pages=ic.num_kv_pages if ic.paged_kv else None)
        return self.kv_cache

    def _rope(self, device) -> Tuple[torch.Tensor, torch.Tensor]:
        if self._rope_tables is None or self._rope_tables[0].device != device:
            cfg = self.config
            self._rope_tables = rope_ops.build_rope_cache(self.inference_config.seq_len, cfg.head_dim, cfg.rope_theta,
                                                          cfg.rope_scaling, device=device)
        return self._rope_tables

    # -- one decoder layer ---------------------------------------------------------------------------------------
    def _qkv_to_cache(self, layer_idx: int, attn, h: torch.Tensor, positions: torch.Tensor, seq_ids):
        """Projection, RoPE at ``positions`` and the write of the new K/V rows into the cache; ``(q, k, v, ck, cv)``."""
        B, T, _ = h.shape
        nq, nkv, D = attn.q_heads_local, attn.kv_heads_local, attn.head_dim
        if self.config.fuse_qkv:
            qkv = attn.qkv_proj(h, return_fused=True)
        else:
            qkv = torch.cat(attn.qkv_proj(h), dim=-1)
        qkv = self._post_qkv(qkv.reshape(B * T, nq + 2 * nkv, D))
        cos, sin = self._rope(h.device)
        qkv = rope_ops.rope_inplace_(qkv.contiguous(), cos, sin, nq + nkv, positions=positions.reshape(-1))
        qkv = qkv.reshape(B, T, nq + 2 * nkv, D)
        q = qkv[:, :, :nq]
        k = qkv[:, :, nq:nq + nkv]
        v = qkv[:, :, nq + nkv:]
        ck, cv = self.kv_cache.layer(layer_idx)
        kw = dict(block_table=self.kv_cache.block_table, page_size=self.kv_cache.page_size) if self.kv_cache.paged else {}
        attn_ops.kv_cache_update(ck, cv, k.permute(0, 2, 1, 3), v.permute(0, 2, 1, 3), positions[:, 0], seq_ids=seq_ids,
                                 **kw)
        return q, k, v, ck, cv

    def _attention_with_past(self, layer_idx: int, attn, h: torch.Tensor, positions: torch.Tensor,
                             seq_ids: torch.Tensor) -> torch.Tensor:
        """A few NEW tokens per sequence attending to the cache AND causally to each other (speculation verify):
        every (sequence, token) pair is one row of the decode-attention kernel with its own length."""
        B, T, _ = h.shape
        nq, D = attn.q_heads_local, attn.head_dim
        q, _, _, ck, cv = self._qkv_to_cache(layer_idx, attn, h, positions, seq_ids)
        rows = seq_ids.long().repeat_interleave(T)
        lengths = (positions + 1).reshape(-1).to(torch.int32)
        q2 = q.reshape(B * T, nq, D).contiguous()
        # every row reads the cache line of its sequence in place (T rows share a line: no gathered copies)
        o = self._decode_attention(q2, ck, cv, rows, lengths)
        return attn.o_proj(o.reshape(B, T, nq * D))

    def _attention(self, layer_idx: int, attn, h: torch.Tensor, positions: torch.Tensor, seq_ids: torch.Tensor,
                   lengths: torch.Tensor, prefill: bool) -> torch.Tensor:
        B, T, _ = h.shape
        nq, D = attn.q_heads_local, attn.head_dim
        q, k, v, ck, cv = self._qkv_to_cache(layer_idx, attn, h, positions, seq_ids)
        if prefill:
            o = attn_ops.flash_attn_func(q.permute(0, 2, 1, 3), k.permute(0, 2, 1, 3), v.permute(0, 2, 1, 3), causal=True)
            o = o.permute(0, 2, 1, 3).reshape(B, T, nq * D)
        else:
            # continuous batching: rows of the step map to arbitrary cache lines, which the kernel follows itself
            o = self._decode_attention(q[:, 0].contiguous(), ck, cv, seq_ids, lengths).reshape(B, 1, nq * D)
        return attn.o_proj(o)

    def _add_norm(self, norm, hidden: torch.Tensor, residual: Optional[torch.Tensor]):
        """``(norm(hidden + residual), hidden + residual)`` -- RMSNorm through the fused add + norm kernel; model
        families with another norm override this."""
        eps = norm.variance_epsilon
        if residual is None:
            return norm_ops.rms_norm(hidden, norm.weight, eps), hidden
        return norm_ops.fused_add_rms_norm(hidden, residual, norm.weight, eps)

    def _post_qkv(self, qkv: torch.Tensor) -> torch.Tensor:
        """Hook on the fused projection output ``[B*T, heads, D]`` before RoPE (DBRX clamps it)."""
        return qkv

    # -- fused token-generation layer (6 launches) -------------------------------------------------------------------
    def _fused_decode_ok(self, layer, hidden: torch.Tensor, batched: bool = False) -> bool:
        """The fused decode kernels apply to the plain Llama layer shape: RMSNorm, bias-free fused QKV, SwiGLU MLP,
        bf16 weights, at most 8 sequences, native backends selected, and (tensor parallel) the one-shot all-reduce
        context for the row-parallel projections.  Model families that override the hooks take the generic path."""
        from ..ops import backend, decode_fused
        from ..parallel.layers import RowParallelLinear

        cached = getattr(layer, "_nxd_fused_decode", None)
        if cached is None:
            cls = type(self)
            attn, mlp = getattr(layer, "self_attn", None), getattr(layer, "mlp", None)
            qkv = getattr(attn, "qkv_proj", None)
            w = getattr(qkv, "weight_qkv", None)
            cached = bool(
                cls._post_qkv is KVCacheInferenceMixin._post_qkv and cls._add_norm is KVCacheInferenceMixin._add_norm
                and cls._mlp is KVCacheInferenceMixin._mlp and getattr(self.config, "fuse_qkv", False)
                and isinstance(w, torch.Tensor) and w.dtype == torch.bfloat16 and getattr(qkv, "bias_qkv", None) is None
                and type(getattr(attn, "o_proj", None)) is RowParallelLinear and attn.o_proj.bias is None
                and type(getattr(mlp, "down_proj", None)) is RowParallelLinear and mlp.down_proj.bias is None
                and hasattr(mlp, "gate_up_proj") and type(mlp.gate_up_proj).__name__ == "ColumnParallelLinear"
                and mlp.gate_up_proj.bias is None and mlp.gate_up_proj.weight.dtype == torch.bfloat16
                and hasattr(layer.input_layernorm, "variance_epsilon") and attn.head_dim in (64, 128)
                and qkv.kv_size_multiplier == 1)
            layer._nxd_fused_decode = cached
        if not cached or backend.gemm_backend() != "tcgen05":
            return False
        x = hidden.reshape(hidden.shape[0], -1)
        if not (decode_fused.qualifies_batched(x) if batched else decode_fused.qualifies(x)):
            return False
        if ps.get_tensor_model_parallel_size() > 1:
            return backend.fused_comm_enabled() and ps.get_symm_mem_context("tp") is not None
        return True

    def _decode_i32(self, positions, seq_ids):
        """int32 copies of the positions / cache lines once per forward, not once per layer."""
        memo = getattr(self, "_decode_i32_memo", None)
        if memo is None or memo[0] is not positions or memo[1] is not seq_ids:
            memo = (positions, seq_ids, positions[:, 0].to(torch.int32).contiguous(),
                    seq_ids.to(torch.int32).contiguous() if seq_ids is not None else None)
            self._decode_i32_memo = memo
        return memo[2], memo[3]

    def _decode_attention(self, q, ck, cv, seq_ids, lengths):
        """Attention of one new token per row over its cache line; with continuous batching the kernel follows
        ``seq_ids`` itself (reference ``attention_base.py:141-170`` gathers nothing either: its cache is per line)."""
        kc = self.kv_cache
        return attn_ops.decode_attention(q, ck, cv, lengths, block_table=kc.block_table if kc.paged else None,
                                         page_size=kc.page_size if kc.paged else 0, seq_ids=seq_ids)

    def _layer_batched_decode(self, i: int, layer, hidden, residual, positions, seq_ids, lengths):
        """One token per sequence for 9 .. 128 sequences on the small-M tensor-core kernels (``csrc/gemm_smallm.cu``):
        RMSNorm | QKV GEMM+RoPE+cache write | attention (split, merge) | o_proj+residual | RMSNorm | gate/up+SwiGLU |
        down_proj+residual -- eight launches, no copies between them.  Tensor parallel: the two row-parallel GEMMs are
        followed by the one-shot all-reduce and the residual moves into the fused add+RMSNorm (ten launches)."""
        from ..ops import activation, decode_fused
        from ..ops.fused_comm import oneshot_all_reduce
        from ..ops.gemm import gemm_smallm

        tp = ps.get_tensor_model_parallel_size()
        ctx = ps.get_symm_mem_context("tp") if tp > 1 else None

        def all_reduce(y):
            r = oneshot_all_reduce(ctx, y)
            return r if r is not None else mappings.reduce_from_tensor_model_parallel_region(y)

        B = hidden.shape[0]
        attn, mlp = layer.self_attn, layer.mlp
        nq, nkv, D = attn.q_heads_local, attn.kv_heads_local, attn.head_dim
        cos, sin = self._rope(hidden.device)
        ck, cv = self.kv_cache.layer(i)
        kc = self.kv_cache
        pos, sid = self._decode_i32(positions, seq_ids)
        xn, x = self._add_norm(layer.input_layernorm, hidden.reshape(B, -1),
                               residual.reshape(B, -1) if residual is not None else None)
        q = decode_fused.qkv_rope_cache_batched(xn, attn.qkv_proj.weight_qkv, cos, sin, pos, ck, cv, nq, nkv, D,
                                                seq_ids=sid, block_table=kc.block_table if kc.paged else None,
                                                page_size=kc.page_size if kc.paged else 0)
        o = self._decode_attention(q, ck, cv, seq_ids, lengths).reshape(B, nq * D)
        if tp == 1:
            x2 = gemm_smallm(o, attn.o_proj.weight, residual=x)
            xn2 = norm_ops.rms_norm(x2, layer.post_attention_layernorm.weight,
                                    layer.post_attention_layernorm.variance_epsilon)
        else:
            xn2, x2 = self._add_norm(layer.post_attention_layernorm, all_reduce(gemm_smallm(o, attn.o_proj.weight)), x)
        if B <= decode_fused.BATCHED_SWIGLU_MAX_ROWS:
            act = decode_fused.gate_up_swiglu_batched(xn2, mlp.gate_up_proj.weight)
        else:
            act = activation.swiglu(gemm_smallm(xn2, mlp.gate_up_proj.weight))
        if tp == 1:
            return gemm_smallm(act, mlp.down_proj.weight, residual=x2).reshape(B, 1, -1), None
        return all_reduce(gemm_smallm(act, mlp.down_proj.weight)).reshape(B, 1, -1), x2.reshape(B, 1, -1)

    def _layer_fused_decode(self, i: int, layer, hidden, residual, positions, seq_ids, lengths):
        """One token per sequence: RMSNorm+QKV+RoPE+cache write | attention (split, merge) | o_proj+residual |
        RMSNorm+gate/up+SwiGLU | down_proj+residual.  The residual stream leaves every row-parallel GEMV already
        summed, so the returned residual is ``None`` (the next layer's norm reads the stream directly)."""
        from ..ops import decode_fused
        from ..ops.fused_comm import skinny_gemm

        B = hidden.shape[0]
        x = hidden.reshape(B, -1)
        if residual is not None:
            x = x + residual.reshape(B, -1)
        attn, mlp = layer.self_attn, layer.mlp
        nq, nkv, D = attn.q_heads_local, attn.kv_heads_local, attn.head_dim
        cos, sin = self._rope(x.device)
        ck, cv = self.kv_cache.layer(i)
        kc = self.kv_cache
        pos, sid = self._decode_i32(positions, seq_ids)
        table = kc.block_table if kc.paged else None
        q = decode_fused.qkv_rope_cache(x, layer.input_layernorm.weight, layer.input_layernorm.variance_epsilon,
                                        attn.qkv_proj.weight_qkv, cos, sin, pos, ck, cv, nq, nkv, D, seq_ids=sid,
                                        block_table=table, page_size=kc.page_size if kc.paged else 0)
        o = self._decode_attention(q, ck, cv, seq_ids, lengths)
        ctx = ps.get_symm_mem_context("tp") if ps.get_tensor_model_parallel_size() > 1 else None
        x2 = skinny_gemm(o.reshape(B, nq * D), attn.o_proj.weight, residual=x, ctx=ctx)
        if x2 is None:      # shape outside the skinny kernel: generic projection, then the residual
            x2 = attn.o_proj(o.reshape(B, 1, nq * D)).reshape(B, -1) + x
        act = decode_fused.gate_up_swiglu(x2, layer.post_attention_layernorm.weight,
                                          layer.post_attention_layernorm.variance_epsilon, mlp.gate_up_proj.weight)
        x3 = skinny_gemm(act, mlp.down_proj.weight, residual=x2, ctx=ctx)
        if x3 is None:
            x3 = mlp.down_proj(act.reshape(B, 1, -1)).reshape(B, -1) + x2
        return x3.reshape(B, 1, -1), None

    def _layer(self, i: int, layer, hidden, residual, positions, seq_ids, lengths, prefill):
        if not prefill and hidden.shape[1] == 1:
            if self._fused_decode_ok(layer, hidden):
                return self._layer_fused_decode(i, layer, hidden, residual, positions, seq_ids, lengths)
            if hidden.shape[0] > 8 and self._fused_decode_ok(layer, hidden, batched=True):
                return self._layer_batched_decode(i, layer, hidden, residual, positions, seq_ids, lengths)
        h, residual = self._add_norm(layer.input_layernorm, hidden, residual)
        a = self._attention(i, layer.self_attn, h, positions, seq_ids, lengths, prefill)
        h, residual = self._add_norm(layer.post_attention_layernorm, a, residual)
        return self._mlp(layer, h), residual

    # -- graphs ----------------------------------------------------------------------------------------------------
    @torch.no_grad()
    def forward_infer(self, input_ids: torch.Tensor, position_ids: torch.Tensor, seq_ids: Optional[torch.Tensor] = None,
                      last_token_index: Optional[torch.Tensor] = None, return_logits: bool = False,
                      uniform: Optional[torch.Tensor] = None, return_hidden: bool = False):
        """``input_ids [B, T]``; ``position_ids [B, T]`` absolute positions; ``seq_ids [B]`` cache lines (default
        ``arange(B)``); ``last_token_index [B]`` index (within T) of the last real prompt token (context encoding with
        right padding).  Returns sampled tokens ``[B]`` (and the ``[B, V]`` logits with ``return_logits``)."""
        assert self.kv_cache is not None, "call init_kv_cache() first"
        B, T = input_ids.shape
        prefill = T > 1
        # the int32 position / cache-line copies are shared by the layers of ONE forward only: a captured graph is
        # replayed with the same static tensor objects but new contents, so the conversion has to be part of every
        # forward (and of every capture)
        self._decode_i32_memo = None
        default_lines = seq_ids is None
        hidden = self._embed(input_ids)
        lengths = (position_ids[:, -1] + 1).to(torch.int32)
        residual = None
        # token generation without continuous batching: row b IS cache line b -- never gather cache lines by seq_ids
        # (an index_select of a [B, Hkv, S, D] cache per layer and token costs more than the attention itself)
        identity = default_lines or not self.inference_config.is_continuous_batching
        sid = None if (identity and not prefill) else (
            seq_ids if seq_ids is not None else torch.arange(B, device=input_ids.device, dtype=torch.int32))
        for i, layer in enumerate(self._decoder_layers()):
            hidden, residual = self._layer(i, layer, hidden, residual, position_ids, sid, lengths, prefill)
        if prefill:
            idx = last_token_index if last_token_index is not None else torch.full((B,), T - 1, device=hidden.device)
            rows = torch.arange(B, device=hidden.device)
            hidden, residual = hidden[rows, idx.long()], residual[rows, idx.long()]
        else:
            hidden, residual = hidden[:, 0], (residual[:, 0] if residual is not None else None)
        hidden, _ = self._add_norm(self._final_norm(), hidden.contiguous(),
                                   residual.contiguous() if residual is not None else None)
        logits = self.lm_head(hidden)                                   # [B, V/tp]
        logits = mappings.gather_from_tensor_model_parallel_region(logits)[:, : self.config.vocab_size]
        tokens = self.sampler.sample(logits.float() if logits.dtype not in (torch.float32, torch.bfloat16) else logits,
                                     uniform)
        if return_hidden:                  # final-norm output of the selected position: what extra heads (Medusa) read
            return tokens, logits, hidden
        return (tokens, logits) if return_logits else tokens

    @torch.no_grad()
    def forward_speculation(self, input_ids: torch.Tensor, position_ids: torch.Tensor,
                            seq_ids: Optional[torch.Tensor] = None, return_logits: bool = False,
                            return_hidden: bool = False):
        """The ``speculation_model`` graph of the reference (``model_base.py:853-860``): the TARGET model on ``k + 1``
        tokens per sequence at consecutive positions, on top of the KV cache.  Writes their K/V into the cache (rows
        past the accepted prefix are simply overwritten by the next round) and returns the greedy token after EVERY
        position, ``[B, k + 1]`` -- what ``NeuronSpeculation.target_verify`` needs."""
        assert self.kv_cache is not None, "call init_kv_cache() first"
        B, T = input_ids.shape
        lines = seq_ids if seq_ids is not None else torch.arange(B, device=input_ids.device, dtype=torch.int32)
        hidden, residual = self._embed(input_ids), None
        for i, layer in enumerate(self._decoder_layers()):
            h, residual = self._add_norm(layer.input_layernorm, hidden, residual)
            a = self._attention_with_past(i, layer.self_attn, h, position_ids, lines)
            h, residual = self._add_norm(layer.post_attention_layernorm, a, residual)
            hidden = self._mlp(layer, h)
        H = hidden.shape[-1]
        hidden, _ = self._add_norm(self._final_norm(), hidden.reshape(B * T, H).contiguous(),
                                   residual.reshape(B * T, H).contiguous())
        logits = mappings.gather_from_tensor_model_parallel_region(self.lm_head(hidden))[:, : self.config.vocab_size]
        tokens = logits.argmax(dim=-1).reshape(B, T)
        if return_hidden:
            return tokens, logits.reshape(B, T, -1), hidden.reshape(B, T, H)
        return (tokens, logits.reshape(B, T, -1)) if return_logits else tokens

    def forward(self, input_ids: torch.Tensor, position_ids: torch.Tensor, seq_ids: Optional[torch.Tensor] = None,
                last_token_index: Optional[torch.Tensor] = None):
        """Calling the inference model runs one graph: context encoding (T > 1) or token generation (T == 1)."""
        return self.forward_infer(input_ids, position_ids, seq_ids, last_token_index)

    def _embed(self, input_ids: torch.Tensor) -> torch.Tensor:
        emb = self._embedding()
        tp = ps.get_tensor_model_parallel_size()
        w = emb.weight
        if tp > 1:
            mask = (input_ids < emb.start_index) | (input_ids >= emb.end_index)
            local = (input_ids - emb.start_index).masked_fill(mask, 0)
            out = nn.functional.embedding(local, w).masked_fill(mask.unsqueeze(-1), 0.0)
            return comm_linear.all_reduce_small(out)
        return nn.functional.embedding(input_ids, w)

    # -- simple generation loop (the runner adds batching / bucketing / CUDA graphs on top) ---------------------------
    @torch.no_grad()
    def generate(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None, max_new_tokens: int = 16,
                 eos_token_id: Optional[int] = None) -> torch.Tensor:
        """Greedy / sampled continuation of right-padded prompts ``[B, T]``; returns ``[B, T + max_new_tokens]``."""
        B, T = input_ids.shape
        dev = input_ids.device
        if self.kv_cache is None:
            self.init_kv_cache(device=dev)
        if attention_mask is None:
            attention_mask = torch.ones_like(input_ids)
        lens = attention_mask.sum(dim=1).to(torch.long)
        pos = torch.arange(T, device=dev).unsqueeze(0).expand(B, T)
        dynamic = self.kv_cache.pool is not None
        if dynamic:                                         # the prompt (incl. its padding) and the first new token
            for b in range(B):
                self.kv_cache.free(b)
                self.kv_cache.reserve(b, T + 1)
        tok = self.forward_infer(input_ids, pos, last_token_index=lens - 1)
        out = [tok]
        cur = lens.clone()
        finished = torch.zeros(B, dtype=torch.bool, device=dev)
        for step in range(max_new_tokens - 1):
            if dynamic:
                for b in range(B):
                    self.kv_cache.reserve(b, max(T + 1, int(lens[b]) + step + 2))
            if eos_token_id is not None:
                finished |= tok == eos_token_id
                if bool(finished.all()):
                    break
            tok = self.forward_infer(tok.reshape(B, 1), cur.reshape(B, 1))
            cur = cur + 1
            out.append(tok)
        gen = torch.stack(out, dim=1)
        res = torch.full((B, T + gen.shape[1]), self.inference_config.pad_token_id, dtype=input_ids.dtype, device=dev)
        for b in range(B):
            L = int(lens[b])
            res[b, :L] = input_ids[b, :L]
            res[b, L:L + gen.shape[1]] = gen[b]
        return res


class LlamaInferenceModel(KVCacheInferenceMixin, LlamaForCausalLM):
    def __init__(self, config: LlamaConfig, inference_config: Optional[InferenceConfig] = None, device=None):
        config.sequence_parallel_enabled = False
        config.activation_checkpointing = False
        LlamaForCausalLM.__init__(self, config, device=device)
        self._init_inference(inference_config)


class _MoEDecoderLayout:
    """Where ``MixtralForCausalLM``-shaped models keep their pieces, and how their feed-forward runs at inference."""

    def _decoder_layers(self):
        return self.layers

    def _embedding(self):
        return self.embed_tokens

    def _final_norm(self):
        return self.norm

    def _mlp(self, layer, h: torch.Tensor) -> torch.Tensor:
        if not layer.is_moe:
            return layer.mlp.down_proj(act_ops.swiglu(layer.mlp.gate_up_proj(h)))
        out = layer.mlp(h.transpose(0, 1))[0]          # the MoE block is sequence-first: [T, B, H]
        return out.transpose(0, 1)


class MixtralInferenceModel(_MoEDecoderLayout, KVCacheInferenceMixin, MixtralForCausalLM):
    """Mixtral-style MoE decoder with a KV cache (reference ``examples/inference/mixtral`` + ``modules/moe`` inference
    paths): context encoding runs the capacity-free all-experts path, token generation loads only the selected experts
    (``ExpertMLPs.forward_selective_loading``)."""

    def __init__(self, config: MixtralConfig, inference_config: Optional[InferenceConfig] = None, device=None):
        config.sequence_parallel_enabled = False
        config.activation_checkpointing = False
        MixtralForCausalLM.__init__(self, config, device=device)
        self._init_inference(inference_config)


class DbrxInferenceModel(_MoEDecoderLayout, KVCacheInferenceMixin, DbrxForCausalLM):
    """DBRX with a KV cache (reference ``examples/inference/dbrx``): the Mixtral inference model with weight-only
    LayerNorm in place of RMSNorm; the QKV clamp is a forward hook of the projection and applies here unchanged."""

    def __init__(self, config: DbrxConfig, inference_config: Optional[InferenceConfig] = None, device=None):
        config.sequence_parallel_enabled = False
        config.activation_checkpointing = False
        DbrxForCausalLM.__init__(self, config, device=device)
        self._init_inference(inference_config)

    def _add_norm(self, norm, hidden: torch.Tensor, residual: Optional[torch.Tensor]):
        summed = hidden if residual is None else hidden + residual
        return norm(summed), summed
