"""``InferenceRunner``: build (trace) the two graphs, generate, check accuracy, benchmark (SURVEY C53, §3.4-3.5;
reference ``examples/inference/runner.py:36-655`` and ``modules/model_wrapper.py:60-301``).

Sub-models as in the reference: ``context_encoding_model`` (bucketed prompt lengths, produces the first token = TTFT)
and ``token_generation_model`` (one token per call).  Both are buckets of ONE module sharing weights and KV cache;
each bucket is a CUDA graph.  Prompts are right-padded to the bucket length, batches padded to the compiled batch.
"""
from __future__ import annotations

import json
import os
from typing import Dict, Optional

import torch

from ..models.llama import LlamaConfig
from ..parallel import parallel_state as ps
from .benchmark import LatencyCollector, generate_report
from .config import InferenceConfig
from .model import LlamaInferenceModel
from .model_builder import BaseModelInstance, ModelBuilder, NxDModel

CONTEXT_ENCODING_MODEL = "context_encoding_model"
TOKEN_GENERATION_MODEL = "token_generation_model"
SPECULATION_MODEL = "speculation_model"


class InferenceRunner:
    """``model_cls`` picks the family (``LlamaInferenceModel`` by default, ``MixtralInferenceModel``,
    ``DbrxInferenceModel`` -- the reference has one runner subclass per family, ``examples/inference/{llama3,mixtral,
    dbrx}/*_runner.py``).  With ``inference_config.quantized`` the decoder layers' parallel linears are swapped for the
    int8 weight-only layers and a float checkpoint is quantized on load (reference ``run_llama_quantized.py``)."""

    model_cls = LlamaInferenceModel

    def __init__(self, model_config: LlamaConfig, inference_config: InferenceConfig, checkpoint_loader=None, device=None,
                 model_cls=None):
        self.model_config = model_config
        self.config = inference_config
        if model_cls is not None:
            self.model_cls = model_cls
        self._float_loader = checkpoint_loader
        self.checkpoint_loader = (self._quantized_loader if (inference_config.quantized and checkpoint_loader is not None)
                                  else checkpoint_loader)
        self.device = device or (torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available()
                                 else torch.device("cpu"))
        self.model: Optional[LlamaInferenceModel] = None
        self.nxd_model: Optional[NxDModel] = None

    # -- build --------------------------------------------------------------------------------------------------------
    def _q_config(self) -> dict:
        from ..quantization.quantization_config import (QuantizationType, get_default_custom_qconfig_dict,
                                                        get_default_per_channel_custom_qconfig_dict)

        if QuantizationType(self.config.quantization_type) == QuantizationType.PER_CHANNEL_SYMMETRIC:
            return get_default_per_channel_custom_qconfig_dict()
        return get_default_custom_qconfig_dict()

    def _quantize_layers(self, m) -> None:
        """int8 weight-only layers inside the decoder blocks; embedding, (tied) LM head, QKV and routers stay float."""
        from ..quantization.quantize import convert

        for layer in m._decoder_layers():
            convert(layer, self._q_config(), inplace=True)

    def _make_module(self) -> LlamaInferenceModel:
        if torch.empty(0).device.type == "meta":   # shard_checkpoint builds shape-only replicas under torch.device("meta")
            m = self.model_cls(self.model_config, self.config, device="meta")
            if self.config.quantized:
                self._quantize_layers(m)
            return m
        m = self.model_cls(self.model_config, self.config, device=self.device)
        if self.config.quantized:
            self._quantize_layers(m)
        m.init_kv_cache(device=self.device)
        self.model = m
        return m

    def generate_quantized_state_dict(self, state_dict: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        """Float full state -> the state the quantized model loads: ``<layer>.weight`` int8 + ``<layer>.scale`` for
        exactly the layers ``_quantize_layers`` converts (reference ``model_base.py`` ``generate_quantized_state_dict``)."""
        from ..quantization.quantization_layers import BaseQuantizeParallelLinear
        from ..quantization.quantization_utils import quantize_per_channel_symmetric, quantize_per_tensor_symmetric

        with torch.device("meta"):
            meta = self.model_cls(self.model_config, self.config, device="meta")
            self._quantize_layers(meta)
        out = dict(state_dict)
        for name, mod in meta.named_modules():
            key = name + ".weight"
            if not isinstance(mod, BaseQuantizeParallelLinear) or key not in out or not out[key].is_floating_point():
                continue
            if mod.scale.numel() > 1:       # per channel: scale already shaped to broadcast against the weight
                q, sc = quantize_per_channel_symmetric(out[key].float(), (mod.per_channel_axis or 0) % out[key].dim())
            else:
                q, sc = quantize_per_tensor_symmetric(out[key].float())
            out[key], out[name + ".scale"] = q, sc.to(torch.float32)
        return out

    def _quantized_loader(self) -> Dict[str, torch.Tensor]:
        sd = self._float_loader()
        return sd if any(k.endswith(".scale") for k in sd) else self.generate_quantized_state_dict(sd)

    def trace(self, traced_model_path: Optional[str] = None) -> NxDModel:
        """Compile both sub-models (reference ``runner.trace`` -> ``model.compile`` -> ``ModelBuilder.trace``) and,
        when a path is given, write the per-rank sharded checkpoint + configs next to it."""
        ic = self.config
        builder = ModelBuilder(tp_degree=ic.tp_degree, checkpoint_loader=self.checkpoint_loader,
                               use_cuda_graphs=ic.use_cuda_graphs and self.device.type == "cuda")
        inst = BaseModelInstance(self._make_module)
        dev = self.device

        def ex(b, t):
            return (torch.zeros(b, t, dtype=torch.long, device=dev), torch.zeros(b, t, dtype=torch.long, device=dev),
                    torch.arange(b, dtype=torch.int32, device=dev), torch.zeros(b, dtype=torch.long, device=dev))

        builder.add(CONTEXT_ENCODING_MODEL, inst, [ex(ic.ctx_batch_size, t) for t in ic.context_buckets()],
                    priority_model_idx=0)
        builder.add(TOKEN_GENERATION_MODEL, inst, [ex(ic.tkg_batch_size, 1)])
        if ic.speculation_length > 0:
            # the target's verify graph: k + 1 tokens per sequence on top of the KV cache (same module, other entry point)
            def shared():
                return self.model if (self.model is not None and torch.empty(0).device.type != "meta") else self._make_module()

            k1 = ic.speculation_length + 1
            builder.add(SPECULATION_MODEL, BaseModelInstance(shared, forward_method="forward_speculation"),
                        [ex(ic.tkg_batch_size, k1)[:3]])
        self.nxd_model = builder.trace(initialize_model_weights=self.checkpoint_loader is not None)
        if traced_model_path:
            os.makedirs(traced_model_path, exist_ok=True)
            with open(os.path.join(traced_model_path, "config.json"), "w") as f:
                mc = self.model_config
                json.dump({"model": mc.to_dict() if hasattr(mc, "to_dict") else dict(vars(mc)),
                           "inference": self.config.__dict__}, f, indent=1, default=str)
            if self.checkpoint_loader is not None and ps.is_global_rank_zero():
                builder.shard_checkpoint(os.path.join(traced_model_path, "weights"))
        return self.nxd_model

    def close(self) -> None:
        """Release the CUDA graphs (required before the process group is destroyed, see ``NxDModel.release_graphs``)."""
        if self.nxd_model is not None:
            self.nxd_model.release_graphs()

    # -- generate -------------------------------------------------------------------------------------------------------
    def _reserve_lines(self, n_lines: int, n_tokens: int) -> None:
        """With a dynamic page pool (``num_kv_pages``) the cache lines of a ``generate`` call get their pages up front:
        the loop itself runs captured graphs and never talks to the allocator."""
        cache = self.model.kv_cache if self.model is not None else None
        if cache is None or cache.pool is None:
            return
        for line in range(n_lines):
            cache.free(line)
        for line in range(n_lines):
            cache.reserve(line, min(n_tokens, self.config.seq_len))

    def _bucket_len(self, n: int) -> int:
        if self.nxd_model is not None:
            hit = self.nxd_model.route_padded(CONTEXT_ENCODING_MODEL, self.config.ctx_batch_size, n)
            if hit is not None:
                return hit[1]
        for b in self.config.context_buckets():
            if n <= b:
                return b
        raise ValueError(f"prompt length {n} exceeds max_context_length {self.config.max_context_length}")

    @torch.no_grad()
    def generate(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
                 max_new_tokens: Optional[int] = None, collectors: Optional[Dict[str, LatencyCollector]] = None
                 ) -> torch.Tensor:
        """Right-padded prompts ``[B, T]`` (B <= compiled batch) -> ``[B, T + new]`` token ids."""
        assert self.nxd_model is not None, "call trace() first"
        ic, dev = self.config, self.device
        max_new = max_new_tokens or ic.max_new_tokens
        B, T = input_ids.shape
        if attention_mask is None:
            attention_mask = torch.ones_like(input_ids)
        lens = attention_mask.sum(dim=1).to(torch.long)
        Tb = self._bucket_len(T)
        Bc = ic.ctx_batch_size
        self._reserve_lines(max(Bc, ic.tkg_batch_size), Tb + max_new)
        ids = torch.full((Bc, Tb), ic.pad_token_id, dtype=torch.long, device=dev)
        ids[:B, :T] = input_ids.to(dev)
        pos = torch.arange(Tb, device=dev).unsqueeze(0).expand(Bc, Tb).contiguous()
        last = torch.zeros(Bc, dtype=torch.long, device=dev)
        last[:B] = lens.to(dev) - 1
        seq = torch.arange(Bc, dtype=torch.int32, device=dev)
        cte = collectors.get(CONTEXT_ENCODING_MODEL) if collectors else None
        tkg = collectors.get(TOKEN_GENERATION_MODEL) if collectors else None
        if cte:
            cte.pre_hook()
        tok = self.nxd_model(ids, pos, seq, last).clone()
        if cte:
            cte.hook()
        outs = [tok[:B]]
        cur = torch.zeros(ic.tkg_batch_size, dtype=torch.long, device=dev)
        cur[:B] = lens.to(dev)
        step_ids = torch.zeros(ic.tkg_batch_size, 1, dtype=torch.long, device=dev)
        seq_t = torch.arange(ic.tkg_batch_size, dtype=torch.int32, device=dev)
        zero_last = torch.zeros(ic.tkg_batch_size, dtype=torch.long, device=dev)
        finished = torch.zeros(B, dtype=torch.bool, device=dev)
        for _ in range(max_new - 1):
            if ic.eos_token_id is not None:
                finished |= outs[-1] == ic.eos_token_id
                if bool(finished.all()):
                    break
            step_ids[:B, 0] = outs[-1]
            if tkg:
                tkg.pre_hook()
            tok = self.nxd_model(step_ids, cur.reshape(-1, 1), seq_t, zero_last).clone()
            if tkg:
                tkg.hook()
            cur = cur + 1
            outs.append(tok[:B])
        gen = torch.stack(outs, dim=1)
        res = torch.full((B, T + gen.shape[1]), ic.pad_token_id, dtype=torch.long, device=dev)
        for b in range(B):
            L = int(lens[b])
            res[b, :L] = input_ids[b, :L]
            res[b, L:L + gen.shape[1]] = gen[b]
        return res

    @torch.no_grad()
    def generate_speculative(self, draft: "InferenceRunner", input_ids: torch.Tensor,
                             max_new_tokens: Optional[int] = None) -> torch.Tensor:
        """Assisted decoding (reference ``run_llama_speculative.py`` / ``utils/speculative_decoding.py``): ``draft`` proposes
        ``speculation_length`` tokens with its token-generation graph, this (target) runner verifies them with ONE call of
        its ``speculation_model`` graph.  Batch 1 as in the reference; the result equals ``generate`` of the target."""
        from .speculative import NeuronSpeculation

        ic, dev = self.config, self.device
        assert ic.speculation_length > 0, "trace the target with InferenceConfig(speculation_length=k)"
        assert input_ids.shape[0] == 1 and ic.tkg_batch_size == 1 and draft.config.tkg_batch_size == 1
        max_new = max_new_tokens or ic.max_new_tokens
        T = input_ids.shape[1]
        first = None
        for r in (draft, self):                                  # context encoding fills both KV caches
            Tb = r._bucket_len(T)
            r._reserve_lines(r.config.ctx_batch_size, Tb + max_new + ic.speculation_length + 1)
            ids = torch.full((r.config.ctx_batch_size, Tb), r.config.pad_token_id, dtype=torch.long, device=dev)
            ids[:1, :T] = input_ids.to(dev)
            pos = torch.arange(Tb, device=dev).unsqueeze(0).expand(r.config.ctx_batch_size, Tb).contiguous()
            last = torch.zeros(r.config.ctx_batch_size, dtype=torch.long, device=dev)
            last[0] = T - 1
            first = r.nxd_model(ids, pos, torch.arange(r.config.ctx_batch_size, dtype=torch.int32, device=dev), last).clone()
        seq1 = torch.arange(1, dtype=torch.int32, device=dev)
        zero = torch.zeros(1, dtype=torch.long, device=dev)

        def draft_step(tok, pos):
            return draft.nxd_model(tok.reshape(1, 1), pos.reshape(1, 1), seq1, zero).clone()

        def target_verify(toks, pos):
            return self.nxd_model(toks.contiguous(), pos.contiguous(), seq1).clone()

        spec = NeuronSpeculation(draft_step, target_verify, ic.speculation_length)
        new = spec.generate(first[:1], torch.tensor([T], device=dev), max_new - 1, eos_token_id=ic.eos_token_id)
        self.last_acceptance = list(spec.accepted_hist)
        return torch.cat([input_ids.to(dev)[0], first[:1], new]).unsqueeze(0)

    # -- checks / benchmark ------------------------------------------------------------------------------------------------
    @torch.no_grad()
    def check_accuracy_logits(self, reference_model, input_ids: torch.Tensor, atol: float = 5e-2) -> bool:
        """Compare prefill logits against a reference causal LM returning ``[S, B, V]`` logits (the training model) --
        the reference runner compares against HF on CPU (``runner.py:313-409``)."""
        B, T = input_ids.shape
        pos = torch.arange(T, device=input_ids.device).unsqueeze(0).expand(B, T)
        _, logits = self.model.forward_infer(input_ids, pos, return_logits=True)
        ref = reference_model(input_ids).transpose(0, 1)[:, -1]
        return bool((logits.float() - ref.float()).abs().max() <= atol)

    def torch_profile(self, chrome_trace_path: str = "torch-trace.json", prompt_len: Optional[int] = None,
                      new_tokens: int = 8, warmup: int = 1, with_stack: bool = False) -> str:
        """One ``generate`` under ``torch.profiler`` (CPU + CUDA activities when a GPU is present), exported as a Chrome
        trace (reference ``runner.py:434-458``).  Graph replays show up as one ``cudaGraphLaunch`` each; trace with
        ``use_cuda_graphs=False`` to see the individual kernels.  Numbers taken under the profiler are not benchmarks."""
        from torch.profiler import ProfilerActivity, profile

        ic = self.config
        T = prompt_len or min(ic.max_context_length, 128)
        ids = torch.randint(1, self.model_config.vocab_size, (ic.ctx_batch_size, T), device=self.device)
        for _ in range(warmup):
            self.generate(ids, max_new_tokens=new_tokens)
        acts = [ProfilerActivity.CPU] + ([ProfilerActivity.CUDA] if self.device.type == "cuda" else [])
        with profile(activities=acts, with_stack=with_stack, record_shapes=True) as prof:
            self.generate(ids, max_new_tokens=new_tokens)
            if self.device.type == "cuda":
                torch.cuda.synchronize()
        d = os.path.dirname(os.path.abspath(chrome_trace_path))
        os.makedirs(d, exist_ok=True)
        prof.export_chrome_trace(chrome_trace_path)
        return chrome_trace_path

    def benchmark_sampling(self, prompt_len: Optional[int] = None, new_tokens: Optional[int] = None, n_runs: int = 20,
                           warmup: int = 1) -> Dict:
        """1 warm-up + ``n_runs`` timed ``generate`` calls; e2e, context-encoding (TTFT) and token-generation reports
        (reference ``runner.py:460-535``)."""
        ic = self.config
        T = prompt_len or ic.max_context_length
        new = new_tokens or ic.max_new_tokens
        ids = torch.randint(1, self.model_config.vocab_size, (ic.ctx_batch_size, T), device=self.device)
        for _ in range(warmup):
            self.generate(ids, max_new_tokens=new)
        e2e = LatencyCollector()
        cols = {CONTEXT_ENCODING_MODEL: LatencyCollector(), TOKEN_GENERATION_MODEL: LatencyCollector()}
        for _ in range(n_runs):
            e2e.pre_hook()
            self.generate(ids, max_new_tokens=new, collectors=cols)
            e2e.hook()
        e2e.finish()
        for c in cols.values():
            c.finish()
        total_len = T + new
        report = {"e2e_model": generate_report(e2e.latency_list, total_len, ic.ctx_batch_size, n_runs),
                  CONTEXT_ENCODING_MODEL: generate_report(cols[CONTEXT_ENCODING_MODEL].latency_list, T, ic.ctx_batch_size,
                                                          n_runs),
                  TOKEN_GENERATION_MODEL: generate_report(cols[TOKEN_GENERATION_MODEL].latency_list, 1, ic.tkg_batch_size,
                                                          len(cols[TOKEN_GENERATION_MODEL].latency_list))}
        report["ttft_ms_p50"] = report[CONTEXT_ENCODING_MODEL]["latency_ms_p50"]
        report["decode_tokens_per_s"] = report[TOKEN_GENERATION_MODEL]["throughput"]
        return report


class LlamaRunner(InferenceRunner):
    """Reference ``examples/inference/llama3/llama3_runner.py``."""
    model_cls = LlamaInferenceModel


class MixtralRunner(InferenceRunner):
    """Reference ``examples/inference/mixtral/mixtral_runner.py``."""

    def __init__(self, *a, **kw):
        from .model import MixtralInferenceModel

        kw.setdefault("model_cls", MixtralInferenceModel)
        super().__init__(*a, **kw)


class DbrxRunner(InferenceRunner):
    """Reference ``examples/inference/dbrx/dbrx_runner.py``."""

    def __init__(self, *a, **kw):
        from .model import DbrxInferenceModel

        kw.setdefault("model_cls", DbrxInferenceModel)
        super().__init__(*a, **kw)
