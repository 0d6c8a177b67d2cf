"""Inference configuration (reference ``examples/inference/modules/config.py:4-68``)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional


@dataclass
class InferenceConfig:
    tp_degree: int = 1
    batch_size: int = 1                      # compiled batch of both sub-models unless overridden below
    ctx_batch_size: Optional[int] = None     # context-encoding batch
    tkg_batch_size: Optional[int] = None     # token-generation batch
    max_batch_size: Optional[int] = None     # KV cache lines (continuous batching: > tkg batch)
    max_context_length: int = 2048           # longest prompt
    seq_len: int = 4096                      # KV cache length (= max total length)
    max_new_tokens: int = 256
    padding_side: str = "right"
    enable_bucketing: bool = True
    buckets: Optional[List[int]] = None      # context-length buckets; default powers of two from 128
    on_device_sampling: bool = True
    do_sample: bool = False
    top_k: int = 1
    top_p: float = 1.0
    temperature: float = 1.0
    quantized: bool = False
    quantization_type: str = "per_tensor_symmetric"
    speculation_length: int = 0
    medusa_speculation_length: int = 0
    num_medusa_heads: int = 0
    is_continuous_batching: bool = False
    use_cuda_graphs: bool = True
    paged_kv: bool = False
    page_size: int = 128
    num_kv_pages: Optional[int] = None      # paged cache: size of the page pool (default: lines * pages per line)
    pad_token_id: int = 0
    eos_token_id: Optional[int] = None

    def __post_init__(self):
        self.ctx_batch_size = self.ctx_batch_size or self.batch_size
        self.tkg_batch_size = self.tkg_batch_size or self.batch_size
        self.max_batch_size = self.max_batch_size or max(self.ctx_batch_size, self.tkg_batch_size)
        if self.max_context_length > self.seq_len:
            raise ValueError("max_context_length cannot exceed seq_len")

    def context_buckets(self) -> List[int]:
        """Powers of two from 128 up to ``max_context_length`` (reference ``autobucketing.py:6-54``)."""
        if self.buckets:
            return sorted(self.buckets)
        if not self.enable_bucketing:
            return [self.max_context_length]
        out, b = [], 128
        while b < self.max_context_length:
            out.append(b)
            b *= 2
        out.append(self.max_context_length)
        return out
