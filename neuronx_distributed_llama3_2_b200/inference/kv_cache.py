"""KV cache state for the inference graphs (reference keeps the cache as aliased parameters of the compiled
graph and re-emits the whole cache every step, ``examples/inference/modules/model_base.py:114-125,388-422``).

Here the cache is a plain device allocation per layer that the kernels update IN PLACE:
  contiguous  ``[lines, kv_heads_local, seq_len, head_dim]``
  paged       ``[pages, kv_heads_local, page_size, head_dim]`` + a block table ``[lines, max_pages]``
"""
from __future__ import annotations

from typing import List, Tuple

import torch


class KVCache:
    def __init__(self, num_layers: int, lines: int, kv_heads: int, seq_len: int, head_dim: int,
                 dtype: torch.dtype = torch.bfloat16, device=None, paged: bool = False, page_size: int = 128):
        self.num_layers, self.lines, self.kv_heads, self.seq_len, self.head_dim = num_layers, lines, kv_heads, seq_len, head_dim
        self.paged, self.page_size = paged, page_size
        if paged:
            self.max_pages = (seq_len + page_size - 1) // page_size
            n_pages = lines * self.max_pages
            shape = (n_pages, kv_heads, page_size, head_dim)
            # static identity mapping by default; a scheduler may permute it
            self.block_table = torch.arange(n_pages, dtype=torch.int32, device=device).reshape(lines, self.max_pages)
        else:
            shape = (lines, kv_heads, seq_len, head_dim)
            self.block_table = None
        self.k: List[torch.Tensor] = [torch.zeros(shape, dtype=dtype, device=device) for _ in range(num_layers)]
        self.v: List[torch.Tensor] = [torch.zeros(shape, dtype=dtype, device=device) for _ in range(num_layers)]

    def layer(self, i: int) -> Tuple[torch.Tensor, torch.Tensor]:
        return self.k[i], self.v[i]

    def reset(self) -> None:
        for t in self.k + self.v:
            t.zero_()

    def bytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in self.k + self.v)
