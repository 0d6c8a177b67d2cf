"""KV cache state for the inference graphs (reference keeps the cache as aliased parameters of the compiled
graph and re-emits the whole cache every step, ``examples/inference/modules/model_base.py:114-125,388-422``).

Here the cache is a plain device allocation per layer that the kernels update IN PLACE:
  contiguous  ``[lines, kv_heads_local, seq_len, head_dim]``
  paged       ``[pages, kv_heads_local, page_size, head_dim]`` + a block table ``[lines, max_pages]``

Paged mode has two flavours: the static identity mapping (every line owns ``max_pages`` pages up front) and, with
``num_pages``, a POOL smaller than ``lines * max_pages`` managed by the native page allocator
(``runtime.native.PagedKVPool``): lines get pages as they grow (``reserve``), give them back (``free``), can be forked
copy-on-write (``fork``) and rolled back (``truncate``).  The block table is a device tensor updated in place, so CUDA
graphs that captured it see the new mapping.
"""
from __future__ import annotations

from typing import Iterable, List, Optional, Tuple

import torch


class KVCache:
    def __init__(self, num_layers: int, lines: int, kv_heads: int, seq_len: int, head_dim: int,
                 dtype: torch.dtype = torch.bfloat16, device=None, paged: bool = False, page_size: int = 128,
                 num_pages: Optional[int] = None):
        self.num_layers, self.lines, self.kv_heads, self.seq_len, self.head_dim = num_layers, lines, kv_heads, seq_len, head_dim
        self.paged, self.page_size = paged, page_size
        self.pool = None
        if paged:
            self.max_pages = (seq_len + page_size - 1) // page_size
            if num_pages is None:
                n_pages = lines * self.max_pages
                # static identity mapping: every line owns its pages up front
                self.block_table = torch.arange(n_pages, dtype=torch.int32, device=device).reshape(lines, self.max_pages)
            else:
                from ..runtime.native import PagedKVPool

                n_pages = num_pages
                self.pool = PagedKVPool(n_pages, page_size, lines, self.max_pages)
                self.block_table = torch.zeros(lines, self.max_pages, dtype=torch.int32, device=device)
            shape = (n_pages, kv_heads, page_size, head_dim)
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
        if self.pool is not None:
            for line in range(self.lines):
                self.pool.free_line(line)
            self.block_table.zero_()

    # -- dynamic page pool (no-ops for the contiguous cache and the static mapping) ---------------------------------------
    def sync_lines(self, lines: Iterable[int]) -> None:
        """Copy the allocator's page lists of ``lines`` into the device block table."""
        if self.pool is None:
            return
        rows = list(lines)
        if not rows:
            return
        host = torch.tensor([self.pool.block_row(r)[0] for r in rows], dtype=torch.int32)
        self.block_table[torch.tensor(rows, dtype=torch.long, device=self.block_table.device)] = host.to(
            self.block_table.device, non_blocking=True)

    def reserve(self, line: int, n_tokens: int) -> None:
        """Room for ``n_tokens`` tokens in total on ``line`` (``MemoryError`` when the pool is exhausted)."""
        if self.pool is not None and self.pool.reserve(line, n_tokens) > 0:
            self.sync_lines([line])

    def free(self, line: int) -> None:
        if self.pool is not None:
            self.pool.free_line(line)

    def truncate(self, line: int, n_tokens: int) -> None:
        if self.pool is not None and self.pool.truncate(line, n_tokens) > 0:
            self.sync_lines([line])

    def _copy_page(self, src: int, dst: int) -> None:
        for t in self.k + self.v:
            t[dst].copy_(t[src])

    def fork(self, src: int, dst: int) -> None:
        """``dst`` continues from ``src``'s content: full pages shared, the partial last page copied."""
        assert self.pool is not None, "fork needs the dynamic page pool (num_pages=...)"
        copy = self.pool.fork(src, dst)
        if copy is not None:
            self._copy_page(*copy)
        self.sync_lines([dst])

    def prepare_write(self, line: int, pos: int) -> None:
        """Before writing position ``pos`` of ``line``: give the line a private copy of that page if it is shared."""
        if self.pool is None:
            return
        copy = self.pool.unshare(line, pos)
        if copy is not None:
            self._copy_page(*copy)
            self.sync_lines([line])

    def bytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in self.k + self.v)
