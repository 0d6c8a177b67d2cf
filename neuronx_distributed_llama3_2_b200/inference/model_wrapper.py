"""``ModelWrapper``: fits arbitrary requests to the compiled shape of a sub-model (reference
``examples/inference/modules/model_wrapper.py:60-301``): pads the sequence to the bucket, pads a short batch by
repeating its first row, splits a long batch into compiled-batch chunks and, under continuous batching, reorders rows by
cache line (``seq_ids``) before the call and restores the caller's order afterwards.

The wrapped callable has the sub-model signature ``(input_ids [B, T], position_ids [B, T], seq_ids [B],
last_token_index [B]) -> tokens [B]`` (a ``NxDModel`` bucket router or ``LlamaInferenceModel.forward_infer``).
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence

import torch

CONTEXT_ENCODING_MODEL_TAG = "context_encoding_model"
TOKEN_GENERATION_MODEL_TAG = "token_generation_model"
SPECULATION_MODEL_TAG = "speculation_model"
MEDUSA_MODEL_TAG = "medusa_speculation_model"


class ModelWrapper(torch.nn.Module):
    def __init__(self, fn: Callable, batch_size: int, buckets: Sequence[int], tag: str = CONTEXT_ENCODING_MODEL_TAG,
                 pad_token_id: int = 0, is_continuous_batching: bool = False, padding_side: str = "right"):
        super().__init__()
        self.fn = fn
        self.batch_size = int(batch_size)
        self.buckets = sorted(int(b) for b in buckets)
        self.tag = tag
        self.pad_token_id = pad_token_id
        self.is_continuous_batching = is_continuous_batching
        if padding_side != "right":
            raise NotImplementedError("the KV-cache kernels assume right padding (cache row = position)")

    # -- sequence ---------------------------------------------------------------------------------------------------------
    def _bucket(self, T: int) -> int:
        for b in self.buckets:
            if T <= b:
                return b
        raise ValueError(f"{self.tag}: {T} tokens exceed the largest compiled bucket {self.buckets[-1]}")

    def _pad_sequence(self, input_ids: torch.Tensor, position_ids: torch.Tensor):
        T = input_ids.shape[1]
        Tb = self._bucket(T)
        if Tb == T:
            return input_ids, position_ids
        B, extra = input_ids.shape[0], Tb - T
        ids = torch.cat([input_ids, input_ids.new_full((B, extra), self.pad_token_id)], dim=1)
        # padded positions continue after the last real one: they write cache rows nobody reads (lengths stop before)
        tail = position_ids[:, -1:] + 1 + torch.arange(extra, device=position_ids.device, dtype=position_ids.dtype)
        return ids, torch.cat([position_ids, tail], dim=1)

    # -- batch -------------------------------------------------------------------------------------------------------------
    def _run_chunk(self, ids, pos, seq, last) -> torch.Tensor:
        n = ids.shape[0]
        if n < self.batch_size:     # repeat the first row: it rewrites its own cache line with the same values
            rep = self.batch_size - n
            ids, pos = torch.cat([ids, ids[:1].expand(rep, -1)]), torch.cat([pos, pos[:1].expand(rep, -1)])
            seq, last = torch.cat([seq, seq[:1].expand(rep)]), torch.cat([last, last[:1].expand(rep)])
        out = self.fn(ids.contiguous(), pos.contiguous(), seq.contiguous(), last.contiguous())
        return out[:n]

    def forward(self, input_ids: torch.Tensor, position_ids: torch.Tensor, seq_ids: Optional[torch.Tensor] = None,
                last_token_index: Optional[torch.Tensor] = None) -> torch.Tensor:
        B, T = input_ids.shape
        dev = input_ids.device
        if seq_ids is None:
            seq_ids = torch.arange(B, dtype=torch.int32, device=dev)
        if last_token_index is None:
            last_token_index = torch.full((B,), T - 1, dtype=torch.long, device=dev)
        input_ids, position_ids = self._pad_sequence(input_ids, position_ids)
        order = None
        if self.is_continuous_batching and B > 1:
            order = torch.argsort(seq_ids.to(torch.long), stable=True)
            input_ids, position_ids = input_ids[order], position_ids[order]
            seq_ids, last_token_index = seq_ids[order], last_token_index[order]
        outs: List[torch.Tensor] = []
        for i in range(0, B, self.batch_size):
            sl = slice(i, min(i + self.batch_size, B))
            outs.append(self._run_chunk(input_ids[sl], position_ids[sl], seq_ids[sl], last_token_index[sl]))
        out = torch.cat(outs, dim=0)
        if order is not None:
            restored = torch.empty_like(out)
            restored[order] = out
            out = restored
        return out
