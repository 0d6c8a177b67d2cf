"""Bucket generation and the "bucket kernels" that fit a request into a compiled bucket (reference
``examples/inference/modules/autobucketing.py:6-126``).

The reference compiles one graph per bucket and ships TorchScript "bucket kernels" that pick the bucket from the
attention mask / position ids and slice or pad the inputs.  Here a bucket is a captured CUDA graph (``model_builder``),
the lookup is ``runtime.NativeRouter.route_padded`` and these functions are the host-side fit-to-bucket step.
"""
from __future__ import annotations

from math import log2
from typing import List, Sequence, Tuple

import torch


def generate_buckets(min_length: int, max_length: int) -> List[int]:
    """Powers of two from ``min_length`` up, closed by ``max_length`` itself; the last power of two is dropped when
    it would sit right below ``max_length`` (``round(log2(max))``), e.g. ``(128, 2048) -> [128, 256, 512, 1024, 2048]``,
    ``(128, 1100) -> [128, 256, 512, 1100]``."""
    if min_length == max_length:
        return [max_length]
    lo, hi = int(log2(min_length)), round(log2(max_length))
    return [2 ** i for i in range(lo, hi)] + [max_length]


def select_bucket(buckets: Sequence[int], length: int) -> int:
    """Index of the smallest bucket that holds ``length`` tokens."""
    for i, b in enumerate(buckets):
        if length <= b:
            return i
    raise ValueError(f"length {length} exceeds the largest bucket {buckets[-1]}")


def slice_lhs(tensor: torch.Tensor, bucket: int, dim: int) -> torch.Tensor:
    """First ``bucket`` entries of ``dim`` (right-padded inputs)."""
    return tensor.narrow(dim, 0, bucket)


def slice_rhs(tensor: torch.Tensor, bucket: int, max_idx: int, dim: int) -> torch.Tensor:
    """``bucket`` entries of ``dim`` ending at ``max_idx`` (left-padded inputs)."""
    return tensor.narrow(dim, max_idx - bucket, bucket)


def context_encoder_bk(tensors: List[torch.Tensor], buckets: Sequence[int], padding_side: str = "right",
                       pad_token: int = 0) -> Tuple[List[torch.Tensor], int]:
    """Fit ``[input_ids, attention_mask, position_ids, ...]`` (each ``[B, T, ...]`` on dim 1) into the smallest context
    bucket that holds the longest prompt: slice when the tensors are longer than the bucket, pad otherwise (ids with
    ``pad_token``, masks with 0, positions continue counting).  Returns the fitted tensors and the bucket index."""
    input_ids, attention_mask = tensors[0], tensors[1]
    longest = int(attention_mask.sum(dim=1).max()) if attention_mask.numel() else input_ids.shape[1]
    idx = select_bucket(buckets, max(longest, 1))
    bucket, T = buckets[idx], input_ids.shape[1]
    out = []
    for i, t in enumerate(tensors):
        if t.dim() < 2 or t.shape[1] != T:
            out.append(t)
        elif T >= bucket:
            out.append(slice_lhs(t, bucket, 1) if padding_side == "right" else slice_rhs(t, bucket, T, 1))
        else:
            pad_shape = list(t.shape)
            pad_shape[1] = bucket - T
            if i == 0:
                pad = torch.full(pad_shape, pad_token, dtype=t.dtype, device=t.device)
            elif i == 2:      # position ids keep counting so that padded rows never alias a real cache row
                pad = (t[:, -1:] + 1 + torch.arange(bucket - T, device=t.device, dtype=t.dtype)).expand(pad_shape)
            else:
                pad = torch.zeros(pad_shape, dtype=t.dtype, device=t.device)
            out.append(torch.cat([t, pad] if padding_side == "right" else [pad, t], dim=1))
    return out, idx


def token_generation_bk(tensors: List[torch.Tensor], buckets: Sequence[int], padding_side: str = "right"
                        ) -> Tuple[List[torch.Tensor], int]:
    """Token generation attends to the cache: pick the smallest cache bucket covering the furthest position and cut
    the attention mask ``tensors[1]`` (``[B, S_max]``) down to it."""
    attention_mask, position_ids = tensors[1], tensors[2]
    furthest = int(position_ids.max()) + 1
    idx = select_bucket(buckets, furthest)
    bucket = buckets[idx]
    out = list(tensors)
    if attention_mask.dim() == 2 and attention_mask.shape[1] > bucket:
        S = attention_mask.shape[1]
        out[1] = slice_lhs(attention_mask, bucket, 1) if padding_side == "right" else slice_rhs(attention_mask, bucket, S, 1)
    return out, idx


def get_context_encoder_bk():
    return context_encoder_bk


def get_token_generation_bk():
    return token_generation_bk
