"""GQA sharding strategies for inference (reference ``examples/inference/modules/gqa.py:19-692``).

With ``tp`` ranks every rank needs a whole number of query heads and at least one K/V head.  When the checkpoint's
head counts do not divide, the weights are rewritten once, before sharding:

``REPLICATE_TO_TP_DEGREE``  K/V heads are repeated until there are exactly ``tp`` of them (needs ``tp % kv == 0``) and the
                            query heads of every K/V group are padded with zero heads *inside their group* so that the
                            q -> kv alignment survives the contiguous split over ranks;
``CONVERT_TO_MHA``          K/V heads are repeated until every (padded) query head has its own -- always possible, costs
                            more KV cache.

Zero query heads produce zero attention output rows only after ``o_proj``'s matching input columns are zero as well, so
the output projection is padded with the same pattern.  ``transform_attention_weights`` applies the whole rewrite to
``q/k/v/o`` weights and ``tests/test_inference_cpu.py`` checks that attention with the rewritten weights equals the
original.
"""
from __future__ import annotations

import enum
from typing import Dict, Optional, Tuple

import torch


class GQA(enum.Enum):
    CONVERT_TO_MHA = "convert-to-mha"
    REPLICATE_TO_TP_DEGREE = "replicate-to-tp-degree"


def get_number_of_extra_heads(num_heads: int, tp_degree: int) -> int:
    """Heads to add so that ``num_heads`` becomes a multiple of ``tp_degree``."""
    return (-num_heads) % tp_degree


def determine_sharding_strategy(tp_degree: int, source_key_value_heads: int,
                                desired_sharding_strategy: Optional[GQA] = None) -> GQA:
    """Prefer replication to ``tp`` K/V heads; fall back to MHA conversion when ``tp`` is not a multiple of the K/V
    head count (reference ``:46-54``)."""
    strategy = desired_sharding_strategy or GQA.REPLICATE_TO_TP_DEGREE
    if strategy is GQA.REPLICATE_TO_TP_DEGREE and tp_degree % source_key_value_heads != 0:
        strategy = GQA.CONVERT_TO_MHA
    return strategy


def get_shardable_head_counts(tp_degree: int, num_attention_heads: int, num_key_value_heads: int,
                              sharding_strategy: GQA) -> Tuple[int, int]:
    """``(query heads, K/V heads)`` after padding / replication (reference ``:57-77``)."""
    q = num_attention_heads + get_number_of_extra_heads(num_attention_heads, tp_degree)
    kv = num_key_value_heads
    if num_attention_heads == num_key_value_heads:
        kv = q
    elif num_key_value_heads < tp_degree or num_key_value_heads % tp_degree != 0:
        if sharding_strategy is GQA.REPLICATE_TO_TP_DEGREE:
            if tp_degree % num_key_value_heads != 0:
                raise ValueError("REPLICATE_TO_TP_DEGREE needs tp_degree to be a multiple of num_key_value_heads")
            kv = tp_degree
        else:
            kv = q
    return q, kv


def _split_heads(t: torch.Tensor, dim: int, heads: int) -> torch.Tensor:
    shape = t.shape[:dim] + (heads, t.shape[dim] // heads) + t.shape[dim + 1:]
    return t.reshape(shape)


def _merge_heads(t: torch.Tensor, dim: int) -> torch.Tensor:
    shape = t.shape[:dim] + (t.shape[dim] * t.shape[dim + 1],) + t.shape[dim + 2:]
    return t.reshape(shape)


def pad_heads_interleaved(t: Optional[torch.Tensor], dim: int, source_heads: int, target_heads: int,
                          source_group_size: int) -> Optional[torch.Tensor]:
    """Zero heads appended to EVERY group of ``source_group_size`` heads along ``dim`` so that the total becomes
    ``target_heads`` (groups stay aligned with their K/V head).  ``target - source`` must split evenly over the groups."""
    if t is None or target_heads == source_heads:
        return t
    groups = source_heads // source_group_size
    extra = target_heads - source_heads
    if source_heads % source_group_size or extra % groups:
        raise ValueError(f"cannot pad {source_heads} heads in groups of {source_group_size} to {target_heads}")
    h = _split_heads(t, dim, source_heads)
    g = h.reshape(h.shape[:dim] + (groups, source_group_size) + h.shape[dim + 1:])
    pad_shape = list(g.shape)
    pad_shape[dim + 1] = extra // groups
    g = torch.cat([g, torch.zeros(pad_shape, dtype=t.dtype, device=t.device)], dim=dim + 1)
    h = g.reshape(g.shape[:dim] + (target_heads,) + g.shape[dim + 2:])
    return _merge_heads(h, dim)


def pad_heads_tail(t: Optional[torch.Tensor], dim: int, source_heads: int, target_heads: int) -> Optional[torch.Tensor]:
    """Zero heads appended at the end of ``dim``."""
    if t is None or target_heads == source_heads:
        return t
    h = _split_heads(t, dim, source_heads)
    pad_shape = list(h.shape)
    pad_shape[dim] = target_heads - source_heads
    return _merge_heads(torch.cat([h, torch.zeros(pad_shape, dtype=t.dtype, device=t.device)], dim=dim), dim)


def replicate_kv(t: Optional[torch.Tensor], source_heads: int, repeats: int, head_dim: int = 0) -> Optional[torch.Tensor]:
    """Every K/V head repeated ``repeats`` times in place (``k1 k1 k2 k2 ...``) along ``head_dim``."""
    if t is None or repeats == 1:
        return t
    h = _split_heads(t, head_dim, source_heads)
    return _merge_heads(torch.repeat_interleave(h, repeats, dim=head_dim), head_dim)


def transform_attention_weights(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, o: torch.Tensor, tp_degree: int,
                                num_attention_heads: int, num_key_value_heads: int,
                                desired_sharding_strategy: Optional[GQA] = None) -> Dict[str, object]:
    """Rewrite ``q [Hq*D, hidden]``, ``k / v [Hkv*D, hidden]``, ``o [hidden, Hq*D]`` so that they shard over
    ``tp_degree`` ranks.  Returns the new tensors and head counts."""
    strategy = determine_sharding_strategy(tp_degree, num_key_value_heads, desired_sharding_strategy)
    hq, hkv = get_shardable_head_counts(tp_degree, num_attention_heads, num_key_value_heads, strategy)
    group = num_attention_heads // num_key_value_heads
    if num_attention_heads == num_key_value_heads:                       # MHA: pad all four at the tail
        q2, k2, v2 = (pad_heads_tail(t, 0, num_attention_heads, hq) for t in (q, k, v))
        o2 = pad_heads_tail(o, 1, num_attention_heads, hq)
    elif hkv == num_key_value_heads:                                     # K/V heads already shard: only q may need padding
        q2 = pad_heads_interleaved(q, 0, num_attention_heads, hq, group) if hq != num_attention_heads else q
        o2 = pad_heads_interleaved(o, 1, num_attention_heads, hq, group) if hq != num_attention_heads else o
        k2, v2 = k, v
    elif strategy is GQA.REPLICATE_TO_TP_DEGREE:
        reps = hkv // num_key_value_heads
        # pad inside each K/V group first (group -> a multiple of `reps`), then every new K/V head serves group'/reps q heads
        q2 = pad_heads_interleaved(q, 0, num_attention_heads, hq, group)
        o2 = pad_heads_interleaved(o, 1, num_attention_heads, hq, group)
        k2, v2 = replicate_kv(k, num_key_value_heads, reps), replicate_kv(v, num_key_value_heads, reps)
    else:                                                                # CONVERT_TO_MHA
        k2 = replicate_kv(k, num_key_value_heads, group)
        v2 = replicate_kv(v, num_key_value_heads, group)
        q2 = pad_heads_tail(q, 0, num_attention_heads, hq)
        o2 = pad_heads_tail(o, 1, num_attention_heads, hq)
        k2 = pad_heads_tail(k2, 0, num_attention_heads, hkv)
        v2 = pad_heads_tail(v2, 0, num_attention_heads, hkv)
    return {"q": q2, "k": k2, "v": v2, "o": o2, "num_attention_heads": hq, "num_key_value_heads": hkv,
            "strategy": strategy}
