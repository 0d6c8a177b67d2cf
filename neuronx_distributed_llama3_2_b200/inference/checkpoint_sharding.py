"""Shard an un-sharded checkpoint for one tensor-parallel rank without a process group (reference
``src/neuronx_distributed/trace/trace.py:618-736``: ``get_sharded_checkpoint`` -> ``invoke_preshard_hook`` +
``shard_children`` driven by ``partition_dim`` / ``partition_stride`` / fused-QKV tags; ``_mock_parallel_state``)."""
from __future__ import annotations

import contextlib
from typing import Any, Dict

import torch

from ..parallel import parallel_state as ps
from ..parallel.layers import create_local_weight


@contextlib.contextmanager
def _mock_parallel_state(tp_degree: int, rank: int):
    """Make model code believe it runs as TP rank ``rank`` of ``tp_degree`` (manual overrides, no communication)."""
    ps.set_tensor_model_parallel_size(tp_degree)
    ps.set_tensor_model_parallel_rank(rank)
    try:
        yield
    finally:
        ps.set_tensor_model_parallel_size(None)
        ps.set_tensor_model_parallel_rank(None)


def invoke_preshard_hook(module: torch.nn.Module, state: Dict[str, Any], prefix: str = "") -> None:
    if hasattr(module, "preshard_hook"):
        module.preshard_hook(state, prefix + "weight")
        return
    for name, child in module.named_children():
        invoke_preshard_hook(child, state, prefix + name + ".")


def shard_children(model: torch.nn.Module, state: Dict[str, Any], tp_degree: int, rank: int) -> Dict[str, Any]:
    """Slice every tensor whose parameter carries TP tags.  Fused QKV weights (``fused_qkv`` tag) are split into their
    q / k / v parts first so that each part is sharded on its own (KV parts are tiled ``kv_size_multiplier`` times)."""
    out = dict(state)
    for name, p in model.state_dict(keep_vars=True).items():
        if name not in out or not getattr(p, "tensor_model_parallel", False):
            continue
        full = out[name]
        dim, stride = p.partition_dim, getattr(p, "partition_stride", 1)
        if getattr(p, "fused_qkv", False):
            local_rows = p.shape[0]
            owner = _owner_module(model, name)
            q_sz, kv_sz = owner.q_output_size_per_partition, owner.kv_output_size_per_partition
            m = owner.kv_size_multiplier
            q_full, kv_full = owner.output_sizes
            qp, kp, vp = torch.split(full, [q_full, kv_full, kv_full], dim=0)
            parts = [torch.chunk(qp, tp_degree, 0)[rank],
                     torch.chunk(kp.repeat(m, *([1] * (kp.dim() - 1))), tp_degree, 0)[rank],
                     torch.chunk(vp.repeat(m, *([1] * (vp.dim() - 1))), tp_degree, 0)[rank]]
            out[name] = torch.cat(parts, dim=0).contiguous()
            assert out[name].shape[0] == local_rows == q_sz + 2 * kv_sz
            continue
        mult = getattr(p, "kv_size_multiplier", 1)
        if mult > 1 and full.shape[dim] * mult == p.shape[dim] * tp_degree:
            reps = [1] * full.dim()
            reps[dim] = mult
            full = full.repeat(*reps)
        per = full.shape[dim] // tp_degree
        out[name] = create_local_weight(full, dim, per, stride, rank=rank, world_size=tp_degree).contiguous()
    return out


def _owner_module(model: torch.nn.Module, param_name: str) -> torch.nn.Module:
    mod_name = param_name.rpartition(".")[0]
    return model.get_submodule(mod_name) if mod_name else model


def get_sharded_checkpoint(checkpoint: Dict[str, Any], model: torch.nn.Module, rank: int, tp_degree: int,
                           is_cached: bool = False) -> Dict[str, Any]:
    """``checkpoint`` = full (un-sharded) state dict keyed like ``model``; returns rank ``rank``'s shard."""
    state = dict(checkpoint)
    invoke_preshard_hook(model, state)
    return shard_children(model, state, tp_degree, rank)
