"""Pad attention heads so ``num_heads % tp == 0`` (reference
``src/neuronx_distributed/parallel_layers/pad.py:10-107``)."""
from __future__ import annotations

import torch

from .layers import ColumnParallelLinear, RowParallelLinear
from .parallel_state import get_tensor_model_parallel_size


def get_number_of_extra_heads(n_head: int, tp_degree: int) -> int:
    return 0 if n_head % tp_degree == 0 else tp_degree - n_head % tp_degree


def pad_model(model: torch.nn.Module, tp_degree: int, n_heads: int, wrapped_classes=(), pad_hook_fn=None):
    """Zero-pad every Column-parallel output dim / Row-parallel input dim inside modules of
    ``wrapped_classes`` by the ratio ``tgt_heads / num_heads`` and patch ``num_heads`` attributes.
    Padded heads produce zeros and receive zero gradients, so results are unchanged."""
    num_heads = n_heads
    tgt = num_heads + get_number_of_extra_heads(num_heads, tp_degree)
    if tgt == num_heads:
        return model
    ratio_num, ratio_den = tgt, num_heads
    tp = get_tensor_model_parallel_size()

    def pad_size(n_local: int) -> int:
        full = n_local * tp
        return (full * ratio_num // ratio_den) // tp - n_local

    def pad_module(mod: torch.nn.Module):
        for child in mod.children():
            if isinstance(child, ColumnParallelLinear):
                extra = pad_size(child.weight.shape[0])
                child.weight.data = torch.nn.functional.pad(child.weight.data, (0, 0, 0, extra))
                if child.bias is not None:
                    child.bias.data = torch.nn.functional.pad(child.bias.data, (0, extra))
                child.output_size_per_partition += extra
                child.output_size = child.output_size_per_partition * tp
            elif isinstance(child, RowParallelLinear):
                extra = pad_size(child.weight.shape[1])
                child.weight.data = torch.nn.functional.pad(child.weight.data, (0, extra))
                child.input_size_per_partition += extra
                child.input_size = child.input_size_per_partition * tp
            else:
                pad_module(child)
        if pad_hook_fn is not None:
            pad_hook_fn(mod, tgt, num_heads)
        elif hasattr(mod, "num_heads") and getattr(mod, "num_heads") == num_heads:
            mod.num_heads = tgt

    for m in model.modules():
        if wrapped_classes and isinstance(m, tuple(wrapped_classes)):
            pad_module(m)
    if not wrapped_classes:
        pad_module(model)
    return model
