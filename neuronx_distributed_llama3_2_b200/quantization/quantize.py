"""``convert``: swap float parallel layers for their quantized counterparts (reference ``quantization/quantize.py:13-66``
and ``quantization_mappings.py:11-16``)."""
from __future__ import annotations

from typing import Dict, Optional, Type

import torch

from ..modules.moe.moe_parallel_layers import ExpertFusedColumnParallelLinear, ExpertFusedRowParallelLinear
from ..parallel.layers import ColumnParallelLinear, RowParallelLinear
from .quantization_config import BASE_QCONFIG_DICT_TYPE, get_default_custom_qconfig_dict
from .quantization_layers import (QuantizedColumnParallel, QuantizedExpertFusedColumnParallel,
                                  QuantizedExpertFusedRowParallel, QuantizedRowParallel)


def get_default_quant_module_mappings() -> Dict[Type[torch.nn.Module], Type[torch.nn.Module]]:
    return {
        ColumnParallelLinear: QuantizedColumnParallel,
        RowParallelLinear: QuantizedRowParallel,
        ExpertFusedColumnParallelLinear: QuantizedExpertFusedColumnParallel,
        ExpertFusedRowParallelLinear: QuantizedExpertFusedRowParallel,
    }


def convert(module: torch.nn.Module, q_config: Optional[BASE_QCONFIG_DICT_TYPE] = None, inplace: bool = False,
            mapping: Optional[Dict[Type[torch.nn.Module], Type[torch.nn.Module]]] = None) -> torch.nn.Module:
    """Replace every layer found in ``mapping`` by ``mapping[type].from_float(layer, q_config)``.  Weights are NOT
    quantized here: they are expected to arrive already quantized through ``load_state_dict`` (``*.weight`` int8 +
    ``*.scale``), as in the reference flow."""
    if q_config is None:
        q_config = get_default_custom_qconfig_dict()
    mapping = mapping or get_default_quant_module_mappings()
    if not inplace:
        import copy

        module = copy.deepcopy(module)
    _convert(module, q_config, mapping)
    return module


def _convert(module, q_config, mapping) -> None:
    for name, child in list(module.named_children()):
        if type(child) in mapping:
            setattr(module, name, mapping[type(child)].from_float(child, q_config))
        else:
            _convert(child, q_config, mapping)
