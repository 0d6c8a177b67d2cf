"""Quantization configuration (reference ``quantization/quantization_config.py:19-60``)."""
from __future__ import annotations

import enum
from typing import Any, Dict, Union

import torch


class QuantizationType(enum.Enum):
    PER_TENSOR_SYMMETRIC = "per_tensor_symmetric"
    PER_CHANNEL_SYMMETRIC = "per_channel_symmetric"


class QuantizedDtype(enum.Enum):
    INT8 = torch.int8


BASE_QCONFIG_DICT_TYPE = Dict[str, Union[QuantizationType, QuantizedDtype, int, Any]]

_DEFAULT = {"quantization_type": QuantizationType.PER_TENSOR_SYMMETRIC, "quantized_dtype": QuantizedDtype.INT8}
_DEFAULT_PC = {"quantization_type": QuantizationType.PER_CHANNEL_SYMMETRIC, "quantized_dtype": QuantizedDtype.INT8,
               "quantization_per_channel_axis": 0}


def get_default_custom_qconfig_dict() -> BASE_QCONFIG_DICT_TYPE:
    return dict(_DEFAULT)


def get_default_per_channel_custom_qconfig_dict() -> BASE_QCONFIG_DICT_TYPE:
    return dict(_DEFAULT_PC)
