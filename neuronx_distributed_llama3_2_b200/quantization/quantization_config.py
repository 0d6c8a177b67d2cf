"""Quantization configuration (reference ``quantization/quantization_config.py:19-60``)."""
from __future__ import annotations

import enum
from typing import TypedDict

import torch


class MyEnumMeta(enum.EnumMeta):
    """``value in EnumClass`` for raw values as well as members (reference ``quantization_config.py:9-16``)."""

    def __contains__(cls, item) -> bool:
        if isinstance(item, cls):
            return True
        return any(item == m.value for m in cls)


class QuantizationType(enum.Enum, metaclass=MyEnumMeta):
    PER_TENSOR_SYMMETRIC = "per_tensor_symmetric"
    PER_CHANNEL_SYMMETRIC = "per_channel_symmetric"


class QuantizedDtype(enum.Enum, metaclass=MyEnumMeta):
    INT8 = torch.int8


class BASE_QCONFIG_DICT_TYPE(TypedDict):
    quantization_type: QuantizationType
    quantized_dtype: QuantizedDtype


class PER_CHANNEL_QCONFIG_DICT_TYPE(BASE_QCONFIG_DICT_TYPE):
    quantization_per_channel_axis: int


_DEFAULT = {"quantization_type": QuantizationType.PER_TENSOR_SYMMETRIC, "quantized_dtype": QuantizedDtype.INT8}
_DEFAULT_PC = {"quantization_type": QuantizationType.PER_CHANNEL_SYMMETRIC, "quantized_dtype": QuantizedDtype.INT8,
               "quantization_per_channel_axis": 0}


def get_default_custom_qconfig_dict() -> BASE_QCONFIG_DICT_TYPE:
    return dict(_DEFAULT)


def get_default_per_channel_custom_qconfig_dict() -> PER_CHANNEL_QCONFIG_DICT_TYPE:
    return dict(_DEFAULT_PC)
