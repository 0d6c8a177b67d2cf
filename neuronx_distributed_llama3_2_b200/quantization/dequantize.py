"""Dequantization helpers (reference ``quantization/dequantize.py:3-32``)."""
from __future__ import annotations

import torch


def direct_cast_dequantize(tensor: torch.Tensor, upcast_dtype: torch.dtype) -> torch.Tensor:
    """int8 -> activation dtype without scaling (the scale is applied to the matmul OUTPUT)."""
    return tensor.to(upcast_dtype)


def scale_dequantize(tensor: torch.Tensor, scale: torch.Tensor, upcast_dtype: torch.dtype) -> torch.Tensor:
    """``tensor * scale`` computed in fp32, returned in ``upcast_dtype``."""
    return (tensor.to(torch.float32) * scale.to(torch.float32)).to(upcast_dtype)
