"""Quantize tensors / state dicts on the host (reference ``quantization/quantization_utils.py:55-126``)."""
from __future__ import annotations

from typing import Dict, Iterable, Optional, Tuple

import torch

from .observer import PerChannelAbsMaxObserver


def quantize_per_tensor_symmetric(w: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    scale = torch.clamp(w.detach().abs().max().float() / 127.0, min=torch.finfo(torch.float32).eps)
    q = torch.clamp(torch.round(w.detach().float() / scale), -128, 127).to(torch.int8)
    return q, scale.reshape(1)


def quantize_per_channel_symmetric(w: torch.Tensor, ch_axis: int = 0) -> Tuple[torch.Tensor, torch.Tensor]:
    obs = PerChannelAbsMaxObserver(ch_axis=ch_axis)
    obs(w)
    scale, _ = obs.calculate_qparams()
    shape = [1] * w.dim()
    shape[ch_axis % w.dim()] = -1
    q = torch.clamp(torch.round(w.detach().float() / scale.reshape(shape)), -128, 127).to(torch.int8)
    return q, scale.reshape(shape)


def convert_qint8_to_int8_state_dict(state_dict: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """torch ``qint8`` tensors -> plain ``int8`` + ``<name>.scale`` entries (reference ``:55-86``)."""
    out = {}
    for name, t in state_dict.items():
        if isinstance(t, torch.Tensor) and t.is_quantized:
            out[name] = t.int_repr()
            if t.qscheme() in (torch.per_channel_symmetric, torch.per_channel_affine):
                shape = [1] * t.dim()
                shape[t.q_per_channel_axis()] = -1
                out[name.replace("weight", "scale")] = t.q_per_channel_scales().to(torch.float32).reshape(shape)
            else:
                out[name.replace("weight", "scale")] = torch.tensor([t.q_scale()], dtype=torch.float32)
        else:
            out[name] = t
    return out


def quantize_state_dict(state_dict: Dict[str, torch.Tensor], per_channel: bool = False, ch_axis: int = 0,
                        include: Optional[Iterable[str]] = None) -> Dict[str, torch.Tensor]:
    """Quantize every 2-D+ floating ``*.weight`` (optionally only names containing one of ``include``) to int8 and
    add the matching ``*.scale`` (what the reference obtains from ``torch.quantization.quantize_dynamic`` on an HF model
    followed by ``convert_qint8_to_int8_state_dict``, ``:89-126``)."""
    out = {}
    for name, t in state_dict.items():
        pick = (isinstance(t, torch.Tensor) and t.is_floating_point() and t.dim() >= 2 and name.endswith("weight")
                and (include is None or any(s in name for s in include)))
        if not pick:
            out[name] = t
            continue
        q, s = quantize_per_channel_symmetric(t, ch_axis) if per_channel else quantize_per_tensor_symmetric(t)
        out[name] = q
        out[name[: -len("weight")] + "scale"] = s
    return out
