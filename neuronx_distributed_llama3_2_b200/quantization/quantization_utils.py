"""Quantize tensors / state dicts on the host (reference ``quantization/quantization_utils.py:55-126``)."""
from __future__ import annotations

from typing import Dict, Iterable, Optional, Tuple

import torch

from .observer import PerChannelAbsMaxObserver


def quantize_per_tensor_symmetric(tensor: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    w = tensor
    scale = torch.clamp(w.detach().abs().max().float() / 127.0, min=torch.finfo(torch.float32).eps)
    q = torch.clamp(torch.round(w.detach().float() / scale), -128, 127).to(torch.int8)
    return q, scale.reshape(1)


def quantize_per_channel_symmetric(tensor: torch.Tensor, channel_axis: int = 0) -> Tuple[torch.Tensor, torch.Tensor]:
    w, ch_axis = tensor, channel_axis
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
    packed_suffix = "_packed_params._packed_params"
    for name, t in state_dict.items():
        if name.endswith(packed_suffix):          # torch.ao dynamic-quantized Linear: (qint8 weight, bias) tuple
            prefix = name[: -len(packed_suffix)]
            qw, bias = t
            out[prefix + "weight"] = qw.int_repr()
            out[prefix + "scale"] = extract_q_scale(qw).to(torch.float32)
            if bias is not None:
                out[prefix + "bias"] = bias
            continue
        if name.endswith("_packed_params.dtype") or (name.endswith(".zero_point") and name[:-10] + packed_suffix[:0] + "_packed_params._packed_params" in state_dict):
            continue
        if name.endswith(".scale") and name[:-5] + "_packed_params._packed_params" in state_dict:
            continue                              # activation scale of the dynamic module: recomputed at run time
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


def extract_q_scale_per_tensor(q_tensor: torch.Tensor) -> torch.Tensor:
    """Scale of a per-tensor quantized tensor, shape ``[1]`` (reference ``:11-21``)."""
    assert q_tensor.qscheme() in (torch.per_tensor_affine, torch.per_tensor_symmetric)
    return torch.tensor([q_tensor.q_scale()])


def extract_q_scale_per_channel(q_tensor: torch.Tensor) -> torch.Tensor:
    """Per-channel scales shaped to broadcast against the weight (reference ``:24-43``)."""
    assert q_tensor.qscheme() in (torch.per_channel_affine, torch.per_channel_symmetric)
    shape = [1] * q_tensor.dim()
    shape[q_tensor.q_per_channel_axis()] = q_tensor.shape[q_tensor.q_per_channel_axis()]
    return q_tensor.q_per_channel_scales().to(torch.float32).reshape(shape)


def extract_q_scale(q_tensor: torch.Tensor) -> torch.Tensor:
    if q_tensor.qscheme() in (torch.per_tensor_affine, torch.per_tensor_symmetric):
        return extract_q_scale_per_tensor(q_tensor)
    if q_tensor.qscheme() in (torch.per_channel_affine, torch.per_channel_symmetric):
        return extract_q_scale_per_channel(q_tensor)
    raise ValueError(f"qscheme {q_tensor.qscheme()} is not supported")


def quantize_pytorch_model_per_tensor_symmetric(float_model: torch.nn.Module, inplace: bool = False) -> torch.nn.Module:
    """``torch.ao`` dynamic quantization of every ``nn.Linear`` with one symmetric int8 scale per weight (reference
    ``:89-95``); its state dict feeds ``convert_qint8_to_int8_state_dict``."""
    from torch.ao.quantization import default_dynamic_qconfig, quantize_dynamic

    return quantize_dynamic(float_model, qconfig_spec={torch.nn.Linear: default_dynamic_qconfig}, dtype=torch.qint8,
                            inplace=inplace)


def quantize_pytorch_model_per_channel_symmetric(float_model: torch.nn.Module, inplace: bool = False) -> torch.nn.Module:
    """Same with one scale per output channel, from ``PerChannelAbsMaxObserver`` (reference ``:98-110``)."""
    from torch.ao.quantization import QConfig, default_observer, quantize_dynamic

    from .observer import PerChannelAbsMaxObserver

    qconfig = QConfig(activation=default_observer,
                      weight=PerChannelAbsMaxObserver.with_args(dtype=torch.qint8, qscheme=torch.per_channel_symmetric,
                                                                ch_axis=0))
    return quantize_dynamic(float_model, qconfig_spec={torch.nn.Linear: qconfig}, dtype=torch.qint8, inplace=inplace)
