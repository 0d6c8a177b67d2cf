"""Quantized tensor-parallel linear layers: int8 weights + fp32 scales (reference
``quantization/quantization_layers.py:40-890``).

Forward = the reference's recipe (``:460-470``): cast int8 -> activation dtype (exact), matmul, multiply the OUTPUT by
the scale.  Per-channel scales follow the weight's sharding; per-tensor scales are replicated.  Inference only.
"""
from __future__ import annotations

import os

from typing import Optional

import torch
from torch.nn.parameter import Parameter

from ..modules.moe.moe_parallel_layers import ExpertFusedColumnParallelLinear, ExpertFusedRowParallelLinear
from ..ops import linear as gemm_ops
from ..parallel import mappings
from ..parallel import parallel_state as ps
from ..parallel.layers import ColumnParallelLinear, RowParallelLinear
from ..parallel.utils import divide, set_tensor_model_parallel_attributes
from .dequantize import direct_cast_dequantize, scale_dequantize
from .quantization_config import QuantizationType, QuantizedDtype


class QuantizedParallelLinearLayerStateDictAdaptor:
    """Reads / writes the three tensors of a quantized linear layer in a state dict whatever produced it (reference
    ``quantization_layers.py:252-339``): plain ``{prefix}weight`` (int8) + ``{prefix}scale``, or a torch
    ``{prefix}_packed_params._packed_params`` tuple ``(qint8 weight, bias)`` from ``torch.ao`` dynamic quantization.
    (The reference's ``get_bias_from_state_dict`` returns the weight for packed params -- SURVEY appendix B; here it
    returns the bias.)"""

    @staticmethod
    def _packed(prefix: str, state_dict: dict):
        return state_dict.get(prefix + "_packed_params._packed_params")

    @staticmethod
    def get_weight_from_state_dict(prefix: str, state_dict: dict) -> torch.Tensor:
        packed = QuantizedParallelLinearLayerStateDictAdaptor._packed(prefix, state_dict)
        w = packed[0] if packed is not None else state_dict[prefix + "weight"]
        return w.int_repr() if w.is_quantized else w

    @staticmethod
    def set_weight_to_state_dict(prefix: str, tensor: torch.Tensor, state_dict: dict) -> None:
        state_dict[prefix + "weight"] = tensor
        state_dict.pop(prefix + "_packed_params._packed_params", None)
        state_dict.pop(prefix + "_packed_params.dtype", None)

    @staticmethod
    def get_bias_from_state_dict(prefix: str, state_dict: dict) -> Optional[torch.Tensor]:
        packed = QuantizedParallelLinearLayerStateDictAdaptor._packed(prefix, state_dict)
        if packed is not None:
            return packed[1]
        return state_dict.get(prefix + "bias")

    @staticmethod
    def set_bias_to_state_dict(prefix: str, tensor: Optional[torch.Tensor], state_dict: dict) -> None:
        if tensor is not None:
            state_dict[prefix + "bias"] = tensor

    @staticmethod
    def get_scale_from_state_dict(prefix: str, state_dict: dict) -> torch.Tensor:
        packed = QuantizedParallelLinearLayerStateDictAdaptor._packed(prefix, state_dict)
        if packed is None:
            return state_dict[prefix + "scale"]
        w = packed[0]
        if w.qscheme() in (torch.per_channel_symmetric, torch.per_channel_affine):
            shape = [1] * w.dim()
            shape[w.q_per_channel_axis()] = -1
            return w.q_per_channel_scales().to(torch.float32).reshape(shape)
        return torch.tensor([w.q_scale()], dtype=torch.float32)

    @staticmethod
    def set_scale_to_state_dict(prefix: str, tensor: torch.Tensor, state_dict: dict) -> None:
        state_dict[prefix + "scale"] = tensor


class BaseQuantizeParallelLinear(torch.nn.Module):
    # state-dict accessors every quantized layer offers (reference ``quantization_layers.py:60-250``)
    get_weight_from_state_dict = staticmethod(QuantizedParallelLinearLayerStateDictAdaptor.get_weight_from_state_dict)
    set_weight_to_state_dict = staticmethod(QuantizedParallelLinearLayerStateDictAdaptor.set_weight_to_state_dict)
    get_bias_from_state_dict = staticmethod(QuantizedParallelLinearLayerStateDictAdaptor.get_bias_from_state_dict)
    set_bias_to_state_dict = staticmethod(QuantizedParallelLinearLayerStateDictAdaptor.set_bias_to_state_dict)
    get_scale_from_state_dict = staticmethod(QuantizedParallelLinearLayerStateDictAdaptor.get_scale_from_state_dict)

    def _setup_scale(self, weight_shape, partition_dim: int, quantization_type: QuantizationType,
                     per_channel_axis: Optional[int], device, expert_dims: int = 0) -> None:
        self.quantization_type = quantization_type
        if quantization_type == QuantizationType.PER_TENSOR_SYMMETRIC:
            self.scale = Parameter(torch.ones(1, dtype=torch.float32, device=device), requires_grad=False)
            self.scale_partitioned = False
        else:
            axis = (per_channel_axis if per_channel_axis is not None else 0) % len(weight_shape)
            shape = [1] * len(weight_shape)
            shape[axis] = weight_shape[axis]
            self.scale = Parameter(torch.ones(shape, dtype=torch.float32, device=device), requires_grad=False)
            self.scale_partitioned = axis == partition_dim
            if self.scale_partitioned:
                set_tensor_model_parallel_attributes(self.scale, True, partition_dim, 1)
        self.per_channel_axis = per_channel_axis

    @classmethod
    def from_float(cls, mod, q_config=None):
        raise NotImplementedError


def _int8_linear(x: torch.Tensor, weight_i8: torch.Tensor, scale: torch.Tensor, out_axis_scale: bool) -> torch.Tensor:
    """``x @ W^T`` with int8 ``W [out, in]``: exact cast, GEMM on the tensor cores, then scale the output."""
    x2 = x.reshape(-1, x.shape[-1])
    if (x2.shape[0] <= 8 and x2.is_cuda and x2.dtype == torch.bfloat16 and weight_i8.dim() == 2
            and os.environ.get("NXD_INT8_SKINNY", "1") == "1"):
        # token generation: stream the int8 weight once (half the bytes of bf16) instead of dequantising all of it
        # (csrc/skinny_int8.cu; NXD_INT8_SKINNY=0 takes the dequantise + GEMM path below)
        from ..ops.fused_comm import skinny_gemm_i8

        per_out = out_axis_scale and scale.numel() > 1
        if per_out or scale.numel() == 1:
            y = skinny_gemm_i8(x2.contiguous(), weight_i8, scale if per_out else scale.reshape(-1)[:1])
            if y is not None:
                return y.reshape(x.shape[:-1] + (weight_i8.shape[0],))
    w = direct_cast_dequantize(weight_i8, x.dtype)
    y = gemm_ops.gemm_nt(x2, w)
    s = scale.reshape(1, -1) if (out_axis_scale and scale.numel() > 1) else scale.reshape(-1)[:1]
    y = (y.float() * s.float()).to(x.dtype)
    return y.reshape(x.shape[:-1] + (w.shape[0],))


class QuantizedColumnParallel(BaseQuantizeParallelLinear):
    """int8 ``ColumnParallelLinear`` (weight ``[out/tp, in]``)."""

    def __init__(self, input_size: int, output_size: int, bias: bool = True, gather_output: bool = True,
                 dtype: torch.dtype = torch.float32, device=None,
                 quantization_type: QuantizationType = QuantizationType.PER_TENSOR_SYMMETRIC,
                 quantized_dtype: QuantizedDtype = QuantizedDtype.INT8, quantization_per_channel_axis: Optional[int] = None,
                 sequence_parallel_enabled: bool = False, stride: int = 1, keep_master_weight: bool = False):
        super().__init__()
        if sequence_parallel_enabled:
            raise NotImplementedError("quantized layers are inference-only: sequence parallelism is not supported")
        tp = ps.get_tensor_model_parallel_size()
        self.input_size, self.output_size = input_size, output_size
        self.stride, self.keep_master_weight = stride, keep_master_weight
        self.gather_output = gather_output
        self.dtype = dtype
        self.output_size_per_partition = divide(output_size, tp)
        self.weight = Parameter(torch.zeros(self.output_size_per_partition, input_size, dtype=quantized_dtype.value,
                                            device=device), requires_grad=False)
        set_tensor_model_parallel_attributes(self.weight, True, 0, stride)
        self._setup_scale(self.weight.shape, 0, quantization_type, quantization_per_channel_axis, device)
        if bias:
            self.bias = Parameter(torch.zeros(self.output_size_per_partition, dtype=dtype, device=device),
                                  requires_grad=False)
            set_tensor_model_parallel_attributes(self.bias, True, 0, 1)
        else:
            self.register_parameter("bias", None)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        y = _int8_linear(x, self.weight, self.scale, out_axis_scale=self.scale.dim() > 1 and self.scale.shape[0] > 1)
        if self.bias is not None:
            y = y + self.bias
        if self.gather_output:
            y = mappings.gather_from_tensor_model_parallel_region(y)
        return y

    @classmethod
    def from_float(cls, mod: ColumnParallelLinear, q_config=None):
        q_config = q_config or {}
        new = cls(mod.input_size, mod.output_size, bias=mod.bias is not None, gather_output=mod.gather_output,
                  dtype=mod.dtype, device=mod.weight.device,
                  quantization_type=q_config.get("quantization_type", QuantizationType.PER_TENSOR_SYMMETRIC),
                  quantized_dtype=q_config.get("quantized_dtype", QuantizedDtype.INT8),
                  quantization_per_channel_axis=q_config.get("quantization_per_channel_axis"))
        if mod.bias is not None and mod.bias.device.type != "meta":
            new.bias.data.copy_(mod.bias.data)
        # a strided shard layout (fused gate|up) must survive the swap: checkpoint sharding reads it from the weight
        setattr(new.weight, "partition_stride", getattr(mod.weight, "partition_stride", 1))
        if new.scale_partitioned:
            setattr(new.scale, "partition_stride", getattr(mod.weight, "partition_stride", 1))
        return new


class QuantizedRowParallel(BaseQuantizeParallelLinear):
    """int8 ``RowParallelLinear`` (weight ``[out, in/tp]``); the output scale is applied before the TP all-reduce."""

    def __init__(self, input_size: int, output_size: int, bias: bool = True, input_is_parallel: bool = False,
                 dtype: torch.dtype = torch.float32, device=None,
                 quantization_type: QuantizationType = QuantizationType.PER_TENSOR_SYMMETRIC,
                 quantized_dtype: QuantizedDtype = QuantizedDtype.INT8, quantization_per_channel_axis: Optional[int] = None,
                 sequence_parallel_enabled: bool = False, stride: int = 1, keep_master_weight: bool = False):
        super().__init__()
        if sequence_parallel_enabled:
            raise NotImplementedError("quantized layers are inference-only: sequence parallelism is not supported")
        tp = ps.get_tensor_model_parallel_size()
        self.input_size, self.output_size = input_size, output_size
        self.stride, self.keep_master_weight = stride, keep_master_weight
        self.input_is_parallel = input_is_parallel
        self.dtype = dtype
        self.input_size_per_partition = divide(input_size, tp)
        self.weight = Parameter(torch.zeros(output_size, self.input_size_per_partition, dtype=quantized_dtype.value,
                                            device=device), requires_grad=False)
        set_tensor_model_parallel_attributes(self.weight, True, 1, 1)
        self._setup_scale(self.weight.shape, 1, quantization_type, quantization_per_channel_axis, device)
        if bias:
            self.bias = Parameter(torch.zeros(output_size, dtype=dtype, device=device), requires_grad=False)
        else:
            self.register_parameter("bias", None)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not self.input_is_parallel:
            x = mappings.scatter_to_tensor_model_parallel_region(x)
        y = _int8_linear(x, self.weight, self.scale, out_axis_scale=self.scale.dim() > 1 and self.scale.shape[0] > 1)
        y = mappings.reduce_from_tensor_model_parallel_region(y)
        if self.bias is not None:
            y = y + self.bias
        return y

    @classmethod
    def from_float(cls, mod: RowParallelLinear, q_config=None):
        q_config = q_config or {}
        new = cls(mod.input_size, mod.output_size, bias=mod.bias is not None, input_is_parallel=mod.input_is_parallel,
                  dtype=mod.dtype, device=mod.weight.device,
                  quantization_type=q_config.get("quantization_type", QuantizationType.PER_TENSOR_SYMMETRIC),
                  quantized_dtype=q_config.get("quantized_dtype", QuantizedDtype.INT8),
                  quantization_per_channel_axis=q_config.get("quantization_per_channel_axis"))
        if mod.bias is not None and mod.bias.device.type != "meta":
            new.bias.data.copy_(mod.bias.data)
        return new


class _QuantizedExpertFused(BaseQuantizeParallelLinear):
    """MoE variants dequantize the (selected) expert slabs first (reference ``:736``) and then run the batched GEMM."""

    def _forward(self, x, expert_indices, reduce: bool):
        w = self.weight if expert_indices is None else self.weight[expert_indices]
        s = self.scale
        if s.dim() == w.dim() and s.shape[0] == self.weight.shape[0] and expert_indices is not None:
            s = s[expert_indices]
        wd = scale_dequantize(w, s, x.dtype)
        lead = x.shape[1:-1]
        y = torch.bmm(x.reshape(x.shape[0], -1, x.shape[-1]), wd).reshape((x.shape[0],) + tuple(lead) + (wd.shape[-1],))
        if reduce:
            y = mappings.reduce_from_tensor_model_parallel_region(y)
        return y


class QuantizedExpertFusedColumnParallel(_QuantizedExpertFused):
    def __init__(self, num_experts: int, input_size: int, output_size: int, dtype=torch.float32, device=None,
                 quantization_type=QuantizationType.PER_TENSOR_SYMMETRIC, quantized_dtype=QuantizedDtype.INT8,
                 quantization_per_channel_axis: Optional[int] = None, stride: int = 1, keep_master_weight: bool = False):
        super().__init__()
        self.stride, self.keep_master_weight = stride, keep_master_weight
        tp, ep = ps.get_tensor_model_parallel_size(), ps.get_expert_model_parallel_size()
        self.num_experts, self.input_size, self.output_size = num_experts, input_size, output_size
        self.weight = Parameter(torch.zeros(divide(num_experts, ep), input_size, divide(output_size, tp),
                                            dtype=quantized_dtype.value, device=device), requires_grad=False)
        set_tensor_model_parallel_attributes(self.weight, True, 2, 1)
        setattr(self.weight, "expert_model_parallel", True)
        self._setup_scale(self.weight.shape, 2, quantization_type, quantization_per_channel_axis, device)

    def forward(self, x, expert_indices=None):
        return self._forward(x, expert_indices, reduce=False)

    @classmethod
    def from_float(cls, mod: ExpertFusedColumnParallelLinear, q_config=None):
        q_config = q_config or {}
        return cls(mod.num_experts, mod.input_size, mod.output_size, dtype=mod.weight.dtype, device=mod.weight.device,
                   quantization_type=q_config.get("quantization_type", QuantizationType.PER_TENSOR_SYMMETRIC),
                   quantized_dtype=q_config.get("quantized_dtype", QuantizedDtype.INT8),
                   quantization_per_channel_axis=q_config.get("quantization_per_channel_axis"))


class QuantizedExpertFusedRowParallel(_QuantizedExpertFused):
    def __init__(self, num_experts: int, input_size: int, output_size: int, reduce_output: bool = True,
                 dtype=torch.float32, device=None, quantization_type=QuantizationType.PER_TENSOR_SYMMETRIC,
                 quantized_dtype=QuantizedDtype.INT8, quantization_per_channel_axis: Optional[int] = None,
                 stride: int = 1, keep_master_weight: bool = False):
        super().__init__()
        self.stride, self.keep_master_weight = stride, keep_master_weight
        tp, ep = ps.get_tensor_model_parallel_size(), ps.get_expert_model_parallel_size()
        self.num_experts, self.input_size, self.output_size = num_experts, input_size, output_size
        self.reduce_output = reduce_output
        self.weight = Parameter(torch.zeros(divide(num_experts, ep), divide(input_size, tp), output_size,
                                            dtype=quantized_dtype.value, device=device), requires_grad=False)
        set_tensor_model_parallel_attributes(self.weight, True, 1, 1)
        setattr(self.weight, "expert_model_parallel", True)
        self._setup_scale(self.weight.shape, 1, quantization_type, quantization_per_channel_axis, device)

    def forward(self, x, expert_indices=None):
        return self._forward(x, expert_indices, reduce=self.reduce_output)

    @classmethod
    def from_float(cls, mod: ExpertFusedRowParallelLinear, q_config=None):
        q_config = q_config or {}
        return cls(mod.num_experts, mod.input_size, mod.output_size, reduce_output=mod.reduce_output,
                   dtype=mod.weight.dtype, device=mod.weight.device,
                   quantization_type=q_config.get("quantization_type", QuantizationType.PER_TENSOR_SYMMETRIC),
                   quantized_dtype=q_config.get("quantized_dtype", QuantizedDtype.INT8),
                   quantization_per_channel_axis=q_config.get("quantization_per_channel_axis"))
