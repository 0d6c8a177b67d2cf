"""Expert-fused parallel linear layers and the router projection (reference
``src/neuronx_distributed/modules/moe/moe_parallel_layers.py:52-394``).

Weights are 3-D, one slab per local expert:
  ``ExpertFusedColumnParallelLinear``  ``(E/ep, H, I/tp)``  ``partition_dim = 2``
  ``ExpertFusedRowParallelLinear``     ``(E/ep, I/tp, H)``  ``partition_dim = 1``
All parameters carry ``expert_model_parallel = True`` (also after ``.to()`` and inside ``state_dict``), which is how
ZeRO-1, grad-norm and checkpoint IO tell them apart from dense parameters.
"""
from __future__ import annotations

import math
from typing import Callable, Optional

import torch
import torch.distributed as dist
import torch.nn.init as init
from torch.nn.parameter import Parameter

from ...parallel import mappings
from ...parallel import parallel_state as ps
from ...parallel.utils import divide, set_tensor_model_parallel_attributes


def _use_tc(*ts: torch.Tensor) -> bool:
    """tcgen05 GEMM per expert (bf16 on the GPU, TMA-compatible shapes); otherwise torch.bmm (CPU tests, fp32)."""
    from ...ops import linear as gemm_ops

    if not all(t.is_cuda and t.dtype == torch.bfloat16 for t in ts) or gemm_ops._backend() != "tcgen05":
        return False
    # only the contiguous (last) dims must be 16-byte multiples; row counts (the expert capacity) are free
    return all(t.shape[-1] % 8 == 0 for t in ts) and _lib_ok()


def _lib_ok() -> bool:
    from ...ops import _lib

    return _lib.available()


def _grouped_mm(x: torch.Tensor, w: torch.Tensor, layout: str = "nn") -> torch.Tensor:
    """Per-expert GEMM.  ``nn``: ``(e, c, a) @ (e, a, b) -> (e, c, b)``; ``nt``: ``(e, c, b) @ (e, a, b)^T -> (e, c, a)``;
    ``tn``: ``(e, c, a)^T @ (e, c, b) -> (e, a, b)``.  Extra leading dims of ``x`` are folded into ``c``.  On the GPU all
    experts run in ONE grouped tcgen05 launch (``ops.gemm.gemm_grouped``; the reference contracts over ``e`` in one
    einsum, ``moe_parallel_layers.py:78,94,108``); the capacity ``c`` needs no alignment."""
    e = x.shape[0]
    lead = x.shape[1:-1]
    x3 = x.reshape(e, -1, x.shape[-1])
    if layout == "tn":
        w3 = w.reshape(e, -1, w.shape[-1])
        if _use_tc(x3, w3):
            from ...ops.gemm import gemm_grouped

            return gemm_grouped(x3, w3, "tn")
        return torch.bmm(x3.transpose(1, 2), w3)
    n_out = w.shape[-1] if layout == "nn" else w.shape[-2]
    if _use_tc(x3, w) and n_out % 8 == 0:
        from ...ops.gemm import gemm_grouped

        y = gemm_grouped(x3, w, layout)
    else:
        y = torch.bmm(x3, w if layout == "nn" else w.transpose(1, 2))
    return y.reshape((e,) + tuple(lead) + (n_out,))


class ExpertFusedLinearWithAsyncCommunication(torch.autograd.Function):
    """Batched-over-experts linear: ``e...h,ehi->e...i`` with the TP all-reduce of the input grad (column flavour)
    issued asynchronously and overlapped with the weight-grad batch GEMM (reference ``:52-110``)."""

    @staticmethod
    def forward(ctx, input, weight, async_grad_allreduce: bool):
        ctx.async_grad_allreduce = async_grad_allreduce
        ctx.save_for_backward(input, weight)
        return _grouped_mm(input, weight)

    @staticmethod
    def backward(ctx, grad_output):
        input, weight = ctx.saved_tensors
        grad_input = _grouped_mm(grad_output, weight, "nt")
        handle = None
        if ctx.async_grad_allreduce and ps.get_tensor_model_parallel_size() > 1:
            grad_input = grad_input.contiguous()
            handle = dist.all_reduce(grad_input, group=ps.get_tensor_model_parallel_group(), async_op=True)
        grad_weight = _grouped_mm(input, grad_output, "tn") if weight.requires_grad else None
        if handle is not None:
            handle.wait()
        return grad_input, grad_weight, None


class ExpertFusedLinear(torch.nn.Module):
    """Mixin: marks parameters ``expert_model_parallel`` and keeps the mark through ``.to()`` / ``state_dict``."""

    def _mark_expert_parallel_weights(self, iterable=None):
        for p in (iterable if iterable is not None else self.parameters()):
            setattr(p, "expert_model_parallel", True)

    def _apply(self, fn, *a, **kw):
        out = super()._apply(fn, *a, **kw)
        self._mark_expert_parallel_weights()
        return out

    def _save_to_state_dict(self, destination, prefix, keep_vars):
        super()._save_to_state_dict(destination, prefix, keep_vars)
        for name in list(self._parameters.keys()):
            key = prefix + name
            if key in destination and destination[key] is not None:
                setattr(destination[key], "expert_model_parallel", True)


def _init_expert_weight(weight: torch.Tensor, init_method: Callable, partition_dim: int) -> None:
    """Every expert slab is initialised like a 2-D linear weight; the TP slice is cut from a full-size draw on CPU
    so that all TP ranks agree on the master weight."""
    set_tensor_model_parallel_attributes(weight, True, partition_dim, 1)
    if weight.device.type == "meta":
        return
    tp, rank = ps.get_tensor_model_parallel_size(), ps.get_tensor_model_parallel_rank()
    shape = list(weight.shape)
    shape[partition_dim] *= tp
    with torch.no_grad():
        full = torch.empty(shape, dtype=torch.float32)
        for e in range(shape[0]):
            init_method(full[e])
        per = weight.shape[partition_dim]
        weight.copy_(full.narrow(partition_dim, rank * per, per).to(weight.dtype))


class ExpertFusedColumnParallelLinear(ExpertFusedLinear):
    def __init__(self, num_experts: int, input_size: int, output_size: int, dtype: torch.dtype = torch.float32,
                 device=None, stride: int = 1, init_method: Optional[Callable] = None, keep_master_weight: bool = False):
        super().__init__()
        self.num_experts = num_experts
        self.input_size, self.output_size = input_size, output_size
        tp, ep = ps.get_tensor_model_parallel_size(), ps.get_expert_model_parallel_size()
        self.num_local_experts = divide(num_experts, ep)
        self.output_size_per_partition = divide(output_size, tp)
        self.async_tensor_model_parallel_allreduce = tp > 1
        self.init_method = init_method or (lambda w: init.kaiming_uniform_(w, a=math.sqrt(5)))
        self.set_weight_and_bias_config()
        self.weight = Parameter(torch.empty(self.weight_shape, dtype=dtype, device=device))
        _init_expert_weight(self.weight, self.init_method, self.weight_partition_dim)
        setattr(self.weight, "partition_stride", stride)
        self.register_parameter("bias", None)
        self._mark_expert_parallel_weights()

    def set_weight_and_bias_config(self) -> None:
        """One ``[in, out / tp]`` slab per local expert, column-parallel inside the slab, no bias (reference ``:183-192``)."""
        self.weight_shape = (self.num_local_experts, self.input_size, self.output_size_per_partition)
        self.weight_partition_dim = 2
        self.bias_shape = None

    def forward(self, input_: torch.Tensor, expert_indices: Optional[torch.Tensor] = None) -> torch.Tensor:
        w = self.weight if expert_indices is None else self.weight[expert_indices]
        return ExpertFusedLinearWithAsyncCommunication.apply(input_, w, self.async_tensor_model_parallel_allreduce)


class ExpertFusedRowParallelLinear(ExpertFusedLinear):
    def __init__(self, num_experts: int, input_size: int, output_size: int, reduce_output: bool = True,
                 dtype: torch.dtype = torch.float32, device=None, stride: int = 1,
                 init_method: Optional[Callable] = None, keep_master_weight: bool = False):
        super().__init__()
        self.num_experts = num_experts
        self.input_size, self.output_size = input_size, output_size
        self.reduce_output = reduce_output
        tp, ep = ps.get_tensor_model_parallel_size(), ps.get_expert_model_parallel_size()
        self.num_local_experts = divide(num_experts, ep)
        self.input_size_per_partition = divide(input_size, tp)
        self.init_method = init_method or (lambda w: init.kaiming_uniform_(w, a=math.sqrt(5)))
        self.set_weight_and_bias_config()
        self.weight = Parameter(torch.empty(self.weight_shape, dtype=dtype, device=device))
        _init_expert_weight(self.weight, self.init_method, self.weight_partition_dim)
        setattr(self.weight, "partition_stride", stride)
        self.register_parameter("bias", None)
        self._mark_expert_parallel_weights()

    def set_weight_and_bias_config(self) -> None:
        """One ``[in / tp, out]`` slab per local expert, row-parallel inside the slab, no bias (reference ``:274-283``)."""
        self.weight_shape = (self.num_local_experts, self.input_size_per_partition, self.output_size)
        self.weight_partition_dim = 1
        self.bias_shape = None

    def forward(self, input_: torch.Tensor, expert_indices: Optional[torch.Tensor] = None) -> torch.Tensor:
        w = self.weight if expert_indices is None else self.weight[expert_indices]
        out = ExpertFusedLinearWithAsyncCommunication.apply(input_, w, False)
        if self.reduce_output:
            out = mappings.reduce_from_tensor_model_parallel_region(out)
        return out


class LinearWithWeightGradAR(torch.autograd.Function):
    """Replicated router projection whose WEIGHT grad is all-reduced over TP: the MoE layer defers its output
    reduction, so each TP rank back-propagates only a partial grad into the router (reference ``:319-394``).
    ``reduce_weight_grad`` is False with expert parallelism: there the down projection has already reduced, the
    affinity gradient is complete on every TP rank and summing it again would scale it by tp (reference ``:383-385``).
    The INPUT grad stays partial here; ``MoE`` sums it together with the experts' input grad at its entry."""

    @staticmethod
    def forward(ctx, input, weight, reduce_weight_grad=True):
        ctx.reduce_weight_grad = reduce_weight_grad
        ctx.save_for_backward(input, weight)
        return torch.matmul(input, weight.t())

    @staticmethod
    def backward(ctx, grad_output):
        input, weight = ctx.saved_tensors
        grad_input = torch.matmul(grad_output, weight)
        gw = torch.matmul(grad_output.reshape(-1, grad_output.shape[-1]).t(), input.reshape(-1, input.shape[-1]))
        if ctx.reduce_weight_grad and ps.get_tensor_model_parallel_size() > 1:
            dist.all_reduce(gw, group=ps.get_tensor_model_parallel_group())
        return grad_input, gw, None


class LinearRouter(torch.nn.Module):
    def __init__(self, input_size: int, output_size: int, dtype: torch.dtype = torch.float32, device=None,
                 init_method: Optional[Callable] = None):
        super().__init__()
        self.weight = Parameter(torch.empty(output_size, input_size, dtype=dtype, device=device))
        self.init_method = init_method or (lambda w: init.kaiming_uniform_(w, a=math.sqrt(5)))
        if self.weight.device.type != "meta":
            self.init_weight_cpu()

    def init_weight_cpu(self) -> None:
        """(Re-)draw the replicated router weight (reference ``:393-394``); every rank must hold the same RNG state."""
        with torch.no_grad():
            self.init_method(self.weight)

    def forward(self, input_: torch.Tensor) -> torch.Tensor:
        return LinearWithWeightGradAR.apply(input_, self.weight, ps.get_expert_model_parallel_size() == 1)
