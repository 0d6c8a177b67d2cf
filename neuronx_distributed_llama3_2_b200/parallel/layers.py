"""Tensor-parallel layers.

Reference: ``src/neuronx_distributed/parallel_layers/layers.py:58-1235``.

Activation layout follows the reference: with sequence parallelism the residual
stream is ``[S/tp, B, H]`` (sequence first) and the collective dim is 0.

What is different on B200:
  * ``LinearWithAsyncCommunication`` / ``_RowParallelMatmul`` call
    ``ops.comm_linear`` which runs ONE fused kernel for all-gather->GEMM and
    GEMM->reduce-scatter / all-reduce over NVSwitch peer memory when a symmetric
    memory context exists for the TP group, else NCCL + the tcgen05 GEMM;
  * the gathered input is *stashed* for backward by default (180 GB HBM makes
    the reference's re-gather, ``layers.py:338-346``, a bad trade);
  * wgrad can accumulate straight into an fp32 ``main_grad`` buffer from the GEMM
    epilogue (the reference's ``use_fp32_grad_acc`` hook, ``trainer.py:252-259``).
"""
from __future__ import annotations

import math
import os
from typing import Callable, Optional, Tuple

import torch
import torch.nn.functional as F
import torch.nn.init as init
from torch.nn.parameter import Parameter

from ..ops import comm_linear
from ..ops import linear as gemm_ops
from . import mappings
from .parallel_state import (
    get_tensor_model_parallel_rank,
    get_tensor_model_parallel_size,
)
from .random import get_rng_tracker
from .utils import (
    EmbeddingUtility,
    cast_if_autocast_enabled,
    divide,
    get_padding_length,
    set_tensor_model_parallel_attributes,
)


def _stash_gathered() -> bool:
    return os.environ.get("NXD_SP_STASH_GATHERED", "1") == "1"


# ---------------------------------------------------------------------------
# weight initialisation
# ---------------------------------------------------------------------------

def _initialize_affine_weight(weight: torch.Tensor, init_method: Callable, partition_dim: int, stride: int = 1) -> None:
    """Device-side init under the TP RNG stream so every TP rank draws a different
    slice (reference ``_initialize_affine_weight_neuron`` layers.py:43-55)."""
    set_tensor_model_parallel_attributes(tensor=weight, is_parallel=True, dim=partition_dim, stride=stride)
    with get_rng_tracker().fork():
        init_method(weight)


def create_local_weight(full_weight: torch.Tensor, partition_dim: int, per_partition_size: int, stride: int,
                        out_weight: Optional[torch.Tensor] = None, rank: Optional[int] = None,
                        world_size: Optional[int] = None) -> torch.Tensor:
    """Slice a full weight for this TP rank.  ``stride`` > 1 means the full weight
    is ``stride`` fused matrices (e.g. gate|up): the rank takes its chunk of each
    and concatenates them (reference layers.py:58-66)."""
    rank = get_tensor_model_parallel_rank() if rank is None else rank
    world_size = get_tensor_model_parallel_size() if world_size is None else world_size
    per_stride = divide(per_partition_size, stride)
    pieces = torch.split(full_weight, per_stride, dim=partition_dim)
    mine = pieces[rank::world_size]
    with torch.no_grad():
        res = torch.cat(mine, dim=partition_dim)
        if out_weight is not None:
            out_weight.copy_(res)
            return out_weight
        return res


def _initialize_parameter_cpu(param: torch.Tensor, partition_dim: int, init_method: Callable,
                              return_master_param: bool = False, *, param_dtype=torch.float32,
                              stride: int = 1) -> Optional[torch.Tensor]:
    """Initialise the full fp32 master weight identically on every rank, then keep
    this rank's slice (reference layers.py:71-98)."""
    set_tensor_model_parallel_attributes(tensor=param, is_parallel=True, dim=partition_dim, stride=stride)
    shape = list(param.shape)
    shape[partition_dim] *= get_tensor_model_parallel_size()
    master = torch.empty(shape, dtype=torch.float32, requires_grad=False)
    init_method(master)
    master = master.to(dtype=param_dtype)
    create_local_weight(master, partition_dim, param.shape[partition_dim], stride, out_weight=param.data)
    return master if return_master_param else None


# ---------------------------------------------------------------------------
# embedding
# ---------------------------------------------------------------------------

class ParallelEmbedding(torch.nn.Module):
    """Embedding sharded over the vocabulary (default) or the embedding dim
    (reference layers.py:101-285).

    Vocab-sharded forward: mask ids outside ``[start, end)``, local lookup, zero
    masked rows, all-reduce.  Dim-sharded: local lookup + all-gather on last dim.
    """

    def __init__(self, num_embeddings: int, embedding_dim: int, padding_idx: Optional[int] = None,
                 max_norm: Optional[float] = None, norm_type: float = 2.0, scale_grad_by_freq: bool = False,
                 sparse: bool = False, init_method: Callable = init.normal_, device=None,
                 dtype: torch.dtype = torch.float32, shard_across_embedding: bool = False, pad: bool = False,
                 sequence_parallel_enabled: bool = False):
        super().__init__()
        self.num_embeddings = num_embeddings
        self.embedding_dim = embedding_dim
        self.max_norm = max_norm
        self.norm_type = norm_type
        self.scale_grad_by_freq = scale_grad_by_freq
        self.sparse = sparse
        self.tensor_model_parallel_size = get_tensor_model_parallel_size()
        self.shard_across_embedding = shard_across_embedding
        self.stride = 1
        self.pad = pad
        self.pad_size = 0
        self.padding_idx = padding_idx
        # B200 addition: emit the [S/tp, B, H] sequence-parallel shard directly with
        # a reduce-scatter instead of all-reduce + local slice.
        self.sequence_parallel_enabled = sequence_parallel_enabled
        tp = self.tensor_model_parallel_size
        if shard_across_embedding:
            self.num_embeddings_per_partition = num_embeddings
            if pad:
                self.pad_size = get_padding_length(embedding_dim, tp)
                self.embedding_dim = embedding_dim + self.pad_size
            self.embedding_dim_per_partition = divide(self.embedding_dim, tp)
            self.weight_partition_dim = 1
            shape = (self.num_embeddings, self.embedding_dim_per_partition)
        else:
            if pad:
                self.pad_size = get_padding_length(num_embeddings, tp)
                self.num_embeddings = num_embeddings + self.pad_size
            self.start_index, self.end_index = EmbeddingUtility.range_from_global_vocab_size(
                self.num_embeddings, get_tensor_model_parallel_rank(), tp)
            self.num_embeddings_per_partition = self.end_index - self.start_index
            self.embedding_dim_per_partition = embedding_dim
            self.weight_partition_dim = 0
            shape = (self.num_embeddings_per_partition, self.embedding_dim)
        self.init_method = init_method
        self.weight = Parameter(torch.empty(*shape, device=device, dtype=dtype))
        self.init_weight_cpu_or_device()

    def init_weight_cpu(self) -> None:
        """Master-weight initialisation on the CPU (reference ``layers.py:207-213``)."""
        _initialize_parameter_cpu(self.weight, self.weight_partition_dim, self.init_method, param_dtype=self.weight.dtype,
                                  stride=self.stride)

    def init_weight_cpu_or_device(self):
        w = self.weight
        if w.device.type == "meta":
            set_tensor_model_parallel_attributes(w, True, self.weight_partition_dim, self.stride)
        elif w.device.type == "cpu":
            _initialize_parameter_cpu(w, self.weight_partition_dim, self.init_method, param_dtype=w.dtype,
                                      stride=self.stride)
        else:
            _initialize_affine_weight(w, self.init_method, self.weight_partition_dim, self.stride)

    def forward(self, input_: torch.Tensor) -> torch.Tensor:
        tp = self.tensor_model_parallel_size
        if self.shard_across_embedding:
            out = F.embedding(input_, self.weight, self.padding_idx, self.max_norm, self.norm_type,
                              self.scale_grad_by_freq, self.sparse)
            out = mappings.gather_from_tensor_model_parallel_region(out)
            if self.pad and self.pad_size > 0:
                out = out[..., : self.embedding_dim - self.pad_size]
            return out
        w = self.weight
        if (w.is_cuda and w.dtype == torch.bfloat16 and self.max_norm is None and self.padding_idx is None
                and input_.dim() == 2):
            # sm_100a gather kernel: masks out-of-shard ids and writes [S, B, H] directly when SP is on
            from ..ops.embedding import sp_embedding_push, vocab_parallel_embedding

            if self.sequence_parallel_enabled and tp > 1 and input_.shape[1] % tp == 0 and w.shape[1] % 8 == 0:
                sym = comm_linear._fused_ctx(w, None)
                if sym is not None:
                    # the owner of a vocab row writes it straight into the sequence shard of the rank that needs it
                    return sp_embedding_push(input_, w, self.start_index, sym)
            out = vocab_parallel_embedding(input_, w, self.start_index if tp > 1 else 0,
                                           seq_first_out=self.sequence_parallel_enabled)
            if self.sequence_parallel_enabled:
                return mappings.reduce_scatter_to_sequence_parallel_region(out)
            return mappings.reduce_from_tensor_model_parallel_region(out)
        if tp > 1:
            mask = (input_ < self.start_index) | (input_ >= self.end_index)
            local = (input_ - self.start_index).masked_fill(mask, 0)
        else:
            mask, local = None, input_
        out = F.embedding(local, w, self.padding_idx if tp == 1 else None, self.max_norm, self.norm_type,
                          self.scale_grad_by_freq, self.sparse)
        if mask is not None:
            out = out.masked_fill(mask.unsqueeze(-1), 0.0)
        if self.sequence_parallel_enabled:
            # [B, S, H] -> [S, B, H] -> reduce-scatter on S
            out = out.transpose(0, 1).contiguous()
            return mappings.reduce_scatter_to_sequence_parallel_region(out)
        return mappings.reduce_from_tensor_model_parallel_region(out)

    def preshard_hook(self, model_state_dict: dict, prefix: str) -> None:
        """Pad an unsharded checkpoint weight before slicing (reference :265-285)."""
        if not self.pad or self.pad_size == 0:
            return
        key = prefix
        w = model_state_dict[key]
        if self.shard_across_embedding:
            model_state_dict[key] = F.pad(w, (0, self.pad_size))
        else:
            model_state_dict[key] = F.pad(w, (0, 0, 0, self.pad_size))


# ---------------------------------------------------------------------------
# the core autograd kernel
# ---------------------------------------------------------------------------

def _flatten(x: torch.Tensor) -> torch.Tensor:
    return x.reshape(-1, x.shape[-1])


def _accumulate_wgrad(weight: torch.Tensor, dy2d: torch.Tensor, x2d: torch.Tensor) -> Optional[torch.Tensor]:
    """``dW = dY^T X``.  If the parameter carries an fp32 ``main_grad`` buffer the
    GEMM epilogue accumulates into it and autograd gets ``None``."""
    main_grad = getattr(weight, "main_grad", None)
    if main_grad is not None:
        gemm_ops.gemm_tn(dy2d, x2d, out=main_grad, accumulate=True)
        return None
    flat = getattr(weight, "_nxd_flat_grad", None)
    if (flat is not None and weight.grad is None and not getattr(weight, "_nxd_flat_written", False)
            and flat.dtype == weight.dtype and flat.shape == weight.shape):
        # first gradient after zero_grad(): the GEMM stores straight into the ZeRO-1 flat gradient buffer (no zero
        # fill before, no add after).  A fresh view object is returned so that autograd adopts it as ``.grad``.
        # ``.grad`` stays None until the END of a backward pass, so a weight that is used several times in one graph
        # (several forward passes before one backward, a module called twice) needs the flag: only the first of its
        # contributions may overwrite the buffer, the others are ordinary tensors that autograd adds.
        gemm_ops.gemm_tn(dy2d, x2d, out=flat)
        weight._nxd_flat_written = True
        return flat.view(flat.shape)
    return gemm_ops.gemm_tn(dy2d, x2d, out_dtype=weight.dtype)


class LinearWithAsyncCommunication(torch.autograd.Function):
    """``Y = X W^T (+ b)`` for a column-parallel weight (reference layers.py:288-417).

    fwd: SP -> all-gather X on dim 0 fused with the GEMM.
    bwd: ``dX = dY W`` then reduce-scatter (SP, fused) or all-reduce (async, overlapped
    with wgrad on the comm stream); ``dW = dY^T X_total``.
    """

    @staticmethod
    def forward(ctx, input: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor],
                async_grad_allreduce: bool, sequence_parallel_enabled: bool,
                save_for_backward: bool = True, process_group=None):
        ctx.use_bias = bias is not None and weight.requires_grad
        ctx.async_grad_allreduce = async_grad_allreduce
        ctx.sequence_parallel_enabled = sequence_parallel_enabled
        ctx.compute_weight_gradient = weight.requires_grad
        ctx.process_group = process_group
        ctx.in_shape = input.shape
        if sequence_parallel_enabled:
            out2d, total_input = comm_linear.ag_gemm(_flatten(input), weight, rows_per_rank_outer=input.shape[0],
                                                     inner=int(math.prod(input.shape[1:-1])), group=process_group)
            out_shape = (input.shape[0] * comm_linear.group_size(process_group),) + tuple(input.shape[1:-1])
        else:
            total_input = _flatten(input)
            out2d = gemm_ops.gemm_nt(total_input, weight)
            out_shape = tuple(input.shape[:-1])
        if bias is not None:
            out2d = out2d + bias
        if save_for_backward:
            if ctx.compute_weight_gradient:
                if sequence_parallel_enabled and not _stash_gathered():
                    ctx.save_for_backward(input, weight)
                    ctx.stashed = False
                else:
                    ctx.save_for_backward(total_input, weight)
                    ctx.stashed = True
            else:
                ctx.save_for_backward(weight)
        return out2d.reshape(out_shape + (weight.shape[0],))

    @staticmethod
    def backward(ctx, grad_output: torch.Tensor):
        if ctx.compute_weight_gradient:
            saved_input, weight = ctx.saved_tensors
        else:
            (weight,) = ctx.saved_tensors
            saved_input = None
        dy2d = _flatten(grad_output)
        if not dy2d.is_contiguous():
            dy2d = dy2d.contiguous()
        grad_weight = grad_bias = None
        handle = None
        if ctx.sequence_parallel_enabled:
            # dX_shard = reduce_scatter(dY W): one fused GEMM->RS kernel
            grad_input2d = comm_linear.gemm_rs(dy2d, weight, layout="nn", rows_outer=grad_output.shape[0],
                                               inner=int(math.prod(grad_output.shape[1:-1])),
                                               group=ctx.process_group)
        else:
            grad_input2d = gemm_ops.gemm_nn(dy2d, weight)
            if ctx.async_grad_allreduce:
                handle = comm_linear.all_reduce_async(grad_input2d, group=ctx.process_group)
        if ctx.compute_weight_gradient:
            if ctx.sequence_parallel_enabled and not ctx.stashed:
                total_input = comm_linear.all_gather_rows(_flatten(saved_input), rows_per_rank_outer=ctx.in_shape[0],
                                                          inner=int(math.prod(ctx.in_shape[1:-1])),
                                                          group=ctx.process_group)
            else:
                total_input = saved_input
            grad_weight = _accumulate_wgrad(weight, dy2d, total_input)
            if ctx.use_bias:
                grad_bias = dy2d.sum(dim=0)
        if handle is not None:
            handle.wait()
        grad_input = grad_input2d.reshape(ctx.in_shape)
        return grad_input, grad_weight, grad_bias, None, None, None, None


class _RowParallelMatmul(torch.autograd.Function):
    """``Y = reduce(X_local W_local^T)`` for a row-parallel weight, with the
    reduction (reduce-scatter under SP, else all-reduce) fused into the GEMM, and
    the adjoint all-gather fused into the dgrad GEMM (reference layers.py:785-797
    + mappings ``_ReduceScatterToSequenceParallelRegion``)."""

    @staticmethod
    def forward(ctx, input: torch.Tensor, weight: torch.Tensor, sequence_parallel_enabled: bool,
                reduce_output: bool, process_group=None):
        ctx.sequence_parallel_enabled = sequence_parallel_enabled
        ctx.reduce_output = reduce_output
        ctx.process_group = process_group
        ctx.in_shape = input.shape
        x2d = _flatten(input)
        ctx.save_for_backward(x2d, weight)
        lead = tuple(input.shape[:-1])
        if sequence_parallel_enabled:
            out2d = comm_linear.gemm_rs(x2d, weight, layout="nt", rows_outer=input.shape[0],
                                        inner=int(math.prod(input.shape[1:-1])), group=process_group)
            lead = (input.shape[0] // comm_linear.group_size(process_group),) + tuple(input.shape[1:-1])
        elif reduce_output:
            out2d = comm_linear.gemm_ar(x2d, weight, group=process_group)
        else:
            out2d = gemm_ops.gemm_nt(x2d, weight)
        return out2d.reshape(lead + (weight.shape[0],))

    @staticmethod
    def backward(ctx, grad_output: torch.Tensor):
        x2d, weight = ctx.saved_tensors
        dy2d = _flatten(grad_output)
        if not dy2d.is_contiguous():
            dy2d = dy2d.contiguous()
        if ctx.sequence_parallel_enabled:
            # dX = all_gather(dY_shard) W : fused AG->GEMM, the gathered dY is reused for wgrad
            grad_input2d, dy_total = comm_linear.ag_gemm(dy2d, weight, layout="nn", persistent_total=False,
                                                         rows_per_rank_outer=grad_output.shape[0],
                                                         inner=int(math.prod(grad_output.shape[1:-1])),
                                                         group=ctx.process_group)
        else:
            dy_total = dy2d
            grad_input2d = gemm_ops.gemm_nn(dy2d, weight)
        grad_weight = _accumulate_wgrad(weight, dy_total, x2d) if weight.requires_grad else None
        return grad_input2d.reshape(ctx.in_shape), grad_weight, None, None, None


def linear_with_async_allreduce(input: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor],
                                async_grad_allreduce: bool, sequence_parallel_enabled: bool,
                                autograd_func_class=LinearWithAsyncCommunication, save_for_backward: bool = True,
                                process_group=None) -> torch.Tensor:
    args = cast_if_autocast_enabled(input, weight, bias)
    dev = input.device.type
    with torch.autocast(device_type=dev, enabled=False):
        return autograd_func_class.apply(*args, async_grad_allreduce, sequence_parallel_enabled,
                                         save_for_backward, process_group)


# ---------------------------------------------------------------------------
# Column / Row parallel linear
# ---------------------------------------------------------------------------

class BaseParallelLinear(torch.nn.Module):
    """Template methods of the reference (``layers.py:420-457``): ``set_weight_and_bias_config`` decides shapes and the
    partition dim, ``initialize_weight_and_bias`` creates and initialises the parameters on ``self.device`` (meta: only
    the TP attributes; cpu: master-weight slice, ``init_weight_cpu``; cuda: direct init under the TP RNG tracker).
    Quantized / LoRA / expert-fused layers override the first and reuse the second."""

    def set_weight_and_bias_config(self) -> None:
        raise NotImplementedError

    def _bias_tp_attrs(self, bias: torch.Tensor) -> None:
        """Subclass hook: parallel attributes of the freshly created bias."""

    def init_weight_cpu(self) -> None:
        self.master_weight = _initialize_parameter_cpu(self.weight, self.weight_partition_dim, self.init_method,
                                                       param_dtype=self.weight.dtype, stride=self.stride,
                                                       return_master_param=self.keep_master_weight)

    def initialize_weight_and_bias(self) -> None:
        self.set_weight_and_bias_config()
        self.weight = Parameter(torch.empty(*self.weight_shape, dtype=self.dtype, device=self.device))
        self.device = self.weight.device
        self.master_weight = None
        if self.device.type == "meta":
            set_tensor_model_parallel_attributes(self.weight, True, self.weight_partition_dim, self.stride)
        elif self.device.type == "cpu":
            self.init_weight_cpu()
        else:
            _initialize_affine_weight(self.weight, self.init_method, self.weight_partition_dim, self.stride)
        if self.bias_shape is not None:
            self.bias = Parameter(torch.zeros(*self.bias_shape, dtype=self.dtype, device=self.device))
            self._bias_tp_attrs(self.bias)
        else:
            self.register_parameter("bias", None)

    def _init_weight(self, weight: torch.Tensor, partition_dim: int):
        if weight.device.type == "meta":
            set_tensor_model_parallel_attributes(weight, True, partition_dim, self.stride)
            return None
        if weight.device.type == "cpu":
            return _initialize_parameter_cpu(weight, partition_dim, self.init_method, param_dtype=weight.dtype,
                                             stride=self.stride, return_master_param=self.keep_master_weight)
        _initialize_affine_weight(weight, self.init_method, partition_dim, self.stride)
        return None


class ColumnParallelLinear(BaseParallelLinear):
    """``Y = X A^T + b`` with A split on its output dim: weight ``[out/tp, in]``,
    ``partition_dim=0`` (reference layers.py:460-634)."""

    autograd_func_class = LinearWithAsyncCommunication

    def __init__(self, input_size: int, output_size: int, bias: bool = True, gather_output: bool = True,
                 dtype: torch.dtype = torch.float32, device=None, stride: int = 1,
                 init_method: Optional[Callable] = None, sequence_parallel_enabled: bool = False,
                 keep_master_weight: bool = False, skip_bias_add: bool = False, pad: bool = False,
                 tensor_model_parallel_group=None):
        super().__init__()
        self.input_size = input_size
        self.output_size = output_size
        self.gather_output = gather_output
        self.pad = pad
        self.pad_size = 0
        self.tensor_model_parallel_group = tensor_model_parallel_group
        world = get_tensor_model_parallel_size()
        if pad:
            self.pad_size = get_padding_length(output_size, world)
            self.output_size = output_size + self.pad_size
        self.output_size_per_partition = divide(self.output_size, world)
        self.skip_bias_add = skip_bias_add
        self.stride = stride
        self.keep_master_weight = keep_master_weight
        self.init_method = init_method if init_method is not None else _default_linear_init
        self.sequence_parallel_enabled = sequence_parallel_enabled
        self.async_tensor_model_parallel_allreduce = (not sequence_parallel_enabled) and world > 1
        if self.sequence_parallel_enabled and world <= 1:
            self.sequence_parallel_enabled = False
        self.dtype = dtype
        self.device = device
        self.use_bias = self.add_bias = bias
        self.initialize_weight_and_bias()

    def set_weight_and_bias_config(self) -> None:
        # F.linear computes X A^T + b: the weight is stored transposed, [out / tp, in]
        self.weight_shape = (self.output_size_per_partition, self.input_size)
        self.weight_partition_dim = 0
        self.bias_shape = (self.output_size_per_partition,) if self.add_bias else None

    def _bias_tp_attrs(self, bias: torch.Tensor) -> None:
        set_tensor_model_parallel_attributes(bias, True, 0, stride=self.stride)

    def forward(self, input: torch.Tensor):
        if self.pad and self.training:
            raise RuntimeError("`pad=True` is only supported for inference (reference layers.py:597-598)")
        if self.async_tensor_model_parallel_allreduce or self.sequence_parallel_enabled:
            input_parallel = input
        else:
            input_parallel = mappings.copy_to_tensor_model_parallel_region(input)
        bias = self.bias if (self.use_bias and not self.skip_bias_add) else None
        out = linear_with_async_allreduce(
            input_parallel, self.weight, bias,
            async_grad_allreduce=self.async_tensor_model_parallel_allreduce,
            sequence_parallel_enabled=self.sequence_parallel_enabled,
            autograd_func_class=self.autograd_func_class,
            process_group=self.tensor_model_parallel_group,
        )
        if self.gather_output:
            assert not self.sequence_parallel_enabled
            out = mappings.gather_from_tensor_model_parallel_region(out)
            if self.pad and self.pad_size > 0:
                out = out[..., : self.output_size - self.pad_size]
        if self.skip_bias_add:
            return out, self.bias
        return out

    def preshard_hook(self, model_state_dict: dict, prefix: str) -> None:
        if not self.pad or self.pad_size == 0:
            return
        model_state_dict[prefix] = F.pad(model_state_dict[prefix], (0, 0, 0, self.pad_size))
        bkey = prefix.replace("weight", "bias")
        if self.use_bias and bkey in model_state_dict:
            model_state_dict[bkey] = F.pad(model_state_dict[bkey], (0, self.pad_size))


class RowParallelLinear(BaseParallelLinear):
    """``Y = X A^T + b`` with A split on its input dim: weight ``[out, in/tp]``,
    ``partition_dim=1``; output reduce-scattered (SP) or all-reduced
    (reference layers.py:637-810)."""

    def __init__(self, input_size: int, output_size: int, bias: bool = True, input_is_parallel: bool = False,
                 dtype: torch.dtype = torch.float32, device=None, stride: int = 1,
                 init_method: Optional[Callable] = None, sequence_parallel_enabled: bool = False,
                 keep_master_weight: bool = False, skip_bias_add: bool = False, pad: bool = False,
                 reduce_output: bool = True, tensor_model_parallel_group=None):
        super().__init__()
        self.input_size = input_size
        self.output_size = output_size
        self.input_is_parallel = input_is_parallel
        self.pad = pad
        self.pad_size = 0
        self.reduce_output = reduce_output
        self.tensor_model_parallel_group = tensor_model_parallel_group
        world = get_tensor_model_parallel_size()
        if pad:
            self.pad_size = get_padding_length(input_size, world)
            self.input_size = input_size + self.pad_size
        self.input_size_per_partition = divide(self.input_size, world)
        self.skip_bias_add = skip_bias_add
        self.stride = stride
        self.keep_master_weight = keep_master_weight
        self.init_method = init_method if init_method is not None else _default_linear_init
        self.sequence_parallel_enabled = sequence_parallel_enabled
        if self.sequence_parallel_enabled and not self.input_is_parallel:
            raise RuntimeError("To enable `sequence_parallel_enabled`, `input_is_parallel` must be `True`")
        if self.sequence_parallel_enabled and world <= 1:
            self.sequence_parallel_enabled = False
        self.dtype = dtype
        self.device = device
        self.use_bias = self.add_bias = bias
        self.initialize_weight_and_bias()

    def set_weight_and_bias_config(self) -> None:
        self.weight_shape = (self.output_size, self.input_size_per_partition)
        self.weight_partition_dim = 1
        self.bias_shape = (self.output_size,) if self.add_bias else None

    def _bias_tp_attrs(self, bias: torch.Tensor) -> None:
        # the bias is replicated; under SP every rank adds it to its sequence shard only -> gradient needs the TP sum
        setattr(bias, "sequence_parallel_enabled", self.sequence_parallel_enabled)

    def forward(self, input_: torch.Tensor):
        if self.pad and self.training:
            raise RuntimeError("`pad=True` is only supported for inference")
        if self.input_is_parallel:
            input_parallel = input_
        else:
            if self.pad and self.pad_size > 0:
                input_ = F.pad(input_, (0, self.pad_size))
            input_parallel = mappings.scatter_to_tensor_model_parallel_region(input_)
        x, w = cast_if_autocast_enabled(input_parallel, self.weight)
        with torch.autocast(device_type=x.device.type, enabled=False):
            out = _RowParallelMatmul.apply(x, w, self.sequence_parallel_enabled,
                                           self.reduce_output and get_tensor_model_parallel_size() > 1,
                                           self.tensor_model_parallel_group)
        if self.skip_bias_add:
            return out, self.bias
        if self.bias is not None:
            out = out + self.bias
        return out

    def preshard_hook(self, model_state_dict: dict, prefix: str) -> None:
        if not self.pad or self.pad_size == 0:
            return
        model_state_dict[prefix] = F.pad(model_state_dict[prefix], (0, self.pad_size))


def _default_linear_init(w: torch.Tensor) -> None:
    init.kaiming_uniform_(w, a=math.sqrt(5))


# ---------------------------------------------------------------------------
# Conv2d
# ---------------------------------------------------------------------------

class Conv2dWithInputGradAllReduce(torch.autograd.Function):
    """Conv2d whose input-grad all-reduce over TP is launched asynchronously and
    overlapped with the weight-grad computation (reference layers.py:813-874)."""

    @staticmethod
    def forward(ctx, input, weight, bias, stride, padding, allreduce_input_grad):
        ctx.stride, ctx.padding = stride, padding
        ctx.allreduce_input_grad = allreduce_input_grad
        ctx.use_bias = bias is not None
        ctx.save_for_backward(input, weight)
        return F.conv2d(input, weight, bias, stride, padding)

    @staticmethod
    def backward(ctx, grad_output):
        input, weight = ctx.saved_tensors
        grad_input = torch.nn.grad.conv2d_input(input.shape, weight, grad_output, ctx.stride, ctx.padding)
        handle = None
        if ctx.allreduce_input_grad and get_tensor_model_parallel_size() > 1:
            handle = comm_linear.all_reduce_async(grad_input)
        grad_weight = torch.nn.grad.conv2d_weight(input, weight.shape, grad_output, ctx.stride, ctx.padding)
        grad_bias = grad_output.sum(dim=(0, 2, 3)) if ctx.use_bias else None
        if handle is not None:
            handle.wait()
        return grad_input, grad_weight, grad_bias, None, None, None


class Conv2dWithWeightGradAllReduce(torch.autograd.Function):
    """Conv2d over a TP-sharded INPUT (the weight is replicated along the sharded dim of the activations): the weight /
    bias gradients are partial on every rank and all-reduced over TP (reference ``layers.py:877-930``)."""

    @staticmethod
    def forward(ctx, input, weight, bias, stride, padding, allreduce_weight_grad):
        ctx.stride, ctx.padding = stride, padding
        ctx.allreduce_weight_grad = allreduce_weight_grad
        ctx.use_bias = bias is not None
        ctx.save_for_backward(input, weight)
        return F.conv2d(input, weight, bias, stride, padding)

    @staticmethod
    def backward(ctx, grad_output):
        input, weight = ctx.saved_tensors
        grad_weight = torch.nn.grad.conv2d_weight(input, weight.shape, grad_output, ctx.stride, ctx.padding)
        grad_bias = grad_output.sum(dim=(0, 2, 3)) if ctx.use_bias else None
        handles = []
        if ctx.allreduce_weight_grad and get_tensor_model_parallel_size() > 1:
            handles.append(comm_linear.all_reduce_async(grad_weight))
            if grad_bias is not None:
                handles.append(comm_linear.all_reduce_async(grad_bias))
        grad_input = torch.nn.grad.conv2d_input(input.shape, weight, grad_output, ctx.stride, ctx.padding)
        for h in handles:
            h.wait()
        return grad_input, grad_weight, grad_bias, None, None, None


def conv2d_with_weight_grad_allreduce(input: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor],
                                      stride: Tuple[int, int], padding: Tuple[int, int],
                                      allreduce_weight_grad: bool) -> torch.Tensor:
    args = cast_if_autocast_enabled(input, weight, bias, stride, padding, allreduce_weight_grad)
    with torch.amp.autocast("cuda", enabled=False):
        return Conv2dWithWeightGradAllReduce.apply(*args)


def _pair(v) -> Tuple[int, int]:
    return (v, v) if isinstance(v, int) else tuple(v)


class BaseParallelConv(torch.nn.Module):
    """Shared ctor checks for the two conv flavours (reference layers.py:904-1020)."""

    def __init__(self, in_channels, out_channels, kernel_size, stride, padding, dilation, groups, bias,
                 padding_mode, partition_dim, dtype, device, init_method, keep_master_weight):
        super().__init__()
        if _pair(dilation) != (1, 1):
            raise NotImplementedError("dilation is not supported")
        if groups != 1:
            raise NotImplementedError("groups != 1 is not supported")
        if padding_mode != "zeros":
            raise NotImplementedError("only zero padding is supported")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride, self.padding = _pair(kernel_size), _pair(stride), _pair(padding)
        self.partition_dim = partition_dim
        self.stride_tp = 1
        self.keep_master_weight = keep_master_weight
        self.init_method = init_method if init_method is not None else _default_linear_init
        world = get_tensor_model_parallel_size()
        oc = divide(out_channels, world) if partition_dim == 0 else out_channels
        ic = divide(in_channels, world) if partition_dim == 1 else in_channels
        self.weight = Parameter(torch.empty(oc, ic, *self.kernel_size, dtype=dtype, device=device))
        if self.weight.device.type == "meta":
            set_tensor_model_parallel_attributes(self.weight, True, partition_dim, 1)
            self.master_weight = None
        elif self.weight.device.type == "cpu":
            self.master_weight = _initialize_parameter_cpu(self.weight, partition_dim, self.init_method,
                                                           param_dtype=dtype,
                                                           return_master_param=keep_master_weight)
        else:
            _initialize_affine_weight(self.weight, self.init_method, partition_dim)
            self.master_weight = None
        if bias:
            self.bias = Parameter(torch.zeros(oc if partition_dim == 0 else out_channels, dtype=dtype, device=device))
            if partition_dim == 0:
                set_tensor_model_parallel_attributes(self.bias, True, 0, 1)
        else:
            self.register_parameter("bias", None)


class OutputChannelParallelConv2d(BaseParallelConv):
    """Conv2d split over output channels; optional channel all-gather of the result
    (reference layers.py:1033-1131)."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1,
                 bias=True, padding_mode="zeros", gather_output=True, dtype=torch.float32, device=None,
                 init_method=None, keep_master_weight=False):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, bias,
                         padding_mode, 0, dtype, device, init_method, keep_master_weight)
        self.gather_output = gather_output

    def forward(self, x):
        x, w, b = cast_if_autocast_enabled(x, self.weight, self.bias)
        out = Conv2dWithInputGradAllReduce.apply(x, w, b, self.stride, self.padding, True)
        if self.gather_output:
            out = mappings.gather_from_tensor_model_parallel_region_with_dim(out, 1)
        return out


class InputChannelParallelConv2d(BaseParallelConv):
    """Conv2d split over input channels; result all-reduced, bias added after
    (reference layers.py:1134-1235)."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1,
                 bias=True, padding_mode="zeros", input_is_parallel=False, dtype=torch.float32, device=None,
                 init_method=None, keep_master_weight=False):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, bias,
                         padding_mode, 1, dtype, device, init_method, keep_master_weight)
        self.input_is_parallel = input_is_parallel

    def forward(self, x):
        if not self.input_is_parallel:
            x = mappings.scatter_input_channels_to_tensor_model_parallel_region(x)
        x, w = cast_if_autocast_enabled(x, self.weight)
        out = Conv2dWithInputGradAllReduce.apply(x, w, None, self.stride, self.padding, False)
        out = mappings.reduce_from_tensor_model_parallel_region(out)
        if self.bias is not None:
            out = out + self.bias.view(1, -1, 1, 1)
        return out
