"""Collective mappings as autograd Functions.

Reference: ``src/neuronx_distributed/parallel_layers/mappings.py:24-486``.  Each
Function pairs a forward collective with its adjoint.  These are the *unfused*
NCCL (GPU) / gloo (CPU) versions; the Column/Row parallel layers bypass them on
B200 and call the fused AG->GEMM / GEMM->RS kernels (``ops/fused_comm.py``) when a
symmetric-memory context is available.  They remain the semantic definition (and
the baseline the fused kernels are measured against).

Unlike the XLA reference there is no ``nonzero_partition_dim_swap`` requirement:
NCCL collectives operate on contiguous buffers, so for a non-zero dim we move the
dim to the front with one ``movedim(...).contiguous()`` (a single strided copy).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from . import parallel_state as ps
from .utils import divide

# ---------------------------------------------------------------------------
# primitives
# ---------------------------------------------------------------------------


def _group_size(group) -> int:
    if group is None:
        return 1
    return dist.get_world_size(group=group)


def _tp_group():
    return ps.get_tensor_model_parallel_group()


def _tp_size() -> int:
    return ps.get_tensor_model_parallel_size()


def _reduce(input_: torch.Tensor, group=None, op=dist.ReduceOp.SUM) -> torch.Tensor:
    """In-place all-reduce over ``group`` (TP by default)."""
    if group is None:
        if _tp_size() == 1:
            return input_
        group = _tp_group()
    if _group_size(group) == 1:
        return input_
    if not input_.is_contiguous():
        input_ = input_.contiguous()
    dist.all_reduce(input_, op=op, group=group)
    return input_


def _split_along_dim(input_: torch.Tensor, dim: int, group=None, rank: Optional[int] = None,
                     world: Optional[int] = None) -> torch.Tensor:
    """Keep this rank's slice of ``dim``."""
    if world is None:
        world = _tp_size() if group is None else _group_size(group)
    if world == 1:
        return input_
    if rank is None:
        rank = ps.get_tensor_model_parallel_rank() if group is None else dist.get_rank(group=group)
    size = divide(input_.size(dim), world)
    return input_.narrow(dim, rank * size, size).contiguous()


def _gather_along_dim(input_: torch.Tensor, dim: int, group=None) -> torch.Tensor:
    """All-gather and concatenate along ``dim``."""
    if group is None:
        if _tp_size() == 1:
            return input_
        group = _tp_group()
    world = _group_size(group)
    if world == 1:
        return input_
    dim = dim % input_.dim()
    src = input_.contiguous()
    flat = torch.empty((world * src.shape[0],) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    dist.all_gather_into_tensor(flat, src, group=group)
    if dim == 0:
        return flat
    # [world, d0, .., dk, ..] -> [d0, .., world*dk, ..]
    out = flat.view((world,) + tuple(src.shape)).movedim(0, dim)
    shape = list(src.shape)
    shape[dim] *= world
    return out.reshape(shape)


def _reduce_scatter_along_dim(input_: torch.Tensor, dim: int, group=None) -> torch.Tensor:
    """Reduce-scatter along ``dim``: every rank gets the sum of its slice."""
    if group is None:
        if _tp_size() == 1:
            return input_
        group = _tp_group()
    world = _group_size(group)
    if world == 1:
        return input_
    dim = dim % input_.dim()
    assert input_.size(dim) % world == 0, (
        f"reduce-scatter dim {dim} of size {input_.size(dim)} not divisible by group size {world}"
    )
    if dim != 0:
        shape = list(input_.shape)
        part = shape[dim] // world
        view = input_.reshape(shape[:dim] + [world, part] + shape[dim + 1:]).movedim(dim, 0).contiguous()
        out = torch.empty(view.shape[1:], dtype=input_.dtype, device=input_.device)
        dist.reduce_scatter_tensor(out, view.view((world * view.shape[1],) + tuple(view.shape[2:])), group=group)
        return out
    src = input_.contiguous()
    out = torch.empty((src.shape[0] // world,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    dist.reduce_scatter_tensor(out, src, group=group)
    return out


def _all_to_all_in_expert_parallel_region(x: torch.Tensor, split_dim: int, concat_dim: int) -> torch.Tensor:
    """All-to-all over the EP group: split ``split_dim`` into ep chunks, send chunk
    i to EP rank i, concatenate what arrives along ``concat_dim``
    (reference mappings.py:149-162)."""
    ep = ps.get_expert_model_parallel_size()
    if ep == 1:
        return x
    group = ps.get_expert_model_parallel_group()
    send = torch.stack([c for c in torch.chunk(x, ep, dim=split_dim)], dim=0).contiguous()
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send, group=group)  # one call; works on nccl and gloo
    return torch.cat(list(recv.unbind(0)), dim=concat_dim)


# ---------------------------------------------------------------------------
# autograd Functions
# ---------------------------------------------------------------------------


class _CopyToModelParallelRegion(torch.autograd.Function):
    """fwd identity / bwd all-reduce."""

    @staticmethod
    def forward(ctx, input_):
        return input_

    @staticmethod
    def backward(ctx, grad_output):
        return _reduce(grad_output.contiguous().clone() if _tp_size() > 1 else grad_output)


class _ReduceFromModelParallelRegion(torch.autograd.Function):
    """fwd all-reduce / bwd identity."""

    @staticmethod
    def forward(ctx, input_):
        return _reduce(input_.clone() if _tp_size() > 1 else input_)

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output


class _ScatterToModelParallelRegion(torch.autograd.Function):
    """fwd split on ``dim`` / bwd all-gather."""

    @staticmethod
    def forward(ctx, input_, dim=-1):
        ctx.dim = dim
        return _split_along_dim(input_, dim)

    @staticmethod
    def backward(ctx, grad_output):
        return _gather_along_dim(grad_output, ctx.dim), None


class _GatherFromModelParallelRegion(torch.autograd.Function):
    """fwd all-gather on ``dim`` / bwd split."""

    @staticmethod
    def forward(ctx, input_, dim=-1):
        ctx.dim = dim
        return _gather_along_dim(input_, dim)

    @staticmethod
    def backward(ctx, grad_output):
        return _split_along_dim(grad_output, ctx.dim), None


class _ScatterToSequenceParallelRegion(torch.autograd.Function):
    """fwd split on the sequence dim / bwd all-gather."""

    @staticmethod
    def forward(ctx, input_, dim=0):
        ctx.dim = dim
        return _split_along_dim(input_, dim)

    @staticmethod
    def backward(ctx, grad_output):
        return _gather_along_dim(grad_output, ctx.dim), None


class _GatherFromSequenceParallelRegion(torch.autograd.Function):
    """fwd all-gather on the sequence dim / bwd reduce-scatter (if the consumer is
    tensor-parallel, i.e. produces partial grads) or split."""

    @staticmethod
    def forward(ctx, input_, dim=0, to_model_parallel=True):
        ctx.dim = dim
        ctx.to_model_parallel = to_model_parallel
        return _gather_along_dim(input_, dim)

    @staticmethod
    def backward(ctx, grad_output):
        if ctx.to_model_parallel:
            return _reduce_scatter_along_dim(grad_output, ctx.dim), None, None
        return _split_along_dim(grad_output, ctx.dim), None, None


class _ReduceScatterToSequenceParallelRegion(torch.autograd.Function):
    """fwd reduce-scatter on the sequence dim / bwd all-gather."""

    @staticmethod
    def forward(ctx, input_, dim=0):
        ctx.dim = dim
        return _reduce_scatter_along_dim(input_, dim)

    @staticmethod
    def backward(ctx, grad_output):
        return _gather_along_dim(grad_output, ctx.dim), None


class _AllToAllInExpertParallelRegion(torch.autograd.Function):
    """fwd all-to-all / bwd the inverse all-to-all (reference :311-338)."""

    @staticmethod
    def forward(ctx, input_, split_dim, concat_dim):
        ctx.split_dim, ctx.concat_dim = split_dim, concat_dim
        return _all_to_all_in_expert_parallel_region(input_, split_dim, concat_dim)

    @staticmethod
    def backward(ctx, grad_output):
        return (
            _all_to_all_in_expert_parallel_region(grad_output, ctx.concat_dim, ctx.split_dim),
            None,
            None,
        )


class _ScatterInputChannelsToModelParallelRegion(torch.autograd.Function):
    """Conv2d input-channel split (dim 1) / all-gather in bwd (reference :341-354)."""

    @staticmethod
    def forward(ctx, input_):
        return _split_along_dim(input_, 1)

    @staticmethod
    def backward(ctx, grad_output):
        return _gather_along_dim(grad_output, 1)


class _ExportCtx:
    """Stand-in for the autograd context when a mapping is evaluated outside autograd."""


def _attach_symbolic(fn_cls) -> None:
    """``symbolic(graph, *inputs)`` -- the static hook ``torch.onnx`` / graph exporters call instead of ``forward``
    (the reference defines one per mapping, ``parallel_layers/mappings.py:169-354``).  The exported computation of every
    mapping is its forward collective, so the hook evaluates ``forward`` with a throw-away context."""
    def symbolic(graph, *inputs):
        return fn_cls.forward(_ExportCtx(), *inputs)

    symbolic.__qualname__ = f"{fn_cls.__name__}.symbolic"
    fn_cls.symbolic = staticmethod(symbolic)


for _cls in (_CopyToModelParallelRegion, _ReduceFromModelParallelRegion, _ScatterToModelParallelRegion,
             _GatherFromModelParallelRegion, _ScatterToSequenceParallelRegion, _GatherFromSequenceParallelRegion,
             _ReduceScatterToSequenceParallelRegion, _AllToAllInExpertParallelRegion,
             _ScatterInputChannelsToModelParallelRegion):
    _attach_symbolic(_cls)
del _cls


# ---------------------------------------------------------------------------
# public helpers
# ---------------------------------------------------------------------------


def copy_to_tensor_model_parallel_region(input_: torch.Tensor) -> torch.Tensor:
    return _CopyToModelParallelRegion.apply(input_)


def reduce_from_tensor_model_parallel_region(input_: torch.Tensor) -> torch.Tensor:
    return _ReduceFromModelParallelRegion.apply(input_)


def scatter_to_tensor_model_parallel_region(input_: torch.Tensor) -> torch.Tensor:
    return _ScatterToModelParallelRegion.apply(input_, -1)


def gather_from_tensor_model_parallel_region(input_: torch.Tensor) -> torch.Tensor:
    return _GatherFromModelParallelRegion.apply(input_, -1)


def scatter_to_sequence_parallel_region(input_: torch.Tensor) -> torch.Tensor:
    return _ScatterToSequenceParallelRegion.apply(input_, 0)


def gather_from_sequence_parallel_region(input_: torch.Tensor, to_model_parallel: bool = True) -> torch.Tensor:
    return _GatherFromSequenceParallelRegion.apply(input_, 0, to_model_parallel)


def reduce_scatter_to_sequence_parallel_region(input_: torch.Tensor) -> torch.Tensor:
    return _ReduceScatterToSequenceParallelRegion.apply(input_, 0)


def scatter_to_tensor_model_parallel_region_with_dim(input_: torch.Tensor, partition_dim: int) -> torch.Tensor:
    return _ScatterToModelParallelRegion.apply(input_, partition_dim)


def gather_from_tensor_model_parallel_region_with_dim(input_: torch.Tensor, gather_dim: int) -> torch.Tensor:
    return _GatherFromModelParallelRegion.apply(input_, gather_dim)


def reduce_scatter_to_tensor_model_parallel_region_with_dim(input_: torch.Tensor, partition_dim: int) -> torch.Tensor:
    return _ReduceScatterToSequenceParallelRegion.apply(input_, partition_dim)


def scatter_input_channels_to_tensor_model_parallel_region(input_: torch.Tensor) -> torch.Tensor:
    return _ScatterInputChannelsToModelParallelRegion.apply(input_)


def enter_expert_parallel_region(x: torch.Tensor, scatter_gather: bool = False) -> torch.Tensor:
    """(e, c, h) -> (e/ep, ep, c, h): dispatch tokens to the ranks owning each expert.

    Reference mappings.py:412-449.  With ``scatter_gather`` the capacity dim is
    first split over TP so each TP rank moves 1/tp of the bytes across EP, and
    re-gathered after the all-to-all (the reference's optional SP optimisation).
    """
    ep = ps.get_expert_model_parallel_size()
    if ep == 1:
        return x.unsqueeze(1)
    e, c, h = x.shape
    assert e % ep == 0, f"num experts {e} not divisible by ep {ep}"
    if scatter_gather:
        x = scatter_to_tensor_model_parallel_region_with_dim(x, 1)
    # split experts (dim 0) across EP ranks; concatenate arrivals on a new dim 1
    x = x.reshape(ep, e // ep, x.shape[1], h)  # [dst_rank, e_local, c, h]
    x = _AllToAllInExpertParallelRegion.apply(x, 0, 0)  # [src_rank, e_local, c, h]
    x = x.transpose(0, 1).contiguous()  # [e_local, src_rank(ep), c, h]
    if scatter_gather:
        x = gather_from_tensor_model_parallel_region_with_dim(x, 2)
    return x


def exit_expert_parallel_region(x: torch.Tensor, scatter_gather: bool = False) -> torch.Tensor:
    """(e/ep, ep, c, h) -> (e, c, h): inverse of ``enter_expert_parallel_region``
    (reference mappings.py:452-486)."""
    ep = ps.get_expert_model_parallel_size()
    if ep == 1:
        return x.squeeze(1)
    e_local, ep_, c, h = x.shape
    assert ep_ == ep
    if scatter_gather:
        x = scatter_to_tensor_model_parallel_region_with_dim(x, 2)
    x = x.transpose(0, 1).contiguous()  # [dst_rank, e_local, c, h]
    x = _AllToAllInExpertParallelRegion.apply(x, 0, 0)  # [owner_rank, e_local, c, h]
    x = x.reshape(ep * e_local, x.shape[2], h)
    if scatter_gather:
        x = gather_from_tensor_model_parallel_region_with_dim(x, 1)
    return x


def nonzero_partition_dim_swap(func):
    """Decorator for ``f(x, partition_dim)``: the wrapped function always sees the partition dim at position 0 (pre- and
    post-transpose), callers keep their own dim (reference ``mappings.py:24-39``)."""
    import functools

    @functools.wraps(func)
    def wrapped(x: torch.Tensor, partition_dim: int) -> torch.Tensor:
        xt = x.transpose(0, partition_dim) if partition_dim != 0 else x
        yt = func(xt, partition_dim=0)
        return yt.transpose(0, partition_dim) if partition_dim != 0 else yt

    return wrapped
