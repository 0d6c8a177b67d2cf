"""Collective + GEMM pairs used by the tensor-parallel layers.

Every function works on 2-D row-major views where the sequence-parallel dim is
the row dim (``[S, B, H]`` flattens to ``[S*B, H]`` with a rank's shard being a
contiguous row block), so all-gather / reduce-scatter on dim 0 are plain row-block
concatenation / partition.

Dispatch (per call):
  1. a symmetric-memory context exists for the TP group and the fused kernels
     are enabled (``NXD_FUSED_COMM=1``, default on GPU)  ->  ONE sm_100a kernel
     from ``csrc/fused_comm_gemm.cu`` that moves slabs over NVLink peer memory
     while tcgen05 tiles are computed;
  2. otherwise NCCL (or gloo on CPU) collective + the GEMM from ``ops.linear``
     -- this is also the *baseline arm* measured in ``bench.py``.
"""
from __future__ import annotations

import os
from typing import Tuple

import torch
import torch.distributed as dist

from ..parallel import parallel_state as ps
from . import backend
from . import linear as gemm_ops


def _group(group):
    if group is not None:
        return group
    if ps.get_tensor_model_parallel_size() == 1:
        return None
    return ps.get_tensor_model_parallel_group()


def group_size(group=None) -> int:
    g = _group(group)
    return 1 if g is None else dist.get_world_size(group=g)


def _fused_ctx(x: torch.Tensor, group):
    """Symmetric-memory context for the fused kernels, or None."""
    if not x.is_cuda or group is not None:
        return None
    if not backend.fused_comm_enabled():
        return None
    if x.dtype != torch.bfloat16:
        return None
    return ps.get_symm_mem_context("tp")


def all_gather_rows(x2d: torch.Tensor, group=None, **_) -> torch.Tensor:
    g = _group(group)
    if g is None:
        return x2d
    world = dist.get_world_size(group=g)
    x2d = x2d.contiguous()
    out = torch.empty((world * x2d.shape[0], x2d.shape[1]), dtype=x2d.dtype, device=x2d.device)
    dist.all_gather_into_tensor(out, x2d, group=g)
    return out


def reduce_scatter_rows(x2d: torch.Tensor, group=None) -> torch.Tensor:
    g = _group(group)
    if g is None:
        return x2d
    world = dist.get_world_size(group=g)
    x2d = x2d.contiguous()
    out = torch.empty((x2d.shape[0] // world, x2d.shape[1]), dtype=x2d.dtype, device=x2d.device)
    dist.reduce_scatter_tensor(out, x2d, group=g)
    return out


def _gemm(a, b, layout):
    return gemm_ops.gemm_nt(a, b) if layout == "nt" else gemm_ops.gemm_nn(a, b)


def ag_gemm(x_shard2d: torch.Tensor, weight: torch.Tensor, layout: str = "nt", group=None,
            persistent_total: bool = True, **_) -> Tuple[torch.Tensor, torch.Tensor]:
    """``all_gather_rows(x) @ W^T`` (``layout="nt"``) or ``@ W`` (``"nn"``).
    Returns ``(y, x_total)`` -- the gathered operand is a by-product of the fused
    kernel and is reused by wgrad."""
    g = _group(group)
    if g is None:
        return _gemm(x_shard2d, weight, layout), x_shard2d
    ctx = _fused_ctx(x_shard2d, group)
    if ctx is not None:
        from .fused_comm import fused_ag_gemm

        res = fused_ag_gemm(ctx, x_shard2d, weight, layout, persistent_total=persistent_total)
        if res is not None:
            return res
        backend.note_fallback("ag_gemm", f"shard {tuple(x_shard2d.shape)} x weight {tuple(weight.shape)} {layout} -> NCCL + GEMM")
    total = all_gather_rows(x_shard2d, group=g)
    return _gemm(total, weight, layout), total


def gemm_rs(x2d: torch.Tensor, weight: torch.Tensor, layout: str = "nt", group=None, **_) -> torch.Tensor:
    """``reduce_scatter_rows(x @ W^T)`` / ``(x @ W)``."""
    g = _group(group)
    if g is None:
        return _gemm(x2d, weight, layout)
    ctx = _fused_ctx(x2d, group)
    if ctx is not None:
        from .fused_comm import fused_gemm_rs

        res = fused_gemm_rs(ctx, x2d, weight, layout)
        if res is not None:
            return res
        backend.note_fallback("gemm_rs", f"x {tuple(x2d.shape)} x weight {tuple(weight.shape)} {layout} -> GEMM + NCCL")
    return reduce_scatter_rows(_gemm(x2d, weight, layout), group=g)


def gemm_ar(x2d: torch.Tensor, weight: torch.Tensor, layout: str = "nt", group=None) -> torch.Tensor:
    """``all_reduce(x @ W^T)`` (RowParallel without SP: the decode path)."""
    g = _group(group)
    if g is None:
        return _gemm(x2d, weight, layout)
    ctx = _fused_ctx(x2d, group)
    if ctx is not None:
        from .fused_comm import fused_gemm_ar

        res = fused_gemm_ar(ctx, x2d, weight, layout)
        if res is not None:
            return res
        if x2d.shape[0] * (weight.shape[0] if layout == "nt" else weight.shape[1]) <= (1 << 20):
            # beyond 2 MB NCCL is the intended path (ring / NVLS wins on bandwidth), below it is a fallback
            backend.note_fallback("gemm_ar", f"x {tuple(x2d.shape)} x weight {tuple(weight.shape)} {layout} -> GEMM + NCCL")
    y = _gemm(x2d, weight, layout)
    dist.all_reduce(y, group=g)
    return y


def all_reduce_small(t: torch.Tensor, group=None) -> torch.Tensor:
    """Sum ``t`` over the TP group (no autograd): one-shot NVLink kernel for messages up to 2 MB, NCCL beyond."""
    g = _group(group)
    if g is None:
        return t
    ctx = _fused_ctx(t, group)
    if ctx is not None and t.is_contiguous():
        from .fused_comm import oneshot_all_reduce

        res = oneshot_all_reduce(ctx, t)
        if res is not None:
            return res
    t = t.contiguous()
    dist.all_reduce(t, group=g)
    return t


class _Done:
    def wait(self):
        return None


def all_reduce_async(t: torch.Tensor, group=None):
    """Kick an all-reduce that overlaps with whatever is enqueued next
    (NCCL runs it on its own stream; ``handle.wait()`` joins)."""
    g = _group(group)
    if g is None:
        return _Done()
    assert t.is_contiguous()
    return dist.all_reduce(t, group=g, async_op=True)
