"""Gradient utilities: global grad-norm, clipping, bucketed DP all-reduce, SP grad all-reduce
(reference ``src/neuronx_distributed/parallel_layers/grads.py:29-329``).

Semantics kept: TP-duplicated parameters are counted once (only on tp rank 0, or summed on all
ranks and divided by tp when ``force_spmd``), expert-parallel parameters are reduced over the EP
group as well, PP-shared (tied) weights are counted only on the stage that owns them.

B200: the per-tensor ``torch.norm`` calls are replaced by the two-stage ``sumsq`` kernel
accumulating into one device scalar; clipping multiplies by a coefficient that stays on the
device (``torch.where``-free, no host sync), as the reference's branch-free form (``:222-240``).
"""
from __future__ import annotations

import os
from typing import Iterable, List, Union

import torch
import torch.distributed as dist

from ..ops import optim as optim_ops
from . import parallel_state as ps


def param_is_not_shared(param: torch.Tensor) -> bool:
    return not getattr(param, "shared", False)


def _all_reduce(t: torch.Tensor, group, size: int, op=dist.ReduceOp.SUM) -> None:
    if size > 1 and group is not None:
        dist.all_reduce(t, op=op, group=group)


def get_grad_norm(parameters: Union[Iterable[torch.Tensor], torch.Tensor], norm_type: float = 2.0,
                  zero1_optimizer: bool = False, zero1_optimizer_groups=None, force_spmd: bool = True
                  ) -> torch.Tensor:
    """Global gradient norm across TP / PP / EP (and ZeRO shard groups)."""
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    params = [p for p in parameters if p.grad is not None]
    if not params:
        return torch.zeros((), dtype=torch.float32)
    device = params[0].grad.device
    tp = ps.get_tensor_model_parallel_size()
    tp_rank = ps.get_tensor_model_parallel_rank()
    inf_norm = norm_type == float("inf")

    def accumulate(group_params: List[torch.Tensor]) -> torch.Tensor:
        total = torch.zeros(1, dtype=torch.float32, device=device)
        dup = torch.zeros(1, dtype=torch.float32, device=device)
        for p in group_params:
            if not param_is_not_shared(p):
                continue
            g = p.grad.detach()
            is_tp = getattr(p, "tensor_model_parallel", False)
            if inf_norm:
                v = g.abs().max().float().reshape(1)
                if is_tp or tp_rank == 0 or force_spmd:
                    total = torch.maximum(total, v)
                continue
            if norm_type == 2.0:
                tgt = total if is_tp else dup
                optim_ops.sumsq_(g.contiguous(), tgt, accumulate=True)
            else:
                v = g.float().abs().pow(norm_type).sum().reshape(1)
                if is_tp:
                    total += v
                else:
                    dup += v
        if not inf_norm:
            # duplicated (non-TP) params: every TP rank holds the same grad -> count once
            total = total + (dup / tp if force_spmd else (dup if tp_rank == 0 else 0.0 * dup))
        return total

    ep_params = [p for p in params if getattr(p, "expert_model_parallel", False)]
    non_ep = [p for p in params if not getattr(p, "expert_model_parallel", False)]
    total = accumulate(non_ep)
    op = dist.ReduceOp.MAX if inf_norm else dist.ReduceOp.SUM
    if ep_params:
        ep_total = accumulate(ep_params)
        _all_reduce(ep_total, ps.get_expert_model_parallel_group() if ps.get_expert_model_parallel_size() > 1 else None,
                    ps.get_expert_model_parallel_size(), op)
        total = torch.maximum(total, ep_total) if inf_norm else total + ep_total
    if zero1_optimizer and zero1_optimizer_groups is not None:
        _all_reduce(total, zero1_optimizer_groups, dist.get_world_size(zero1_optimizer_groups) if
                    dist.is_initialized() else 1, op)
    if ps.get_tensor_model_parallel_size() > 1:
        _all_reduce(total, ps.get_tensor_model_parallel_group(), tp, op)
    if ps.get_pipeline_model_parallel_size() > 1:
        _all_reduce(total, ps.get_pipeline_model_parallel_group(), ps.get_pipeline_model_parallel_size(), op)
    if inf_norm:
        return total.reshape(())
    return total.pow(1.0 / norm_type).reshape(())


def clip_grads_with_norm(parameters, max_norm: float, total_norm: torch.Tensor) -> None:
    """Scale grads by ``min(1, max_norm / (norm + 1e-6))`` without a host sync."""
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    coef = torch.clamp(max_norm / (total_norm + 1e-6), max=1.0)
    for p in parameters:
        if p.grad is not None:
            p.grad.detach().mul_(coef.to(device=p.grad.device, dtype=p.grad.dtype))


def clip_grad_norm(parameters, max_norm: float, norm_type: float = 2.0, zero1_optimizer: bool = False,
                   zero1_optimizer_groups=None, force_spmd: bool = True) -> torch.Tensor:
    """Megatron-style global-norm clipping (reference grads.py:180-219).  Returns the pre-clip norm."""
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    parameters = list(parameters)
    total_norm = get_grad_norm(parameters, norm_type, zero1_optimizer, zero1_optimizer_groups, force_spmd)
    clip_grads_with_norm(parameters, max_norm, total_norm)
    return total_norm


def bucket_allreduce_gradients(grads: List[torch.Tensor], reduce_over_ep_group: bool = False) -> None:
    """Average grads over the (expert-)data-parallel group in flat per-dtype buckets of at most
    ``ALLREDUCE_BUCKET_CAP_MB`` (default 512), last-produced grads first (reference :243-310)."""
    # Scaling as the reference (grads.py:243-310): the FIRST pass sums every gradient -- dense and expert -- over the
    # expert-data-parallel group and divides by the FULL data-parallel size; the second pass (dense gradients only)
    # sums over the EP group with divisor 1.  Dense: sum over exp_dp x ep = dp ranks, / dp.  Expert: the token
    # all-to-all already brought every EP rank's tokens to the expert's owner, so its gradient is summed over exp_dp
    # only and still divided by dp.
    if reduce_over_ep_group:
        size = 1
        group = ps.get_expert_model_parallel_group() if ps.get_expert_model_parallel_size() > 1 else None
    else:
        size = ps.get_data_parallel_size()
        group = ps.get_expert_data_parallel_group() if ps.get_expert_data_parallel_size() > 1 else None
    if size == 1 and group is None:
        return
    cap = int(os.environ.get("ALLREDUCE_BUCKET_CAP_MB", 512)) * 1024 * 1024
    by_dtype = {}
    for g in reversed(grads):
        by_dtype.setdefault(g.dtype, []).append(g)
    for _, gs in by_dtype.items():
        bucket, nbytes = [], 0
        def flush():
            if not bucket:
                return
            flat = torch.cat([g.reshape(-1) for g in bucket])
            if size != 1:
                flat.div_(size)
            if group is not None:
                dist.all_reduce(flat, group=group)
            off = 0
            for g in bucket:
                g.copy_(flat[off: off + g.numel()].view_as(g))
                off += g.numel()
        for g in gs:
            sz = g.numel() * g.element_size()
            if bucket and nbytes + sz > cap:
                flush()
                bucket, nbytes = [], 0
            bucket.append(g)
            nbytes += sz
        flush()


def allreduce_sequence_parallel_gradients(optimizer_or_params) -> None:
    """All-reduce over TP the grads of params tagged ``sequence_parallel_enabled`` (norm weights,
    RowParallel bias): under SP each rank saw only its sequence slice (reference :313-329).
    One flat buffer instead of ~2L+1 tiny collectives."""
    if ps.get_tensor_model_parallel_size() == 1:
        return
    if hasattr(optimizer_or_params, "param_groups"):
        params = [p for g in optimizer_or_params.param_groups for p in g["params"]]
    else:
        params = list(optimizer_or_params)
    grads = []
    for p in params:
        if getattr(p, "sequence_parallel_enabled", False):
            g = getattr(p, "main_grad", None) if p.grad is None else p.grad
            if g is not None:
                grads.append(g)
    if not grads:
        return
    by_dtype = {}
    for g in grads:
        by_dtype.setdefault(g.dtype, []).append(g)
    for gs in by_dtype.values():
        flat = torch.cat([g.reshape(-1) for g in gs])
        dist.all_reduce(flat, group=ps.get_tensor_model_parallel_group())
        off = 0
        for g in gs:
            g.copy_(flat[off: off + g.numel()].view_as(g))
            off += g.numel()
