"""``GQAQKVColumnParallelLinear``: Q/K/V projection for grouped-query attention when the number of
KV heads is smaller than the TP degree.

Reference: ``src/neuronx_distributed/modules/qkv_linear.py:34-772``.

KV heads are replicated ``kv_size_multiplier`` times so that every TP rank owns (a replica of) at
least one KV head.  The K/V master weight ``[kv_heads*D, H]`` is tiled ``m`` times along dim 0 and
then sharded over TP, so rank ``r`` holds KV head ``r % (tp/m)`` and the ranks
``{j, j + tp/m, ...}`` ("KV-shared group") hold replicas of the same head.

Backward keeps the replicas identical: dK/dV (output grads) are all-reduced over the KV-shared
group before the weight-grad GEMM, and their contribution to the input grad is scaled by ``1/m``
because the following TP reduction would otherwise count it ``m`` times (reference ``:250-273``).

B200: one AG->GEMM fused kernel for the first (or the fused) projection; the gathered input is a
by-product reused by the other projections and by wgrad.
"""
from __future__ import annotations

import math
from typing import Callable, Optional

import torch
import torch.distributed as dist
import torch.nn.init as init
from torch.nn.parameter import Parameter

from ..ops import comm_linear
from ..ops import linear as gemm_ops
from ..parallel import mappings
from ..parallel import parallel_state as ps
from ..parallel.layers import _accumulate_wgrad, _flatten, _initialize_affine_weight, _stash_gathered
from ..parallel.utils import cast_if_autocast_enabled, divide, set_tensor_model_parallel_attributes


def get_kv_shared_group(kv_size_multiplier: int, as_list: bool = False):
    return ps.get_kv_shared_group(kv_size_multiplier, as_list=as_list)


def _init_replicated_cpu(param: torch.Tensor, full_out: int, in_size: int, multiplier: int, init_method: Callable,
                         stride: int = 1) -> Optional[torch.Tensor]:
    """Master weight ``[full_out, in]`` -> tiled ``multiplier`` times along dim 0 -> this rank's slice."""
    tp, rank = ps.get_tensor_model_parallel_size(), ps.get_tensor_model_parallel_rank()
    master = torch.empty(full_out, in_size, dtype=torch.float32)
    init_method(master)
    tiled = master.repeat(multiplier, 1).to(param.dtype)
    per = tiled.shape[0] // tp
    with torch.no_grad():
        param.copy_(tiled[rank * per:(rank + 1) * per])
    return master


class GQAQKVLinearWithAsyncCommunication(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input, weight_q, weight_k, weight_v, bias_q, bias_k, bias_v, async_grad_allreduce,
                sequence_parallel_enabled, kv_size_multiplier, weight_qkv=None, bias_qkv=None, fuse_qkv=False,
                output_sizes=None):
        ctx.async_grad_allreduce = async_grad_allreduce
        ctx.sequence_parallel_enabled = sequence_parallel_enabled
        ctx.kv_size_multiplier = kv_size_multiplier
        ctx.fuse_qkv = fuse_qkv
        ctx.in_shape = input.shape
        ctx.use_bias = (bias_qkv is not None) if fuse_qkv else (bias_q is not None)
        first_w = weight_qkv if fuse_qkv else weight_q
        ctx.compute_weight_gradient = first_w.requires_grad
        x2d = _flatten(input)
        if sequence_parallel_enabled:
            first, total = comm_linear.ag_gemm(x2d, first_w, layout="nt")
            lead = (input.shape[0] * ps.get_tensor_model_parallel_size(),) + tuple(input.shape[1:-1])
        else:
            total = x2d
            first = gemm_ops.gemm_nt(total, first_w)
            lead = tuple(input.shape[:-1])
        if fuse_qkv:
            if bias_qkv is not None:
                first = first + bias_qkv
            q, k, v = torch.split(first, output_sizes, dim=-1)
            ctx.output_sizes = output_sizes
            saved_w = (weight_qkv,)
        else:
            q = first
            k = gemm_ops.gemm_nt(total, weight_k)
            v = gemm_ops.gemm_nt(total, weight_v)
            if bias_q is not None:
                q, k, v = q + bias_q, k + bias_k, v + bias_v
            saved_w = (weight_q, weight_k, weight_v)
        if ctx.compute_weight_gradient:
            if sequence_parallel_enabled and not _stash_gathered():
                ctx.save_for_backward(x2d, *saved_w)
                ctx.stashed = False
            else:
                ctx.save_for_backward(total, *saved_w)
                ctx.stashed = True
        else:
            ctx.save_for_backward(*saved_w)
        return (q.reshape(lead + (q.shape[-1],)), k.reshape(lead + (k.shape[-1],)),
                v.reshape(lead + (v.shape[-1],)))

    @staticmethod
    def backward(ctx, grad_q, grad_k, grad_v):
        m = ctx.kv_size_multiplier
        if ctx.compute_weight_gradient:
            saved_input, *weights = ctx.saved_tensors
        else:
            weights = list(ctx.saved_tensors)
            saved_input = None
        gq, gk, gv = (_flatten(g).contiguous() for g in (grad_q, grad_k, grad_v))
        if m > 1:
            group = get_kv_shared_group(m)
            dist.all_reduce(gk, group=group)
            dist.all_reduce(gv, group=group)
        sp = ctx.sequence_parallel_enabled
        if ctx.fuse_qkv:
            (w_qkv,) = weights
            g_scaled = torch.cat([gq, gk / m, gv / m], dim=-1) if m > 1 else torch.cat([gq, gk, gv], dim=-1)
            if sp:
                grad_input2d = comm_linear.gemm_rs(g_scaled, w_qkv, layout="nn")
            else:
                grad_input2d = gemm_ops.gemm_nn(g_scaled, w_qkv)
        else:
            w_q, w_k, w_v = weights
            gi = gemm_ops.gemm_nn(gq, w_q)
            gkv = gemm_ops.gemm_nn(gk, w_k) + gemm_ops.gemm_nn(gv, w_v)
            gi = gi + (gkv / m if m > 1 else gkv)
            grad_input2d = comm_linear.reduce_scatter_rows(gi) if sp else gi
        handle = None
        if (not sp) and ctx.async_grad_allreduce:
            handle = comm_linear.all_reduce_async(grad_input2d)
        grads_w = [None, None, None]
        grads_b = [None, None, None]
        grad_w_qkv = grad_b_qkv = None
        if ctx.compute_weight_gradient:
            if sp and not ctx.stashed:
                total = comm_linear.all_gather_rows(saved_input)
            else:
                total = saved_input
            if ctx.fuse_qkv:
                g_full = torch.cat([gq, gk, gv], dim=-1)
                grad_w_qkv = _accumulate_wgrad(weights[0], g_full, total)
                if ctx.use_bias:
                    grad_b_qkv = g_full.sum(0)
            else:
                for i, (g, w) in enumerate(zip((gq, gk, gv), weights)):
                    grads_w[i] = _accumulate_wgrad(w, g, total)
                    if ctx.use_bias:
                        grads_b[i] = g.sum(0)
        if handle is not None:
            handle.wait()
        in_shape = ctx.in_shape
        if sp:
            in_shape = (in_shape[0],) + tuple(in_shape[1:])
        return (grad_input2d.reshape(in_shape), grads_w[0], grads_w[1], grads_w[2], grads_b[0], grads_b[1],
                grads_b[2], None, None, None, grad_w_qkv, grad_b_qkv, None, None)


class FusedQKVLinearWithAsyncCommunication(torch.autograd.Function):
    """Single-weight variant: ``qkv = X W_qkv^T`` with ``W_qkv = [W_q; W_k; W_v]`` (local shards).
    Returns the fused ``[..., q + 2 kv]`` tensor so RoPE can run in place on it and attention can
    read q/k/v as strided views (no split/concat copies)."""

    @staticmethod
    def forward(ctx, input, weight_qkv, bias_qkv, async_grad_allreduce, sequence_parallel_enabled,
                kv_size_multiplier, output_sizes):
        ctx.async_grad_allreduce = async_grad_allreduce
        ctx.sequence_parallel_enabled = sequence_parallel_enabled
        ctx.kv_size_multiplier = kv_size_multiplier
        ctx.output_sizes = output_sizes
        ctx.in_shape = input.shape
        ctx.use_bias = bias_qkv is not None
        ctx.compute_weight_gradient = weight_qkv.requires_grad
        x2d = _flatten(input)
        if sequence_parallel_enabled:
            out, total = comm_linear.ag_gemm(x2d, weight_qkv, layout="nt")
            lead = (input.shape[0] * ps.get_tensor_model_parallel_size(),) + tuple(input.shape[1:-1])
        else:
            total = x2d
            out = gemm_ops.gemm_nt(total, weight_qkv)
            lead = tuple(input.shape[:-1])
        if bias_qkv is not None:
            out = out + bias_qkv
        if ctx.compute_weight_gradient:
            if sequence_parallel_enabled and not _stash_gathered():
                ctx.save_for_backward(x2d, weight_qkv)
                ctx.stashed = False
            else:
                ctx.save_for_backward(total, weight_qkv)
                ctx.stashed = True
        else:
            ctx.save_for_backward(weight_qkv)
        return out.reshape(lead + (out.shape[-1],))

    @staticmethod
    def backward(ctx, grad_qkv):
        m = ctx.kv_size_multiplier
        if ctx.compute_weight_gradient:
            saved_input, w_qkv = ctx.saved_tensors
        else:
            (w_qkv,) = ctx.saved_tensors
            saved_input = None
        g = _flatten(grad_qkv)
        if not g.is_contiguous():
            g = g.contiguous()
        sp = ctx.sequence_parallel_enabled
        g_for_input = g
        if m > 1:
            q_sz, kv_sz, _ = ctx.output_sizes
            group = get_kv_shared_group(m)
            gkv = g[:, q_sz:].contiguous()
            dist.all_reduce(gkv, group=group)
            g = torch.cat([g[:, :q_sz], gkv], dim=-1)
            g_for_input = torch.cat([g[:, :q_sz], gkv / m], dim=-1)
        if sp:
            grad_input2d = comm_linear.gemm_rs(g_for_input, w_qkv, layout="nn")
        else:
            grad_input2d = gemm_ops.gemm_nn(g_for_input, w_qkv)
        handle = None
        if (not sp) and ctx.async_grad_allreduce:
            handle = comm_linear.all_reduce_async(grad_input2d)
        grad_w = grad_b = None
        if ctx.compute_weight_gradient:
            total = comm_linear.all_gather_rows(saved_input) if (sp and not ctx.stashed) else saved_input
            grad_w = _accumulate_wgrad(w_qkv, g, total)
            if ctx.use_bias:
                grad_b = g.sum(0)
        if handle is not None:
            handle.wait()
        return grad_input2d.reshape(ctx.in_shape), grad_w, grad_b, None, None, None, None


def gqa_qkv_linear_with_async_allreduce(input, weight_q, weight_k, weight_v, bias_q, bias_k, bias_v,
                                        async_grad_allreduce, sequence_parallel_enabled, kv_size_multiplier=1,
                                        weight_qkv=None, bias_qkv=None, fuse_qkv=False, output_sizes=None):
    args = cast_if_autocast_enabled(input, weight_q, weight_k, weight_v, bias_q, bias_k, bias_v)
    wqkv, bqkv = cast_if_autocast_enabled(weight_qkv, bias_qkv)
    with torch.autocast(device_type=input.device.type, enabled=False):
        return GQAQKVLinearWithAsyncCommunication.apply(*args, async_grad_allreduce, sequence_parallel_enabled,
                                                        kv_size_multiplier, wqkv, bqkv, fuse_qkv, output_sizes)


class GQAQKVColumnParallelLinear(torch.nn.Module):
    """Args mirror the reference (``qkv_linear.py:454-552``): ``output_sizes = [q_out, kv_out]``
    are the *global* (unsharded, un-replicated) output sizes."""

    def __init__(self, input_size: int, output_sizes, bias: bool = True, gather_output: bool = True,
                 dtype: torch.dtype = torch.float32, device=None, init_method: Optional[Callable] = None,
                 sequence_parallel_enabled: bool = False, keep_master_weight: bool = False,
                 kv_size_multiplier: int = 1, fuse_qkv: bool = False):
        super().__init__()
        self.input_size = input_size
        self.output_sizes = list(output_sizes)
        self.gather_output = gather_output
        self.kv_size_multiplier = kv_size_multiplier
        self.fuse_qkv = fuse_qkv
        self.dtype = dtype
        self.use_bias = bias
        self.keep_master_weight = keep_master_weight
        self.init_method = init_method or (lambda w: init.kaiming_uniform_(w, a=math.sqrt(5)))
        tp = ps.get_tensor_model_parallel_size()
        self.sequence_parallel_enabled = sequence_parallel_enabled and tp > 1
        self.async_tensor_model_parallel_allreduce = (not self.sequence_parallel_enabled) and tp > 1
        self.q_output_size_per_partition = divide(self.output_sizes[0], tp)
        self.kv_output_size_per_partition = divide(self.output_sizes[1] * kv_size_multiplier, tp)
        if tp > 1 and kv_size_multiplier > 1:
            get_kv_shared_group(kv_size_multiplier)  # create the groups collectively at construction
        q_sz, kv_sz = self.q_output_size_per_partition, self.kv_output_size_per_partition
        if fuse_qkv:
            self.weight_qkv = Parameter(torch.empty(q_sz + 2 * kv_sz, input_size, dtype=dtype, device=device))
            self._init_fused()
            if bias:
                self.bias_qkv = Parameter(torch.zeros(q_sz + 2 * kv_sz, dtype=dtype, device=device))
                set_tensor_model_parallel_attributes(self.bias_qkv, True, 0, 1)
            else:
                self.register_parameter("bias_qkv", None)
        else:
            self.weight_q = Parameter(torch.empty(q_sz, input_size, dtype=dtype, device=device))
            self.weight_k = Parameter(torch.empty(kv_sz, input_size, dtype=dtype, device=device))
            self.weight_v = Parameter(torch.empty(kv_sz, input_size, dtype=dtype, device=device))
            self._init_split()
            if bias:
                self.bias_q = Parameter(torch.zeros(q_sz, dtype=dtype, device=device))
                self.bias_k = Parameter(torch.zeros(kv_sz, dtype=dtype, device=device))
                self.bias_v = Parameter(torch.zeros(kv_sz, dtype=dtype, device=device))
                for b in (self.bias_q, self.bias_k, self.bias_v):
                    set_tensor_model_parallel_attributes(b, True, 0, 1)
            else:
                for n in ("bias_q", "bias_k", "bias_v"):
                    self.register_parameter(n, None)

    # -- init ----------------------------------------------------------------------------------
    def _tag(self, w, fused: bool = False):
        if not hasattr(w, "tensor_model_parallel"):
            set_tensor_model_parallel_attributes(w, True, 0, 1)
        if fused:
            setattr(w, "fused_qkv", True)
            setattr(w, "num_attention_heads", None)
        setattr(w, "kv_size_multiplier", self.kv_size_multiplier)

    def _init_split(self):
        ws = (self.weight_q, self.weight_k, self.weight_v)
        outs = (self.output_sizes[0], self.output_sizes[1], self.output_sizes[1])
        mults = (1, self.kv_size_multiplier, self.kv_size_multiplier)
        for w, out, m in zip(ws, outs, mults):
            self._tag(w)
            if w.device.type == "meta":
                continue
            if w.device.type == "cpu" or m > 1:
                tmp = torch.empty(w.shape, dtype=w.dtype)
                _init_replicated_cpu(tmp, out, self.input_size, m, self.init_method)
                with torch.no_grad():
                    w.copy_(tmp)
            else:
                for tag in ("tensor_model_parallel", "partition_dim", "partition_stride"):
                    delattr(w, tag)
                _initialize_affine_weight(w, self.init_method, 0, 1)

    def _init_fused(self):
        w = self.weight_qkv
        self._tag(w, fused=True)
        if w.device.type == "meta":
            return
        q_sz, kv_sz = self.q_output_size_per_partition, self.kv_output_size_per_partition
        parts = []
        for out, m, sz in ((self.output_sizes[0], 1, q_sz), (self.output_sizes[1], self.kv_size_multiplier, kv_sz),
                           (self.output_sizes[1], self.kv_size_multiplier, kv_sz)):
            tmp = torch.empty(sz, self.input_size, dtype=w.dtype)
            _init_replicated_cpu(tmp, out, self.input_size, m, self.init_method)
            parts.append(tmp)
        with torch.no_grad():
            w.copy_(torch.cat(parts, dim=0))

    def initialize_weight_biases(self) -> None:
        """(Re-)draw the weights with ``init_method`` and zero the biases -- the reference's post-construction init
        entry point (``qkv_linear.py:590-650``); the constructor already ran it once."""
        if self.fuse_qkv:
            self._init_fused()
            biases = (self.bias_qkv,)
        else:
            self._init_split()
            biases = (self.bias_q, self.bias_k, self.bias_v)
        with torch.no_grad():
            for b in biases:
                if b is not None and b.device.type != "meta":
                    b.zero_()

    # -- forward -------------------------------------------------------------------------------
    def forward(self, input: torch.Tensor, return_fused: bool = False):
        if self.async_tensor_model_parallel_allreduce or self.sequence_parallel_enabled:
            x = input
        else:
            x = mappings.copy_to_tensor_model_parallel_region(input)
        q_sz, kv_sz = self.q_output_size_per_partition, self.kv_output_size_per_partition
        if self.fuse_qkv:
            xc, wc, bc = cast_if_autocast_enabled(x, self.weight_qkv, self.bias_qkv)
            with torch.autocast(device_type=x.device.type, enabled=False):
                qkv = FusedQKVLinearWithAsyncCommunication.apply(
                    xc, wc, bc, self.async_tensor_model_parallel_allreduce, self.sequence_parallel_enabled,
                    self.kv_size_multiplier, [q_sz, kv_sz, kv_sz])
            if return_fused:
                assert not self.gather_output
                return qkv
            q, k, v = torch.split(qkv, [q_sz, kv_sz, kv_sz], dim=-1)
        else:
            q, k, v = gqa_qkv_linear_with_async_allreduce(
                x, self.weight_q, self.weight_k, self.weight_v, self.bias_q, self.bias_k, self.bias_v,
                self.async_tensor_model_parallel_allreduce, self.sequence_parallel_enabled,
                self.kv_size_multiplier)
        if self.gather_output:
            assert not self.sequence_parallel_enabled
            q = mappings.gather_from_tensor_model_parallel_region(q)
            k = mappings.gather_from_tensor_model_parallel_region(k)
            v = mappings.gather_from_tensor_model_parallel_region(v)
        return q, k, v


def destroy_kv_group() -> None:
    """Forget the cached KV-sharing groups (reference ``qkv_linear.py:81-85``); they live in ``parallel_state`` here."""
    ps.destroy_kv_shared_groups() if hasattr(ps, "destroy_kv_shared_groups") else None


def check_requires_grad(weight_qkv, fuse_qkv: bool, weight_q) -> bool:
    return bool(weight_qkv.requires_grad if fuse_qkv else weight_q.requires_grad)


def check_use_bias(weight_qkv, fuse_qkv: bool, weight_q, bias_q, bias_qkv) -> bool:
    has_bias = (bias_qkv is not None) if fuse_qkv else (bias_q is not None)
    return has_bias and check_requires_grad(weight_qkv, fuse_qkv, weight_q)
