"""``Experts``: the expert MLPs over expert-aligned tokens ``(E, C, H)`` (reference
``src/neuronx_distributed/modules/moe/experts.py:17-163``).

With expert parallelism the ``(E, C, H)`` batch enters the EP region through an all-to-all
(``(E, C, H) -> (E/ep, ep*C, H)``: every rank keeps its local experts and receives their tokens from all EP peers),
runs the local experts, and leaves through the inverse all-to-all.
"""
from __future__ import annotations

from typing import Callable, Optional

import torch

from ...parallel import mappings
from ...parallel import parallel_state as ps
from .moe_parallel_layers import ExpertFusedColumnParallelLinear, ExpertFusedRowParallelLinear


def _expert_wgrad(x: torch.Tensor, g: torch.Tensor) -> torch.Tensor:
    """``dW[e] = sum_s x[s, e]^T g[s, e]`` for ``x [ep, E_loc, C, a]``, ``g [ep, E_loc, C, b]``."""
    from .moe_parallel_layers import _grouped_mm

    ep, el, c, a = x.shape
    xe = x.transpose(0, 1).reshape(el, ep * c, a)
    ge = g.transpose(0, 1).reshape(el, ep * c, g.shape[-1])
    return _grouped_mm(xe, ge, "tn")


class _DispatchGemm(torch.autograd.Function):
    """Expert-parallel dispatch all-to-all fused with the first expert GEMM (``nxd_a2a_gemm_bf16``):
    ``x [ep(dst), E_loc, C, H]`` -> ``y [ep(src), E_loc, C, N]``.  Backward: dgrad GEMM fused with the inverse
    all-to-all (``nxd_gemm_a2a_bf16``); the received tokens, a by-product of the forward, feed the weight gradient."""

    @staticmethod
    def forward(ctx, x, weight, sym):
        from ...ops.fused_comm import fused_a2a_gemm

        y, x_total = fused_a2a_gemm(sym, x, weight, "nn")
        ctx.save_for_backward(x_total, weight)
        ctx.sym = sym
        return y

    @staticmethod
    def backward(ctx, dy):
        from ...ops.fused_comm import fused_gemm_a2a

        x_total, weight = ctx.saved_tensors
        dy = dy.contiguous()
        dx = fused_gemm_a2a(ctx.sym, dy, weight, "nt")                       # [E, C, H], back on the token owners
        dx = dx.reshape(x_total.shape)       # partial over TP: the copy_to_tensor_model_parallel_region before sums it
        dw = _expert_wgrad(x_total, dy) if weight.requires_grad else None
        return dx, dw, None


class _GemmCombine(torch.autograd.Function):
    """Second expert GEMM fused with the combine all-to-all (``nxd_gemm_a2a_bf16``): ``x [ep(src), E_loc, C, I]`` ->
    ``[E, C, H]`` on the rank the tokens came from.  Backward: dispatch of dOut fused with the dgrad GEMM."""

    @staticmethod
    def forward(ctx, x, weight, sym):
        from ...ops.fused_comm import fused_gemm_a2a

        out = fused_gemm_a2a(sym, x, weight, "nn")
        ctx.save_for_backward(x, weight)
        ctx.sym = sym
        return out

    @staticmethod
    def backward(ctx, dout):
        from ...ops.fused_comm import fused_a2a_gemm

        x, weight = ctx.saved_tensors
        ep, el, c, _ = x.shape
        dx, dout_total = fused_a2a_gemm(ctx.sym, dout.contiguous().reshape(ep, el, c, dout.shape[-1]), weight, "nt")
        dw = _expert_wgrad(x, dout_total) if weight.requires_grad else None
        return dx, dw, None


class Experts(torch.nn.Module):
    def __init__(self, num_experts: int, hidden_size: int, intermediate_size: int, glu: bool,
                 activation_fn: Callable[[torch.Tensor], torch.Tensor], dtype: torch.dtype = torch.float32, device=None,
                 input_layer_init_method=None, output_layer_init_method=None):
        super().__init__()
        self.num_experts = num_experts
        self.glu = glu
        self.activation_fn = activation_fn
        self.intermediate_size = intermediate_size
        ep = ps.get_expert_model_parallel_size()
        if glu:  # gate | up fused on the output dim, sharded with stride 2 like the dense MLP
            self.gate_up_proj = ExpertFusedColumnParallelLinear(num_experts, hidden_size, 2 * intermediate_size,
                                                                dtype=dtype, device=device, stride=2,
                                                                init_method=input_layer_init_method)
        else:
            self.up_proj = ExpertFusedColumnParallelLinear(num_experts, hidden_size, intermediate_size, dtype=dtype,
                                                           device=device, init_method=input_layer_init_method)
        # with EP the TP reduction is deferred to the MoE exit (reference ``reduce_output = ep > 1``)
        self.down_proj = ExpertFusedRowParallelLinear(num_experts, intermediate_size, hidden_size,
                                                      reduce_output=ep > 1, dtype=dtype, device=device,
                                                      init_method=output_layer_init_method)

    def forward(self, hidden_states: torch.Tensor, expert_indices: Optional[torch.Tensor] = None) -> torch.Tensor:
        ep = ps.get_expert_model_parallel_size()
        if ep > 1:
            if expert_indices is not None:
                raise NotImplementedError("selective loading is not supported with expert parallelism")
            sym = self._fused_context(hidden_states)
            if sym is not None:
                return self._forward_fused(hidden_states, sym)
            hidden_states = mappings.enter_expert_parallel_region(hidden_states)   # (E/ep, ep, C, H)
        if self.glu:
            gu = self.gate_up_proj(hidden_states, expert_indices)
            gate, up = torch.chunk(gu, 2, dim=-1)
            inter = self.activation_fn(gate) * up
        else:
            inter = self.activation_fn(self.up_proj(hidden_states, expert_indices))
        out = self.down_proj(inter, expert_indices)
        if ep > 1:
            out = mappings.exit_expert_parallel_region(out)                        # back to (E, C, H)
        return out

    # -- fused dispatch / combine (one kernel per all-to-all + GEMM pair) -------------------------------------------------
    def _fused_context(self, hidden_states: torch.Tensor):
        """Symmetric-memory context of the EP group when the fused kernels apply: bf16 on the GPU, capacity a multiple
        of the 256-row tile, TMA-compatible feature sizes, at most 8 EP ranks (``NXD_FUSED_MOE=0`` disables)."""
        from ...ops import backend

        if (not hidden_states.is_cuda or hidden_states.dtype != torch.bfloat16 or not backend.fused_moe_enabled()):
            return None
        e, c, h = hidden_states.shape
        w1 = (self.gate_up_proj if self.glu else self.up_proj).weight
        w2 = self.down_proj.weight
        ep = ps.get_expert_model_parallel_size()
        el = e // ep
        if (c % 256 or h % 64 or w1.shape[2] % 64 or w2.shape[1] % 64 or ep > 8 or el * c // 256 > 64
                or w1.dtype != torch.bfloat16):
            backend.note_fallback("moe_a2a_gemm", f"E={e} C={c} H={h} ep={ep}: shape outside the fused dispatch/combine "
                                  "kernels -> NCCL all-to-all + grouped GEMM")
            return None
        try:
            return ps.get_symm_mem_context("ep")
        except Exception:  # noqa: BLE001 - EP ranks that cannot map each other's memory (other node)
            return None

    def _forward_fused(self, hidden_states: torch.Tensor, sym) -> torch.Tensor:
        e, c, h = hidden_states.shape
        ep = ps.get_expert_model_parallel_size()
        x4 = hidden_states.reshape(ep, e // ep, c, h)                      # slab d = tokens for the experts of rank d
        if ps.get_tensor_model_parallel_size() > 1:
            x4 = mappings.copy_to_tensor_model_parallel_region(x4)         # dgrad is partial over TP
        w1 = (self.gate_up_proj if self.glu else self.up_proj).weight
        y = _DispatchGemm.apply(x4, w1, sym)                               # [ep(src), E_loc, C, N]
        if self.glu:
            gate, up = torch.chunk(y, 2, dim=-1)
            inter = self.activation_fn(gate) * up
        else:
            inter = self.activation_fn(y)
        out = _GemmCombine.apply(inter.contiguous(), self.down_proj.weight, sym)   # [E, C, H]
        if ps.get_tensor_model_parallel_size() > 1:
            out = mappings.reduce_from_tensor_model_parallel_region(out)
        return out
