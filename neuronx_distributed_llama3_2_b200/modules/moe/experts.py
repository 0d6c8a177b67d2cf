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
