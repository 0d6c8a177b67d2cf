"""``MoE``: router + expert MLPs inside the tensor/sequence/expert-parallel regions (reference
``src/neuronx_distributed/modules/moe/model.py:7-160``)."""
from __future__ import annotations

import torch

from ...parallel import mappings
from ...parallel import parallel_state as ps


class MoE(torch.nn.Module):
    """``forward(hidden_states (S, B, H) [SP: (S/tp, B, H)])`` -> ``(output, [bias], [router_logits], [expert_index])``.

    Exit collectives as the reference (``:141-150``): SP -> reduce-scatter on the sequence dim, otherwise all-reduce
    over TP (the expert row-parallel projections deliver partial sums when ``ep == 1``); with EP the projections have
    already reduced and SP just scatters.
    """

    def __init__(self, router, expert_mlps, return_router_logits: bool = False, return_expert_index: bool = False,
                 sequence_parallel_enabled: bool = False):
        super().__init__()
        for attr in ("num_experts", "top_k", "hidden_size"):
            if getattr(router, attr) != getattr(expert_mlps, attr):
                raise ValueError(f"Inconsistent {attr} across the router and expert_mlps")
        self.router = router
        self.expert_mlps = expert_mlps
        self.return_router_logits = return_router_logits
        self.return_expert_index = return_expert_index
        self.sequence_parallel_enabled = sequence_parallel_enabled
        self.ep_enabled = ps.get_expert_model_parallel_size() > 1
        # TP without EP: the expert down projections deliver PARTIAL sums and this module reduces once at its exit
        # (reference :141-150).  Everything between entry and exit then back-propagates per-rank partial gradients --
        # the expert input grads and, through the affinities, the router's -- so the entry op sums them over TP in ONE
        # collective (reduce-scatter under SP, all-reduce otherwise) and the expert column projections skip theirs.
        # The reference all-reduces only the expert part (moe_parallel_layers.py:96-101) and drops the router part of
        # the input gradient; summing at the entry is exact and saves a collective.
        self.deferred_reduction = ps.get_tensor_model_parallel_size() > 1 and not self.ep_enabled
        if self.deferred_reduction:
            from .moe_parallel_layers import ExpertFusedColumnParallelLinear

            for m in self.expert_mlps.modules():
                if isinstance(m, ExpertFusedColumnParallelLinear):
                    m.async_tensor_model_parallel_allreduce = False

    def forward(self, hidden_states: torch.Tensor):
        tp = ps.get_tensor_model_parallel_size()
        sp = self.sequence_parallel_enabled and tp > 1
        if sp and not self.training:
            raise NotImplementedError("sequence parallelism is not supported for MoE inference (reference :108-110)")
        if sp:  # deferred reduction: backward = reduce-scatter of the partial input grads; with EP they are complete: split
            full = mappings.gather_from_sequence_parallel_region(hidden_states,
                                                                 to_model_parallel=self.deferred_reduction)
        elif self.deferred_reduction:
            full = mappings.copy_to_tensor_model_parallel_region(hidden_states)
        else:
            full = hidden_states
        shape = full.shape
        seq_len = shape[0]
        flat = full.reshape(-1, shape[-1])
        router_logits, expert_affinities, expert_index = self.router(flat)
        out = self.expert_mlps(flat, expert_affinities, expert_index, seq_len).reshape(shape)
        if sp and self.ep_enabled:
            out = mappings.scatter_to_sequence_parallel_region(out)
        elif sp:
            out = mappings.reduce_scatter_to_sequence_parallel_region(out)
        elif tp > 1 and not self.ep_enabled:
            out = mappings.reduce_from_tensor_model_parallel_region(out)
        ret = [out]
        if getattr(self.expert_mlps, "return_bias", False):
            ret.append(None)
        if self.return_router_logits:
            ret.append(router_logits)
        if self.return_expert_index:
            ret.append(expert_index)
        return tuple(ret)
