"""``ExpertMLPs``: dispatch tokens to experts, run them, combine with the affinities (reference
``src/neuronx_distributed/modules/moe/expert_mlps.py:36-357``).

Three execution modes with the reference's semantics:
  * ``forward_capacity_factor``  every expert takes at most ``C = ceil(T k cf / E)`` tokens, surplus tokens are
    DROPPED in token order (position in expert = running count over the token dim);
  * ``forward_all_experts``      full capacity: result equals "every token through every expert, weighted by the
    masked affinities".  Computed DROPLESS here: tokens are sorted by expert and each expert multiplies only its own
    segment (the reference really runs every token through every expert, ``:139-167``) -- same result, 1/E·k of the FLOPs;
  * ``forward_selective_loading`` token generation: only the chosen experts' weights are touched.
The dispatch bookkeeping is scatter/sort based; the reference's matmul-cumsum and arange-compare one-hots exist only
because of XLA.
"""
from __future__ import annotations

import math
from typing import Any, Callable, Optional, Union

import torch
import torch.nn.functional as F

from ...parallel import parallel_state as ps
from ...utils.model_utils import ACT2FN
from .experts import Experts

# token-gen batches with T * top_k / E below this take the selective-loading path (reference ``:36``)
SELECTIVE_LOADING_THRESHOLD = 1.0


class ExpertMLPs(torch.nn.Module):
    def __init__(self, num_experts: int, top_k: int, hidden_size: int, intermediate_size: int, hidden_act: str,
                 glu_mlp: bool, capacity_factor: Union[None, float], normalize_top_k_affinities: bool = False,
                 return_bias: bool = False, init_method: Optional[Callable[..., Any]] = torch.nn.init.kaiming_uniform_,
                 output_layer_init_method: Optional[Callable[..., Any]] = torch.nn.init.kaiming_uniform_,
                 dtype: torch.dtype = torch.float32, device: torch.device = torch.device("cpu")):
        super().__init__()
        self.num_experts = num_experts
        if not (0 < top_k <= num_experts):
            raise ValueError(f"Invalid top_k={top_k} for num_experts={num_experts}")
        self.top_k = top_k
        self.hidden_size, self.intermediate_size = hidden_size, intermediate_size
        if hidden_act not in ACT2FN:
            raise ValueError(f"Unknown activation: {hidden_act} ; Supported: {list(ACT2FN.keys())}")
        self.act_fn = ACT2FN[hidden_act]
        self.glu_mlp = glu_mlp
        if capacity_factor is None or capacity_factor >= num_experts / top_k:
            capacity_factor = None  # full capacity
        self.capacity_factor = capacity_factor
        if normalize_top_k_affinities and top_k == 1:
            raise ValueError("top_k must be greater than 1 for normalizing top-k expert affinities")
        self.normalize_top_k_affinities = normalize_top_k_affinities
        if return_bias:
            raise NotImplementedError("bias is currently unsupported for MoE")
        self.return_bias = return_bias
        self.mlp_op = Experts(num_experts, hidden_size, intermediate_size, glu_mlp, self.act_fn, dtype=dtype,
                              device=device, input_layer_init_method=init_method,
                              output_layer_init_method=output_layer_init_method)
        self.dtype, self.device = dtype, device

    # -- helpers ---------------------------------------------------------------------------------------------
    def get_expert_mask(self, expert_index: torch.Tensor) -> torch.Tensor:
        """(T, E) top_k-hot mask of the chosen experts."""
        mask = torch.zeros(expert_index.shape[0], self.num_experts, device=expert_index.device, dtype=torch.float32)
        mask.scatter_(1, expert_index, 1.0)
        return mask

    def get_expert_affinities_masked(self, expert_affinities: torch.Tensor, expert_mask: torch.Tensor) -> torch.Tensor:
        aff = expert_affinities.masked_fill(expert_mask == 0, 0)
        if self.normalize_top_k_affinities:
            aff = F.normalize(aff, p=1.0, dim=1)
        return aff

    # -- full capacity, dropless -----------------------------------------------------------------------------
    def forward_all_experts(self, hidden_states, expert_affinities, expert_index):
        if ps.get_expert_model_parallel_size() > 1:
            # the reference refuses this combination (:145-146).  With experts spread over ranks the dispatch needs a
            # fixed per-expert buffer, so run the capacity path with room for EVERY token (capacity = T): nothing is
            # dropped, the result equals the dropless one.
            return self.forward_capacity_factor(hidden_states, expert_affinities, expert_index, full_capacity=True)
        T, H = hidden_states.shape
        mask = self.get_expert_mask(expert_index)
        aff = self.get_expert_affinities_masked(expert_affinities, mask)                 # (T, E)
        flat_expert = expert_index.reshape(-1)                                            # (T*k,)
        order = torch.argsort(flat_expert, stable=True)
        token_of = (order // self.top_k)
        sorted_expert = flat_expert[order]
        counts = torch.bincount(sorted_expert, minlength=self.num_experts)
        x_sorted = hidden_states[token_of]                                                # (T*k, H)
        w_sorted = aff[token_of, sorted_expert].unsqueeze(1).to(hidden_states.dtype)
        out_sorted = torch.empty_like(x_sorted)
        start = 0
        for e, n in enumerate(counts.tolist()):
            if n == 0:
                continue
            seg = x_sorted[start:start + n].unsqueeze(0)                                  # (1, n, H)
            out_sorted[start:start + n] = self.mlp_op(seg, expert_indices=torch.tensor([e], device=seg.device))[0]
            start += n
        output = torch.zeros(T, H, device=hidden_states.device, dtype=hidden_states.dtype)
        output.index_add_(0, token_of, out_sorted * w_sorted)
        return output

    # -- fixed capacity with token dropping ------------------------------------------------------------------
    def forward_capacity_factor(self, hidden_states, expert_affinities, expert_index, full_capacity: bool = False):
        T, H = hidden_states.shape
        E, k = self.num_experts, self.top_k
        capacity = T if full_capacity else min(math.ceil(T * k * self.capacity_factor / E), T)
        mask = self.get_expert_mask(expert_index)                                         # (T, E)
        position = torch.cumsum(mask, dim=0)                                              # 1-based slot in expert
        mask = mask.masked_fill(position > capacity, 0)                                   # drop the overflow
        aff = self.get_expert_affinities_masked(expert_affinities, mask)                  # (T, E)
        tok = torch.arange(T, device=hidden_states.device).unsqueeze(1).expand(T, k)
        kept = mask.gather(1, expert_index) > 0                                           # (T, k)
        slot = (position.gather(1, expert_index).long() - 1).clamp(min=0)                 # (T, k)
        flat_pos = expert_index * capacity + slot                                         # index into (E*C)
        from ...ops import moe_permute

        if moe_permute.usable(hidden_states):
            # two kernels (+ the grouped expert GEMMs) instead of indexing chains: gather by token_assignments, weighted
            # scatter-combine (csrc/moe_permute.cu); empty slots are zero rows, dropped pairs carry weight 0
            slot_token = torch.full((E * capacity,), -1, dtype=torch.int32, device=hidden_states.device)
            slot_token[flat_pos[kept]] = tok[kept].to(torch.int32)
            slot_of = flat_pos.clamp(max=E * capacity - 1).to(torch.int32).contiguous()
            aligned = moe_permute.dispatch(hidden_states, slot_token, slot_of, kept).reshape(E, capacity, H)
            aligned_out = self.mlp_op(aligned).reshape(E * capacity, H)
            w = aff.gather(1, expert_index) * kept
            return moe_permute.combine(aligned_out, w, slot_of, slot_token, kept)
        # expert-aligned gather table; empty slots point at token 0 and are zeroed through the affinity
        assign = torch.zeros(E * capacity, dtype=torch.long, device=hidden_states.device)
        assign[flat_pos[kept]] = tok[kept]
        aligned = hidden_states[assign].reshape(E, capacity, H)
        aligned_out = self.mlp_op(aligned).reshape(E * capacity, H)
        w = aff.gather(1, expert_index) * kept                                            # (T, k)
        gathered = aligned_out[flat_pos.clamp(max=E * capacity - 1)]                      # (T, k, H)
        return (gathered * w.unsqueeze(-1).to(gathered.dtype)).sum(dim=1)

    # -- token generation ------------------------------------------------------------------------------------
    def forward_selective_loading(self, hidden_states, expert_affinities, expert_index):
        T = hidden_states.shape[0]
        chosen = expert_affinities.gather(1, expert_index)
        if self.normalize_top_k_affinities:
            chosen = F.normalize(chosen, p=1.0, dim=1)
        outs = []
        for t in range(T):
            y = self.mlp_op(hidden_states[t].reshape(1, 1, -1).expand(self.top_k, 1, -1), expert_indices=expert_index[t])
            outs.append((y.squeeze(1) * chosen[t].unsqueeze(1).to(y.dtype)).sum(dim=0))
        return torch.stack(outs, dim=0)

    def forward(self, hidden_states, expert_affinities, expert_index, seq_len):
        """``hidden_states (T, H)``, ``expert_affinities (T, E)``, ``expert_index (T, top_k)``; ``seq_len`` tells
        context encoding (> 1) from token generation (== 1) in eval (reference ``:298-357``)."""
        if self.training:
            if self.capacity_factor is None:
                return self.forward_all_experts(hidden_states, expert_affinities, expert_index)
            return self.forward_capacity_factor(hidden_states, expert_affinities, expert_index)
        if seq_len == 1:
            if ps.get_expert_model_parallel_size() > 1:
                raise NotImplementedError("expert parallelism is not supported for token generation")
            T = hidden_states.shape[0]
            if T * self.top_k / self.num_experts < SELECTIVE_LOADING_THRESHOLD:
                return self.forward_selective_loading(hidden_states, expert_affinities, expert_index)
            return self.forward_all_experts(hidden_states, expert_affinities, expert_index)
        if self.capacity_factor is None:
            return self.forward_all_experts(hidden_states, expert_affinities, expert_index)
        return self.forward_capacity_factor(hidden_states, expert_affinities, expert_index)
