"""Switch-Transformer load-balancing auxiliary loss (reference
``src/neuronx_distributed/modules/moe/loss_function.py:5-48``).

``loss = E / k * sum_e(fraction_of_tokens_routed_to_e * mean_router_prob_e)`` computed over all layers' router logits
concatenated on the token dim.  The reference builds the one-hot with an arange-compare in fp64 (an XLA workaround);
here it is a plain ``scatter_add`` in fp32.
"""
from __future__ import annotations

from typing import Sequence, Union

import torch


def load_balancing_loss_func(router_logits: Union[torch.Tensor, Sequence[torch.Tensor]], num_experts: int,
                             top_k: int) -> torch.Tensor:
    if not isinstance(router_logits, torch.Tensor):
        router_logits = torch.cat([g.reshape(-1, g.shape[-1]) for g in router_logits
                                              if g is not None], dim=0)
    logits = router_logits.reshape(-1, num_experts).float()
    probs = torch.softmax(logits, dim=-1)
    _, selected = torch.topk(probs, top_k, dim=-1)
    counts = torch.zeros(num_experts, dtype=torch.float32, device=logits.device)
    counts.scatter_add_(0, selected.reshape(-1), torch.ones(selected.numel(), dtype=torch.float32, device=logits.device))
    tokens_per_expert = counts / logits.shape[0]          # sums to top_k
    router_prob_per_expert = probs.mean(dim=0)
    return (tokens_per_expert * router_prob_per_expert).sum() * (num_experts / top_k) * 1.0
