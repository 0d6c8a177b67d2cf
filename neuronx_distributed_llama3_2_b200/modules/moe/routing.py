"""Token routers (reference ``src/neuronx_distributed/modules/moe/routing.py:9-218``).

``forward(hidden_states)`` takes ``(S, B, H)`` or ``(T, H)`` and returns
``(router_logits (T, E), expert_affinities (T, E), expert_index (T, top_k))``.  Affinities are computed in fp32
(the reference uses fp64 to dodge XLA's auto-downcast) and cast back to the activation dtype.
"""
from __future__ import annotations

from abc import ABC, abstractmethod

import torch
import torch.nn.functional as F

from .moe_parallel_layers import LinearRouter


class RouterBase(torch.nn.Module, ABC):
    def __init__(self, num_experts: int, top_k: int, hidden_size: int, act_fn: str, dtype: torch.dtype = torch.float32,
                 device: torch.device = torch.device("cpu")):
        super().__init__()
        self.num_experts = num_experts
        if not (0 < top_k <= num_experts):
            raise ValueError(f"Invalid top_k={top_k} for num_experts={num_experts}")
        self.top_k = top_k
        self.hidden_size = hidden_size
        if act_fn not in ("sigmoid", "softmax"):
            raise ValueError("act_fn must be either 'sigmoid' or 'softmax'")
        self.act_fn = act_fn
        self.linear_router = LinearRouter(hidden_size, num_experts, dtype=dtype, device=device)

    def get_router_logits(self, hidden_states: torch.Tensor) -> torch.Tensor:
        return self.linear_router(hidden_states).reshape(-1, self.num_experts)

    def apply_activation_fn(self, router_logits: torch.Tensor) -> torch.Tensor:
        x = router_logits.float()
        aff = torch.sigmoid(x) if self.act_fn == "sigmoid" else F.softmax(x, dim=1)
        return aff

    def get_router_logits_and_expert_affinities(self, hidden_states: torch.Tensor):
        """``(router_logits (T, E), expert_affinities (T, E) in the dtype of hidden_states)`` -- the activation runs in
        fp32 (the reference upcasts to fp64 only to dodge XLA's auto-downcast; ``routing.py:50-67``)."""
        router_logits = self.get_router_logits(hidden_states)
        return router_logits, self.apply_activation_fn(router_logits).to(hidden_states.dtype)

    @abstractmethod
    def forward(self, hidden_states):
        ...


class RouterTopK(RouterBase):
    """Softmax + top-k; no balancing inside -- pair it with ``load_balancing_loss_func`` (reference ``:89-120``)."""

    def __init__(self, num_experts: int, top_k: int, hidden_size: int, dtype: torch.dtype = torch.float32,
                 device: torch.device = torch.device("cpu")):
        super().__init__(num_experts, top_k, hidden_size, "softmax", dtype, device)

    def forward(self, hidden_states):
        router_logits = self.get_router_logits(hidden_states)
        expert_affinities = self.apply_activation_fn(router_logits)
        _, expert_index = torch.topk(router_logits, self.top_k, dim=1)
        return router_logits, expert_affinities.to(hidden_states.dtype), expert_index.detach().to(torch.long)


class RouterSinkhorn(RouterBase):
    """Top-1 routing; during training the assignment is balanced with a fixed number of Sinkhorn iterations on the
    (detached, fp32) logits, in eval it is the plain argmax (reference ``:123-218``)."""

    DEFAULT_SINKHORN_ITERS = 30

    def __init__(self, num_experts: int, top_k: int, hidden_size: int, act_fn: str = "sigmoid",
                 dtype: torch.dtype = torch.float32, device: torch.device = torch.device("cpu"),
                 sinkhorn_iterations: int = None, sinkhorn_tol: float = None):
        if top_k != 1:
            raise NotImplementedError("RouterSinkhorn only supports Top-1 routing")
        super().__init__(num_experts, top_k, hidden_size, act_fn, dtype, device)
        self.sinkhorn_iterations = sinkhorn_iterations if sinkhorn_iterations is not None else self.DEFAULT_SINKHORN_ITERS
        self.sinkhorn_tol = sinkhorn_tol

    @staticmethod
    def _sinkhorn(cost: torch.Tensor, iters: int, tol=None) -> torch.Tensor:
        cost = torch.exp(cost)
        d0 = torch.ones(cost.shape[0], device=cost.device, dtype=cost.dtype)
        d1 = torch.ones(cost.shape[1], device=cost.device, dtype=cost.dtype)
        eps = 1e-8
        d1_old = d1
        for _ in range(iters):
            d0 = (1.0 / d0.numel()) / (torch.sum(d1 * cost, dim=1) + eps)
            d1 = (1.0 / d1.numel()) / (torch.sum(d0.unsqueeze(1) * cost, dim=0) + eps)
            if tol is not None and torch.mean(torch.abs(d1_old - d1)) <= tol:
                break
            d1_old = d1
        return d1 * cost * d0.unsqueeze(1)

    def forward(self, hidden_states):
        router_logits = self.get_router_logits(hidden_states)
        expert_affinities = self.apply_activation_fn(router_logits)
        if self.training:
            with torch.no_grad():
                balanced = self._sinkhorn(router_logits.detach().float(), self.sinkhorn_iterations, self.sinkhorn_tol)
                expert_index = torch.argmax(balanced, dim=1, keepdim=True)
        else:
            expert_index = torch.argmax(router_logits, dim=1, keepdim=True)
        return router_logits, expert_affinities.to(hidden_states.dtype), expert_index.detach().to(torch.long)
