"""LayerNorm / RMSNorm whose parameters are tagged for sequence parallelism
(reference ``parallel_layers/layer_norm.py:17-37``).

Under SP every TP rank sees a different sequence slice, so the weight/bias grads
are partial sums that ``allreduce_sequence_parallel_gradients`` (grads.py)
all-reduces over TP; the ``sequence_parallel_enabled`` tag is how they are found.
"""
from __future__ import annotations

import torch

from ..ops import norm as norm_ops


class LayerNorm(torch.nn.LayerNorm):
    def __init__(self, normalized_shape, eps: float = 1e-5, elementwise_affine: bool = True,
                 sequence_parallel_enabled: bool = False, dtype=torch.float32, device=None):
        super().__init__(normalized_shape, eps=eps, elementwise_affine=elementwise_affine, dtype=dtype, device=device)
        self.sequence_parallel_enabled = sequence_parallel_enabled
        if elementwise_affine:
            setattr(self.weight, "sequence_parallel_enabled", sequence_parallel_enabled)
            setattr(self.bias, "sequence_parallel_enabled", sequence_parallel_enabled)


class RMSNorm(torch.nn.Module):
    """``w * x * rsqrt(mean(x^2) + eps)`` (Llama norm; reference training model
    ``examples/training/llama/modeling_llama_nxd.py:142-149`` does it in fp64 to
    dodge XLA auto-downcast; here: fp32 statistics inside one sm_100a kernel)."""

    def __init__(self, hidden_size: int, eps: float = 1e-6, sequence_parallel_enabled: bool = False,
                 dtype=torch.float32, device=None):
        super().__init__()
        self.weight = torch.nn.Parameter(torch.ones(hidden_size, dtype=dtype, device=device))
        self.variance_epsilon = eps
        self.sequence_parallel_enabled = sequence_parallel_enabled
        setattr(self.weight, "sequence_parallel_enabled", sequence_parallel_enabled)

    def forward(self, hidden_states: torch.Tensor, residual: torch.Tensor = None) -> torch.Tensor:
        """``rmsnorm(hidden_states)``, or with ``residual``: ``rmsnorm(hidden_states + residual)`` in one kernel."""
        if residual is not None:
            return norm_ops.fused_add_rms_norm(hidden_states, residual, self.weight, self.variance_epsilon)[0]
        return norm_ops.rms_norm(hidden_states, self.weight, self.variance_epsilon)
