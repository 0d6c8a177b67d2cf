"""On-device sampling (reference ``src/neuronx_distributed/utils/sampling.py:27-77``): greedy when
``top_k == 1``, otherwise top-k (+ top-p / temperature) multinomial by inverse CDF against one uniform per row -- one
fused sm_100a kernel per batch (``csrc/sampling.cu``), no ``torch.multinomial``."""
from __future__ import annotations

from typing import Optional

import torch

from ..ops import sampling as ops


class Sampler:
    def __init__(self, config=None, top_k: Optional[int] = None, top_p: float = 1.0, temperature: float = 1.0,
                 do_sample: Optional[bool] = None, seed: int = 0):
        self.top_k = top_k if top_k is not None else getattr(config, "top_k", 1)
        self.top_p = getattr(config, "top_p", top_p) if config is not None else top_p
        self.temperature = getattr(config, "temperature", temperature) if config is not None else temperature
        ds = do_sample if do_sample is not None else getattr(config, "do_sample", self.top_k > 1)
        self.is_greedy = (not ds) or self.top_k == 1
        self.is_medusa = bool(getattr(config, "is_medusa", False))
        self.on_device_sampling = bool(getattr(config, "on_device_sampling", True))
        self.generator: Optional[torch.Generator] = None
        self.seed = seed
        self.sampling_method = self.multinomial

    def _uniform(self, n: int, device) -> torch.Tensor:
        if self.generator is None or self.generator.device != torch.device(device):
            self.generator = torch.Generator(device=device)
            self.generator.manual_seed(self.seed)
        return torch.rand(n, device=device, dtype=torch.float32, generator=self.generator)

    def sample(self, logits: torch.Tensor, uniform: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``logits [B, V]`` -> token ids ``[B]``."""
        if self.is_greedy:
            return ops.argmax(logits)
        u = uniform if uniform is not None else self._uniform(logits.shape[0], logits.device)
        return ops.topk_sample(logits, self.top_k, self.top_p, self.temperature, u)

    def multinomial(self, token_logits: torch.Tensor, uniform: Optional[torch.Tensor] = None) -> torch.Tensor:
        """The reference's sampling method name (``utils/sampling.py:27-77``): greedy for ``top_k == 1``, else one draw
        from the soft-maxed top-k.  Medusa needs the candidate set itself, so it gets the top-k indices ``[B, k]``."""
        if self.is_medusa and not self.is_greedy:
            return torch.topk(token_logits, self.top_k, dim=1).indices
        return self.sample(token_logits, uniform)
