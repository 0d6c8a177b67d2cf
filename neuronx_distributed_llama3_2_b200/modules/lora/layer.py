"""LoRA adapter layers for plain torch modules (reference
``src/neuronx_distributed/modules/lora/layer.py:20-424``).

``y = base(x) + scale * B(A(dropout(x)))``; ``B`` starts at zero so the wrapped model is unchanged at step 0.
``merge()`` folds ``scale * B A`` into the base weight, ``unmerge()`` removes it again.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F


class LoraLayer:
    """Mixin with the adapter bookkeeping shared by every wrapped layer type."""

    adapter_layer_names = ("lora_A", "lora_B", "lora_embedding_A", "lora_embedding_B")

    def init_lora(self, base_layer: nn.Module, lora_config) -> None:
        self.base_layer = base_layer
        self.lora_config = lora_config
        self.lora_rank = lora_config.lora_rank
        self.lora_alpha = lora_config.lora_alpha
        self.scaling = (self.lora_alpha / math.sqrt(self.lora_rank)) if lora_config.use_rslora \
            else (self.lora_alpha / self.lora_rank)
        self.lora_dropout = nn.Dropout(p=lora_config.lora_dropout) if lora_config.lora_dropout > 0.0 else nn.Identity()
        self.merged = False

    def get_base_layer(self) -> nn.Module:
        b = self.base_layer
        while hasattr(b, "base_layer"):
            b = b.base_layer
        return b

    def reset_lora_parameters(self, a: torch.Tensor, b: torch.Tensor, embedding: bool = False) -> None:
        mode = self.lora_config.init_lora_weights
        with torch.no_grad():
            if embedding:
                nn.init.zeros_(a)
                nn.init.normal_(b)
                return
            if mode == "gaussian":
                nn.init.normal_(a, std=1.0 / self.lora_rank)
            else:
                nn.init.kaiming_uniform_(a, a=math.sqrt(5))
            nn.init.zeros_(b)

    def get_delta_weight(self) -> torch.Tensor:
        raise NotImplementedError

    # -- reference method names (``layer.py:55-110``) -----------------------------------------------------------------
    def _adapter_pairs(self):
        """(A, B, is_embedding) weight pairs of this wrapper; TP wrappers with several B matrices override."""
        if hasattr(self, "lora_embedding_A"):
            return [(self.lora_embedding_A, self.lora_embedding_B, True)]
        return [(self.lora_A.weight, self.lora_B.weight, False)]

    def init_lora_parameters(self, init_lora_weights=None) -> None:
        """(Re-)initialise the adapter: ``"default"`` = Kaiming-uniform A, zero B; ``"gaussian"`` = N(0, 1/r) A, zero B
        (embeddings: zero A, N(0, 1) B).  ``None`` keeps the mode of the stored configuration."""
        if init_lora_weights is not None:
            self.lora_config.init_lora_weights = init_lora_weights
        for a, b, emb in self._adapter_pairs():
            self.reset_lora_parameters(a, b, embedding=emb)

    def update_layer(self, lora_config=None) -> None:
        """Rebuild the adapter for a (new) configuration -- rank, alpha, dropout, init mode -- around the same base
        layer.  A merged adapter is unmerged first so that the base weight is not left modified."""
        if self.merged:
            self.unmerge()
        base = self.base_layer
        for name in list(self._modules):
            if name.startswith("lora_"):
                del self._modules[name]
        for name in list(self._parameters):
            if name.startswith("lora_"):
                del self._parameters[name]
        type(self).__init__(self, base, lora_config or self.lora_config)

    def transpose(self, weight: torch.Tensor) -> torch.Tensor:
        """``weight.T`` when the base layer stores ``(in, out)`` (``fan_in_fan_out``, e.g. GPT-2 Conv1D)."""
        return weight.T if getattr(self.lora_config, "fan_in_fan_out", False) else weight

    def merge(self) -> None:
        if self.merged:
            return
        w = self.get_base_layer().weight
        with torch.no_grad():
            w.add_(self.get_delta_weight().to(w.dtype))
        self.merged = True

    def unmerge(self) -> None:
        if not self.merged:
            return
        w = self.get_base_layer().weight
        with torch.no_grad():
            w.sub_(self.get_delta_weight().to(w.dtype))
        self.merged = False


class LoraLinear(nn.Module, LoraLayer):
    def __init__(self, base_layer: nn.Linear, lora_config):
        super().__init__()
        self.init_lora(base_layer, lora_config)
        dt, dev = base_layer.weight.dtype, base_layer.weight.device
        self.lora_A = nn.Linear(base_layer.in_features, self.lora_rank, bias=False, dtype=dt, device=dev)
        self.lora_B = nn.Linear(self.lora_rank, base_layer.out_features, bias=False, dtype=dt, device=dev)
        self.reset_lora_parameters(self.lora_A.weight, self.lora_B.weight)

    def get_delta_weight(self) -> torch.Tensor:
        return (self.lora_B.weight.float() @ self.lora_A.weight.float()) * self.scaling

    def forward(self, x, *a, **kw):
        out = self.base_layer(x, *a, **kw)
        if self.merged:
            return out
        return out + self.lora_B(self.lora_A(self.lora_dropout(x))) * self.scaling


class LoraEmbedding(nn.Module, LoraLayer):
    def __init__(self, base_layer: nn.Embedding, lora_config):
        super().__init__()
        self.init_lora(base_layer, lora_config)
        dt, dev = base_layer.weight.dtype, base_layer.weight.device
        self.lora_embedding_A = nn.Parameter(torch.empty(self.lora_rank, base_layer.num_embeddings, dtype=dt, device=dev))
        self.lora_embedding_B = nn.Parameter(torch.empty(base_layer.embedding_dim, self.lora_rank, dtype=dt, device=dev))
        self.reset_lora_parameters(self.lora_embedding_A, self.lora_embedding_B, embedding=True)

    def get_delta_weight(self) -> torch.Tensor:
        return (self.lora_embedding_B.float() @ self.lora_embedding_A.float()).t() * self.scaling

    def forward(self, x, *a, **kw):
        out = self.base_layer(x, *a, **kw)
        if self.merged:
            return out
        b = self.get_base_layer()
        after_a = F.embedding(x, self.lora_embedding_A.t(), b.padding_idx, b.max_norm, b.norm_type,
                              b.scale_grad_by_freq, b.sparse)
        return out + (after_a @ self.lora_embedding_B.t()) * self.scaling


class LoraConv2d(nn.Module, LoraLayer):
    def __init__(self, base_layer: nn.Conv2d, lora_config):
        super().__init__()
        self.init_lora(base_layer, lora_config)
        dt, dev = base_layer.weight.dtype, base_layer.weight.device
        self.lora_A = nn.Conv2d(base_layer.in_channels, self.lora_rank, base_layer.kernel_size, base_layer.stride,
                                base_layer.padding, bias=False, dtype=dt, device=dev)
        self.lora_B = nn.Conv2d(self.lora_rank, base_layer.out_channels, (1, 1), (1, 1), bias=False, dtype=dt, device=dev)
        self.reset_lora_parameters(self.lora_A.weight, self.lora_B.weight)

    def get_delta_weight(self) -> torch.Tensor:
        a, b = self.lora_A.weight.float(), self.lora_B.weight.float()
        if self.get_base_layer().weight.shape[2:4] == (1, 1):
            d = (b.squeeze(3).squeeze(2) @ a.squeeze(3).squeeze(2)).unsqueeze(2).unsqueeze(3)
        else:
            d = F.conv2d(a.permute(1, 0, 2, 3), b).permute(1, 0, 2, 3)
        return d * self.scaling

    def forward(self, x, *a, **kw):
        out = self.base_layer(x, *a, **kw)
        if self.merged:
            return out
        return out + self.lora_B(self.lora_A(self.lora_dropout(x))) * self.scaling
