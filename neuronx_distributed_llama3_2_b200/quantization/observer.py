"""Per-channel abs-max observer (reference ``quantization/observer.py:12-166``): tracks the running max |x| of every
channel along ``ch_axis`` and derives symmetric int8 scales (zero point 0)."""
from __future__ import annotations

from typing import Tuple

import torch


class PerChannelAbsMaxObserver(torch.nn.Module):
    def __init__(self, ch_axis: int = 0, dtype: torch.dtype = torch.qint8, qscheme=torch.per_channel_symmetric,
                 quant_min: int = -128, quant_max: int = 127, eps: float = torch.finfo(torch.float32).eps,
                 reduce_range: bool = False, factory_kwargs=None, **_):
        # ``reduce_range`` / ``factory_kwargs`` belong to the torch.ao observer protocol the reference derives from
        super().__init__()
        if reduce_range:
            raise NotImplementedError("reduce_range is not supported for symmetric int8 weights")
        if qscheme != torch.per_channel_symmetric:
            raise NotImplementedError("PerChannelAbsMaxObserver only supports per_channel_symmetric")
        self.ch_axis, self.dtype, self.qscheme = ch_axis, dtype, qscheme
        self.quant_min, self.quant_max, self.eps = quant_min, quant_max, eps
        self.register_buffer("max_val", torch.tensor([]))

    def forward(self, x_orig: torch.Tensor) -> torch.Tensor:
        if x_orig.numel() == 0:
            return x_orig
        x = x_orig.detach().to(torch.float32)
        axis = self.ch_axis % x.dim()
        y = x.movedim(axis, 0).reshape(x.shape[axis], -1)
        cur = y.abs().amax(dim=1)
        self.max_val = cur if self.max_val.numel() == 0 else torch.maximum(self.max_val.to(cur.device), cur)
        return x_orig

    @torch.jit.export
    def calculate_qparams(self) -> Tuple[torch.Tensor, torch.Tensor]:
        if self.max_val.numel() == 0:
            return torch.tensor([1.0]), torch.tensor([0])
        scale = torch.clamp(self.max_val / float(self.quant_max), min=self.eps)
        return scale, torch.zeros_like(scale, dtype=torch.int64)

    @torch.jit.export
    def reset_min_max_vals(self) -> None:
        """Forget the running statistics (reference ``observer.py:158-166``)."""
        self.max_val = torch.tensor([], device=self.max_val.device)

    @classmethod
    def with_args(cls, **kwargs):
        """``torch.ao`` observer-factory protocol: ``QConfig(weight=PerChannelAbsMaxObserver.with_args(...))``."""
        from torch.ao.quantization.observer import _with_args

        return _with_args(cls, **kwargs)

    def extra_repr(self) -> str:
        return f"max_val={self.max_val}"
