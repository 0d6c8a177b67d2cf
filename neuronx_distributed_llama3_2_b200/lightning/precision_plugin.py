"""Precision plugin (reference ``lightning/precision_plugin.py:11-26`` toggles XLA bf16 flags): bf16 parameters with
fp32 master weights in the ZeRO-1 optimizer need no autocast; ``mode='autocast'`` wraps the step in ``torch.autocast``."""
from __future__ import annotations

import contextlib

import torch

from ._compat import PrecisionBase


class NeuronXLAPrecisionPlugin(PrecisionBase):
    def __init__(self, mode: str = "bf16-params", dtype: torch.dtype = torch.bfloat16,
                 mixed_precision_enabled: bool = False):
        """``mixed_precision_enabled`` is the reference's only argument (``precision_plugin.py:12-16``): fp32 master
        weights live in the optimizer, the forward runs in ``dtype`` parameters -- the default ``bf16-params`` mode."""
        super().__init__()
        self.mode, self.dtype, self.mixed_precision_enabled = mode, dtype, mixed_precision_enabled

    def forward_context(self):
        if self.mode == "autocast":
            dev = "cuda" if torch.cuda.is_available() else "cpu"
            return torch.autocast(device_type=dev, dtype=self.dtype)
        return contextlib.nullcontext()

    def optimizer_step(self, optimizer, model=None, closure=None, **kwargs):
        if closure is not None:
            closure()
        return optimizer.step()
