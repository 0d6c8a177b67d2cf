"""Activation checkpointing by module class.

Reference: ``src/neuronx_distributed/utils/activation_checkpoint.py:20-83`` wraps matching sub-modules in a
reentrant XLA checkpoint wrapper and strips the wrapper prefix from parameter names.  Here the wrapper patches
``forward`` in place (non-reentrant ``torch.utils.checkpoint``), so module paths and ``state_dict`` keys do not
change at all and nothing needs stripping.
"""
from __future__ import annotations

import functools
from typing import Callable

import torch
from torch.utils.checkpoint import checkpoint

_WRAPPED = "_nxd_checkpoint_wrapped"


def checkpoint_wrapper(module: torch.nn.Module) -> torch.nn.Module:
    if getattr(module, _WRAPPED, False):
        return module
    inner = module.forward

    @functools.wraps(inner)
    def forward(*args, **kwargs):
        if not (module.training and torch.is_grad_enabled()):
            return inner(*args, **kwargs)
        return checkpoint(inner, *args, use_reentrant=False, **kwargs)

    module.forward = forward
    setattr(module, _WRAPPED, True)
    return module


def apply_activation_checkpointing(model: torch.nn.Module, check_fn: Callable[[torch.nn.Module], bool] = lambda _: True
                                   ) -> None:
    """Checkpoint every sub-module for which ``check_fn`` is true (outermost match wins)."""
    root = model.module if hasattr(model, "module") and isinstance(model.module, torch.nn.Module) else model
    if hasattr(root, "local_modules"):
        candidates = list(root.local_modules())
    else:
        candidates = list(root.modules())
    wrapped = []
    for m in candidates:
        if m is root or not check_fn(m):
            continue
        if any(m is not w and any(m is c for c in w.modules()) for w in wrapped):
            continue
        checkpoint_wrapper(m)
        wrapped.append(m)
