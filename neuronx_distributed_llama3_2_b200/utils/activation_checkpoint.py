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


def apply_activation_checkpointing(model: torch.nn.Module, checkpoint_wrapper_fn: Callable = None,
                                   check_fn: Callable[[torch.nn.Module], bool] = lambda _: True) -> None:
    """Checkpoint every sub-module for which ``check_fn`` is true (outermost match wins).  Argument order as the
    reference (``activation_checkpoint.py:55-59``); ``checkpoint_wrapper_fn(module)`` defaults to ``checkpoint_wrapper``,
    which patches ``forward`` in place -- a custom wrapper that RETURNS a new module is put in the parent's place."""
    wrap = checkpoint_wrapper_fn or checkpoint_wrapper
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
        holders = [(parent, name) for parent in root.modules() for name, child in parent.named_children() if child is m]
        new = wrap(m)
        if new is not None and new is not m:
            for parent, name in holders:
                setattr(parent, name, new)
        wrapped.append(m)


class NxDCheckpointWrapper(torch.nn.Module):
    """Module form of ``checkpoint_wrapper`` (reference ``utils/activation_checkpoint.py``: a
    ``CheckpointWrapper`` subclass whose state-dict keys do not carry the wrapper prefix): parameters and
    ``state_dict`` of the wrapped module appear under their ORIGINAL names, so checkpoints are interchangeable between
    runs with and without activation checkpointing."""

    def __init__(self, module: torch.nn.Module):
        super().__init__()
        self._checkpoint_wrapped_module = module

    def forward(self, *args, **kwargs):
        inner = self._checkpoint_wrapped_module
        if not (self.training and torch.is_grad_enabled()):
            return inner(*args, **kwargs)
        return checkpoint(inner, *args, use_reentrant=False, **kwargs)

    def __getattr__(self, name):
        try:
            return super().__getattr__(name)
        except AttributeError:
            return getattr(self._checkpoint_wrapped_module, name)

    def named_parameters(self, prefix: str = "", recurse: bool = True, remove_duplicate: bool = True):
        for n, p in super().named_parameters(prefix=prefix, recurse=recurse, remove_duplicate=remove_duplicate):
            yield n.replace("_checkpoint_wrapped_module.", ""), p

    def state_dict(self, *args, destination=None, prefix: str = "", keep_vars: bool = False):
        """Keys without the wrapper's attribute name -- also when this wrapper sits inside a larger model whose
        ``state_dict`` recursion lands here with a ``destination`` and a ``prefix``."""
        if args:                                   # deprecated positional form (destination, prefix, keep_vars)
            destination = args[0] if len(args) > 0 else destination
            prefix = args[1] if len(args) > 1 else prefix
            keep_vars = args[2] if len(args) > 2 else keep_vars
        local = super().state_dict(prefix=prefix, keep_vars=keep_vars)
        out = destination if destination is not None else type(local)()
        for k, v in local.items():
            out[k.replace("_checkpoint_wrapped_module.", "")] = v
        return out

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        inner_prefix = prefix + "_checkpoint_wrapped_module."
        for k in [k for k in state_dict if k.startswith(prefix) and not k.startswith(inner_prefix)]:
            state_dict[inner_prefix + k[len(prefix):]] = state_dict.pop(k)
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)
