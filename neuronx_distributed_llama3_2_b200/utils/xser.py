"""Tensor-per-file checkpoint serialisation, on-disk compatible with ``torch_xla.utils.serialization``
(SURVEY 2.2 "xser", 7.4 item 8).

Format (what the reference writes through ``xser.save``, see reference
``src/neuronx_distributed/parallel_layers/checkpointing.py:80-113`` and
``trainer/checkpoint.py:426-504``):

    <path>                      torch.save of the object with every tensor replaced by
                                ``TensorReference(tid)``
    <path>.tensors/tensor_<tid>.pt   one ``torch.save``d CPU tensor per reference

``TensorReference`` is pickled under the module path ``torch_xla.utils.serialization`` so files
written here load with real torch_xla and vice versa; when torch_xla is absent (always, on B200)
a stub module is registered in ``sys.modules`` so unpickling resolves.
"""
from __future__ import annotations

import os
import sys
import types
from typing import Any, Callable, List, Tuple

import torch


def _install_shim():
    name = "torch_xla.utils.serialization"
    if name in sys.modules and hasattr(sys.modules[name], "TensorReference"):
        return sys.modules[name].TensorReference
    try:  # a real torch_xla wins if present
        import importlib

        mod = importlib.import_module(name)
        return mod.TensorReference
    except Exception:
        pass

    class TensorReference(object):
        def __init__(self, tid):
            self.tid = tid

        def __repr__(self):
            return f"TensorReference({self.tid})"

    TensorReference.__module__ = name
    TensorReference.__qualname__ = "TensorReference"
    for parent in ("torch_xla", "torch_xla.utils"):
        if parent not in sys.modules:
            m = types.ModuleType(parent)
            m.__path__ = []  # mark as package
            m.__nxd_b200_shim__ = True
            sys.modules[parent] = m
    mod = types.ModuleType(name)
    mod.TensorReference = TensorReference
    mod.__nxd_b200_shim__ = True
    sys.modules[name] = mod
    sys.modules["torch_xla.utils"].serialization = mod
    sys.modules["torch_xla"].utils = sys.modules["torch_xla.utils"]
    return TensorReference


TensorReference = _install_shim()


def rewrite_data(data: Any, select_fn: Callable[[Any], bool], convert_fn: Callable[[List[Any]], List[Any]]) -> Any:
    """Replace every leaf matching ``select_fn`` inside nested dict/list/tuple with the element
    ``convert_fn`` returns for it (all selected leaves are converted in one call)."""
    selected: List[Any] = []

    def collect(x):
        if select_fn(x):
            selected.append(x)
        elif isinstance(x, dict):
            for v in x.values():
                collect(v)
        elif isinstance(x, (list, tuple)):
            for v in x:
                collect(v)

    collect(data)
    converted = convert_fn(selected)
    it = iter(converted)

    def rebuild(x):
        if select_fn(x):
            return next(it)
        if isinstance(x, dict):
            return type(x)((k, rebuild(v)) for k, v in x.items())
        if isinstance(x, list):
            return [rebuild(v) for v in x]
        if isinstance(x, tuple):
            return tuple(rebuild(v) for v in x)
        return x

    return rebuild(data)


def tensors_dir(path: str) -> str:
    return path + ".tensors"


def split_tensors(data: Any) -> Tuple[Any, List[torch.Tensor]]:
    """(object with TensorReference leaves, list of tensors indexed by tid)."""
    tensors: List[torch.Tensor] = []

    def convert(ts):
        out = []
        for t in ts:
            out.append(TensorReference(len(tensors)))
            tensors.append(t)
        return out

    ref = rewrite_data(data, lambda x: isinstance(x, torch.Tensor), convert)
    return ref, tensors


def save(data: Any, path: str, master_only: bool = True, global_master: bool = False) -> None:
    """Write ``data`` in tensor-per-file form (every tensor moved to CPU first)."""
    ref, tensors = split_tensors(data)
    os.makedirs(tensors_dir(path), exist_ok=True)
    for tid, t in enumerate(tensors):
        torch.save(t.detach().cpu(), os.path.join(tensors_dir(path), f"tensor_{tid}.pt"))
    torch.save(ref, path)


def load(path: str, map_location="cpu") -> Any:
    ref = torch.load(path, weights_only=False)

    def convert(refs):
        return [torch.load(os.path.join(tensors_dir(path), f"tensor_{r.tid}.pt"), map_location=map_location,
                           weights_only=False) for r in refs]

    return rewrite_data(ref, lambda x: isinstance(x, TensorReference), convert)
