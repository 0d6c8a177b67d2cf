"""FX tracing of the user model for pipeline partitioning.

Reference: ``src/neuronx_distributed/pipeline/trace.py:77-211``: the transformer layer class and the parallel
modules are leaves, parallel functions are auto-wrapped, unused inputs become concrete ``None`` args, the traced
graph is DCE'd and linted.  The HF tracer is used when the model is a ``transformers.PreTrainedModel``.
"""
from __future__ import annotations

import contextlib

import inspect
from typing import Any, Callable, Dict, Iterable, List, Optional, Sequence, Type

import torch
import torch.fx as fx

from ..utils.logger import get_logger

logger = get_logger()


def _default_leaf_classes() -> tuple:
    from ..modules.qkv_linear import GQAQKVColumnParallelLinear
    from ..parallel import PARALLEL_MODULES
    from ..parallel.layer_norm import LayerNorm, RMSNorm

    return tuple(PARALLEL_MODULES) + (GQAQKVColumnParallelLinear, RMSNorm, LayerNorm)


class NxDTracer(fx.Tracer):
    """torch.fx tracer whose leaves are the pipeline-atomic modules."""

    def __init__(self, leaf_classes: Sequence[Type[torch.nn.Module]] = (), autowrap_functions: Iterable[Callable] = (),
                 autowrap_modules: Iterable[Any] = ()):
        import math

        super().__init__(autowrap_modules=(math, *autowrap_modules), autowrap_functions=tuple(autowrap_functions))
        # leaves are given as classes or, as in the reference (``leaf_modules=["MyModule"]``), as class names
        self.leaf_classes = tuple(c for c in leaf_classes if isinstance(c, type))
        self.leaf_names = frozenset(c for c in leaf_classes if isinstance(c, str))

    def is_leaf_module(self, m: torch.nn.Module, module_qualified_name: str) -> bool:
        if isinstance(m, self.leaf_classes) or getattr(m, "_nxd_fx_leaf", False):
            return True
        if self.leaf_names and any(c.__name__ in self.leaf_names for c in type(m).__mro__):
            return True
        return super().is_leaf_module(m, module_qualified_name)


def get_concrete_args(model: torch.nn.Module, input_names: Optional[List[str]]) -> Dict[str, Any]:
    """Every ``forward`` parameter that is not a traced input is fixed to its default (reference ``:77-104``)."""
    sig = inspect.signature(model.forward)
    if input_names is None:
        return {}
    concrete = {}
    for name, prm in sig.parameters.items():
        if name in input_names or prm.kind in (prm.VAR_POSITIONAL, prm.VAR_KEYWORD):
            continue
        concrete[name] = None if prm.default is inspect.Parameter.empty else prm.default
    return concrete


def trace_model(model: torch.nn.Module, input_names: Optional[List[str]] = None,
                leaf_modules: Optional[Sequence[Type[torch.nn.Module]]] = None,
                autowrap_functions: Optional[Iterable[Callable]] = None, autowrap_modules: Optional[Iterable[Any]] = None,
                autowrap_obj_methods: Optional[Dict[Any, List[str]]] = None, tracer_cls=None) -> fx.GraphModule:
    """Symbolically trace ``model`` keeping ``leaf_modules`` (+ the parallel layers) opaque."""
    from ..parallel import PARALLEL_FUNCTIONS

    leaves = tuple(leaf_modules or ()) + _default_leaf_classes()
    wrap_fns = list(PARALLEL_FUNCTIONS) + list(autowrap_functions or [])
    wrap_fns += list(getattr(model, "_fx_wrap_functions", []))
    if tracer_cls is None or isinstance(tracer_cls, str):        # None / "torch" / "hf" as in the reference's pipeline_config
        tracer_cls = get_tracer_class(model, tracer_cls)
    # model code reaches the collectives as ``mappings.gather_from_...(x)``: a module attribute, which only
    # ``autowrap_modules`` keeps opaque (``autowrap_functions`` patches bare global names)
    from ..parallel import mappings as _mappings

    tracer = tracer_cls(leaf_classes=leaves, autowrap_functions=wrap_fns,
                        autowrap_modules=(_mappings,) + tuple(autowrap_modules or ()))
    patched = []
    if autowrap_obj_methods:  # reference ``:125-148``: make bound methods opaque call_function nodes
        for obj, names in autowrap_obj_methods.items():
            for n in names:
                orig = getattr(obj, n)
                fx.wrap(orig) if inspect.isfunction(orig) else None
                patched.append((obj, n, orig))
    try:
        graph = tracer.trace(model, concrete_args=get_concrete_args(model, input_names))
    finally:
        for obj, n, orig in patched:
            setattr(obj, n, orig)
    gm = fx.GraphModule(tracer.root, graph, class_name=model.__class__.__name__)
    gm.graph.eliminate_dead_code()
    gm.graph.lint()
    gm.recompile()
    return gm


class TorchTracerWrapper(NxDTracer):
    """Name of the reference's torch.fx tracer (``pipeline/trace.py:57-76``): ``NxDTracer`` with keyword configuration."""

    def __init__(self, **config):
        super().__init__(leaf_classes=tuple(config.get("leaf_modules", config.get("leaf_classes", ()))),
                         autowrap_functions=tuple(config.get("autowrap_functions", ())),
                         autowrap_modules=tuple(config.get("autowrap_modules", ())))


class HFTracerWrapper(TorchTracerWrapper):
    """Tracer selected for Hugging Face ``PreTrainedModel``s (reference ``pipeline/trace.py:24-55`` wraps transformers'
    ``HFTracer``).  ``transformers.utils.fx`` was removed in transformers 5 and its models' decorated ``forward`` is not
    symbolically traceable by plain torch.fx either, so this class exists for the selection logic (``tracer_cls="hf"``)
    and traces models written in torch.fx-friendly style -- the model families of this package, which load HF
    checkpoints through the converters -- with the same leaf rules as ``TorchTracerWrapper``."""


def _is_hf_model(model) -> bool:
    try:
        from transformers import PreTrainedModel

        return isinstance(model, PreTrainedModel)
    except Exception:  # noqa: BLE001
        return False


def get_tracer_class(model: torch.nn.Module = None, tracer_cls=None):
    """``"torch"`` / ``"hf"`` pick the wrapper, ``None`` picks by model type, a class is returned unchanged, any other
    string is an error (reference ``:107-122``)."""
    if tracer_cls is None:
        return HFTracerWrapper if _is_hf_model(model) else TorchTracerWrapper
    if tracer_cls == "torch":
        return TorchTracerWrapper
    if tracer_cls == "hf":
        return HFTracerWrapper
    if isinstance(tracer_cls, str):
        raise ValueError(f"Unsupported tracer_cls {tracer_cls}")
    return tracer_cls


@contextlib.contextmanager
def patch_obj_method(autowrap_obj_methods: Optional[Dict[Any, List[str]]]):
    """Inside the context the listed bound methods are opaque to ``torch.fx`` symbolic tracing: a call on a Proxy is
    recorded as ONE ``call_function`` node instead of being traced through (reference ``:126-148``)."""
    originals = []
    try:
        for obj, names in (autowrap_obj_methods or {}).items():
            assert isinstance(names, list), f"autowrap_obj_methods values must be lists, got {type(names)}"
            for name in names:
                if not hasattr(obj, name):
                    raise ValueError(f"{type(obj)} has no method {name}")
                fn = getattr(obj, name)
                originals.append((obj, name, fn))
                setattr(obj, name, fx._symbolic_trace._create_wrapped_func(fn))
        yield
    finally:
        for obj, name, fn in reversed(originals):
            setattr(obj, name, fn)
