"""Legacy tracing API kept for source compatibility (reference ``src/neuronx_distributed/trace/trace.py``:
``parallel_model_trace`` / ``parallel_model_save`` / ``parallel_model_load``).  The reference spawns ``tp`` processes and
compiles one TorchScript per rank; here every rank already is a process, "tracing" captures a CUDA graph."""
from __future__ import annotations

import os
from typing import Any, Callable, Sequence

import torch

from ..parallel import parallel_state as ps
from .model_builder import BaseModelInstance, ModelBuilder, NxDModel


class TensorParallelModel(torch.nn.Module):
    """What ``parallel_model_trace`` returns: this rank's captured executable + module."""

    def __init__(self, nxd_model: NxDModel, key: str = "model"):
        super().__init__()
        self.nxd_model = nxd_model
        self.key = key

    def forward(self, *args):
        return self.nxd_model(*args)

    @property
    def module(self) -> torch.nn.Module:
        return self.nxd_model.models[self.key]


def parallel_model_trace(func: Callable[[], Any], example_inputs: Sequence[torch.Tensor], tp_degree: int = 1,
                         use_cuda_graphs: bool = True, **_) -> TensorParallelModel:
    """``func()`` returns the module (or ``(module, aliases)`` as in the reference)."""
    def build():
        r = func()
        return r[0] if isinstance(r, tuple) else r

    builder = ModelBuilder(tp_degree=tp_degree, use_cuda_graphs=use_cuda_graphs)
    builder.add("model", BaseModelInstance(build), [tuple(example_inputs)])
    return TensorParallelModel(builder.trace(initialize_model_weights=False))


def parallel_model_save(model: TensorParallelModel, save_dir: str) -> None:
    os.makedirs(save_dir, exist_ok=True)
    rank = ps.get_tensor_model_parallel_rank()
    torch.save(model.module.state_dict(), os.path.join(save_dir, f"tp_{rank}.pt"))


def parallel_model_load(model_or_func, load_dir: str, example_inputs: Sequence[torch.Tensor] = None, tp_degree: int = 1,
                        **kw) -> TensorParallelModel:
    """Load this rank's weights into a traced model (pass the traced model, or the builder function + example inputs)."""
    model = model_or_func if isinstance(model_or_func, TensorParallelModel) else parallel_model_trace(
        model_or_func, example_inputs, tp_degree, **kw)
    rank = ps.get_tensor_model_parallel_rank()
    sd = torch.load(os.path.join(load_dir, f"tp_{rank}.pt"), map_location="cpu", weights_only=False)
    model.module.load_state_dict(sd)
    return model
