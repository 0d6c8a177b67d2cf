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


# ---- names of the reference's trace module (``trace/trace.py``) -----------------------------------------------------
from .checkpoint_sharding import (_mock_parallel_state, get_sharded_checkpoint, invoke_preshard_hook,  # noqa: E402,F401
                                  shard_children)

ParallelModel = TensorParallelModel
TensorParallelNeuronModel = TensorParallelModel     # reference: thread pool over per-rank NEFFs in ONE process; here one
                                                    # process per GPU runs its own rank


def generate_ranked_folder(tp_rank: int, bucket_rank: int, bucket_degree: int) -> str:
    """Folder suffix of one (rank, bucket) artefact (reference ``:162-163``)."""
    return f"_tp{tp_rank}" + (f"_bk{bucket_rank}" if bucket_degree > 1 else "")


def find_unique_dtypes(model: torch.nn.Module) -> dict:
    """``{dtype: number of state-dict entries}`` (reference ``:372-377``)."""
    out: dict = {}
    for v in model.state_dict().values():
        out[v.dtype] = out.get(v.dtype, 0) + 1
    return out


def create_local_weight(rank: int, world_size: int, full_weight: torch.Tensor, partition_dim: int, per_partition_size: int,
                        stride: int, out_weight: torch.Tensor = None) -> torch.Tensor:
    """Slice of ``full_weight`` owned by ``rank`` (strided column groups honoured), reference argument order
    (``:643-669``)."""
    from ..parallel.layers import create_local_weight as _clw

    w = _clw(full_weight, partition_dim, per_partition_size, stride, rank=rank, world_size=world_size)
    if out_weight is not None:
        out_weight.copy_(w)
        return out_weight
    return w


def create_local_weight_qkv(rank: int, world_size: int, full_weight: torch.Tensor, partition_dim: int, q_len: int,
                            kv_len: int, out_weight: torch.Tensor = None) -> torch.Tensor:
    """Fused ``[q | k | v]`` weight: every part is sharded on its own, then the rank's three slices are re-fused
    (reference ``:629-641``)."""
    q, k, v = torch.split(full_weight, [q_len, kv_len, kv_len], dim=partition_dim)
    parts = [torch.chunk(t, world_size, dim=partition_dim)[rank] for t in (q, k, v)]
    w = torch.cat(parts, dim=partition_dim)
    if out_weight is not None:
        out_weight.copy_(w)
        return out_weight
    return w


def collect_tp_neuron_models(models: list, results, bucket_config=None, tp_degree: int = 1) -> None:
    """Gather per-rank artefacts produced by tracing workers into ``models[tp_rank]`` (reference ``:115-120`` drains a
    multiprocessing queue of NEFFs; here ``results`` is any iterable of ``(tp_rank, model)``)."""
    for tp_rank, model in results:
        models[tp_rank] = model


def collect_tp_bucket_neuron_models(models: list, results, bucket_config=None, tp_degree: int = 1) -> None:
    """Same for bucketed models: ``results`` yields ``(tp_rank, bucket_rank, model)``; ``models[tp_rank]`` becomes the
    list of that rank's buckets in bucket order."""
    per_rank: dict = {}
    for tp_rank, bucket_rank, model in results:
        per_rank.setdefault(tp_rank, {})[bucket_rank] = model
    for tp_rank, buckets in per_rank.items():
        models[tp_rank] = [buckets[b] for b in sorted(buckets)]
