"""Tensor-parallel parameter attributes and small helpers.

Reference: ``src/neuronx_distributed/parallel_layers/utils.py:25-266``.  The
parameter-attribute contract (``tensor_model_parallel`` / ``partition_dim`` /
``partition_stride`` / ``sequence_parallel_enabled`` / ``expert_model_parallel`` /
``shared``) is what checkpoint sharding, grad-norm and ZeRO-1 key on, so it is kept
verbatim; everything XLA-specific (``mark_step`` syncs, fp64 tricks) is gone.
"""
from __future__ import annotations

import os
from typing import Any, Sequence

import torch

_MODEL_PARALLEL_ATTRIBUTE_DEFAULTS = {
    "tensor_model_parallel": False,
    "partition_dim": -1,
    "partition_stride": 1,
}
_EXTRA_PARAM_TAGS = ("sequence_parallel_enabled", "expert_model_parallel", "shared")


class EmbeddingUtility:
    """Vocab range owned by a TP rank: ``[first, last)`` (reference utils.py:25-39)."""

    @staticmethod
    def range_from_per_partition_vocab_size(per_partition_vocab_size: int, rank: int, world_size: int):
        first = rank * per_partition_vocab_size
        return first, first + per_partition_vocab_size

    @staticmethod
    def range_from_global_vocab_size(global_vocab_size: int, rank: int, world_size: int):
        return EmbeddingUtility.range_from_per_partition_vocab_size(
            divide(global_vocab_size, world_size), rank, world_size
        )


def param_is_not_tensor_parallel_duplicate(param: torch.Tensor) -> bool:
    from . import parallel_state as ps

    return getattr(param, "tensor_model_parallel", False) or ps.get_tensor_model_parallel_rank() == 0


def set_tensor_model_parallel_attributes(tensor: torch.Tensor, is_parallel: bool, dim: int, stride: int = 1,
                                         num_partitions: int = -1) -> None:
    for attr in _MODEL_PARALLEL_ATTRIBUTE_DEFAULTS:
        assert not hasattr(tensor, attr), f"{attr} already set"
    setattr(tensor, "tensor_model_parallel", is_parallel)
    setattr(tensor, "partition_dim", dim)
    setattr(tensor, "partition_stride", stride)
    if num_partitions > 0:
        setattr(tensor, "num_partitions", num_partitions)


def set_defaults_if_not_set_tensor_model_parallel_attributes(tensor: torch.Tensor) -> None:
    for attr, default in _MODEL_PARALLEL_ATTRIBUTE_DEFAULTS.items():
        if not hasattr(tensor, attr):
            setattr(tensor, attr, default)


def copy_tensor_model_parallel_attributes(destination_tensor: torch.Tensor, source_tensor: torch.Tensor) -> None:
    for attr in list(_MODEL_PARALLEL_ATTRIBUTE_DEFAULTS) + list(_EXTRA_PARAM_TAGS) + ["num_partitions"]:
        if hasattr(source_tensor, attr):
            setattr(destination_tensor, attr, getattr(source_tensor, attr))


def divide(numerator: int, denominator: int) -> int:
    assert numerator % denominator == 0, f"{numerator} is not divisible by {denominator}"
    return numerator // denominator


def split_tensor_along_last_dim(tensor: torch.Tensor, num_partitions: int,
                                contiguous_split_chunks: bool = False) -> Sequence[torch.Tensor]:
    return split_tensor_along_dim(tensor, tensor.dim() - 1, num_partitions, contiguous_split_chunks)


def split_tensor_along_second_dim(tensor, num_partitions, contiguous_split_chunks: bool = False):
    return split_tensor_along_dim(tensor, 1, num_partitions, contiguous_split_chunks)


def split_tensor_along_dim(tensor: torch.Tensor, dim: int, num_partitions: int,
                           contiguous_split_chunks: bool = False) -> Sequence[torch.Tensor]:
    size = divide(tensor.size(dim), num_partitions)
    chunks = torch.split(tensor, size, dim=dim)
    if contiguous_split_chunks:
        return tuple(c.contiguous() for c in chunks)
    return chunks


def get_padding_length(numerator: int, denominator: int) -> int:
    """How much to add to ``numerator`` so it divides by ``denominator``."""
    return (denominator - numerator % denominator) % denominator


def cast_tensor(tensor, from_dtype=torch.float32, to_dtype=torch.bfloat16):
    return tensor.to(dtype=to_dtype) if tensor.dtype == from_dtype else tensor


def cast_all(state: Any, from_dtype=torch.float32, to_dtype=torch.bfloat16) -> Any:
    """Recursively cast tensors of ``from_dtype`` inside nested containers."""
    if isinstance(state, torch.Tensor):
        return cast_tensor(state, from_dtype, to_dtype)
    if isinstance(state, dict):
        return {k: cast_all(v, from_dtype, to_dtype) for k, v in state.items()}
    if isinstance(state, (list, tuple)):
        return type(state)(cast_all(v, from_dtype, to_dtype) for v in state)
    return state


def cast_if_autocast_enabled(*args):
    """Cast floating tensors to the autocast dtype of the device of the first
    tensor (reference utils.py:185-210, minus the XLA special-casing)."""
    dev = next((a.device.type for a in args if isinstance(a, torch.Tensor)), "cpu")
    if not torch.is_autocast_enabled(dev):
        return args
    dt = torch.get_autocast_dtype(dev)
    return tuple(a.to(dt) if isinstance(a, torch.Tensor) and a.is_floating_point() else a for a in args)


def move_all_tensor_to_cpu(data: Any, convert: bool = True) -> Any:
    """Nested-container D2H copy (reference utils.py:213-226)."""
    if isinstance(data, torch.Tensor):
        return data.detach().to("cpu") if convert else data
    if isinstance(data, dict):
        return {k: move_all_tensor_to_cpu(v, convert) for k, v in data.items()}
    if isinstance(data, (list, tuple)):
        return type(data)(move_all_tensor_to_cpu(v, convert) for v in data)
    return data


def get_local_world_size() -> int:
    if "LOCAL_WORLD_SIZE" in os.environ:
        return int(os.environ["LOCAL_WORLD_SIZE"])
    if torch.cuda.is_available():
        return max(torch.cuda.device_count(), 1)
    return 1


def move_model_to_device(model: torch.nn.Module, device) -> torch.nn.Module:
    from ..utils.model_utils import move_model_to_device as _impl

    return _impl(model, device)


def is_torch_version_greater_than_2() -> bool:
    return int(torch.__version__.split(".")[0]) >= 2


def requires_init_pg_override() -> bool:
    return False


def add_barrier(name: str = "nxd-barrier") -> None:
    """``xm.rendezvous`` replacement: a world barrier when distributed is up."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        if os.environ.get("NXD_SKIP_RENDEZVOUS", "0") == "1":
            return
        dist.barrier()


def ensure_divisibility(numerator: int, denominator: int) -> None:
    """Assert that ``numerator`` is divisible by ``denominator`` (reference ``utils.py:78-80``)."""
    assert numerator % denominator == 0, f"{numerator} is not divisible by {denominator}"


def is_pjrt_device() -> bool:
    """The reference asks whether torch_xla runs on the PJRT runtime (``utils.py``); there is no XLA here."""
    return False


def verify_casted_dtype(value, _device_type=None) -> None:
    """Under autocast, check that every floating tensor inside ``value`` already has the autocast dtype (reference
    ``utils.py:249-266``; used after ``cast_if_autocast_enabled``).  Autocast is looked up for the device type of the
    tensor itself (CUDA and CPU autocast are separate switches in torch)."""
    import collections.abc

    if isinstance(value, torch.Tensor):
        dev = _device_type or value.device.type
        if not torch.is_autocast_enabled(dev) or not value.is_floating_point():
            return
        want = torch.get_autocast_dtype(dev)
        assert value.dtype == want, f"Datatype of tensor is expected to be {want}, got {value.dtype} instead"
    elif isinstance(value, collections.abc.Mapping):
        for k, v in value.items():
            verify_casted_dtype(k, _device_type)
            verify_casted_dtype(v, _device_type)
    elif isinstance(value, (list, tuple, set)):
        for v in value:
            verify_casted_dtype(v, _device_type)
