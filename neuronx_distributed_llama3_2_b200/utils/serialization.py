"""Tensor <-> container (de)composition used by the pipeline engine.

Reference: ``src/neuronx_distributed/utils/serialization.py:14-253`` -- ``SerializationManager`` strips tensors out of
nested python containers into ``(stub, tensors, TensorMeta[])`` and rebuilds them on the receiving stage;
``find_loss_from_output_and_spec`` locates the loss inside a model output.
"""
from __future__ import annotations

import base64
import pickle
from dataclasses import dataclass
from typing import Any, List, NamedTuple, Optional, Tuple

import torch


def pickle_to_str(obj: Any) -> str:
    return base64.b64encode(pickle.dumps(obj)).decode("ascii")


def unpickle_from_str(s: str) -> Any:
    return pickle.loads(base64.b64decode(s.encode("ascii")))


class TensorMeta(NamedTuple):
    """What the receiving side needs to allocate the buffer of one tensor that travels separately from the pickled
    structure (reference ``utils/serialization.py:73-83``)."""

    tensor_index: int
    dtype: torch.dtype
    shape: torch.Size
    requires_grad: bool
    device: Optional[torch.device] = None


class _Stub:
    __slots__ = ("idx",)

    def __init__(self, idx):
        self.idx = idx


class SerializationManager:
    """``serialize(obj) -> (stub_obj, tensors, metas)``; ``deserialize(stub_obj, tensors) -> obj``."""

    def serialize(self, obj: Any):
        tensors: List[torch.Tensor] = []
        metas: List[TensorMeta] = []

        def walk(x):
            if isinstance(x, torch.Tensor):
                idx = len(tensors)
                tensors.append(x)
                metas.append(TensorMeta(idx, x.dtype, x.shape, x.requires_grad, x.device))
                return _Stub(idx)
            if isinstance(x, dict):
                return type(x)((k, walk(v)) for k, v in x.items())
            if isinstance(x, tuple) and hasattr(x, "_fields"):
                return type(x)(*[walk(v) for v in x])
            if isinstance(x, (list, tuple)):
                return type(x)(walk(v) for v in x)
            if isinstance(x, slice):
                return slice(walk(x.start), walk(x.stop), walk(x.step))
            return x

        return walk(obj), tensors, metas

    MAX_OBJECT_BYTES = 64 * 1024 * 1024

    def catch_and_raise_for_large_object(self, obj: Any) -> None:
        """The python structure around the tensors travels pickled through an object channel: refuse anything that is
        large once the tensors are taken out (it is almost always a tensor hidden in an unsupported container)."""
        import pickle

        stub, _, _ = self.serialize(obj)
        n = len(pickle.dumps(stub))
        if n > self.MAX_OBJECT_BYTES:
            raise ValueError(f"pipeline stage output serialises to {n} bytes of python objects (limit "
                             f"{self.MAX_OBJECT_BYTES}): pass tensors in lists / tuples / dicts / namedtuples")

    def extract_stubs(self, stub_obj: Any) -> List[Any]:
        """The tensor placeholders of a serialised structure, in tensor-index order."""
        found: List[_Stub] = []

        def walk(x):
            if isinstance(x, _Stub):
                found.append(x)
            elif isinstance(x, dict):
                for v in x.values():
                    walk(v)
            elif isinstance(x, (list, tuple)):
                for v in x:
                    walk(v)
            elif isinstance(x, slice):
                for v in (x.start, x.stop, x.step):
                    walk(v)

        walk(stub_obj)
        return sorted(found, key=lambda t: t.idx)

    def deserialize(self, stub_obj: Any, tensors: List[torch.Tensor]) -> Any:
        def walk(x):
            if isinstance(x, _Stub):
                return tensors[x.idx]
            if isinstance(x, dict):
                return type(x)((k, walk(v)) for k, v in x.items())
            if isinstance(x, tuple) and hasattr(x, "_fields"):
                return type(x)(*[walk(v) for v in x])
            if isinstance(x, (list, tuple)):
                return type(x)(walk(v) for v in x)
            if isinstance(x, slice):
                return slice(walk(x.start), walk(x.stop), walk(x.step))
            return x

        return walk(stub_obj)


def find_loss_from_output_and_spec(output_val: Any, spec_val: Any):
    """``spec`` mirrors the output structure with ``True`` at the loss position (reference ``:36-70``); ``None``
    / ``True`` alone means the output itself is the loss."""
    if spec_val is None or spec_val is True:
        if isinstance(output_val, torch.Tensor):
            return output_val
        if isinstance(output_val, (tuple, list)):
            return output_val[0]
        if isinstance(output_val, dict):
            return output_val["loss"]
        return getattr(output_val, "loss")
    if spec_val is False:
        return None
    if isinstance(spec_val, (tuple, list)):
        for o, s in zip(output_val, spec_val):
            r = find_loss_from_output_and_spec(o, s)
            if r is not None:
                return r
        return None
    if isinstance(spec_val, dict):
        for k, s in spec_val.items():
            o = output_val[k] if isinstance(output_val, dict) else getattr(output_val, k)
            r = find_loss_from_output_and_spec(o, s)
            if r is not None:
                return r
        return None
    raise RuntimeError(f"unsupported loss spec {spec_val!r}")


# names of the reference module (``utils/serialization.py:14-33``): base64 text transport of pickled objects between
# pipeline stages, and the namedtuple test used when tensors inside user structures are replaced by metadata
compress_to_string = pickle_to_str
uncompress_from_string = unpickle_from_str


def is_instance_namedtuple(obj: Any) -> bool:
    return isinstance(obj, tuple) and hasattr(obj, "_fields") and hasattr(obj, "_asdict")
