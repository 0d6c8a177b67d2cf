"""Pipeline point-to-point communication.

Reference: ``src/neuronx_distributed/pipeline/comm.py:38-197`` emulates send/recv with a 2-rank all-gather (XLA has no
p2p) and ships python metadata through the TCP store or a padded gloo buffer.  Here tensors travel with real
``isend`` / ``irecv`` (NCCL p2p on its own stream, overlapping compute; gloo on CPU) and metadata with
``send_object_list`` / ``recv_object_list`` -- only during the first (shape-inference) micro-batch.
"""
from __future__ import annotations

from typing import Any, List, Optional

import torch
import torch.distributed as dist

from ..parallel import parallel_state as ps
from ..utils.serialization import TensorMeta


_GRAD_GROUP = None
_TAG_ACT, _TAG_GRAD = 11, 12


def init_p2p_groups() -> None:
    """Activations and gradients are two logical channels.  With pp == 2 (or wrap-around traffic of the interleaved
    schedule) both directions connect the SAME pair of ranks, and p2p matching is FIFO per pair: gloo separates the
    channels by tag, NCCL (no tags) gets a second communicator for the gradient channel.  Collective: every rank
    must call this once after ``initialize_model_parallel``."""
    global _GRAD_GROUP
    if _GRAD_GROUP is not None or not dist.is_initialized():
        return
    if dist.get_backend(ps.get_pipeline_model_parallel_group()) != "nccl":
        _GRAD_GROUP = ps.get_pipeline_model_parallel_group()
        return
    mine = None
    for row in ps.get_pipeline_model_parallel_group(as_list=True):
        g = dist.new_group(ranks=row)
        if dist.get_rank() in row:
            mine = g
    _GRAD_GROUP = mine


def reset_p2p_groups() -> None:
    global _GRAD_GROUP
    _GRAD_GROUP = None


def _pp_group(grad: bool = False):
    if grad and _GRAD_GROUP is not None:
        return _GRAD_GROUP
    return ps.get_pipeline_model_parallel_group()


def _next_rank() -> int:
    return ps.get_pipeline_model_parallel_next_rank()


def _prev_rank() -> int:
    return ps.get_pipeline_model_parallel_prev_rank()


class _Pending:
    """Outstanding isends: buffers must stay alive until the work completes."""

    def __init__(self):
        self.items = []

    def add(self, work, tensor):
        self.items.append((work, tensor))
        if len(self.items) > 64:
            self.drain(keep=32)

    def drain(self, keep: int = 0):
        while len(self.items) > keep:
            w, _ = self.items.pop(0)
            w.wait()


_pending = _Pending()


def send(tensor: torch.Tensor, send_next: bool = True, all_reduce_send_recv: bool = False) -> None:
    """Asynchronous send to the next (or previous) pipeline stage."""
    dst = _next_rank() if send_next else _prev_rank()
    t = tensor.detach().contiguous()
    grad = not send_next   # activations flow to the next stage, gradients to the previous one
    work = dist.isend(t, dst=dst, group=_pp_group(grad), tag=_TAG_GRAD if grad else _TAG_ACT)
    _pending.add(work, t)


def recv_from(meta: TensorMeta, recv_prev: bool = True, device=None, all_reduce_send_recv: bool = False) -> torch.Tensor:
    """Blocking (stream-ordered on CUDA) receive of a tensor described by ``meta``."""
    src = _prev_rank() if recv_prev else _next_rank()
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(_pp_group()) == "nccl" \
            else torch.device("cpu")
    buf = torch.empty(tuple(meta.shape), dtype=meta.dtype, device=device)
    grad = not recv_prev
    dist.irecv(buf, src=src, group=_pp_group(grad), tag=_TAG_GRAD if grad else _TAG_ACT).wait()
    return buf


def flush_sends() -> None:
    _pending.drain(0)


def send_python_object(obj: Any, send_next: bool = True, method: Optional[str] = None) -> None:
    dst = _next_rank() if send_next else _prev_rank()
    dist.send_object_list([obj], dst=dst, group=_pp_group())


def recv_python_object(recv_prev: bool = True, method: Optional[str] = None) -> Any:
    src = _prev_rank() if recv_prev else _next_rank()
    out: List[Any] = [None]
    dist.recv_object_list(out, src=src, group=_pp_group())
    return out[0]
