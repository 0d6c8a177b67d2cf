"""Legacy TP/PP-sharded ``save`` / ``load`` (reference
``src/neuronx_distributed/parallel_layers/checkpointing.py:23-269``).

Layout: ``<dir>/tp_rank_XX_pp_rank_XX[_dp_rank_XX]/checkpoint.pt`` (plain) or
``<dir>/tp_rank_XX_pp_rank_XX[_dp_rank_XX]`` + ``.tensors/`` (xser mode).  Only the DP-rank-0
replica writes unless ``master_dp_only=False``.  ``load`` can also shard an unsharded checkpoint
on the fly via ``get_sharded_model_dict`` (+ ``preshard_hook``).
"""
from __future__ import annotations

import gc
import os
from typing import Any, Dict, Optional

import torch

from ..utils import xser
from ..utils.logger import get_logger
from . import parallel_state as ps
from .utils import add_barrier, cast_all, move_all_tensor_to_cpu

logger = get_logger()
NXD_SKIP_RENDEZVOUS = "NXD_SKIP_RENDEZVOUS"


def ensure_directory_exists(filename: str) -> None:
    os.makedirs(os.path.dirname(filename), exist_ok=True)


def _invoke_preshard_hook(module: Optional[torch.nn.Module], model_state_dict: Dict[str, Any], prefix: str = "") -> None:
    if module is None:
        return
    if hasattr(module, "preshard_hook"):
        module.preshard_hook(model_state_dict, prefix + "weight")
        return
    for name, child in module.named_children():
        _invoke_preshard_hook(child, model_state_dict, prefix + name + ".")


def get_sharded_model_dict(model: torch.nn.Module, model_state_dict: Dict[str, Any]) -> Dict[str, Any]:
    """Slice a full (unsharded) state dict for this TP rank using each parameter's
    ``partition_dim`` / ``partition_stride`` tags (reference :51-67)."""
    from .layers import create_local_weight

    inner = getattr(model, "original_torch_module", model)
    tp = ps.get_tensor_model_parallel_size()
    _invoke_preshard_hook(inner, model_state_dict)
    for name, param in inner.state_dict(keep_vars=True).items():
        if getattr(param, "tensor_model_parallel", False) and name in model_state_dict:
            if param.partition_dim not in (0, 1, 2):
                raise Exception(f"unsupported partition_dim {param.partition_dim}")
            full = model_state_dict[name]
            per = full.shape[param.partition_dim] // tp
            model_state_dict[name] = create_local_weight(full, param.partition_dim, per, param.partition_stride)
    return model_state_dict


def _chkpt_dir(output_dir: str, dp: bool) -> str:
    name = "tp_rank_{:02d}_pp_rank_{:02d}".format(ps.get_tensor_model_parallel_rank(),
                                                  ps.get_pipeline_model_parallel_rank())
    if dp:
        name += "_dp_rank_{:02d}".format(ps.get_data_parallel_rank())
    return os.path.join(output_dir, name)


def save(checkpoint: dict, output_dir: str, save_serially: bool = True, save_xser: bool = False,
         down_cast_bf16: bool = False, master_dp_only: bool = True) -> None:
    """Save a model/optimizer checkpoint, one file (tree) per (tp, pp[, dp]) rank."""
    base = _chkpt_dir(output_dir, dp=not master_dp_only)
    path = base if save_xser else os.path.join(base, "checkpoint.pt")
    should_write = (not master_dp_only) or ps.get_data_parallel_rank() == 0
    if down_cast_bf16:
        checkpoint = cast_all(checkpoint, torch.float32, torch.bfloat16)
    if should_write:
        ensure_directory_exists(path)
        if save_xser:
            xser.save(checkpoint, path)   # <path> (reference file) + <path>.tensors/tensor_N.pt
        elif save_serially:
            # stagger writers TP-rank by TP-rank to bound host memory (reference :114-134)
            cpu = move_all_tensor_to_cpu(checkpoint)
            torch.save(cpu, path)
            del cpu
            gc.collect()
        else:
            torch.save(move_all_tensor_to_cpu(checkpoint), path)
    add_barrier("nxd-save-checkpoint")


def load(output_dir: str, model_or_optimizer: Any = None, model_key: Optional[str] = "model", load_xser: bool = False,
         sharded: bool = True, strict: bool = True, master_dp_only: bool = True, weights_only: bool = False) -> Any:
    """Load what ``save`` wrote.  With ``sharded=False`` ``output_dir`` is a single full checkpoint
    file that is sliced for this rank.  Returns the loaded dict when no target object is given."""
    if sharded:
        base = _chkpt_dir(output_dir, dp=not master_dp_only)
        path = base if load_xser else os.path.join(base, "checkpoint.pt")
    else:
        path = output_dir
    if load_xser and sharded:
        ckpt = xser.load(path)
    else:
        ckpt = torch.load(path, map_location="cpu", weights_only=weights_only)
    if model_or_optimizer is None:
        add_barrier("nxd-load-checkpoint")
        return ckpt
    state = ckpt[model_key] if (model_key and isinstance(ckpt, dict) and model_key in ckpt) else ckpt
    if isinstance(model_or_optimizer, torch.nn.Module):
        if not sharded:
            state = get_sharded_model_dict(model_or_optimizer, state)
        model_or_optimizer.load_state_dict(state, strict=strict)
    else:
        model_or_optimizer.load_state_dict(state)
    del ckpt
    gc.collect()
    add_barrier("nxd-load-checkpoint")
    return None
