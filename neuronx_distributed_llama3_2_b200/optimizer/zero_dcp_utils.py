"""ZeRO-1 optimizer state through ``torch.distributed.checkpoint`` (reference
``src/neuronx_distributed/optimizer/zero_dcp_utils.py:53-519``).

The reference avoids collectives by letting every rank compute all ranks' save plans and polling the file system for
``write_results.tmp.N``.  With NCCL/gloo available the stock DCP planner does the coordination; what is kept is the
naming contract: every tensor is stored under ``<fqn>|dp-DDDD|pp-PPPP|tp-TTTT`` so that shards of different ranks never
collide inside the one ``optim/`` DCP directory (``.metadata`` + ``__R_0.distcp``).
"""
from __future__ import annotations

from typing import Any, Dict

import torch
import torch.distributed as dist
import torch.distributed.checkpoint as dcp

from ..parallel import parallel_state as ps


def _suffix() -> str:
    if ps.model_parallel_is_initialized():
        return "|dp-{:04d}|pp-{:04d}|tp-{:04d}".format(ps.get_data_parallel_rank(),
                                                       ps.get_pipeline_model_parallel_rank(),
                                                       ps.get_tensor_model_parallel_rank())
    return "|dp-0000|pp-0000|tp-0000"


def _flatten(sd: Dict[str, Any]) -> Dict[str, torch.Tensor]:
    flat: Dict[str, torch.Tensor] = {}
    suf = _suffix()
    for i, t in sd.get("base_state", {}).items():
        flat[f"base_state.{i}{suf}"] = t
    for i, st in sd.get("state", {}).items():
        for k, v in st.items():
            if isinstance(v, torch.Tensor):
                flat[f"state.{i}.{k}{suf}"] = v.reshape(-1) if v.dim() == 0 else v
    return flat


def save_optim_state_dict(path: str, state_dict: Dict[str, Any]) -> None:
    flat = _flatten(state_dict)
    flat = {k: v for k, v in flat.items() if v.numel() > 0}
    meta = {k: v for k, v in state_dict.items() if k not in ("base_state", "state")}
    dcp.save(flat, storage_writer=dcp.FileSystemWriter(path), no_dist=not dist.is_initialized())
    if (not dist.is_initialized()) or dist.get_rank() == 0:
        torch.save(meta, f"{path}/nxd_meta.pt")
    torch.save({k: v for k, v in state_dict.items() if k in ("layout", "shape_info", "param_groups")},
               f"{path}/nxd_meta{_suffix().replace('|', '_')}.pt")


def load_optim_state_dict(path: str, template: Dict[str, Any]) -> Dict[str, Any]:
    """Fill ``template`` (a fresh ``optimizer.state_dict()`` with the right shapes) from the DCP directory."""
    flat = _flatten(template)
    nonempty = {k: v for k, v in flat.items() if v.numel() > 0}
    dcp.load(nonempty, storage_reader=dcp.FileSystemReader(path), no_dist=not dist.is_initialized())
    try:
        extra = torch.load(f"{path}/nxd_meta{_suffix().replace('|', '_')}.pt", map_location="cpu", weights_only=False)
        template["layout"]["step"] = extra["layout"].get("step", 0)
    except FileNotFoundError:
        pass
    return template


def get_dcp_aux_infos(model: torch.nn.Module, optim: torch.optim.Optimizer) -> Dict[str, Any]:
    """What a DCP save needs to know besides the tensors (reference ``:374-382``): optimizer parameter ids -> parameters
    and -> names, and the un-sharded shape of every parameter."""
    names = {id(p): n for n, p in model.named_parameters()}
    pid_to_params, pid_to_names, pid = {}, {}, 0
    for group in optim.param_groups:
        for p in group["params"]:
            pid_to_params[pid] = p
            pid_to_names[pid] = names.get(id(p), f"param_{pid}")
            pid += 1
    sd = optim.state_dict()
    return {"optim_pid_to_params": pid_to_params, "optim_pid_to_pnames": pid_to_names,
            "shape_info": sd.get("shape_info", {i: tuple(p.shape) for i, p in pid_to_params.items()})}
