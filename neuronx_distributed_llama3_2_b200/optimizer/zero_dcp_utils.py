"""ZeRO-1 optimizer state through ``torch.distributed.checkpoint`` WITHOUT collectives (reference
``src/neuronx_distributed/optimizer/zero_dcp_utils.py:53-519``).

Why no collectives: the save runs on the asynchronous checkpoint thread (``trainer/checkpoint.py::CheckpointIOState``)
while the training thread keeps issuing NCCL work on the same communicators; a collective from a second thread would
interleave with it.  The reference gets there by letting every rank plan for every other rank and by having the
coordinator poll the file system for per-rank result files (``:289-370,436-466``).  The same protocol is used here, with
one difference that the flat-buffer ZeRO-1 engine needs: a rank cannot derive another rank's plan (pipeline stages own
different parameters, and a flat 1/dp slice cuts parameters at rank-dependent offsets), so each rank ships its OWN plan
next to its write results and the coordinator merges plans and results into ``.metadata``:

  every rank   plan locally -> write ``__<rank>_0.distcp`` -> pickle (plan, write results) to ``write_results.tmp.<rank>``
               (written as ``.tmp.tmp.<rank>`` and renamed, so a visible file is complete)
  coordinator  poll until all ``write_results.tmp.*`` exist (``MAX_RETRY`` seconds) -> global metadata from the plans ->
               ``.metadata`` -> delete the result files

Every tensor is stored under ``<key>|dp-DDDD|pp-PPPP|tp-TTTT`` (reference ``:166-212``), so shards of different ranks
never collide inside the one ``optim/`` directory.  Loading is rank-local as well: read ``.metadata``, plan for the
own keys, read.
"""
from __future__ import annotations

import dataclasses
import os
import pickle
import time
from typing import Any, Dict, List, Optional

import torch
import torch.distributed as dist
import torch.distributed.checkpoint as dcp
from torch.distributed.checkpoint.default_planner import (DefaultLoadPlanner, DefaultSavePlanner,
                                                          create_default_global_save_plan)
from torch.distributed.checkpoint.filesystem import _StoragePrefix
from torch.distributed.checkpoint.metadata import Metadata

from ..parallel import parallel_state as ps

MAX_RETRY = 100          # seconds the coordinator waits for the other ranks' result files (reference ``:53``)
_RESULT = "write_results.tmp.{}"


def _suffix() -> str:
    if ps.model_parallel_is_initialized():
        return "|dp-{:04d}|pp-{:04d}|tp-{:04d}".format(ps.get_data_parallel_rank(),
                                                       ps.get_pipeline_model_parallel_rank(),
                                                       ps.get_tensor_model_parallel_rank())
    return "|dp-0000|pp-0000|tp-0000"


def _rank_world():
    if dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


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


def _local_save_plan(flat: Dict[str, torch.Tensor], rank: int):
    planner = DefaultSavePlanner()
    planner.set_up_planner(flat, is_coordinator=False)
    plan = planner.create_local_plan()
    # what FileSystemWriter.prepare_global_plan would assign to the rank-th plan: data goes to ``__<rank>_0.distcp``
    plan = dataclasses.replace(plan, storage_data=_StoragePrefix(f"__{rank}_"))
    return planner, planner.finish_plan(plan)


def save_optim_state_dict(path: str, state_dict: Dict[str, Any], aux_infos: Optional[Dict[str, Any]] = None,
                          dedup: bool = False) -> Optional[Metadata]:
    """Write this rank's ZeRO-1 shards into the DCP directory ``path``; no collective is issued, so the call is safe on
    the asynchronous checkpoint thread.  Returns the global metadata on the coordinator (global rank 0), else None."""
    rank, world = _rank_world()
    os.makedirs(path, exist_ok=True)
    flat = {k: v.detach().cpu() for k, v in _flatten(state_dict).items() if v.numel() > 0}
    writer = dcp.FileSystemWriter(path)
    writer.set_up_storage_writer(rank == 0)
    planner, plan = _local_save_plan(flat, rank)
    fut = writer.write_data(plan, planner)
    fut.wait()
    results = fut.value()

    # rank-private side file: the non-tensor part of the state dict (param groups, flat layout, shapes)
    torch.save({k: v for k, v in state_dict.items() if k in ("layout", "shape_info", "param_groups")},
               os.path.join(path, f"nxd_meta{_suffix().replace('|', '_')}.pt"))

    # the plan carries tensor metadata only (no data): ship it with the results; rename = atomic publication
    tmp = os.path.join(path, "write_results.tmp.tmp.{}".format(rank))
    with open(tmp, "wb") as f:
        pickle.dump({"plan": dataclasses.replace(plan, planner_data=None), "results": results}, f)
        f.flush()
        os.fsync(f.fileno())
    os.rename(tmp, os.path.join(path, _RESULT.format(rank)))
    if rank != 0:
        return None

    files = [os.path.join(path, _RESULT.format(r)) for r in range(world)]
    waited = 0
    while not all(os.path.exists(p) for p in files):
        if waited >= MAX_RETRY:
            missing = [r for r, p in enumerate(files) if not os.path.exists(p)]
            raise RuntimeError(f"zero1 DCP save: no result file from ranks {missing} after {MAX_RETRY} s under {path}")
        time.sleep(1.0)
        waited += 1
    plans, all_results = [], []
    for p in files:
        with open(p, "rb") as f:
            rec = pickle.load(f)
        plans.append(rec["plan"])
        all_results.append(rec["results"])
    _, metadata = create_default_global_save_plan(plans)
    writer.finish(metadata=metadata, results=all_results)
    for p in files:
        os.remove(p)
    torch.save({k: v for k, v in state_dict.items() if k not in ("base_state", "state")},
               os.path.join(path, "nxd_meta.pt"))
    return metadata


def wait_for_metadata(path: str, timeout_s: float = 2.0 * MAX_RETRY) -> None:
    """Block until the coordinator has published ``.metadata`` (file-system poll, no collective)."""
    t0 = time.time()
    while not os.path.exists(os.path.join(path, ".metadata")):
        if time.time() - t0 > timeout_s:
            raise RuntimeError(f"zero1 DCP: {path}/.metadata did not appear within {timeout_s:.0f} s")
        time.sleep(0.2)


def load_optim_state_dict(path: str, template: Dict[str, Any], aux_infos: Optional[Dict[str, Any]] = None,
                          dedup: bool = False) -> Dict[str, Any]:
    """Fill ``template`` (a fresh ``optimizer.state_dict()`` with this rank's shard shapes) from the DCP directory.
    Rank-local: metadata read, local plan, reads."""
    wait_for_metadata(path, timeout_s=5.0)
    flat = {k: v for k, v in _flatten(template).items() if v.numel() > 0}
    staging = {k: torch.empty_like(v, device="cpu") for k, v in flat.items()}
    reader = dcp.FileSystemReader(path)
    metadata = reader.read_metadata()
    missing = [k for k in staging if k not in metadata.state_dict_metadata]
    if missing:
        raise KeyError(f"zero1 DCP load: {len(missing)} keys of this rank are not in {path}/.metadata, e.g. {missing[:3]} "
                       "(saved with another dp/tp/pp layout? convert with nxd_convert_zero_checkpoints)")
    planner = DefaultLoadPlanner()
    planner.set_up_planner(staging, metadata, is_coordinator=False)
    reader.set_up_storage_reader(metadata, is_coordinator=False)
    plan = reader.prepare_local_plan(planner.create_local_plan())
    plan = reader.prepare_global_plan(planner.create_global_plan([plan]))[0]
    reader.read_data(planner.finish_plan(plan), planner).wait()
    for k, v in flat.items():
        v.copy_(staging[k].to(v.device))
    try:
        extra = torch.load(os.path.join(path, f"nxd_meta{_suffix().replace('|', '_')}.pt"), map_location="cpu",
                           weights_only=False)
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
