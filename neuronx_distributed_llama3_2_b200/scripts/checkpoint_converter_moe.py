"""Full checkpoint <-> (tp, pp, ep)-sharded checkpoint for ANY model of this package, driven by the model's own parameter
tags instead of a name table (reference ``examples/training/mixtral/convert_checkpoints.py``:
``CheckpointConverterMixtral`` stacks the per-expert HF weights and shards ``gate_up_proj`` on dim 2, ``down_proj`` on dim 1,
experts on dim 0).

A shape-only replica of the model is built on the ``meta`` device under a mocked parallel state for every target
coordinate; its ``tensor_model_parallel / partition_dim / partition_stride / fused_qkv / expert_model_parallel`` tags say
how each tensor is cut (``inference.checkpoint_sharding.shard_children`` does the TP part, expert slabs are chunked on
dim 0 over EP, layers are dealt to PP stages as ``NxDPPModel`` auto-partitioning would).  Merging reads the same tags.

File layout is the trainer's (``trainer/checkpoint.py``): ``<out>/model/dp_rank_00_tp_rank_TT_pp_rank_PP.pt`` and, with
expert parallelism, ``dp_rank_00_ep_rank_EE_tp_rank_TT_pp_rank_PP.pt``.  The trainer writes one COMPLETE file per EP rank
(dense tensors included, the reference layout); the converter's own output keeps the dense remainder once per (tp, pp) in
the ep-less file, and both layouts are read back (``load_coord``); ``checkpoint`` / ``done`` markers.
"""
from __future__ import annotations

import argparse
import contextlib
import json
import os
import re
from typing import Any, Callable, Dict, List, Optional

import torch

from ..inference.checkpoint_sharding import shard_children
from ..parallel import parallel_state as ps
from ..pipeline.partition import create_partitions
from ..utils import xser


@contextlib.contextmanager
def mock_parallel_state(tp: int, tp_rank: int, ep: int = 1, ep_rank: int = 0):
    ps.set_tensor_model_parallel_size(tp)
    ps.set_tensor_model_parallel_rank(tp_rank)
    ps.set_expert_model_parallel_size(ep)
    ps.set_expert_model_parallel_rank(ep_rank)
    try:
        yield
    finally:
        for fn in (ps.set_tensor_model_parallel_size, ps.set_tensor_model_parallel_rank,
                   ps.set_expert_model_parallel_size, ps.set_expert_model_parallel_rank):
            fn(None)


class ModelDrivenCheckpointConverter:
    """``model_fn()`` builds the model (it is called under ``torch.device('meta')``); ``to_model_names`` / ``to_full_names``
    translate a full external (e.g. HF) state dict to / from the model's own un-sharded names."""

    layer_pattern = re.compile(r"(?:^|\.)layers\.(\d+)\.")

    def __init__(self, model_fn: Callable[[], torch.nn.Module], num_layers: int, tp_size: int = 1, pp_size: int = 1,
                 ep_size: int = 1, to_model_names: Optional[Callable[[Dict], Dict]] = None,
                 to_full_names: Optional[Callable[[Dict], Dict]] = None, tied_output: Optional[str] = None,
                 tied_input: Optional[str] = None):
        self.model_fn, self.L = model_fn, num_layers
        self.tp, self.pp, self.ep = tp_size, pp_size, ep_size
        self.to_model_names = to_model_names or (lambda sd: sd)
        self.to_full_names = to_full_names or (lambda sd: sd)
        self.tied_output, self.tied_input = tied_output, tied_input       # e.g. "lm_head.weight" <- "embed_tokens.weight"

    # -- helpers ---------------------------------------------------------------------------------------------------
    def _meta_model(self, tp_rank: int, ep_rank: int) -> torch.nn.Module:
        with mock_parallel_state(self.tp, tp_rank, self.ep, ep_rank), torch.device("meta"):
            return self.model_fn()

    def stage_of_key(self, key: str) -> List[int]:
        """Layers are split evenly (remainder to the later stages); what precedes the layers lives on the first stage,
        what follows on the last; a tied output weight lives on both ends."""
        if self.pp == 1:
            return [0]
        m = self.layer_pattern.search(key)
        if m:
            cuts = create_partitions(self.pp, self.L)
            layer = int(m.group(1))
            return [next((s for s, c in enumerate(cuts) if layer <= c), self.pp - 1)]
        if key == self.tied_input and self.tied_output:
            return [0, self.pp - 1]
        first = any(t in key for t in ("embed", "wte", "embeddings"))
        return [0] if first else [self.pp - 1]

    @staticmethod
    def shard_file(out_dir: str, tp: int, pp: int, ep: Optional[int] = None) -> str:
        ep_s = "" if ep is None else f"_ep_rank_{ep:02d}"
        return os.path.join(out_dir, "model", f"dp_rank_00{ep_s}_tp_rank_{tp:02d}_pp_rank_{pp:02d}.pt")

    def _ep_arg(self, e: int) -> Optional[int]:
        return e if self.ep > 1 else None

    def _expert_keys(self, tp_rank: int, ep_rank: int) -> set:
        return {n for n, p in self._meta_model(tp_rank, ep_rank).state_dict(keep_vars=True).items()
                if getattr(p, "expert_model_parallel", False)}

    # -- full -> sharded ---------------------------------------------------------------------------------------------
    def shard(self, full: Dict[str, torch.Tensor], tp_rank: int, ep_rank: int = 0) -> Dict[str, torch.Tensor]:
        """One (tp, ep) coordinate of the model-named full state (all PP stages)."""
        model = self._meta_model(tp_rank, ep_rank)
        state = dict(full)
        if self.ep > 1:
            for name, p in model.state_dict(keep_vars=True).items():
                if getattr(p, "expert_model_parallel", False) and name in state:
                    state[name] = state[name].chunk(self.ep, dim=0)[ep_rank]
        with mock_parallel_state(self.tp, tp_rank, self.ep, ep_rank):
            return shard_children(model, state, self.tp, tp_rank)

    def convert_from_full_state(self, external_state: Dict[str, torch.Tensor], out_dir: str, use_xser: bool = False) -> None:
        full = self.to_model_names(dict(external_state))
        if self.tied_output and self.tied_output not in full and self.tied_input in full:
            full[self.tied_output] = full[self.tied_input]
        os.makedirs(os.path.join(out_dir, "model"), exist_ok=True)
        save = xser.save if use_xser else torch.save
        for e in range(self.ep):
            for t in range(self.tp):
                shard = self.shard(full, t, e)
                expert_keys = self._expert_keys(t, e) if self.ep > 1 else set()
                for s in range(self.pp):
                    part = {k: v.contiguous().clone() for k, v in shard.items() if s in self.stage_of_key(k)}
                    if self.ep > 1:
                        save({k: v for k, v in part.items() if k in expert_keys}, self.shard_file(out_dir, t, s, e))
                        if e == 0:
                            save({k: v for k, v in part.items() if k not in expert_keys}, self.shard_file(out_dir, t, s))
                    else:
                        save(part, self.shard_file(out_dir, t, s))
        for marker in ("checkpoint", "done"):
            open(os.path.join(out_dir, marker), "w").write("1")

    # -- sharded -> full ---------------------------------------------------------------------------------------------
    def _merge_tp(self, model: torch.nn.Module, shards: List[Dict[str, torch.Tensor]]) -> Dict[str, torch.Tensor]:
        out: Dict[str, torch.Tensor] = {}
        params = model.state_dict(keep_vars=True)
        for name in shards[0]:
            vs = [s[name] for s in shards]
            p = params.get(name)
            if p is None or not getattr(p, "tensor_model_parallel", False) or self.tp == 1:
                out[name] = vs[0]
                continue
            dim, stride = p.partition_dim, getattr(p, "partition_stride", 1)
            if getattr(p, "fused_qkv", False):
                owner = model.get_submodule(name.rpartition(".")[0])
                q_sz, kv_sz = owner.q_output_size_per_partition, owner.kv_output_size_per_partition
                parts = [torch.split(v, [q_sz, kv_sz, kv_sz], dim=0) for v in vs]
                kv_full = owner.output_sizes[1]
                out[name] = torch.cat([torch.cat([x[0] for x in parts], 0), torch.cat([x[1] for x in parts], 0)[:kv_full],
                                       torch.cat([x[2] for x in parts], 0)[:kv_full]], 0)
            elif stride > 1:          # rank r holds its chunk of each of the ``stride`` fused matrices
                pieces = [v.chunk(stride, dim) for v in vs]
                out[name] = torch.cat([torch.cat([pc[j] for pc in pieces], dim) for j in range(stride)], dim)
            else:
                merged = torch.cat(vs, dim)
                mult = getattr(p, "kv_size_multiplier", 1)
                if mult > 1:
                    merged = merged.narrow(dim, 0, merged.shape[dim] // mult)
                out[name] = merged
        return out

    def convert_to_full_state(self, in_dir: str, load_xser: Optional[bool] = None) -> Dict[str, torch.Tensor]:
        def load(path):
            is_x = os.path.isdir(path + ".tensors") if load_xser is None else load_xser
            return xser.load(path) if is_x else torch.load(path, map_location="cpu", weights_only=False)

        def load_coord(t, s, e):
            """Everything rank (tp t, pp s, ep e) holds: its expert file plus the dense file shared by the EP ranks."""
            if self.ep == 1:
                return load(self.shard_file(in_dir, t, s))
            state = dict(load(self.shard_file(in_dir, t, s, e)))
            dense = self.shard_file(in_dir, t, s)
            if os.path.exists(dense) or os.path.isdir(dense + ".tensors"):
                state.update(load(dense))
            return state

        per_ep: List[Dict[str, torch.Tensor]] = []
        for e in range(self.ep):
            model = self._meta_model(0, e)
            merged: Dict[str, torch.Tensor] = {}
            for s in range(self.pp):
                merged.update(self._merge_tp(model, [load_coord(t, s, e) for t in range(self.tp)]))
            per_ep.append(merged)
        full = dict(per_ep[0])
        if self.ep > 1:
            params = self._meta_model(0, 0).state_dict(keep_vars=True)
            for name, p in params.items():
                if getattr(p, "expert_model_parallel", False) and name in full:
                    full[name] = torch.cat([m[name] for m in per_ep], dim=0)
        if self.tied_output and self.tied_input in full:
            full.pop(self.tied_output, None)
        return self.to_full_names(full)


class CheckpointConverterMixtral(ModelDrivenCheckpointConverter):
    """HF Mixtral checkpoint (``block_sparse_moe.experts.N.w1/w2/w3``) <-> shards of ``MixtralForCausalLM``."""

    def __init__(self, config, tp_size: int = 1, pp_size: int = 1, ep_size: int = 1):
        from ..models.mixtral import (MixtralForCausalLM, convert_mixtral_hf_state_dict,
                                      convert_mixtral_state_dict_to_hf)

        self.config = config
        super().__init__(lambda: MixtralForCausalLM(config, device="meta"), config.num_hidden_layers, tp_size, pp_size,
                         ep_size, to_model_names=lambda sd: convert_mixtral_hf_state_dict(sd, config),
                         to_full_names=lambda sd: convert_mixtral_state_dict_to_hf(sd, config))

    @staticmethod
    def get_arg_parser() -> argparse.ArgumentParser:
        ap = argparse.ArgumentParser("mixtral checkpoint converter")
        ap.add_argument("--input_dir", required=True)
        ap.add_argument("--output_dir", required=True)
        ap.add_argument("--config", required=True, help="HF config.json of the model")
        ap.add_argument("--tp_size", type=int, default=1)
        ap.add_argument("--pp_size", type=int, default=1)
        ap.add_argument("--ep_size", type=int, default=1)
        ap.add_argument("--convert_from_full_state", action="store_true")
        ap.add_argument("--convert_to_full_state", action="store_true")
        ap.add_argument("--load_xser", action="store_true")
        ap.add_argument("--save_xser", action="store_true")
        return ap

    @classmethod
    def run(cls, argv=None) -> None:
        from dataclasses import fields

        from ..inference.checkpoint import load_state_dict
        from ..models.mixtral import MixtralConfig

        a = cls.get_arg_parser().parse_args(argv)
        raw = json.load(open(a.config))
        known = {f.name for f in fields(MixtralConfig)}
        kw: Dict[str, Any] = {k: v for k, v in raw.items() if k in known and k != "torch_dtype"}
        kw.setdefault("rope_scaling", None)
        cfg = MixtralConfig(**kw)
        conv = cls(cfg, a.tp_size, a.pp_size, a.ep_size)
        if a.convert_from_full_state:
            src = load_state_dict(a.input_dir) if os.path.isdir(a.input_dir) else torch.load(a.input_dir, map_location="cpu",
                                                                                              weights_only=False)
            conv.convert_from_full_state(src, a.output_dir, use_xser=a.save_xser)
        elif a.convert_to_full_state:
            os.makedirs(a.output_dir, exist_ok=True)
            torch.save(conv.convert_to_full_state(a.input_dir, load_xser=a.load_xser or None),
                       os.path.join(a.output_dir, "checkpoint.pt"))
        else:
            raise SystemExit("pass --convert_from_full_state or --convert_to_full_state")


if __name__ == "__main__":
    CheckpointConverterMixtral.run()
