"""HF / Megatron-style FULL checkpoint <-> NxD (tp, pp)-sharded checkpoint (reference
``src/neuronx_distributed/scripts/checkpoint_converter.py:22-739``).

Direction ``convert_from_full_state``: HF names -> our model names, q/k/v fused into ``qkv_proj.weight_qkv`` (KV heads
replicated ``kv_size_multiplier`` times, Q heads / o_proj columns re-ordered so that every TP rank's query heads sit
next to the KV head it holds), gate/up fused into the stride-2 ``gate_up_proj``; then sliced per TP rank by
``partition_dim`` and distributed over PP stages (even split, remainder to later stages), and written as
``<out>/model/dp_rank_00_tp_rank_TT_pp_rank_PP.pt`` (optionally xser).  ``convert_to_full_state`` is the inverse.
"""
from __future__ import annotations

import argparse
import json
import os
import re
from typing import Any, Dict, List, Optional

import torch

from ..pipeline.partition import create_partitions
from ..utils import xser

# name -> partition dim of the TP-sharded parameters of the Llama family (reference ``:22-46``)
PARTITION_DIMS = {
    "embed_tokens.weight": 0, "lm_head.weight": 0, "qkv_proj.weight_qkv": 0, "qkv_proj.weight_q": 0,
    "qkv_proj.weight_k": 0, "qkv_proj.weight_v": 0, "o_proj.weight": 1, "gate_up_proj.weight": 0, "down_proj.weight": 1,
}


def q_head_permutation(n_q: int, n_kv: int, tp: int, kv_mult: int) -> List[int]:
    """Order of the query heads such that a contiguous TP split gives rank ``r`` the query heads that attend with KV
    head ``r % n_kv`` -- the head its K/V shard holds after tiling the KV heads ``kv_mult`` times.  Identity when no
    replication is needed."""
    if kv_mult == 1:
        return list(range(n_q))
    g = n_q // n_kv
    per = g // kv_mult
    order: List[int] = []
    ranks_per_head = tp // (n_kv * kv_mult) if tp >= n_kv * kv_mult else 1
    for r in range(n_kv * kv_mult):
        h, j = r % n_kv, r // n_kv
        order += list(range(h * g + j * per, h * g + (j + 1) * per))
    assert sorted(order) == list(range(n_q)) and ranks_per_head >= 1
    return order


class CheckpointConverterBase:
    def __init__(self, num_layers: int, num_heads: int, num_kv_heads: int, head_dim: int, tp_size: int = 1,
                 pp_size: int = 1, kv_size_multiplier: int = 1, fuse_qkv: bool = True, tie_word_embeddings: bool = True):
        self.L, self.nq, self.nkv, self.D = num_layers, num_heads, num_kv_heads, head_dim
        self.tp, self.pp, self.m = tp_size, pp_size, kv_size_multiplier
        self.fuse_qkv, self.tied = fuse_qkv, tie_word_embeddings

    # -- name maps -------------------------------------------------------------------------------------------------
    def hf_to_nxd_full(self, hf: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        """HF-named full state -> our names, still un-sharded (QKV / gate-up fused, heads permuted for KV replication)."""
        out: Dict[str, torch.Tensor] = {}
        perm = q_head_permutation(self.nq, self.nkv, self.tp, self.m)
        rows = torch.tensor([h * self.D + d for h in perm for d in range(self.D)])
        for k, v in hf.items():
            m = re.match(r"(model\.layers\.\d+\.)self_attn\.([qkvo])_proj\.weight$", k)
            if m:
                continue
            m = re.match(r"(model\.layers\.\d+\.)mlp\.(gate|up)_proj\.weight$", k)
            if m:
                continue
            out[k] = v
        for i in range(self.L):
            p = f"model.layers.{i}."
            q, k_, v_, o = (hf[p + f"self_attn.{n}_proj.weight"] for n in "qkvo")
            q = q[rows]
            o = o[:, rows]
            if self.fuse_qkv:
                out[p + "self_attn.qkv_proj.weight_qkv"] = torch.cat([q, k_, v_], dim=0)
            else:
                out[p + "self_attn.qkv_proj.weight_q"], out[p + "self_attn.qkv_proj.weight_k"] = q, k_
                out[p + "self_attn.qkv_proj.weight_v"] = v_
            out[p + "self_attn.o_proj.weight"] = o
            out[p + "mlp.gate_up_proj.weight"] = torch.cat([hf[p + "mlp.gate_proj.weight"], hf[p + "mlp.up_proj.weight"]], 0)
        if self.tied and "lm_head.weight" not in out and "model.embed_tokens.weight" in out:
            out["lm_head.weight"] = out["model.embed_tokens.weight"]
        return out

    def nxd_full_to_hf(self, full: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        out: Dict[str, torch.Tensor] = {}
        perm = q_head_permutation(self.nq, self.nkv, self.tp, self.m)
        inv = [0] * len(perm)
        for new, old in enumerate(perm):
            inv[old] = new
        rows = torch.tensor([h * self.D + d for h in inv for d in range(self.D)])
        qs, kvs = self.nq * self.D, self.nkv * self.D
        for k, v in full.items():
            if k.endswith("qkv_proj.weight_qkv"):
                p = k[: -len("qkv_proj.weight_qkv")]
                q, k_, v_ = torch.split(v, [qs, kvs, kvs], dim=0)
                out[p + "q_proj.weight"], out[p + "k_proj.weight"], out[p + "v_proj.weight"] = q[rows], k_, v_
            elif k.endswith("qkv_proj.weight_q"):
                out[k.replace("qkv_proj.weight_q", "q_proj.weight")] = v[rows]
            elif k.endswith("qkv_proj.weight_k"):
                out[k.replace("qkv_proj.weight_k", "k_proj.weight")] = v
            elif k.endswith("qkv_proj.weight_v"):
                out[k.replace("qkv_proj.weight_v", "v_proj.weight")] = v
            elif k.endswith("o_proj.weight"):
                out[k] = v[:, rows]
            elif k.endswith("gate_up_proj.weight"):
                g, u = v.chunk(2, dim=0)
                out[k.replace("gate_up_proj", "gate_proj")], out[k.replace("gate_up_proj", "up_proj")] = g, u
            else:
                out[k] = v
        if self.tied:
            out.pop("lm_head.weight", None)
        return out

    # -- sharding ------------------------------------------------------------------------------------------------
    def _partition_dim(self, name: str) -> Optional[int]:
        for suffix, dim in PARTITION_DIMS.items():
            if name.endswith(suffix):
                return dim
        return None

    def shard_tp(self, full: Dict[str, torch.Tensor], rank: int) -> Dict[str, torch.Tensor]:
        out = {}
        qs, kvs = self.nq * self.D, self.nkv * self.D
        for k, v in full.items():
            dim = self._partition_dim(k)
            if dim is None:
                out[k] = v
            elif k.endswith("weight_qkv"):
                q, k_, v_ = torch.split(v, [qs, kvs, kvs], dim=0)
                rep = [self.m] + [1] * (v.dim() - 1)
                out[k] = torch.cat([q.chunk(self.tp, 0)[rank], k_.repeat(*rep).chunk(self.tp, 0)[rank],
                                    v_.repeat(*rep).chunk(self.tp, 0)[rank]], dim=0).contiguous()
            elif k.endswith("weight_k") or k.endswith("weight_v"):
                out[k] = v.repeat(self.m, 1).chunk(self.tp, 0)[rank].contiguous()
            elif k.endswith("gate_up_proj.weight"):
                g, u = v.chunk(2, dim=0)
                out[k] = torch.cat([g.chunk(self.tp, 0)[rank], u.chunk(self.tp, 0)[rank]], dim=0).contiguous()
            else:
                out[k] = v.chunk(self.tp, dim)[rank].contiguous()
        return out

    def merge_tp(self, shards: List[Dict[str, torch.Tensor]]) -> Dict[str, torch.Tensor]:
        out = {}
        for k in shards[0]:
            dim = self._partition_dim(k)
            vs = [s[k] for s in shards]
            if dim is None:
                out[k] = vs[0]
            elif k.endswith("weight_qkv"):
                q_sz = self.nq * self.D // self.tp
                kv_sz = self.nkv * self.D * self.m // self.tp
                parts = [torch.split(v, [q_sz, kv_sz, kv_sz], dim=0) for v in vs]
                q = torch.cat([p[0] for p in parts], 0)
                k_ = torch.cat([p[1] for p in parts], 0)[: self.nkv * self.D]   # drop the replicas
                v_ = torch.cat([p[2] for p in parts], 0)[: self.nkv * self.D]
                out[k] = torch.cat([q, k_, v_], 0)
            elif k.endswith("weight_k") or k.endswith("weight_v"):
                out[k] = torch.cat(vs, 0)[: self.nkv * self.D]
            elif k.endswith("gate_up_proj.weight"):
                gs, us = zip(*[v.chunk(2, 0) for v in vs])
                out[k] = torch.cat(list(gs) + list(us), 0)
            else:
                out[k] = torch.cat(vs, dim)
        return out

    def stage_of_key(self, key: str) -> List[int]:
        """PP stage(s) that own ``key`` (tied embedding lives on the first AND the last stage)."""
        if self.pp == 1:
            return [0]
        cuts = create_partitions(self.pp, self.L)
        m = re.match(r"model\.layers\.(\d+)\.", key)
        if m:
            layer = int(m.group(1))
            for s, c in enumerate(cuts):
                if layer <= c:
                    return [s]
            return [self.pp - 1]
        if key.startswith("model.embed_tokens"):
            return [0]
        return [self.pp - 1]

    # -- directory level -----------------------------------------------------------------------------------------
    @staticmethod
    def shard_file(out_dir: str, tp: int, pp: int, ep: Optional[int] = None) -> str:
        ep_s = f"_ep_rank_{ep:02d}" if ep is not None else ""
        return os.path.join(out_dir, "model", f"dp_rank_00{ep_s}_tp_rank_{tp:02d}_pp_rank_{pp:02d}.pt")

    def convert_from_full_state(self, hf_state: Dict[str, torch.Tensor], out_dir: str, use_xser: bool = False) -> None:
        full = self.hf_to_nxd_full(hf_state)
        os.makedirs(os.path.join(out_dir, "model"), exist_ok=True)
        for t in range(self.tp):
            shard = self.shard_tp(full, t)
            for s in range(self.pp):
                part = {k: v for k, v in shard.items() if s in self.stage_of_key(k)}
                if self.tied and self.pp > 1 and s == self.pp - 1 and "lm_head.weight" not in part:
                    part["lm_head.weight"] = shard["model.embed_tokens.weight"]
                path = self.shard_file(out_dir, t, s)
                if use_xser:
                    xser.save(part, path)
                else:
                    torch.save(part, path)
        for marker in ("checkpoint", "done"):
            open(os.path.join(out_dir, marker), "w").write("1")

    def convert_to_full_state(self, in_dir: str, load_xser: Optional[bool] = None) -> Dict[str, torch.Tensor]:
        merged_stages: Dict[str, torch.Tensor] = {}
        for s in range(self.pp):
            shards = []
            for t in range(self.tp):
                path = self.shard_file(in_dir, t, s)
                is_x = os.path.isdir(path + ".tensors") if load_xser is None else load_xser
                shards.append(xser.load(path) if is_x else torch.load(path, map_location="cpu", weights_only=False))
            merged_stages.update(self.merge_tp(shards))
        return self.nxd_full_to_hf(merged_stages)

    # -- reference method surface (``scripts/checkpoint_converter.py:22-637``): small overridable steps -----------------
    def get_partition_dim(self, name: str) -> Optional[int]:
        """Dimension a parameter is sharded on over TP, ``None`` for replicated ones (override to add layer types)."""
        return self._partition_dim(name)

    def get_hf_to_nxd_model_keys(self, qkv_linear: bool = True, is_gqa: bool = True) -> Dict[str, str]:
        """HF sub-key -> framework sub-key (reference ``:49-69``)."""
        keys = {"mlp.gate_proj.weight": "mlp.gate_up_proj.weight", "mlp.up_proj.weight": "mlp.gate_up_proj.weight"}
        if qkv_linear:
            tgt = "qkv_proj.weight_qkv" if self.fuse_qkv else None
            for n in "qkv":
                keys[f"self_attn.{n}_proj.weight"] = f"self_attn.{tgt or 'qkv_proj.weight_' + n}"
        return keys

    def get_fused_qkv_key(self, prefix: str = "") -> str:
        return prefix + "self_attn.qkv_proj.weight_qkv"

    @staticmethod
    def is_qkv_weight(name: str) -> bool:
        return bool(re.search(r"self_attn\.(q_proj|k_proj|v_proj|qkv_proj)\.", name))

    @staticmethod
    def get_weight_key(prefix: str, name: str) -> str:
        return f"{prefix}{name}.weight"

    def coalesce_qkv(self, state_dict: Dict[str, torch.Tensor], config: Optional[dict] = None, tp_degree: Optional[int] = None
                     ) -> Dict[str, torch.Tensor]:
        """Fuse every layer's ``q/k/v_proj.weight`` into ``qkv_proj.weight_qkv`` in place (reference ``:78-93``)."""
        for i in range(self.L):
            p = f"model.layers.{i}.self_attn."
            if p + "q_proj.weight" in state_dict:
                parts = [state_dict.pop(p + f"{n}_proj.weight") for n in "qkv"]
                state_dict[p + "qkv_proj.weight_qkv"] = torch.cat(parts, dim=0)
        return state_dict

    def convert_partial_state_to_fused_qkv(self, partial: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        """One TP shard with separate ``weight_q / weight_k / weight_v`` -> fused ``weight_qkv``."""
        for k in [k for k in partial if k.endswith("qkv_proj.weight_q")]:
            base = k[: -len("weight_q")]
            partial[base + "weight_qkv"] = torch.cat([partial.pop(base + "weight_q"), partial.pop(base + "weight_k"),
                                                      partial.pop(base + "weight_v")], dim=0)
        return partial

    def convert_partial_state_to_non_fused_qkv(self, partial: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        q_sz = self.nq * self.D // self.tp
        kv_sz = self.nkv * self.D * self.m // self.tp
        for k in [k for k in partial if k.endswith("qkv_proj.weight_qkv")]:
            base = k[: -len("weight_qkv")]
            q, k_, v_ = torch.split(partial.pop(k), [q_sz, kv_sz, kv_sz], dim=0)
            partial[base + "weight_q"], partial[base + "weight_k"], partial[base + "weight_v"] = q, k_, v_
        return partial

    # -- Megatron (NxDT) naming, ``--model_style megatron`` (reference ``:102-196``) ---------------------------------------
    # whole-key renames first, then fragments; the pairs are (megatron, hf)
    _MEGATRON_WHOLE = (("language_model.embedding.word_embeddings.weight", "model.embed_tokens.weight"),
                       ("language_model.encoder.final_layernorm.weight", "model.norm.weight"),
                       ("language_model.output_layer.weight", "lm_head.weight"))
    _MEGATRON_PARTS = (("language_model.encoder.", "model."), ("self_attention.", "self_attn."),
                       ("core_attention.rotary_emb.inv_freq", "rotary_emb.inv_freq"),
                       ("dense_h_to_4h.weight", "gate_up_proj.weight"), ("dense_4h_to_h.weight", "down_proj.weight"),
                       ("query_key_value.weight", "qkv_proj.weight"), ("dense.weight", "o_proj.weight"))

    def rename_keys_for_megatron(self, key: str, model_style: str = "megatron", hf_to_nxdt: bool = False) -> str:
        """One key between Megatron / NxDT names and HF-style names; untouched unless ``model_style == 'megatron'``."""
        if model_style != "megatron":
            return key
        a, b = (1, 0) if hf_to_nxdt else (0, 1)
        for pair in self._MEGATRON_WHOLE:
            if key == pair[a]:
                return pair[b]
        for pair in self._MEGATRON_PARTS:
            key = key.replace(pair[a], pair[b])
        return key

    def modify_qkv_for_megatron(self, partial_state: Dict[str, torch.Tensor], args) -> None:
        """In place.  Megatron keeps attention as ``query_key_value`` (MHA: rows q|k|v) or, with ``--qkv_linear`` (GQA),
        as ``query`` + ``key_value`` (rows k|v); the framework side has ``q_proj/k_proj/v_proj`` (no qkv_linear) or
        ``qkv_proj.weight_q/_k/_v``.  Direction follows ``args.convert_from_full_state`` the way the reference's does:
        set -> the state is being turned INTO Megatron form, unset -> out of it."""
        if getattr(args, "model_style", "hf") != "megatron":
            return
        to_megatron = bool(getattr(args, "convert_from_full_state", False))
        keys = list(partial_state.keys())
        if not getattr(args, "qkv_linear", False):
            if to_megatron:
                for k in keys:
                    if "q_proj" in k:
                        parts = [partial_state.pop(k.replace("q_proj", n)) for n in ("q_proj", "k_proj", "v_proj")]
                        partial_state[k.replace("q_proj", "query_key_value")] = torch.cat(parts, dim=0)
            return
        for k in keys:
            if k not in partial_state:
                continue
            if to_megatron:
                if k.endswith("query_key_value.weight_q"):
                    partial_state[k.replace("query_key_value.weight_q", "query.weight")] = partial_state.pop(k)
                elif k.endswith("query_key_value.weight_k"):
                    kv = [partial_state.pop(k), partial_state.pop(k[:-1] + "v")]
                    partial_state[k.replace("query_key_value.weight_k", "key_value.weight")] = torch.cat(kv, dim=0)
            else:
                if "query.weight" in k:
                    partial_state[k.replace("query.weight", "qkv_proj.weight_q")] = partial_state.pop(k)
                elif "key_value.weight" in k:
                    kk, vv = partial_state.pop(k).chunk(2, dim=0)
                    partial_state[k.replace("key_value.weight", "qkv_proj.weight_k")] = kk.contiguous()
                    partial_state[k.replace("key_value.weight", "qkv_proj.weight_v")] = vv.contiguous()

    def is_q_or_o_for_megatron(self, args, name: str) -> bool:
        """Megatron checkpoints have no KV replication, so Q and O shards are concatenated in plain rank order (no head
        shuffle) -- true for Q / O projection weights in that style."""
        if getattr(args, "model_style", "hf") != "megatron":
            return False
        return bool(re.search(r"(weight_q|q_proj|query\.weight|o_proj|dense\.weight)", name))

    def megatron_full_to_hf(self, state: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        """Full Megatron/NxDT-named state -> HF names with separate q/k/v and gate/up (what ``hf_to_nxd_full`` takes)."""
        qs, kvs = self.nq * self.D, self.nkv * self.D
        out: Dict[str, torch.Tensor] = {}
        for k, v in state.items():
            if k.endswith("self_attention.query_key_value.weight"):
                base = self.rename_keys_for_megatron(k[: -len("query_key_value.weight")])
                q, k_, v_ = torch.split(v, [qs, kvs, kvs], dim=0)
                out[base + "q_proj.weight"], out[base + "k_proj.weight"], out[base + "v_proj.weight"] = q, k_, v_
            elif k.endswith("self_attention.query.weight"):
                out[self.rename_keys_for_megatron(k[: -len("query.weight")]) + "q_proj.weight"] = v
            elif k.endswith("self_attention.key_value.weight"):
                base = self.rename_keys_for_megatron(k[: -len("key_value.weight")])
                out[base + "k_proj.weight"], out[base + "v_proj.weight"] = v.chunk(2, dim=0)
            elif k.endswith("mlp.dense_h_to_4h.weight"):
                base = self.rename_keys_for_megatron(k[: -len("dense_h_to_4h.weight")])
                out[base + "gate_proj.weight"], out[base + "up_proj.weight"] = v.chunk(2, dim=0)
            else:
                out[self.rename_keys_for_megatron(k)] = v
        return out

    def hf_full_to_megatron(self, state: Dict[str, torch.Tensor], qkv_linear: bool = False) -> Dict[str, torch.Tensor]:
        """Inverse of ``megatron_full_to_hf``: ``qkv_linear`` picks ``query`` + ``key_value`` over ``query_key_value``."""
        out: Dict[str, torch.Tensor] = {}
        for k, v in state.items():
            m = re.match(r"(.*self_attn\.)q_proj\.weight$", k)
            if m:
                base = self.rename_keys_for_megatron(m.group(1), hf_to_nxdt=True)
                kk, vv = state[m.group(1) + "k_proj.weight"], state[m.group(1) + "v_proj.weight"]
                if qkv_linear:
                    out[base + "query.weight"], out[base + "key_value.weight"] = v, torch.cat([kk, vv], dim=0)
                else:
                    out[base + "query_key_value.weight"] = torch.cat([v, kk, vv], dim=0)
            elif re.search(r"self_attn\.[kv]_proj\.weight$", k) or k.endswith("mlp.up_proj.weight"):
                continue
            elif k.endswith("mlp.gate_proj.weight"):
                base = self.rename_keys_for_megatron(k[: -len("gate_proj.weight")], hf_to_nxdt=True)
                out[base + "dense_h_to_4h.weight"] = torch.cat([v, state[k.replace("gate_proj", "up_proj")]], dim=0)
            else:
                out[self.rename_keys_for_megatron(k, hf_to_nxdt=True)] = v
        return out

    def pre_process_full_state_before_tp_conversion(self, state_dict: Dict[str, torch.Tensor], args=None):
        """Hook: edit the full state before it is renamed and sharded.  Megatron-style input is brought to HF names
        here; otherwise nothing."""
        if getattr(args, "model_style", "hf") == "megatron":
            return self.megatron_full_to_hf(state_dict)
        return state_dict

    def post_process_full_state_after_tp_conversion(self, state_dict: Dict[str, torch.Tensor], args=None):
        """Hook: edit the merged full (HF-named) state before it is written.  ``--model_style megatron`` writes
        Megatron names; otherwise nothing."""
        if getattr(args, "model_style", "hf") == "megatron":
            return self.hf_full_to_megatron(state_dict, qkv_linear=self.nkv != self.nq)
        return state_dict

    @staticmethod
    def find_size(obj) -> int:
        """Bytes held by the tensors inside a (nested) state."""
        if isinstance(obj, torch.Tensor):
            return obj.numel() * obj.element_size()
        if isinstance(obj, dict):
            return sum(CheckpointConverterBase.find_size(v) for v in obj.values())
        if isinstance(obj, (list, tuple)):
            return sum(CheckpointConverterBase.find_size(v) for v in obj)
        return 0

    @staticmethod
    def prune_state(state: Dict[str, Any]) -> Dict[str, Any]:
        """Drop everything that is not part of the model state: trainer checkpoints keep it under ``"model"``."""
        return state["model"] if isinstance(state, dict) and "model" in state and isinstance(state["model"], dict) else state

    def get_input_filename(self, args, tp: int, pp: int, ep: Optional[int] = None, xser: Optional[bool] = None) -> str:
        return self.shard_file(args.input_dir, tp, pp, ep)

    def get_output_filename(self, args, tp: int, pp: int, ep: Optional[int] = None, xser: Optional[bool] = None) -> str:
        return self.shard_file(args.output_dir, tp, pp, ep)

    def load_partial_xser(self, args, tp: int, pp: int, ep: Optional[int] = None) -> Dict[str, Any]:
        return self.prune_state(xser.load(self.get_input_filename(args, tp, pp, ep, True)))

    def load_partial_no_xser(self, args, tp: int, pp: int, ep: Optional[int] = None) -> Dict[str, Any]:
        return self.prune_state(torch.load(self.get_input_filename(args, tp, pp, ep, False), map_location="cpu",
                                           weights_only=False))

    def save_partial_xser(self, args, state: Dict[str, Any], tp: int, pp: int, ep: Optional[int] = None) -> None:
        path = self.get_output_filename(args, tp, pp, ep, True)
        os.makedirs(os.path.dirname(path), exist_ok=True)
        xser.save(state, path)

    def save_partial_no_xser(self, args, state: Dict[str, Any], tp: int, pp: int, ep: Optional[int] = None) -> None:
        path = self.get_output_filename(args, tp, pp, ep, False)
        os.makedirs(os.path.dirname(path), exist_ok=True)
        torch.save(state, path)

    def load_full_state(self, args) -> Dict[str, torch.Tensor]:
        """A full HF-style checkpoint: a directory (safetensors / bin, sharded or not) or a single ``torch.save`` file."""
        path = args.input_dir
        if os.path.isdir(path):
            from ..inference.checkpoint import load_state_dict

            try:
                return load_state_dict(path)
            except FileNotFoundError:
                state: Dict[str, Any] = {}
                for f in sorted(os.listdir(path)):
                    if f.endswith(".safetensors"):
                        from safetensors.torch import load_file

                        state.update(load_file(os.path.join(path, f)))
                    elif f.endswith((".bin", ".pt")):
                        state.update(torch.load(os.path.join(path, f), map_location="cpu", weights_only=False))
                return state
        return self.prune_state(torch.load(path, map_location="cpu", weights_only=False))

    def save_full(self, args, full_state: Dict[str, torch.Tensor]) -> None:
        os.makedirs(args.output_dir, exist_ok=True)
        torch.save(full_state, os.path.join(args.output_dir, "checkpoint.pt"))

    def convert_full_state_to_tp(self, args) -> None:
        state = self.pre_process_full_state_before_tp_conversion(self.load_full_state(args), args)
        self.convert_from_full_state(state, args.output_dir, use_xser=getattr(args, "save_xser", False))

    def merge_tp_checkpoints(self, args) -> Dict[str, torch.Tensor]:
        full = self.convert_to_full_state(args.input_dir, load_xser=getattr(args, "load_xser", False) or None)
        return self.post_process_full_state_after_tp_conversion(full, args)

    def _all_ranks(self):
        return [(t, s) for t in range(self.tp) for s in range(self.pp)]

    def convert_from_xser(self, args) -> None:
        """Every ``(tp, pp)`` shard: tensor-per-file (xser) layout -> one ``torch.save`` file (reference ``:627-631``)."""
        for t, s in self._all_ranks():
            self.save_partial_no_xser(args, self.load_partial_xser(args, t, s), t, s)

    def convert_to_xser(self, args) -> None:
        for t, s in self._all_ranks():
            self.save_partial_xser(args, self.load_partial_no_xser(args, t, s), t, s)

    # -- CLI -------------------------------------------------------------------------------------------------------
    @staticmethod
    def get_arg_parser() -> argparse.ArgumentParser:
        ap = argparse.ArgumentParser("nxd checkpoint converter")
        ap.add_argument("--input_dir", required=True)
        ap.add_argument("--output_dir", required=True)
        ap.add_argument("--config", required=True, help="json with num_hidden_layers / num_attention_heads / ...")
        ap.add_argument("--tp_size", type=int, default=1)
        ap.add_argument("--pp_size", type=int, default=1)
        ap.add_argument("--kv_size_multiplier", type=int, default=1)
        ap.add_argument("--qkv_linear", type=lambda s: s.lower() != "false", default=True)
        ap.add_argument("--model_style", choices=["hf", "megatron"], default="hf",
                        help="naming of the FULL checkpoint (input of --convert_from_full_state, output of --convert_to_full_state)")
        ap.add_argument("--convert_from_full_state", action="store_true")
        ap.add_argument("--convert_to_full_state", action="store_true")
        ap.add_argument("--load_xser", action="store_true")
        ap.add_argument("--save_xser", action="store_true")
        ap.add_argument("--convert_from_xser", action="store_true", help="sharded xser -> sharded torch.save")
        ap.add_argument("--convert_to_xser", action="store_true", help="sharded torch.save -> sharded xser")
        return ap

    @classmethod
    def run(cls, argv=None) -> None:
        a = cls.get_arg_parser().parse_args(argv)
        cfg = json.load(open(a.config))
        heads = cfg["num_attention_heads"]
        conv = cls(cfg["num_hidden_layers"], heads, cfg.get("num_key_value_heads", heads),
                   cfg.get("head_dim", cfg["hidden_size"] // heads), a.tp_size, a.pp_size, a.kv_size_multiplier,
                   tie_word_embeddings=cfg.get("tie_word_embeddings", False))
        if a.convert_from_full_state:
            conv.convert_full_state_to_tp(a)
        elif a.convert_to_full_state:
            conv.save_full(a, conv.merge_tp_checkpoints(a))
        elif a.convert_from_xser:
            conv.convert_from_xser(a)
        elif a.convert_to_xser:
            conv.convert_to_xser(a)
        else:
            raise SystemExit("pass --convert_from_full_state, --convert_to_full_state, --convert_from_xser or --convert_to_xser")


if __name__ == "__main__":
    CheckpointConverterBase.run()
