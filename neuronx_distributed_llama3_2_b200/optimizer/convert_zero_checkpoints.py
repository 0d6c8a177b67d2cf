"""Offline converter for ZeRO-1 optimizer checkpoints: sharded -> full, full -> sharded, sharded -> sharded at another DP
size (reference ``src/neuronx_distributed/optimizer/convert_zero_checkpoints.py:54-227``, CLI
``nxd_convert_zero_checkpoints``).

Shards are the files ``optim/dp_rank_DD_tp_rank_TT_pp_rank_PP.pt`` written by ``save_checkpoint``.  Each holds the
reference's top-level keys; with our flat-buffer engine every per-parameter entry is the 1-D slice of that parameter
which falls into the DP rank's contiguous shard (``layout`` records the slice).  "Full" = one file per (tp, pp) with
every parameter's fp32 master weight and Adam moments in its original shape.
"""
from __future__ import annotations

import argparse
import glob
import os
import re
from concurrent.futures import ProcessPoolExecutor
from typing import Any, Dict, List

import torch

_ALIGN = 128


def _load(path: str) -> Dict[str, Any]:
    return torch.load(path, map_location="cpu", weights_only=False)


def merge_shards(shards: List[Dict[str, Any]]) -> Dict[str, Any]:
    """DP shards (ordered by dp rank) -> full state: ``{"state": {i: {exp_avg, exp_avg_sq, step}}, "base_state": {i: w},
    "shape_info", "param_groups"}`` with tensors in original shapes."""
    first = shards[0]
    n_params = len(first["shape_info"])
    full = {"state": {}, "base_state": {}, "shape_info": dict(first["shape_info"]),
            "param_groups": first["param_groups"], "step": first["layout"].get("step", 0)}
    for i in range(n_params):
        shape = tuple(first["shape_info"][i])
        numel = 1
        for s in shape:
            numel *= s
        base = torch.zeros(numel, dtype=torch.float32)
        m = torch.zeros(numel, dtype=torch.float32)
        v = torch.zeros(numel, dtype=torch.float32)
        have_state = False
        for sd in shards:
            lay = sd["layout"]["params"][i]
            if lay["len"] == 0:
                continue
            sl = slice(lay["begin"], lay["begin"] + lay["len"])
            base[sl] = sd["base_state"][i].float()
            st = sd["state"].get(i)
            if st is not None:
                have_state = True
                m[sl] = st["exp_avg"].float()
                v[sl] = st["exp_avg_sq"].float()
        full["base_state"][i] = base.reshape(shape)
        if have_state:
            full["state"][i] = {"step": torch.tensor(float(full["step"])), "exp_avg": m.reshape(shape),
                                "exp_avg_sq": v.reshape(shape)}
    return full


def split_full(full: Dict[str, Any], dp_size: int, template_layout: Dict[str, Any]) -> List[Dict[str, Any]]:
    """Full state -> ``dp_size`` shards in the flat-contiguous layout (bucket offsets come from ``template_layout``,
    i.e. the layout record of any existing shard of the same model)."""
    params = template_layout["params"]
    by_bucket: Dict[int, List[int]] = {}
    for i, p in enumerate(params):
        by_bucket.setdefault(p["bucket"], []).append(i)
    out = []
    for r in range(dp_size):
        sd = {"state": {}, "base_state": {}, "shape_info": dict(full["shape_info"]), "param_groups": full["param_groups"],
              "layout": {"kind": "flat_contiguous", "params": [None] * len(params), "dp_size": dp_size, "dp_rank": r,
                         "step": full.get("step", 0)}}
        out.append(sd)
    for bucket, idxs in by_bucket.items():
        used = max(params[i]["offset"] + ((params[i]["numel"] + 7) // 8) * 8 for i in idxs)
        quantum = dp_size * _ALIGN
        total = (used + quantum - 1) // quantum * quantum
        shard = total // dp_size
        for r in range(dp_size):
            lo, hi = r * shard, (r + 1) * shard
            for i in idxs:
                o, n = params[i]["offset"], params[i]["numel"]
                a, b = max(o, lo), min(o + n, hi)
                ln = max(b - a, 0)
                begin = a - o if ln else 0
                flat_w = full["base_state"][i].reshape(-1)
                out[r]["base_state"][i] = flat_w[begin:begin + ln].clone()
                st = full["state"].get(i)
                if st is not None:
                    out[r]["state"][i] = {"step": st["step"].clone(),
                                          "exp_avg": st["exp_avg"].reshape(-1)[begin:begin + ln].clone(),
                                          "exp_avg_sq": st["exp_avg_sq"].reshape(-1)[begin:begin + ln].clone()}
                out[r]["layout"]["params"][i] = {"bucket": bucket, "offset": o, "numel": n, "begin": begin, "len": ln}
    return out


_PAT = re.compile(r"dp_rank_(\d+)_tp_rank_(\d+)_pp_rank_(\d+)\.pt$")


def _group_files(input_dir: str) -> Dict[Any, List[str]]:
    groups: Dict[Any, List[Any]] = {}
    for f in glob.glob(os.path.join(input_dir, "dp_rank_*_tp_rank_*_pp_rank_*.pt")):
        m = _PAT.search(f)
        if m:
            dp, tp, pp = (int(x) for x in m.groups())
            groups.setdefault((tp, pp), []).append((dp, f))
    return {k: [f for _, f in sorted(v)] for k, v in groups.items()}


def _convert_one(args) -> str:
    (tp, pp), files, output_dir, convert_to_full, dp_size = args
    shards = [_load(f) for f in files]
    if "layout" in shards[0] and shards[0]["layout"].get("kind") == "flat_contiguous":
        full = merge_shards(shards)
        template = shards[0]["layout"]
    else:  # already full
        full, template = shards[0], shards[0]["template_layout"]
    os.makedirs(output_dir, exist_ok=True)
    if convert_to_full:
        full["template_layout"] = template
        torch.save(full, os.path.join(output_dir, f"dp_rank_00_tp_rank_{tp:02d}_pp_rank_{pp:02d}.pt"))
    else:
        for r, sd in enumerate(split_full(full, dp_size, template)):
            torch.save(sd, os.path.join(output_dir, f"dp_rank_{r:02d}_tp_rank_{tp:02d}_pp_rank_{pp:02d}.pt"))
    return f"tp{tp} pp{pp}: {len(files)} file(s) -> {'full' if convert_to_full else f'{dp_size} shard(s)'}"


def convert(input_dir: str, output_dir: str, convert_to_full: bool = False, convert_to_sharded: bool = False,
            dp_size: int = 1, num_workers: int = 4) -> List[str]:
    assert convert_to_full != convert_to_sharded or not convert_to_full, "choose one direction"
    groups = _group_files(input_dir)
    if not groups:
        raise FileNotFoundError(f"no optimizer shards found under {input_dir}")
    jobs = [((tp, pp), files, output_dir, convert_to_full, dp_size) for (tp, pp), files in sorted(groups.items())]
    if num_workers <= 1 or len(jobs) == 1:
        return [_convert_one(j) for j in jobs]
    with ProcessPoolExecutor(max_workers=num_workers) as ex:
        return list(ex.map(_convert_one, jobs))


# -- the reference's helper names (``convert_zero_checkpoints.py:14-145``), taking its ``args`` namespace ---------------------
def _optim_dir(args) -> str:
    """The reference passes ``<ckpt>/<tag>`` and looks into ``optim/``; this tool also accepts the ``optim`` directory itself."""
    sub = os.path.join(args.input_dir, "optim")
    return sub if os.path.isdir(sub) else args.input_dir


def is_full(args) -> bool:
    """True when the input holds merged (un-sharded) optimizer state: one file per (tp, pp) carrying ``template_layout``."""
    d = _optim_dir(args)
    names = os.listdir(d)
    if any(n.startswith("full") for n in names):
        return True
    groups = _group_files(d)
    if not groups or any(len(v) != 1 for v in groups.values()):
        return False
    return "template_layout" in _load(next(iter(groups.values()))[0])


def is_xser(args) -> bool:
    """True when shards were written tensor-by-tensor (``*.tensors`` side directories)."""
    d = _optim_dir(args)
    return any(n.endswith(".tensors") and os.path.isdir(os.path.join(d, n)) for n in os.listdir(d))


def get_parallel_info(args):
    """``(dp_size, tp_size, pp_size)`` from the shard file names (dp_size 0 for a full checkpoint)."""
    d = _optim_dir(args)
    dps, tps, pps = set(), set(), set()
    for f in os.listdir(d):
        m = _PAT.search(f)
        if m:
            dp, tp, pp = (int(x) for x in m.groups())
            dps.add(dp), tps.add(tp), pps.add(pp)
    if not tps:
        raise FileNotFoundError(f"no optimizer shards found under {d}")
    full = is_full(args)
    return (0 if full else max(dps) + 1), max(tps) + 1, max(pps) + 1


def merge_optim_dp_checkpoints(args, tp_rank: int, pp_rank: int) -> Dict[str, Any]:
    """All DP shards of one (tp, pp) coordinate -> full fp32 state (``merge_shards``)."""
    files = _group_files(_optim_dir(args))[(tp_rank, pp_rank)]
    shards = [_load(f) for f in files]
    if "template_layout" in shards[0]:
        return shards[0]
    full = merge_shards(shards)
    full["template_layout"] = shards[0]["layout"]
    return full


def split_and_save_ckpts(args, merged_ckpt: Dict[str, Any], tp_rank: int, pp_rank: int) -> List[str]:
    """Full state -> ``args.dp_size`` shard files under ``args.output_dir[/optim]``."""
    out_dir = args.output_dir
    os.makedirs(out_dir, exist_ok=True)
    paths = []
    for r, sd in enumerate(split_full(merged_ckpt, int(args.dp_size), merged_ckpt["template_layout"])):
        paths.append(os.path.join(out_dir, f"dp_rank_{r:02d}_tp_rank_{tp_rank:02d}_pp_rank_{pp_rank:02d}.pt"))
        torch.save(sd, paths[-1])
    return paths


def main(argv=None) -> None:
    ap = argparse.ArgumentParser("nxd_convert_zero_checkpoints")
    ap.add_argument("--input_dir", required=True, help="<ckpt>/<tag>/optim")
    ap.add_argument("--output_dir", required=True)
    ap.add_argument("--convert_to_full", action="store_true")
    ap.add_argument("--convert_to_sharded", action="store_true")
    ap.add_argument("--dp_size", type=int, default=1, help="target DP size for --convert_to_sharded")
    ap.add_argument("--num_workers", type=int, default=4)
    a = ap.parse_args(argv)
    if a.convert_to_full == a.convert_to_sharded:
        ap.error("pass exactly one of --convert_to_full / --convert_to_sharded")
    for line in convert(a.input_dir, a.output_dir, a.convert_to_full, a.convert_to_sharded, a.dp_size, a.num_workers):
        print(line)


if __name__ == "__main__":
    main()
