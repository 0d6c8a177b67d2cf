"""Weight-layout transformation between compiled graphs (reference ``src/neuronx_distributed/trace/hlo_utils.py:19-283``).

What the reference does: the PRIORITY graph is compiled first and the compiler may pick another memory layout for some
weights; that choice is pulled out of its HLO as (a) a map ``weight -> layout-transform computation``, which is appended
to every other graph so that they expect the same layout, and (b) one stand-alone "weight layout transformation" program
that takes ALL weights in checkpoint layout and returns them in the chosen layout -- run once at load time.

Here a compiled graph is a CUDA-graph bucket over ONE shared set of parameters, and GEMM operand layout is a descriptor
bit of the tcgen05 kernels (K-major or MN-major), so the dense layers never need a re-layout.  What remains useful is the
mechanism itself -- e.g. expert weights stored ``[E, in, out]`` that a decode-priority graph wants as ``[E, out, in]``
(rows contiguous along K for the skinny GEMM).  So the same steps are kept, on a small JSON-serialisable description
of a graph instead of an HLO protobuf:

    graph  = {"name", "id", "frontend_attributes": {str: str}, "weights": {name: {"idx", "shape", "dtype"}},
              "computations": [computation, ...]}
    computation = {"id", "name", "weight_idx", "perm": [..], "dtype": optional target dtype, "instructions": [...]}

The layout choice ("the compiler") is ``suggest_layout_transforms``: a module may tag a parameter with
``preferred_layout_perm`` (a permutation of its dims), or a policy callable decides per weight.
"""
from __future__ import annotations

import copy
import json
from typing import Any, Callable, Dict, List, Optional

import torch

TRANSPOSABLE_WEIGHT_IDX = "transposable_weight_idx"
REQUIRE_TRANSPOSE_WEIGHT_IDX = "require_transpose_weight_idx"
REQUIRE_TRANSPOSE_CUSTOM_CALL = "require_transpose_custom_call"

Graph = Dict[str, Any]
Computation = Dict[str, Any]


def new_graph(name: str, module: torch.nn.Module, graph_id: int = 0) -> Graph:
    """Describe a graph over ``module``'s parameters; weight indices follow ``named_parameters`` order."""
    weights = {n: {"idx": i, "shape": list(p.shape), "dtype": str(p.dtype).replace("torch.", "")}
               for i, (n, p) in enumerate(module.named_parameters())}
    return {"name": name, "id": graph_id, "frontend_attributes": {}, "weights": weights, "computations": []}


def weight_name_to_idx_of(graph: Graph) -> Dict[str, int]:
    return {n: w["idx"] for n, w in graph["weights"].items()}


def read_hlo(hlo_path: str) -> Graph:
    """Load a graph description written by ``write_hlo`` (reference ``:19-24`` parses an ``HloModuleProto``)."""
    with open(hlo_path) as f:
        return json.load(f)


def write_hlo(hlo: Graph, hlo_path: str) -> None:
    with open(hlo_path, "w") as f:
        json.dump(hlo, f, indent=1)


def add_weight_idx_attr_to_hlo(hlo: Graph, weight_name_to_idx: Dict[str, int]) -> Graph:
    """Record which inputs are weights whose layout may be changed (reference ``:27-34``)."""
    hlo["frontend_attributes"][TRANSPOSABLE_WEIGHT_IDX] = ",".join(str(i) for i in sorted(weight_name_to_idx.values()))
    return hlo


def suggest_layout_transforms(hlo: Graph, module: torch.nn.Module,
                              policy: Optional[Callable[[str, torch.nn.Parameter], Optional[List[int]]]] = None) -> Graph:
    """The compiler's part: give the priority graph one layout-transform computation per weight that wants another
    layout.  ``policy(name, param) -> perm | None``; default: the parameter's ``preferred_layout_perm`` tag."""
    allowed = {int(i) for i in hlo["frontend_attributes"].get(TRANSPOSABLE_WEIGHT_IDX, "").split(",") if i != ""}
    next_id = max([hlo["id"]] + [c["id"] for c in hlo["computations"]]) + 1
    for name, p in module.named_parameters():
        idx = hlo["weights"][name]["idx"]
        if idx not in allowed:
            continue
        perm = policy(name, p) if policy is not None else getattr(p, "preferred_layout_perm", None)
        if perm is None or list(perm) == list(range(p.dim())):
            continue
        assert sorted(perm) == list(range(p.dim())), f"{name}: {perm} is not a permutation of {p.dim()} dims"
        hlo["computations"].append({
            "id": next_id, "name": f"layout_{idx}", "weight_idx": idx, "perm": list(perm), "dtype": None,
            "frontend_attributes": {TRANSPOSABLE_WEIGHT_IDX: str(idx)},
            "instructions": [{"id": next_id + 1, "name": "p0", "opcode": "parameter", "operand_ids": []},
                             {"id": next_id + 2, "name": "t0", "opcode": "transpose", "operand_ids": [next_id + 1]}],
            "root_id": next_id + 2})
        next_id += 3
    return hlo


def get_layout_transform_map(hlo_stub: Graph, weight_name_to_idx: Dict[str, int]) -> Dict[str, Computation]:
    """``{weight name: transform computation}`` for the weights the priority graph re-laid-out; weights already in
    their best layout are absent (reference ``:37-60``)."""
    idx_to_name = {i: n for n, i in weight_name_to_idx.items()}
    out: Dict[str, Computation] = {}
    for cpt in hlo_stub["computations"]:
        tag = cpt.get("frontend_attributes", {}).get(TRANSPOSABLE_WEIGHT_IDX)
        if tag is not None and int(tag) in idx_to_name:
            out[idx_to_name[int(tag)]] = cpt
    return out


def update_computation_id_and_name(src_cpt: Computation, start_id: int, name_prefix: str) -> Computation:
    """A copy of the computation with every id shifted by ``start_id`` and every name prefixed, so that it can be put
    into another graph without clashes (reference ``:63-93``)."""
    cpt = copy.deepcopy(src_cpt)
    cpt["id"] += start_id
    cpt["root_id"] += start_id
    cpt["name"] = name_prefix + cpt["name"]
    for ins in cpt["instructions"]:
        ins["id"] += start_id
        ins["operand_ids"] = [o + start_id for o in ins["operand_ids"]]
        ins["name"] = name_prefix + ins["name"]
    return cpt


def append_layout_computation_to_hlo(hlo: Graph, weight_name_to_transform_cpt: Dict[str, Computation]) -> Graph:
    """Tell a non-priority graph which of ITS weights arrive re-laid-out and by which computation (reference
    ``:96-128``; the reference passes an ``HloArtifacts``, whose ``weight_name_to_idx`` is ``graph["weights"]`` here)."""
    idxs, names = [], []
    top = max([hlo["id"]] + [c["root_id"] for c in hlo["computations"]] + [0])
    for wname, w in hlo["weights"].items():
        if wname in weight_name_to_transform_cpt:
            cpt = update_computation_id_and_name(weight_name_to_transform_cpt[wname], start_id=top + 1, name_prefix="wlt_")
            cpt["weight_idx"] = w["idx"]
            cpt["frontend_attributes"] = {TRANSPOSABLE_WEIGHT_IDX: str(w["idx"])}
            hlo["computations"].append(cpt)
            idxs.append(w["idx"])
            names.append(cpt["name"])
            top = cpt["root_id"]
    hlo["frontend_attributes"][REQUIRE_TRANSPOSE_WEIGHT_IDX] = ",".join(str(i) for i in idxs)
    hlo["frontend_attributes"][REQUIRE_TRANSPOSE_CUSTOM_CALL] = ",".join(names)
    return hlo


def extract_weight_layout_transform_hlo(hlo_stub: Graph, weight_name_to_idx: Dict[str, int]) -> Graph:
    """The stand-alone load-time program: inputs = ALL weights in checkpoint layout in ascending index order, outputs =
    the same weights in the same order, those with a transform re-laid-out, the rest passed through (reference
    ``:131-250``)."""
    by_idx = {c["weight_idx"]: c for c in hlo_stub["computations"]
              if TRANSPOSABLE_WEIGHT_IDX in c.get("frontend_attributes", {})}
    wlt = {"name": hlo_stub["name"] + "_wlt", "id": hlo_stub["id"], "frontend_attributes": {}, "weights": {},
           "computations": [], "outputs": []}
    for name, idx in sorted(weight_name_to_idx.items(), key=lambda kv: kv[1]):
        w = hlo_stub["weights"][name]
        wlt["weights"][name] = dict(w, idx=len(wlt["weights"]))
        cpt = by_idx.get(idx)
        shape = list(w["shape"]) if cpt is None else [w["shape"][d] for d in cpt["perm"]]
        dtype = w["dtype"] if cpt is None or not cpt.get("dtype") else cpt["dtype"]
        if cpt is not None:
            wlt["computations"].append(dict(copy.deepcopy(cpt), weight_idx=wlt["weights"][name]["idx"]))
        wlt["outputs"].append({"name": name, "shape": shape, "dtype": dtype, "transformed": cpt is not None})
    return wlt


def prepare_metaneff_for_wlt_hlo(wlt_hlo: Graph, weight_name_to_idx: Dict[str, int]) -> Dict[str, List[Dict[str, Any]]]:
    """IO description of the transform program: ``input{i}`` / ``output{i}`` with shape, dtype and the checkpoint key
    they are bound to (reference ``:253-283`` builds a MetaNeff proto)."""
    names = [n.replace("->", ".") for n, _ in sorted(weight_name_to_idx.items(), key=lambda kv: kv[1])]
    ins, outs = [], []
    for i, n in enumerate(names):
        w = wlt_hlo["weights"][n]
        ins.append({"name": f"input{i}", "shape": list(w["shape"]), "data_type": w["dtype"], "type": "INPUT_WEIGHT",
                    "checkpoint_key": n})
        o = wlt_hlo["outputs"][i]
        outs.append({"name": f"output{i}", "shape": list(o["shape"]), "data_type": o["dtype"], "checkpoint_key": n})
    return {"input_tensors": ins, "output_tensors": outs}


def run_weight_layout_transform(wlt_hlo: Graph, weights: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Execute the transform program on a (sharded) checkpoint: what ``NxDModel.initialize`` runs before binding the
    weights.  Permute + contiguous copy (+ optional dtype cast) on whatever device the tensors live on."""
    by_idx = {c["weight_idx"]: c for c in wlt_hlo["computations"]}
    out = dict(weights)
    for name, w in wlt_hlo["weights"].items():
        cpt = by_idx.get(w["idx"])
        if cpt is None or name not in weights:
            continue
        t = weights[name].permute(*cpt["perm"]).contiguous()
        if cpt.get("dtype"):
            t = t.to(getattr(torch, cpt["dtype"]))
        out[name] = t
    return out


__all__ = ["REQUIRE_TRANSPOSE_CUSTOM_CALL", "REQUIRE_TRANSPOSE_WEIGHT_IDX", "TRANSPOSABLE_WEIGHT_IDX",
           "add_weight_idx_attr_to_hlo", "append_layout_computation_to_hlo", "extract_weight_layout_transform_hlo",
           "get_layout_transform_map", "new_graph", "prepare_metaneff_for_wlt_hlo", "read_hlo",
           "run_weight_layout_transform", "suggest_layout_transforms", "update_computation_id_and_name",
           "weight_name_to_idx_of", "write_hlo"]
