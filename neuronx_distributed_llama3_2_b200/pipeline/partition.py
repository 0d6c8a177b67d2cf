"""Cut a traced model into pipeline stages and derive each stage's IO.

Reference: ``src/neuronx_distributed/pipeline/partition.py:18-303``.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.fx as fx
from torch.fx.passes.split_module import split_module


@dataclass
class PipelineIO:
    """One value crossing a stage boundary: ``name`` is the FX node name in the split graph."""
    name: str
    input_idx: Optional[int] = None   # position among the consuming stage's placeholders
    output_idx: Optional[int] = None  # position among the producing stage's outputs
    metadata: Optional[list] = None
    obj: Optional[object] = None

    def __repr__(self):
        return f"PipelineIO({self.name}, in={self.input_idx}, out={self.output_idx})"


@dataclass
class StageIO:
    model_inputs: List[Tuple[str, int]] = field(default_factory=list)   # (model input name, placeholder index)
    recv: List[PipelineIO] = field(default_factory=list)                # values received from the previous stage
    send: List[PipelineIO] = field(default_factory=list)                # values sent to the next stage
    num_placeholders: int = 0
    returns_model_output: bool = False


def create_partitions(num_stages: int, num_layers) -> List:
    """Even split of ``num_layers`` over ``num_stages``, remainder to the LATER stages; returns the 0-based index of
    the last layer of every stage but the final one (reference ``:280-303``).  The reference's call form
    ``create_partitions(pipeline_parallel_size, model_layer_names)`` -- a LIST of layer names -- returns the names of
    those layers (the pipeline cuts) instead."""
    if not isinstance(num_layers, int):
        names = list(num_layers)
        return [names[i] for i in create_partitions(num_stages, len(names))]
    if num_stages > num_layers:
        raise ValueError(f"cannot cut {num_layers} layers into {num_stages} stages")
    base, rem = divmod(num_layers, num_stages)
    sizes = [base + (1 if s >= num_stages - rem else 0) for s in range(num_stages)]
    cuts, acc = [], 0
    for s in sizes[:-1]:
        acc += s
        cuts.append(acc - 1)
    return cuts


def stage_to_pipeline_parallel_rank(stage: int, pp_size: int) -> int:
    """Virtual stages are dealt round-robin (reference ``:226-229``)."""
    return stage % pp_size


def partition_traced_model(traced: fx.GraphModule, cut_names: Sequence[str]) -> fx.GraphModule:
    """A new stage starts right AFTER each node whose name / call_module target is in ``cut_names``."""
    cut_set = set(cut_names)
    found = set()
    stage_of: Dict[fx.Node, int] = {}
    cur = 0
    for node in traced.graph.nodes:
        if node.op in ("placeholder", "output"):
            continue
        stage_of[node] = cur
        key = node.target if node.op == "call_module" else node.name
        hit = (isinstance(key, str) and key in cut_set) or node.name in cut_set
        if hit:
            found.add(key if isinstance(key, str) and key in cut_set else node.name)
            cur += 1
    missing = cut_set - found
    if missing:
        raise ValueError(f"pipeline cuts {sorted(missing)} not found in the traced graph; candidates: "
                         f"{[n.target for n in traced.graph.nodes if n.op == 'call_module'][:50]}")
    # getitem nodes that unpack a cut node's tuple output belong to the NEXT stage (they were assigned there already
    # because they come after the cut node in topological order)
    split = split_module(traced, None, lambda n: stage_of[n])
    _sink_root_attributes(split)
    return split


def _sink_root_attributes(split: fx.GraphModule) -> None:
    """Parameters / buffers that the model's own ``forward`` reads directly (``self.bias``, a norm that is traced through)
    are ``get_attr`` nodes of the ROOT graph after ``split_module`` and reach the stages as call arguments.  A stage must
    own what it uses -- local parameter lists, optimizer, checkpoint names all go by stage -- so each such tensor is
    registered on the consuming stage module (the same object on every stage that uses it: a shared weight) and read there."""
    changed = False
    for node in list(split.graph.nodes):
        if node.op != "get_attr":
            continue
        obj = split
        for part in node.target.split("."):
            obj = getattr(obj, part)
        if not isinstance(obj, torch.Tensor):
            continue
        for user in list(node.users):
            if user.op != "call_module" or not str(user.target).startswith("submod_"):
                continue
            sub = getattr(split, user.target)
            local = "_root_" + node.target.replace(".", "_")
            if not hasattr(sub, local):
                if isinstance(obj, torch.nn.Parameter):
                    sub.register_parameter(local, obj)
                else:
                    sub.register_buffer(local, obj)
            while node in user.args:
                pos = list(user.args).index(node)
                ph = [n for n in sub.graph.nodes if n.op == "placeholder"][pos]
                with sub.graph.inserting_after(ph):
                    inner = sub.graph.get_attr(local)
                ph.replace_all_uses_with(inner)
                sub.graph.erase_node(ph)
                user.args = tuple(a for i, a in enumerate(user.args) if i != pos)
            sub.graph.lint()
            sub.recompile()
            changed = True
        if not node.users:
            split.graph.erase_node(node)
    if changed:
        split.graph.lint()
        split.recompile()


def analyze_pipeline_module(split: fx.GraphModule) -> Tuple[List[str], List[StageIO]]:
    """Derive, for every ``submod_k``, which placeholders are model inputs, which values it receives from stage k-1
    and which it sends to stage k+1 -- including PASS-THROUGH values that a later stage needs (reference
    ``:76-107,132-223``)."""
    model_inputs = [n.name for n in split.graph.nodes if n.op == "placeholder"]
    calls = [n for n in split.graph.nodes if n.op == "call_module"]
    calls.sort(key=lambda n: int(n.target.split("_")[-1]))
    n_stages = len(calls)
    producer: Dict[str, int] = {}
    aliases: Dict[str, Tuple[str, Optional[int]]] = {}
    for k, c in enumerate(calls):
        producer[c.name] = k
    for n in split.graph.nodes:   # getitem on a multi-output submodule
        if n.op == "call_function" and getattr(n.target, "__name__", "") == "getitem":
            src, idx = n.args
            aliases[n.name] = (src.name, idx)
            producer[n.name] = producer[src.name]
    out_node = next(n for n in split.graph.nodes if n.op == "output")
    final_vals = set()

    def collect(a):
        if isinstance(a, fx.Node):
            final_vals.add(a.name)
        elif isinstance(a, (tuple, list)):
            for x in a:
                collect(x)
        elif isinstance(a, dict):
            for x in a.values():
                collect(x)

    collect(out_node.args)
    ios = [StageIO() for _ in range(n_stages)]
    needed_at: Dict[str, int] = {}    # value -> last stage that consumes it
    for k, c in enumerate(calls):
        ios[k].num_placeholders = len(c.args)
        for pos, a in enumerate(c.args):
            assert isinstance(a, fx.Node), "stage inputs must be graph values"
            if a.op == "placeholder":
                ios[k].model_inputs.append((a.name, pos))
            else:
                needed_at[a.name] = max(needed_at.get(a.name, -1), k)
    # values travel stage by stage from their producer to their last consumer
    for name, last in needed_at.items():
        src = producer[name]
        for hop in range(src, last):
            ios[hop].send.append(PipelineIO(name))
            ios[hop + 1].recv.append(PipelineIO(name))
    for k, c in enumerate(calls):
        for pos, a in enumerate(c.args):
            if a.op != "placeholder":
                for io in ios[k].recv:
                    if io.name == a.name:
                        io.input_idx = pos
        # output index of values this stage produced itself
        for io in ios[k].send:
            if producer[io.name] == k:
                io.output_idx = aliases[io.name][1] if io.name in aliases else None
        ios[k].returns_model_output = any(producer.get(v) == k for v in final_vals)
    return model_inputs, ios


def iterate_graph_model_outputs(output_node_args):
    """Yield the values an FX ``output`` node returns: its ``args`` is a 1-tuple holding either one node or a tuple of
    them (reference ``:108-129``)."""
    if not (isinstance(output_node_args, tuple) and len(output_node_args) == 1):
        raise AssertionError(f"unsupported output args {output_node_args}")
    inner = output_node_args[0]
    yield from (inner if isinstance(inner, (tuple, list)) else (inner,))


def adding_live_obj_for_previous_stages(stage_id_to_IO_input_names: Sequence[Dict[str, PipelineIO]],
                                        stage_id_to_IO_output_names: Sequence[Dict[str, PipelineIO]], obj_name: str,
                                        current_stage: int) -> None:
    """Make ``obj_name`` -- needed by the stage after ``current_stage`` -- travel there from whichever earlier stage
    holds it: every stage on the way that neither consumes nor produces it gets it as a pass-through input AND output;
    the stage where it is found gets it as an output (reference ``:76-105``; ``analyze_pipeline_module`` does the same
    bookkeeping with hop lists).  Walks backwards instead of recursing."""
    passthrough = []
    k = current_stage
    while True:
        if k < 0:
            raise RuntimeError(f"{obj_name} is not available in any stage up to {current_stage}; outputs: "
                               f"{[sorted(d) for d in stage_id_to_IO_output_names]}")
        if obj_name in stage_id_to_IO_input_names[k] or obj_name in stage_id_to_IO_output_names[k]:
            break
        passthrough.append(k)
        k -= 1
    stage_id_to_IO_output_names[k].setdefault(obj_name, PipelineIO(obj_name))
    for hop in passthrough:
        stage_id_to_IO_input_names[hop][obj_name] = PipelineIO(obj_name)
        stage_id_to_IO_output_names[hop][obj_name] = PipelineIO(obj_name)


def analyze_shared_weights_across_stages(split: fx.GraphModule) -> List[List[Tuple[str, int]]]:
    """Parameters (same tensor object) that live in more than one stage, e.g. tied embeddings
    (reference ``:232-277``).  Returns ``[[(local_name, stage), ...], ...]``."""
    seen: Dict[int, List[Tuple[str, int]]] = {}
    for cname, child in split.named_children():
        if not cname.startswith("submod_"):
            continue
        stage = int(cname.split("_")[-1])
        for pname, p in child.named_parameters(remove_duplicate=False):
            entry = seen.setdefault(id(p), [])
            if all(s != stage for _, s in entry):
                entry.append((pname, stage))
    return [v for v in seen.values() if len(v) > 1]
