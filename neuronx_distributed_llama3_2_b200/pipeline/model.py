"""``NxDPPModel``: pipeline-parallel wrapper with 1F1B / interleaved execution.

Reference: ``src/neuronx_distributed/pipeline/model.py:54-1641``.  Same constructor surface and the same observable
behaviour (FX trace -> cut at transformer layers -> one ``submod_k`` per stage -> schedule-driven execution -> loss =
mean over micro-batches; ``local_*`` accessors and state dicts under the ORIGINAL module names so checkpoints are
PP-agnostic; tied weights living on two stages are synchronised and their grads all-reduced).

What is different on B200 / eager CUDA:
  * no dry-run "shape inference" graph: tensor metadata crosses the boundary once, with the first micro-batch, through
    ``send_object_list``; afterwards only raw tensors move (NCCL p2p on its own stream, gloo on CPU);
  * no ``mark_step`` graph cuts -- tasks are plain eager calls on the compute stream;
  * every rank receives the whole batch (as in the reference) and each stage picks the model inputs it needs by
    name, so labels never travel through the pipe.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Any, Dict, Iterator, List, Optional, Tuple

import torch
import torch.distributed as dist
import torch.nn as nn

from ..parallel import parallel_state as ps
from ..utils.logger import get_logger
from ..utils.serialization import SerializationManager, TensorMeta, find_loss_from_output_and_spec
from . import comm
from . import scheduler as sch
from .partition import (StageIO, analyze_pipeline_module, analyze_shared_weights_across_stages, create_partitions,
                        partition_traced_model, stage_to_pipeline_parallel_rank)
from .timeline import PPTimeline
from .trace import trace_model

logger = get_logger()


class NxDPPModel(nn.Module):
    def __init__(self, module: nn.Module, transformer_layer_cls: Optional[type] = None, num_microbatches: int = 1,
                 virtual_pipeline_size: int = 1, output_loss_value_spec: Any = None, return_mb_loss: bool = False,
                 broadcast_and_average_loss: bool = False, pipeline_cuts: Optional[List[str]] = None,
                 input_names: Optional[List[str]] = None, leaf_module_cls: Optional[List[Any]] = None,
                 autowrap_functions: Optional[Tuple[Any, ...]] = None, autowrap_modules: Optional[Tuple[Any, ...]] = None,
                 autowrap_obj_methods: Optional[Dict[Any, List[str]]] = None, tracer_cls=None, param_init_fn=None,
                 trace_file_path: Optional[str] = None, use_zero1_optimizer: bool = False, auto_partition: bool = False,
                 deallocate_pipeline_outputs: bool = False, use_model_wrapper: bool = False,
                 use_optimizer_wrapper: bool = False, return_loss_on_cpu: bool = True, fuse_microbatches: bool = False,
                 _turn_off_odd_even_scheduler: bool = False, _all_reduce_send_recv: bool = False,
                 _fused_send_recv: bool = False, _fused_fwd_bwd: bool = False, _deallocate_send_tensors: bool = True,
                 _use_gloo_for_metadata_comm: bool = False, _debug_mode: bool = False, _debug_pp_size: int = 8,
                 _debug_pp_rank: int = 0, _delay_tracing: bool = False, _mark_step_before_pp_runtime: bool = True,
                 _mark_step_after_pp_runtime: bool = True):
        # the two ``_mark_step_*`` switches cut the XLA graph around the pipeline runtime in the reference; eager CUDA
        # execution has no graph to cut, they are accepted so that reference configs load unchanged
        super().__init__()
        self.original_torch_module = module
        self.transformer_layer_cls = transformer_layer_cls
        self.num_microbatches = num_microbatches
        self.virtual_pipeline_size = virtual_pipeline_size
        self.output_loss_value_spec = output_loss_value_spec
        self.return_mb_loss = return_mb_loss
        self.broadcast_and_average_loss = broadcast_and_average_loss
        self.pipeline_cuts = list(pipeline_cuts or [])
        self.input_names = input_names
        self.leaf_module_cls = list(leaf_module_cls or [])
        self.autowrap_functions = autowrap_functions
        self.autowrap_modules = autowrap_modules
        self.autowrap_obj_methods = autowrap_obj_methods
        self.tracer_cls = tracer_cls
        self.param_init_fn = param_init_fn
        self.use_zero1_optimizer = use_zero1_optimizer
        self.auto_partition = auto_partition
        self.deallocate_pipeline_outputs = deallocate_pipeline_outputs
        self.use_model_wrapper = use_model_wrapper
        self.use_optimizer_wrapper = use_optimizer_wrapper
        self.return_loss_on_cpu = return_loss_on_cpu
        self._debug_mode = _debug_mode
        self._fused_send_recv = _fused_send_recv
        self._fused_fwd_bwd = _fused_fwd_bwd
        self._turn_off_odd_even_scheduler = _turn_off_odd_even_scheduler
        if _debug_mode:
            self.pipeline_parallel_size, self.pipeline_parallel_rank = _debug_pp_size, _debug_pp_rank
        else:
            self.pipeline_parallel_size = ps.get_pipeline_model_parallel_size()
            self.pipeline_parallel_rank = ps.get_pipeline_model_parallel_rank()
        if transformer_layer_cls is not None and transformer_layer_cls not in self.leaf_module_cls:
            self.leaf_module_cls.append(transformer_layer_cls)
        self.serialization_manager = SerializationManager()
        self.timeline = PPTimeline(trace_file_path, self.pipeline_parallel_rank)
        self.partitioned = False
        self.shared_weights_name_to_pg: Dict[str, Any] = {}
        self._shape_known: Dict[Tuple[str, int], Any] = {}
        self.training_mode = True
        self.losses: List[torch.Tensor] = []
        self.trace(input_names=input_names)
        if self.auto_partition or self.pipeline_cuts:
            self.partition()

    # ------------------------------------------------------------------------------------------------------
    # trace / cut / partition
    # ------------------------------------------------------------------------------------------------------
    def trace(self, args=None, kwargs=None, input_names=None, leaf_modules=None, autowrap_functions=None,
              autowrap_modules=None, autowrap_obj_methods=None, tracer_cls=None):
        if input_names is None and kwargs:
            input_names = list(kwargs.keys())
        leaves = list(self.leaf_module_cls) + list(leaf_modules or [])
        self.traced_model = trace_model(self.original_torch_module, input_names=input_names, leaf_modules=leaves,
                                        autowrap_functions=autowrap_functions or self.autowrap_functions,
                                        autowrap_modules=autowrap_modules or self.autowrap_modules,
                                        autowrap_obj_methods=autowrap_obj_methods or self.autowrap_obj_methods,
                                        tracer_cls=tracer_cls or self.tracer_cls)
        self.input_names = input_names

    def get_transformer_layer_names(self) -> List[str]:
        names = []
        for node in self.traced_model.graph.nodes:
            if node.op == "call_module" and self.transformer_layer_cls is not None:
                if isinstance(self.traced_model.get_submodule(node.target), self.transformer_layer_cls):
                    names.append(node.target)
        return names

    def cut_pipeline_stage(self, cut_point: str) -> None:
        assert not self.partitioned, "cuts must be added before partition()"
        if cut_point not in self.pipeline_cuts:
            self.pipeline_cuts.append(cut_point)

    def _auto_cuts(self) -> List[str]:
        layers = self.get_transformer_layer_names()
        n_stages = self.pipeline_parallel_size * self.virtual_pipeline_size
        if not layers:
            raise ValueError("auto_partition needs `transformer_layer_cls` layers in the traced model")
        return [layers[i] for i in create_partitions(n_stages, len(layers))]

    def partition(self) -> None:
        if self.partitioned:
            return
        n_stages = self.pipeline_parallel_size * self.virtual_pipeline_size
        if self.auto_partition and not self.pipeline_cuts:
            self.pipeline_cuts = self._auto_cuts()
        if len(self.pipeline_cuts) != n_stages - 1:
            raise ValueError(f"need {n_stages - 1} pipeline cuts for {n_stages} stages, got {self.pipeline_cuts}")
        self.split = partition_traced_model(self.traced_model, self.pipeline_cuts)
        self.model_input_names, self.stage_io = analyze_pipeline_module(self.split)
        self.num_stages = len(self.stage_io)
        assert self.num_stages == n_stages, f"partition produced {self.num_stages} stages, expected {n_stages}"
        self.stage_modules: List[nn.Module] = [getattr(self.split, f"submod_{k}") for k in range(self.num_stages)]
        self.local_stage_ids = [k for k in range(self.num_stages)
                                if stage_to_pipeline_parallel_rank(k, self.pipeline_parallel_size)
                                == self.pipeline_parallel_rank]
        self.local_stage_modules = nn.ModuleList([self.stage_modules[k] for k in self.local_stage_ids])
        self._build_name_maps()
        self._mark_shared_weights()
        self._drop_non_local_parameters()
        self._create_schedules()
        self.partitioned = True

    # -- names: local (split graph) <-> original --------------------------------------------------------------
    def _build_name_maps(self) -> None:
        orig_mod_name = {id(m): n for n, m in self.original_torch_module.named_modules()}
        orig_param_name = {}
        for n, p in self.original_torch_module.named_parameters(remove_duplicate=False):
            orig_param_name.setdefault(id(p), []).append(n)
        orig_buf_name = {id(b): n for n, b in self.original_torch_module.named_buffers(remove_duplicate=False)}
        self.local_name_to_original_name: Dict[str, str] = {}
        self.original_name_to_local_name: Dict[str, str] = {}
        self._stage_of_original_param: Dict[str, int] = {}
        for k, sm in enumerate(self.stage_modules):
            prefix = f"submod_{k}."
            child_map = {}
            for cname, child in sm.named_modules():
                if cname and id(child) in orig_mod_name:
                    child_map[cname] = orig_mod_name[id(child)]
            def translate(local: str, tensor, table) -> Optional[str]:
                # longest module prefix that is an original module
                parts = local.split(".")
                for cut in range(len(parts) - 1, 0, -1):
                    head = ".".join(parts[:cut])
                    if head in child_map:
                        base = child_map[head]
                        return (base + "." if base else "") + ".".join(parts[cut:])
                names = table.get(id(tensor))
                if isinstance(names, list):
                    return names[0]
                return names
            for pname, p in sm.named_parameters(remove_duplicate=False):
                orig = translate(pname, p, orig_param_name)
                if orig is None:
                    continue
                # a tied parameter has several original names: keep the one that belongs to this module path
                cands = orig_param_name.get(id(p), [orig])
                if orig not in cands and cands:
                    orig = cands[0]
                self.local_name_to_original_name[prefix + pname] = orig
                self.original_name_to_local_name.setdefault(orig, prefix + pname)
                self._stage_of_original_param.setdefault(orig, k)
            for bname, b in sm.named_buffers(remove_duplicate=False):
                orig = translate(bname, b, orig_buf_name)
                if orig is not None:
                    self.local_name_to_original_name[prefix + bname] = orig
                    self.original_name_to_local_name.setdefault(orig, prefix + bname)

    def _mark_shared_weights(self) -> None:
        """Tied weights used by several stages: tag ``p.shared`` and (when the stages sit on different ranks) build
        the process group used for the initial sync and the grad all-reduce (reference ``model.py:591-641``)."""
        self.shared_weights: List[Dict[str, Any]] = []
        for group in analyze_shared_weights_across_stages(self.split):
            stages = sorted({s for _, s in group})
            pp_ranks = sorted({stage_to_pipeline_parallel_rank(s, self.pipeline_parallel_size) for s in stages})
            local = [(n, s) for n, s in group if s in self.local_stage_ids]
            param = None
            if local:
                param = dict(self.stage_modules[local[0][1]].named_parameters(remove_duplicate=False))[local[0][0]]
            entry = {"group": group, "pp_ranks": pp_ranks, "param": param, "pg": None}
            if param is not None and len(pp_ranks) > 1:
                # the tag says "skip me in the global grad norm": after the grad all-reduce every owner holds the same
                # summed gradient, so it is counted on the FIRST owning pp rank only (reference partition.py:240-268)
                setattr(param, "shared", self.pipeline_parallel_rank != pp_ranks[0])
            self.shared_weights.append(entry)
        if self._debug_mode or self.pipeline_parallel_size == 1 or not dist.is_initialized():
            return
        comm.init_p2p_groups()
        pp_global = ps.get_pipeline_model_parallel_ranks()
        for entry in self.shared_weights:
            if len(entry["pp_ranks"]) < 2:
                continue
            ranks = [pp_global[r] for r in entry["pp_ranks"]]
            mine = ranks if self.pipeline_parallel_rank in entry["pp_ranks"] else []
            # every rank must take part in group creation; ranks outside contribute their own PP column's list
            entry["pg"] = ps.create_pg_with_ranks(ranks)
            if not mine:
                entry["pg"] = None

    def _sync_shared_weights(self) -> None:
        for entry in self.shared_weights:
            if entry["pg"] is not None and entry["param"] is not None:
                src = dist.get_process_group_ranks(entry["pg"])[0]
                dist.broadcast(entry["param"].data, src=src, group=entry["pg"])

    def _reduce_shared_weights(self) -> None:
        for entry in self.shared_weights:
            p = entry["param"]
            if entry["pg"] is not None and p is not None:
                g = p.grad if p.grad is not None else getattr(p, "main_grad", None)
                if g is None:
                    g = torch.zeros_like(p)
                    p.grad = g
                dist.all_reduce(g, group=entry["pg"])

    def _drop_non_local_parameters(self) -> None:
        local_ids = {id(p) for m in self.local_stage_modules for p in m.parameters()}
        local_ids |= {id(b) for m in self.local_stage_modules for b in m.buffers()}
        for k, sm in enumerate(self.stage_modules):
            if k in self.local_stage_ids:
                continue
            for p in sm.parameters():
                if id(p) not in local_ids:
                    p.requires_grad_(False)
                    p.data = torch.empty(0, dtype=p.dtype, device=p.device)
            for b in sm.buffers():
                if id(b) not in local_ids:
                    b.data = torch.empty(0, dtype=b.dtype, device=b.device)

    def _create_schedules(self) -> None:
        pp, r, vp = self.pipeline_parallel_size, self.pipeline_parallel_rank, self.virtual_pipeline_size
        if vp > 1:
            self.train_scheduler = sch.TrainInterleavedSchedule(
                self.num_microbatches, vp, pp, r, fused_send_recv=self._fused_send_recv,
                fused_fwd_bwd=self._fused_fwd_bwd, use_odd_even_scheduler=not self._turn_off_odd_even_scheduler)
        else:
            self.train_scheduler = sch.Train1F1BSchedule(self.num_microbatches, pp, r)
        self.eval_scheduler = sch.InferenceSchedule(self.num_microbatches, pp, r, num_model_chunks=vp)

    # ------------------------------------------------------------------------------------------------------
    # execution
    # ------------------------------------------------------------------------------------------------------
    def forward(self, *a, **kw):
        raise RuntimeError("NxDPPModel: call run_train() / run_eval() (reference model.py:1410-1429)")

    def run_train(self, **kwargs):
        self.training_mode = True
        return self._run(self.train_scheduler, kwargs)

    def run_eval(self, **kwargs):
        self.training_mode = False
        with torch.no_grad():
            return self._run(self.eval_scheduler, kwargs)

    def _stage_id(self, chunk: int) -> int:
        return self.local_stage_ids[chunk]

    def _split_microbatches(self, kwargs: Dict[str, Any]) -> List[Dict[str, Any]]:
        n = self.num_microbatches
        out = [dict() for _ in range(n)]
        for name, val in kwargs.items():
            if isinstance(val, torch.Tensor) and val.dim() > 0:
                if val.shape[0] % n != 0:
                    raise ValueError(f"batch dim of `{name}` ({val.shape[0]}) is not divisible by num_microbatches={n}")
                for i, c in enumerate(torch.chunk(val, n, dim=0)):
                    out[i][name] = c
            else:
                for i in range(n):
                    out[i][name] = val
        return out

    def _device(self) -> torch.device:
        for p in self.local_stage_modules.parameters():
            return p.device
        return torch.device("cpu")

    def _run(self, schedule, kwargs):
        if not self.partitioned:
            self.partition()
        if not getattr(self, "_synced", False):
            self._sync_shared_weights()
            self._synced = True
        dev = self._device()
        kwargs = {k: (v.to(dev, non_blocking=True) if isinstance(v, torch.Tensor) else v) for k, v in kwargs.items()}
        self.microbatches = self._split_microbatches(kwargs)
        self.recv_values: Dict[Tuple[int, int], Dict[str, Any]] = {}
        self.recv_tensors: Dict[Tuple[int, int], List[torch.Tensor]] = {}
        self.stage_outputs: Dict[Tuple[int, int], Any] = {}
        self.sent_tensors: Dict[Tuple[int, int], List[torch.Tensor]] = {}
        self.recv_grads: Dict[Tuple[int, int], List[Optional[torch.Tensor]]] = {}
        self.losses = []
        self.eval_outputs = []
        for step in schedule.steps():
            for task in step:
                self._dispatch(task)
        comm.flush_sends()
        self.timeline.mark_step_end(ps.get_pp_gloo_group() if False else None)
        return self._process_loss() if self.training_mode else self._process_eval_outputs()

    def _dispatch(self, task) -> None:
        name = type(task).__name__
        self.timeline.mark_event_start(f"{name}_mb{task.mb}")
        if isinstance(task, sch.ForwardPreprocessTask):
            self._fwd_preprocess(task.mb, task.model_chunk)
        elif isinstance(task, sch.ForwardStepTask):
            self._fwd_step(task.mb, task.model_chunk)
        elif isinstance(task, sch.ForwardPostprocessTask):
            self._fwd_postprocess(task.mb, task.model_chunk)
        elif isinstance(task, sch.BackwardPreprocessTask):
            self._bwd_preprocess(task.mb, task.model_chunk)
        elif isinstance(task, sch.BackwardStepTask):
            self._bwd_step(task.mb, task.model_chunk)
        elif isinstance(task, sch.BackwardPostprocessTask):
            self._bwd_postprocess(task.mb, task.model_chunk)
        elif isinstance(task, sch.ReduceGradsTask):
            self._reduce_shared_weights()
        else:
            raise RuntimeError(f"unknown pipeline task {task}")
        self.timeline.mark_event_end(f"{name}_mb{task.mb}")

    # -- forward ---------------------------------------------------------------------------------------------
    def _fwd_preprocess(self, mb: int, chunk: int) -> None:
        stage = self._stage_id(chunk)
        io: StageIO = self.stage_io[stage]
        if stage == 0 or not io.recv:
            self.recv_values[(mb, chunk)] = {}
            self.recv_tensors[(mb, chunk)] = []
            return
        key = ("fwd", stage)
        if key not in self._shape_known:  # first micro-batch ever: learn structure + tensor metadata
            self._shape_known[key] = comm.recv_python_object(recv_prev=True)
        stubs, metas = self._shape_known[key]
        tensors = []
        for m in metas:
            t = comm.recv_from(m, recv_prev=True, device=self._device())
            if m.requires_grad and self.training_mode and t.is_floating_point():
                t.requires_grad_(True)
            tensors.append(t)
        values = self.serialization_manager.deserialize(stubs, tensors)
        self.recv_values[(mb, chunk)] = values
        self.recv_tensors[(mb, chunk)] = tensors

    def _fwd_step(self, mb: int, chunk: int) -> None:
        stage = self._stage_id(chunk)
        io = self.stage_io[stage]
        args: List[Any] = [None] * io.num_placeholders
        batch = self.microbatches[mb]
        for name, pos in io.model_inputs:
            if name not in batch:
                raise KeyError(f"stage {stage} needs model input `{name}`; run_train got {list(batch)}")
            args[pos] = batch[name]
        values = self.recv_values.get((mb, chunk), {})
        for r in io.recv:
            if r.input_idx is not None:
                args[r.input_idx] = values[r.name]
        out = self.stage_modules[stage](*args)
        self.stage_outputs[(mb, chunk)] = out
        if stage == self.num_stages - 1:
            if self.training_mode:
                loss = find_loss_from_output_and_spec(out, self.output_loss_value_spec)
                if loss is None:
                    raise RuntimeError("could not find the loss in the last stage's output (output_loss_value_spec)")
                self.losses.append(loss)
            else:
                self.eval_outputs.append(out)

    def _values_to_send(self, mb: int, chunk: int) -> Dict[str, Any]:
        stage = self._stage_id(chunk)
        io = self.stage_io[stage]
        out = self.stage_outputs[(mb, chunk)]
        received = self.recv_values.get((mb, chunk), {})
        vals = OrderedDict()
        for s in io.send:
            if s.name in received:               # pass-through value from an earlier stage
                vals[s.name] = received[s.name]
            elif s.output_idx is None:
                vals[s.name] = out
            else:
                vals[s.name] = out[s.output_idx]
        return vals

    def _fwd_postprocess(self, mb: int, chunk: int) -> None:
        stage = self._stage_id(chunk)
        if stage == self.num_stages - 1:
            return
        vals = self._values_to_send(mb, chunk)
        stubs, tensors, metas = self.serialization_manager.serialize(vals)
        key = ("fwd_sent", stage)
        if key not in self._shape_known:
            comm.send_python_object((stubs, metas), send_next=True)
            self._shape_known[key] = True
        for t in tensors:
            comm.send(t, send_next=True)
        self.sent_tensors[(mb, chunk)] = tensors
        if not self.training_mode:
            self.stage_outputs.pop((mb, chunk), None)

    # -- backward --------------------------------------------------------------------------------------------
    def _bwd_preprocess(self, mb: int, chunk: int) -> None:
        stage = self._stage_id(chunk)
        if stage == self.num_stages - 1:
            return
        grads: List[Optional[torch.Tensor]] = []
        for t in self.sent_tensors[(mb, chunk)]:
            if t.requires_grad:
                m = TensorMeta(0, t.dtype, tuple(t.shape), False)
                grads.append(comm.recv_from(m, recv_prev=False, device=t.device))
            else:
                grads.append(None)
        self.recv_grads[(mb, chunk)] = grads

    def _bwd_step(self, mb: int, chunk: int) -> None:
        stage = self._stage_id(chunk)
        if stage == self.num_stages - 1:
            loss = self.losses[mb]
            (loss / self.num_microbatches).backward()
            self.losses[mb] = loss.detach()
        else:
            outs, gs = [], []
            for t, g in zip(self.sent_tensors.pop((mb, chunk)), self.recv_grads.pop((mb, chunk))):
                # pass-through tensors were received, not produced, here: their grad flows straight back
                if g is not None and t.grad_fn is not None:
                    outs.append(t)
                    gs.append(g)
                elif g is not None and t.requires_grad:
                    t.grad = g if t.grad is None else t.grad + g
            if outs:
                torch.autograd.backward(outs, gs)
        self.stage_outputs.pop((mb, chunk), None)

    def _bwd_postprocess(self, mb: int, chunk: int) -> None:
        stage = self._stage_id(chunk)
        if stage == 0:
            return
        for t in self.recv_tensors.pop((mb, chunk)):
            if t.requires_grad:
                g = t.grad if t.grad is not None else torch.zeros_like(t)
                comm.send(g, send_next=False)
        self.recv_values.pop((mb, chunk), None)

    # -- results ---------------------------------------------------------------------------------------------
    def _process_loss(self):
        """Mean of the micro-batch losses on the last PP rank; optional DP average + PP broadcast
        (reference ``model.py:1611-1641``)."""
        last = self.pipeline_parallel_rank == self.pipeline_parallel_size - 1
        loss = None
        if last and self.losses:
            stacked = torch.stack([l.detach().float() for l in self.losses])
            loss = stacked if self.return_mb_loss else stacked.mean()
        if self.broadcast_and_average_loss and dist.is_initialized() and not self._debug_mode:
            dev = self._device()
            shape = (self.num_microbatches,) if self.return_mb_loss else ()
            buf = loss.to(dev) if loss is not None else torch.zeros(shape, device=dev)
            if ps.get_data_parallel_size() > 1 and last:
                dist.all_reduce(buf, group=ps.get_data_parallel_group())
                buf = buf / ps.get_data_parallel_size()
            if self.pipeline_parallel_size > 1:
                src = ps.get_pipeline_model_parallel_ranks()[-1]
                dist.broadcast(buf, src=src, group=ps.get_pipeline_model_parallel_group())
            loss = buf
        return loss

    def _process_eval_outputs(self):
        if self.pipeline_parallel_rank != self.pipeline_parallel_size - 1:
            return None
        return self.eval_outputs

    # ------------------------------------------------------------------------------------------------------
    # local views (reference model.py:1447-1571)
    # ------------------------------------------------------------------------------------------------------
    def _local_prefix_items(self, getter: str, *a, **kw):
        for k in self.local_stage_ids:
            for n, v in getattr(self.stage_modules[k], getter)(*a, **kw):
                yield f"submod_{k}.{n}", v

    def local_named_parameters(self, *a, **kw) -> Iterator[Tuple[str, nn.Parameter]]:
        seen = set()
        for n, p in self._local_prefix_items("named_parameters", *a, **kw):
            if id(p) in seen:
                continue
            seen.add(id(p))
            yield self.local_name_to_original_name.get(n, n), p

    def local_parameters(self, *a, **kw) -> Iterator[nn.Parameter]:
        for _, p in self.local_named_parameters(*a, **kw):
            yield p

    def local_named_buffers(self, *a, **kw):
        for n, b in self._local_prefix_items("named_buffers", *a, **kw):
            yield self.local_name_to_original_name.get(n, n), b

    def local_buffers(self, *a, **kw):
        for _, b in self.local_named_buffers(*a, **kw):
            yield b

    def local_named_modules(self, *a, **kw):
        for k in self.local_stage_ids:
            for n, m in self.stage_modules[k].named_modules(*a, **kw):
                yield f"submod_{k}.{n}" if n else f"submod_{k}", m

    def local_modules(self):
        for _, m in self.local_named_modules():
            yield m

    def local_named_children(self):
        for k in self.local_stage_ids:
            yield f"submod_{k}", self.stage_modules[k]

    def local_children(self):
        for _, m in self.local_named_children():
            yield m

    def local_state_dict(self, *a, **kw) -> Dict[str, torch.Tensor]:
        """State of the local stages keyed by ORIGINAL names."""
        out = OrderedDict()
        for k in self.local_stage_ids:
            for n, v in self.stage_modules[k].state_dict(*a, **kw).items():
                full = f"submod_{k}.{n}"
                out[self.local_name_to_original_name.get(full, full)] = v
        return out

    def load_local_state_dict(self, state_dict: Dict[str, torch.Tensor], strict: bool = True):
        missing, used = [], set()
        for k in self.local_stage_ids:
            sm = self.stage_modules[k]
            local_sd = {}
            for n in sm.state_dict().keys():
                orig = self.local_name_to_original_name.get(f"submod_{k}.{n}", f"submod_{k}.{n}")
                if orig in state_dict:
                    local_sd[n] = state_dict[orig]
                    used.add(orig)
                else:
                    missing.append(orig)
            sm.load_state_dict(local_sd, strict=False)
        if strict and missing:
            raise RuntimeError(f"missing keys for local pipeline stages: {missing}")
        return missing

    def parameters(self, recurse: bool = True):
        return self.local_parameters(recurse=recurse) if self.partitioned else super().parameters(recurse)

    def named_parameters(self, *a, **kw):
        return self.local_named_parameters(*a, **kw) if self.partitioned else super().named_parameters(*a, **kw)

    def state_dict(self, *a, **kw):
        return self.local_state_dict(*a, **kw) if self.partitioned else super().state_dict(*a, **kw)

    def get_model_layer(self, chunk: int = 0) -> nn.Module:
        return self.stage_modules[self._stage_id(chunk)]

    # -- reference method names (``pipeline/model.py``) ----------------------------------------------------------------------
    def get_model_layers(self, module: Optional[nn.Module] = None, layer_cls: Optional[type] = None) -> List[nn.Module]:
        """Transformer layers (instances of ``transformer_layer_cls``) inside ``module`` -- by default the local pipeline
        stages, in execution order."""
        cls = layer_cls or self.transformer_layer_cls
        roots = [module] if module is not None else (list(self.local_stage_modules) if self.partitioned
                                                     else [self.original_torch_module])
        return [m for r in roots for m in r.modules() if cls is not None and isinstance(m, cls)]

    def register_shared_weights(self) -> None:
        """Tag tied weights that several stages use and build the group that keeps them in sync."""
        self._mark_shared_weights()

    def create_schedule(self) -> None:
        self._create_schedules()

    def get_current_stage(self, model_chunk_id: int = 0) -> int:
        """Global stage index executed by this rank for ``model_chunk_id`` (interleaved schedule: several)."""
        return self._stage_id(model_chunk_id)

    def is_last_pp_rank_last_model_chunk(self, model_chunk_id: int = 0) -> bool:
        return self._stage_id(model_chunk_id) == self.num_stages - 1

    def get_batch_iterator(self, batch_kwargs: Dict[str, Any]):
        """Micro-batches of a batch (dim 0 split in ``num_microbatches`` equal parts)."""
        return iter(self._split_microbatches(batch_kwargs))

    def clear_minibatch_state(self) -> None:
        """Drop everything that belongs to the mini-batch that just ran (received activations, stage outputs kept for
        the backward, gradients in flight, losses)."""
        for name in ("microbatches", "losses", "eval_outputs"):
            setattr(self, name, [])
        for name in ("recv_values", "recv_tensors", "stage_outputs", "sent_tensors", "recv_grads"):
            setattr(self, name, {})

    def maybe_deallocate_output_tensor(self, tensor: torch.Tensor) -> None:
        """After an activation has been sent to the next stage only its autograd node is needed: release the storage
        (reference ``deallocate_pipeline_outputs``).  Views of other tensors are left alone."""
        if self.deallocate_pipeline_outputs and isinstance(tensor, torch.Tensor) and tensor._base is None:
            tensor.data = torch.empty((1,), device=tensor.device, dtype=tensor.dtype)

    @staticmethod
    def custom_backward(outputs: List[torch.Tensor], grad_outputs: List[Optional[torch.Tensor]]) -> None:
        """Backward through stage outputs whose storage may have been released (the shapes of output and gradient no
        longer match, which ``torch.autograd.backward`` checks): call the engine directly."""
        pairs = [(o, g) for o, g in zip(outputs, grad_outputs) if isinstance(o, torch.Tensor) and o.requires_grad]
        if not pairs:
            return
        outs, gs = zip(*pairs)
        try:
            torch.autograd.backward(outs, gs)
        except RuntimeError:
            from torch.autograd import Variable

            Variable._execution_engine.run_backward(tuple(outs), tuple(gs), False, False, tuple(), True, True)

    def translate_origin_state_dict_to_local_state_dict(self, state_dict: Dict[str, Any]) -> Dict[str, Any]:
        """Original (un-partitioned) key names -> ``submod_k.*`` names of the local stages; keys of other stages are
        dropped."""
        local_prefixes = tuple(f"submod_{k}." for k in self.local_stage_ids)
        out = OrderedDict()
        for k, v in state_dict.items():
            loc = self.original_name_to_local_name.get(k)
            if loc is not None and loc.startswith(local_prefixes):
                out[loc] = v
        return out

    def translate_local_state_dict_to_origin_state_dict(self, state_dict: Dict[str, Any]) -> Dict[str, Any]:
        return OrderedDict((self.local_name_to_original_name.get(k, k), v) for k, v in state_dict.items())

    def construct_state_dict_per_model_chunk(self, state_dict: Dict[str, Any], strict: bool = True) -> List[Dict[str, Any]]:
        """One state dict per local model chunk (keys relative to the stage module), from original-name keys."""
        chunks = []
        for k in self.local_stage_ids:
            part = {}
            for n in self.stage_modules[k].state_dict().keys():
                orig = self.local_name_to_original_name.get(f"submod_{k}.{n}", f"submod_{k}.{n}")
                if orig in state_dict:
                    part[n] = state_dict[orig]
                elif strict:
                    raise RuntimeError(f"missing key {orig} for pipeline stage {k}")
            chunks.append(part)
        return chunks

    def move_model_to_device(self, device=None) -> None:
        """Move the LOCAL stages to this rank's device (parameters of other stages were dropped at partition time)."""
        device = device or (torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu"))
        if any(p.device.type == "meta" for p in self.local_stage_modules.parameters()):
            self.maybe_materialize_local_module(device)
        self.local_stage_modules.to(device)

    def maybe_materialize_local_module(self, device=None) -> None:
        """Materialise local stages that were built on the meta device (delayed initialisation)."""
        from ..utils.model_utils import maybe_materalize_model

        for m in self.local_stage_modules:
            maybe_materalize_model(m, device)

    def perform_delayed_tracing_and_partition(self, args=None, kwargs=None) -> None:
        """Trace + partition on first use when construction was asked to wait for real inputs (``trace_file_path`` /
        delayed tracing of the reference; hard-disabled there, kept for API parity)."""
        if self.traced_model is None:
            self.trace(args=args, kwargs=kwargs, input_names=self.input_names)
        if not self.partitioned:
            self.partition()

    @property
    def config(self):
        return getattr(self.original_torch_module, "config", None)
