"""Ahead-of-time inference builder: shape-routed, bucketed, CUDA-graph-captured execution.

Reference: ``src/neuronx_distributed/trace/model_builder.py:37-586`` and ``trace/spmd.py:9-187``: ``ModelBuilder.add()``
registers a model instance with example inputs per bucket, ``trace()`` spawns ``tp`` processes, emits one HLO per
(key, bucket), compiles NEFFs (propagating the priority model's weight layout through HLO surgery,
``hlo_utils.py``), and assembles an ``NxDModel`` that routes a call by its input-shape string to the right bucket;
weights are shared across graphs, state (KV cache) is aliased; ``shard_checkpoint`` writes
``tp{rank}_sharded_checkpoint.safetensors``.

On B200 the same surface maps to:
  * one process per GPU (the caller's torchrun), no spawn; "compile" = capture a CUDA graph of the module's forward for
    every (key, bucket) with static input buffers -- NCCL collectives and our ctypes-launched kernels are all
    stream-ordered and capturable; on CPU (tests) the graph is the eager callable;
  * weight-layout optimisation / HLO surgery has no counterpart: the tcgen05 kernels read row-major weights through
    UMMA descriptor major bits, so there is nothing to transform;
  * state aliasing is real aliasing: the KV cache tensors live in the module and are updated in place.
"""
from __future__ import annotations

import os
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple, Union

import torch

from ..parallel import parallel_state as ps
from ..utils.logger import get_logger
from .checkpoint_sharding import _mock_parallel_state, get_sharded_checkpoint

logger = get_logger()


class BaseModelInstance:
    """How to build the module for one sub-model (reference ``model_builder.py:37-47``)."""

    def __init__(self, module_cls: Callable[[], torch.nn.Module], input_output_aliases: Optional[Dict] = None,
                 forward_method: Optional[str] = None):
        self.module_cls = module_cls
        self.module: Optional[torch.nn.Module] = None
        self.input_output_aliases = input_output_aliases or {}
        # entry point of this sub-model when it is not the module's ``forward`` (e.g. ``forward_speculation``): several
        # sub-models of one module then differ by what they run, not only by their input shapes
        self.forward_method = forward_method

    def entry_point(self) -> Callable:
        return getattr(self.module, self.forward_method) if self.forward_method else self.module

    def load_module(self) -> None:
        self.module = self.module_cls()

    def get(self, bucket_rank: int, **kwargs) -> Tuple[torch.nn.Module, Dict]:
        return self.module, self.input_output_aliases


def _shape_key(args: Sequence[torch.Tensor]) -> str:
    return "|".join(f"{tuple(a.shape)}:{str(a.dtype).replace('torch.', '')}" for a in args)


class _Bucket:
    """One (key, bucket) executable: static inputs + (optionally) a captured CUDA graph."""

    def __init__(self, fn: Callable, example_inputs: Tuple[torch.Tensor, ...], use_graph: bool):
        self.fn = fn
        self.static_in = [t.clone() for t in example_inputs]
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.static_out = None
        self.use_graph = use_graph and all(t.is_cuda for t in example_inputs)

    @property
    def first_input_shape(self) -> Optional[Tuple[int, ...]]:
        return tuple(self.static_in[0].shape) if self.static_in else None

    def capture(self) -> None:
        if not self.use_graph:
            return
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(2):   # warm-up outside capture (allocator, lazy inits, symmetric-memory setup)
                self.fn(*self.static_in)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.static_out = self.fn(*self.static_in)

    def __call__(self, *args: torch.Tensor):
        if self.graph is None:
            return self.fn(*args)
        for dst, src in zip(self.static_in, args):
            dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.static_out


class NxDModel(torch.nn.Module):
    """Routes a call by input shapes to a compiled bucket (reference ``spmd.py:82-175``)."""

    def __init__(self, buckets: Dict[str, Dict[str, _Bucket]], modules: Dict[str, torch.nn.Module]):
        super().__init__()
        self._buckets = buckets
        self.models = torch.nn.ModuleDict(modules)
        self._route: Dict[str, Tuple[str, _Bucket]] = {}
        self._graphs: List[Tuple[str, str, _Bucket]] = []
        for key, bs in buckets.items():
            for shape_key, b in bs.items():
                self._route[shape_key] = (key, b)
                self._graphs.append((key, shape_key, b))
        self.weights_initialized = True
        self._native = self._build_native_router()

    def _build_native_router(self):
        """(batch, seq) of the first input -> bucket in the C++ runtime (``csrc/runtime/nxd_runtime.cpp``); the Python
        dict stays as the general path (arbitrary argument shapes, no compiler on the box)."""
        try:
            from ..runtime import NativeRouter, available

            if not available():
                return None
            r = NativeRouter()
            for i, (key, _, b) in enumerate(self._graphs):
                shp = b.first_input_shape
                if shp is None or len(shp) != 2:
                    return None
                r.add(key, int(shp[0]), int(shp[1]), i)
            return r
        except Exception:  # noqa: BLE001 - e.g. two models with the same first-input shape: keep the dict router
            return None

    def router(self, args: Sequence[torch.Tensor]) -> Tuple[str, _Bucket]:
        if self._native is not None and args and args[0].dim() == 2:
            g = self._native.route_exact(int(args[0].shape[0]), int(args[0].shape[1]))
            if g >= 0:
                key, shape_key, b = self._graphs[g]
                if shape_key == _shape_key(args):
                    return key, b
        k = _shape_key(args)
        if k not in self._route:
            raise RuntimeError(f"no compiled bucket for input shapes {k}; have {list(self._route)}")
        return self._route[k]

    def route_padded(self, key: str, batch: int, seq: int) -> Optional[Tuple[int, int]]:
        """Smallest compiled bucket ``(batch', seq')`` of sub-model ``key`` that covers a ``[batch, seq]`` request
        (the runner pads prompts up to it), or ``None``."""
        if self._native is not None:
            g, b, s = self._native.route_padded(key, batch, seq)
            return (b, s) if g >= 0 else None
        best = None
        for k, _, bk in self._graphs:
            shp = bk.first_input_shape
            if k != key or shp is None or len(shp) != 2 or shp[0] < batch or shp[1] < seq:
                continue
            if best is None or shp[0] * shp[1] < best[0] * best[1]:
                best = (int(shp[0]), int(shp[1]))
        return best

    def initialize(self, state_dict: Dict[str, torch.Tensor], strict: bool = False) -> None:
        """Load (already sharded) weights; they are shared by every graph of a key."""
        for m in self.models.values():
            m.load_state_dict(state_dict, strict=strict)

    def load_sharded_checkpoint(self, serialize_path: str, rank: Optional[int] = None, strict: bool = False) -> None:
        """Load ``tp{rank}_sharded_checkpoint.safetensors`` written by ``ModelBuilder.shard_checkpoint`` (reference
        ``model_builder.py:302-321``).  The file is read by the C++ runtime (mmap + a pool of copy threads into pinned
        memory) and moved to the device with non-blocking copies; falls back to the safetensors package."""
        rank = ps.get_tensor_model_parallel_rank() if rank is None else rank
        path = os.path.join(serialize_path, f"tp{rank}_sharded_checkpoint.safetensors")
        dev = next(iter(self.models.values())).parameters().__next__().device
        try:
            from ..runtime import available, load_safetensors

            if not available():
                raise RuntimeError("no native runtime")
            sd = load_safetensors(path, device=dev)
        except (RuntimeError, OSError, KeyError):
            from safetensors.torch import load_file

            sd = load_file(path, device=str(dev))
        self.initialize(sd, strict=strict)

    def forward(self, *args: torch.Tensor):
        _, bucket = self.router(args)
        return bucket(*args)

    def initialize_spmd_models(self, states=None, weights: Optional[Dict[str, torch.Tensor]] = None) -> None:
        """Hand (already sharded) weights to every sub-model; ``states`` (KV caches) are owned by the modules here."""
        if weights is not None:
            self.initialize(weights)

    def initialize_with_saved_weights(self, serialize_path: Optional[str] = None, rank: Optional[int] = None) -> None:
        """Load this rank's ``tp{rank}_sharded_checkpoint.safetensors`` written by ``ModelBuilder.shard_checkpoint``
        (the reference re-uses weights kept inside the saved TorchScript artefact)."""
        path = serialize_path or getattr(self, "serialize_path", None)
        if path is None:
            raise ValueError("initialize_with_saved_weights needs the directory of the sharded checkpoint")
        self.load_sharded_checkpoint(path, rank=rank)

    def mock_initialization(self, mock: bool = True) -> None:
        """Mark the weights as initialised without loading anything (random weights: shape / latency runs)."""
        self.weights_initialized = bool(mock)

    def release_graphs(self) -> None:
        """Destroy the captured CUDA graphs (buckets fall back to eager execution).  Call this -- or drop the model --
        BEFORE ``torch.distributed.destroy_process_group()``: NCCL keeps a communicator alive while a CUDA graph that
        captured one of its collectives exists, and the destroy call then waits forever."""
        for _, _, b in self._graphs:
            b.graph = None
            b.static_out = None

    def keys(self) -> List[str]:
        return list(self._buckets)


class ModelBuilder:
    def __init__(self, router=None, tp_degree: int = 1, checkpoint_loader: Optional[Callable[[], Dict[str, Any]]] = None,
                 compiler_workdir: Optional[str] = None, debug: bool = False, use_cuda_graphs: bool = True,
                 master_proc_env_vars: Optional[Dict[str, str]] = None):
        # ``master_proc_env_vars``: environment of the reference's compile sub-processes; every rank is its own
        # process here, so the variables are simply exported
        for k, v in (master_proc_env_vars or {}).items():
            os.environ[str(k)] = str(v)
        self.router = router
        self.tp_degree = tp_degree
        self.checkpoint_loader = checkpoint_loader
        self.compiler_workdir = compiler_workdir
        self.debug = debug
        self.use_cuda_graphs = use_cuda_graphs
        self.model_collection: Dict[str, Dict[str, Any]] = {}

    def add(self, key: str, model_instance: BaseModelInstance, example_inputs: List[Tuple[torch.Tensor, ...]],
            compiler_args: Union[str, List[str], None] = None, bucket_config=None,
            priority_model_idx: Optional[int] = None) -> "ModelBuilder":
        """Register a sub-model; ``example_inputs`` holds one tuple of example tensors per bucket."""
        if key in self.model_collection:
            raise ValueError(f"model key {key!r} already added")
        self.model_collection[key] = {"instance": model_instance, "example_inputs": list(example_inputs),
                                      "compiler_args": compiler_args, "bucket_config": bucket_config,
                                      "priority_model_idx": priority_model_idx}
        return self

    def trace(self, tp_degree: Optional[int] = None, initialize_model_weights: bool = True) -> NxDModel:
        """Build every module on this rank's device and capture one executable per (key, bucket)."""
        tp = tp_degree or self.tp_degree
        have = ps.get_tensor_model_parallel_size()
        if have != tp:
            raise RuntimeError(f"ModelBuilder(tp_degree={tp}) but the process group has tp={have}: launch one process per "
                               "GPU (torchrun) and call initialize_model_parallel first")
        weights = self.checkpoint_loader() if (initialize_model_weights and self.checkpoint_loader is not None) else None
        rank = ps.get_tensor_model_parallel_rank()
        modules: Dict[str, torch.nn.Module] = {}
        # priority model first (kept for parity with the reference's compile order)
        order = sorted(self.model_collection, key=lambda k: self.model_collection[k]["priority_model_idx"] is None)
        for key in order:
            inst: BaseModelInstance = self.model_collection[key]["instance"]
            if inst.module is None:
                inst.load_module()
            if weights is not None:
                inst.module.load_state_dict(self.shard_weights(rank, self.model_collection[key], weights), strict=False)
            modules[key] = inst.module
        return self.build_nxd_model(modules, order)

    # -- the build steps under the reference's names (``model_builder.py:263-451``) ---------------------------------------
    def shard_weights(self, rank: int, model_container, checkpoint: Optional[Dict[str, Any]] = None) -> Dict[str, Any]:
        """This rank's slice of the un-sharded checkpoint for one registered sub-model (preshard hooks, partition dim /
        stride, fused-QKV aware)."""
        inst = model_container["instance"] if isinstance(model_container, dict) else model_container.model_instance
        checkpoint = checkpoint if checkpoint is not None else self.checkpoint_loader()
        module = inst.module
        if module is None:
            with torch.device("meta"):
                module = inst.module_cls()
        return get_sharded_checkpoint(checkpoint, module, rank, self.tp_degree)

    def build_state_initializer(self, modules: Optional[Dict[str, torch.nn.Module]] = None) -> "StateInitializer":
        """Shapes / dtypes of the per-rank state (the KV cache) of the registered models, as a ``StateInitializer``."""
        shapes, dtypes = {}, {}
        for key, entry in self.model_collection.items():
            m = (modules or {}).get(key) or entry["instance"].module
            cache = getattr(m, "kv_cache", None)
            if cache is not None and getattr(cache, "k", None):
                for i, (k, v) in enumerate(zip(cache.k, cache.v)):          # one (K, V) pair per layer
                    shapes[f"{key}.layers.{i}.kv_cache.k"], dtypes[f"{key}.layers.{i}.kv_cache.k"] = tuple(k.shape), k.dtype
                    shapes[f"{key}.layers.{i}.kv_cache.v"], dtypes[f"{key}.layers.{i}.kv_cache.v"] = tuple(v.shape), v.dtype
        return StateInitializer(shapes, dtypes)

    def build_flattener_map(self) -> Dict[str, Callable]:
        """Per sub-model: nested inputs -> flat tensor list (the reference builds torch_neuronx Flatteners)."""
        def flatten(inputs):
            out = []
            for x in (inputs if isinstance(inputs, (list, tuple)) else [inputs]):
                out.extend(flatten(x) if isinstance(x, (list, tuple)) else [x])
            return out
        return {key: flatten for key in self.model_collection}

    def build_packer(self) -> Callable:
        """Flat result list -> user structure; the sub-models here return one tensor (or a flat tuple) already."""
        return lambda result: result[0] if isinstance(result, (list, tuple)) and len(result) == 1 else result

    def build_nxd_model(self, modules: Dict[str, torch.nn.Module], order: Optional[List[str]] = None) -> NxDModel:
        """Capture one executable per (key, bucket) over the loaded modules and assemble the shape-routed ``NxDModel``."""
        buckets: Dict[str, Dict[str, _Bucket]] = {}
        for key in (order or list(self.model_collection)):
            entry = self.model_collection[key]
            inst: BaseModelInstance = entry["instance"]
            buckets[key] = {}
            for rank_idx, ex in enumerate(entry["example_inputs"]):
                mod, _aliases = inst.get(rank_idx)
                fn = getattr(mod, inst.forward_method) if getattr(inst, "forward_method", None) else mod
                b = _Bucket(fn, tuple(ex), self.use_cuda_graphs)
                b.capture()
                buckets[key][_shape_key(ex)] = b
        return NxDModel(buckets, modules)

    def shard_checkpoint(self, serialize_path: str) -> None:
        """Write ``tp{rank}_sharded_checkpoint.safetensors`` for every TP rank from the un-sharded checkpoint; runs in
        a single process with mocked parallel state and meta-device modules (reference ``:263-331``)."""
        from safetensors.torch import save_file

        assert self.checkpoint_loader is not None, "shard_checkpoint needs a checkpoint_loader"
        os.makedirs(serialize_path, exist_ok=True)
        full = self.checkpoint_loader()
        for rank in range(self.tp_degree):
            merged: Dict[str, torch.Tensor] = {}
            with _mock_parallel_state(self.tp_degree, rank):
                for key, entry in self.model_collection.items():
                    inst: BaseModelInstance = entry["instance"]
                    with torch.device("meta"):
                        module = inst.module_cls()
                    merged.update(get_sharded_checkpoint(full, module, rank, self.tp_degree))
            tensors = {k: v.contiguous() for k, v in merged.items() if isinstance(v, torch.Tensor)}
            # safetensors refuses aliased storage (tied lm_head / embedding): materialise copies
            seen = {}
            for k, v in list(tensors.items()):
                if v.data_ptr() in seen:
                    tensors[k] = v.clone()
                seen[v.data_ptr()] = k
            save_file(tensors, os.path.join(serialize_path, f"tp{rank}_sharded_checkpoint.safetensors"))


# ---- names of the reference's SPMD runtime classes (``trace/spmd.py``, ``trace/model_builder.py``) ---------------------
# SPMDBucketModel = one compiled bucket (NEFF there, captured CUDA graph here); NxDModelExecutor = the object that runs
# the routed bucket; StateInitializer allocated the aliased KV state per rank (``KVCache`` does that here).
NxDModelExecutor = NxDModel


class SPMDBucketModelScript(torch.nn.Module):
    """The compiled buckets of ONE model key, picked by index (reference ``spmd.py:9-30``; a bucket there is an
    ``SPMDModel`` running a NEFF on every rank, here a ``_Bucket`` replaying its CUDA graph on this rank's GPU)."""

    def __init__(self, compiled_models: Sequence[_Bucket]):
        super().__init__()
        self.models = list(compiled_models)

    def forward(self, inputs: List[torch.Tensor], bucket_idx_tensor: torch.Tensor):
        idx = int(bucket_idx_tensor)
        if not 0 <= idx < len(self.models):
            raise ValueError(f"bucket index {idx} out of range: {len(self.models)} bucket(s) compiled")
        return self.models[idx](*inputs)


class SPMDBucketModel(torch.nn.Module):
    """``bucket_kernel(inputs, *constant_args) -> (inputs padded to a bucket, bucket index)`` in front of the bucket
    executor (reference ``spmd.py:32-61``).  Weights are not registered here: all buckets close over the one module."""

    def __init__(self, bucket_kernel: Callable, bucket_kernel_constant_args: Sequence[Any],
                 bucket_model_executor: SPMDBucketModelScript):
        super().__init__()
        self.bucket_kernel = bucket_kernel
        self.bucket_kernel_constant_args = tuple(bucket_kernel_constant_args)
        self.bucket_model_executor = bucket_model_executor

    def forward(self, inputs: List[torch.Tensor]):
        padded, idx = self.bucket_kernel(inputs, *self.bucket_kernel_constant_args)
        return self.bucket_model_executor(padded, idx)


def default_bucket_kernel(inputs: List[torch.Tensor]):
    """Single-bucket "kernel" of the reference (``spmd.py``): inputs unchanged, bucket 0."""
    return inputs, torch.tensor(0)


class StateInitializer(torch.nn.Module):
    """Allocates the per-rank state tensors of a traced model from ``{name: (shape, dtype)}`` (reference
    ``spmd.py:63-80`` builds them from the metaneff); used for KV caches that live outside ``KVCache``."""

    def __init__(self, shapes: Dict[str, Any], dtypes: Dict[str, torch.dtype], local_ranks_size: int = 1,
                 tp_degree: Optional[int] = None):
        super().__init__()
        local_ranks_size = tp_degree if tp_degree is not None else local_ranks_size     # reference argument name
        self.shapes, self.dtypes, self.local_ranks_size = shapes, dtypes, local_ranks_size

    def forward(self, device=None) -> Dict[str, torch.Tensor]:
        return {k: torch.zeros(tuple(shape), dtype=self.dtypes[k], device=device) for k, shape in self.shapes.items()}


class ModelContainer:
    """One registered sub-model of ``ModelBuilder`` (reference ``model_builder.py``)."""

    def __init__(self, model_instance: "BaseModelInstance", example_inputs, compiler_args=None, bucket_config=None,
                 priority_model_idx=None):
        self.model_instance, self.example_inputs = model_instance, example_inputs
        self.compiler_args, self.bucket_config, self.priority_model_idx = compiler_args, bucket_config, priority_model_idx


class JITWrapper(torch.nn.Module):
    """Makes a function traceable as a module (reference ``model_builder.py``: ``torch.jit.trace`` needs an
    ``nn.Module``); here it only forwards."""

    def __init__(self, func: Callable, is_torchscript_model: bool = False):
        super().__init__()
        self.func, self.is_torchscript_model = func, is_torchscript_model

    def forward(self, *inputs):
        return self.func(*inputs)
