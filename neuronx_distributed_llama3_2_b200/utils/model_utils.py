"""Model construction helpers (reference ``src/neuronx_distributed/utils/model_utils.py:4-348``):
meta-device init that preserves tied weights and TP attributes, staggered materialisation,
shared-parameter analysis."""
from __future__ import annotations

import contextlib
from typing import Any, Callable, Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from .logger import get_logger

logger = get_logger()

_PARAM_TAGS = ("tensor_model_parallel", "partition_dim", "partition_stride", "num_partitions",
               "sequence_parallel_enabled", "expert_model_parallel", "shared")


# activation registry used by the MoE modules (reference ``utils/model_utils.py:4-11``)
ACT2FN = {
    "gelu": torch.nn.functional.gelu,
    "leaky_relu": torch.nn.functional.leaky_relu,
    "relu": torch.nn.functional.relu,
    "sigmoid": torch.sigmoid,
    "silu": torch.nn.functional.silu,
    "tanh": torch.tanh,
}


def analyze_shared_parameters(module: nn.Module) -> List[List[Tuple[nn.Module, str]]]:
    """Groups of (module, attr-name) that alias the same Parameter (tied weights)."""
    seen: Dict[int, List[Tuple[nn.Module, str]]] = {}
    for mod in module.modules():
        for name, p in mod._parameters.items():
            if p is not None:
                seen.setdefault(id(p), []).append((mod, name))
    return [v for v in seen.values() if len(v) > 1]


def retie_shared_weights(shared_groups: List[List[Tuple[nn.Module, str]]]) -> None:
    for group in shared_groups:
        src_mod, src_name = group[0]
        src = src_mod._parameters[src_name]
        for mod, name in group[1:]:
            mod._parameters[name] = src


def recursive_filter(item: Any, predicate: Callable[[Any], bool]) -> Any:
    """Keep only leaves for which ``predicate`` holds, preserving container structure."""
    if isinstance(item, dict):
        return {k: recursive_filter(v, predicate) for k, v in item.items()
                if isinstance(v, (dict, list, tuple)) or predicate(v)}
    if isinstance(item, (list, tuple)):
        return type(item)(recursive_filter(v, predicate) for v in item
                          if isinstance(v, (dict, list, tuple)) or predicate(v))
    return item if predicate(item) else None


@contextlib.contextmanager
def init_on_device(device: torch.device, include_buffers: bool = False, force_custom_init_on_device: bool = False):
    """Construct modules with parameters on ``device`` (typically ``meta``) while preserving
    parameter attributes and aliasing (reference :244-317).  Parallel layers see the target
    device through ``torch.device`` context and skip real initialisation on ``meta``."""
    old_register = nn.Module.register_parameter

    def register_parameter(module, name, param):
        old_register(module, name, param)
        if param is not None and param.device != torch.device(device):
            cur = module._parameters[name]
            new = nn.Parameter(cur.to(device), requires_grad=cur.requires_grad)
            for tag in _PARAM_TAGS:
                if hasattr(cur, tag):
                    setattr(new, tag, getattr(cur, tag))
            module._parameters[name] = new

    nn.Module.register_parameter = register_parameter
    try:
        with torch.device(device):
            yield
    finally:
        nn.Module.register_parameter = old_register


def _materialize_param(p: nn.Parameter, device, init_fn=None) -> nn.Parameter:
    new = nn.Parameter(torch.empty_like(p, device=device), requires_grad=p.requires_grad)
    for tag in _PARAM_TAGS:
        if hasattr(p, tag):
            setattr(new, tag, getattr(p, tag))
    return new


def reinit_model(model: nn.Module, device, param_init_fn: Optional[Callable[[nn.Module, Any], None]] = None) -> None:
    """Materialise a meta-device model on ``device`` and run each module's init
    (``param_init_fn(module, device)``, else ``module.init_weight_cpu_or_device`` / ``reset_parameters``),
    keeping tags and tied weights (reference :137-191)."""
    shared = analyze_shared_parameters(model)
    for mod in model.modules():
        for name, p in list(mod._parameters.items()):
            if p is not None and p.device.type == "meta":
                mod._parameters[name] = _materialize_param(p, device)
        for name, b in list(mod._buffers.items()):
            if b is not None and b.device.type == "meta":
                mod._buffers[name] = torch.empty_like(b, device=device)
    retie_shared_weights(shared)
    for mod in model.modules():
        (param_init_fn or default_param_init_fn)(mod, device)


def default_param_init_fn(mod: nn.Module, device=None) -> None:
    """Initialise ONE module's own (already materialised) parameters: the ``param_init_fn`` to put into
    ``model_init_config`` with ``meta_device_init`` when the model has no special needs (the reference's examples carry
    their own ``init_weights(module, device)``, e.g. ``run_llama_nxd.py``).  TP-aware inits draw the full-size master
    weight and keep this rank's slice, norms get ones / zeros, anything else falls back to ``reset_parameters``."""
    own = [p for p in mod.parameters(recurse=False)]
    if not own or any(p.device.type == "meta" for p in own) or all(p.numel() == 0 for p in own):
        return          # nothing here, not materialised yet, or released (another pipeline stage owns this module)
    name = type(mod).__name__

    def strip(w):      # the TP-aware initialisers tag the tensor themselves and refuse one that is tagged already
        for tag in ("tensor_model_parallel", "partition_dim", "partition_stride", "num_partitions"):
            if hasattr(w, tag):
                delattr(w, tag)

    if hasattr(mod, "init_weight_cpu_or_device"):
        strip(mod.weight)
        mod.init_weight_cpu_or_device()
    elif hasattr(mod, "initialize_weight_biases"):            # GQAQKVColumnParallelLinear
        mod.initialize_weight_biases()
    elif hasattr(mod, "_init_weight") and hasattr(mod, "weight_partition_dim"):
        # re-run the TP-aware init on the now-real tensor; tags are already set
        w = mod.weight
        strip(w)
        mod._init_weight(w, mod.weight_partition_dim)
        if getattr(mod, "bias", None) is not None:
            with torch.no_grad():
                mod.bias.zero_()
    elif name.endswith("Norm") and not hasattr(mod, "reset_parameters"):
        with torch.no_grad():
            if getattr(mod, "weight", None) is not None:
                mod.weight.fill_(1.0)
            if getattr(mod, "bias", None) is not None:
                mod.bias.zero_()
    elif hasattr(mod, "init_weight_cpu") and name.endswith("Router"):
        mod.init_weight_cpu()
    elif hasattr(mod, "reset_parameters"):
        mod.reset_parameters()


def move_model_to_device(model: nn.Module, device) -> nn.Module:
    """``model.to(device)`` that keeps TP/EP/SP/shared tags and tied weights (reference :193-224)."""
    shared = analyze_shared_parameters(model)
    tags: Dict[int, Dict[str, Any]] = {}
    for p in model.parameters():
        tags[id(p)] = {t: getattr(p, t) for t in _PARAM_TAGS if hasattr(p, t)}
    names = {n: id(p) for n, p in model.named_parameters(remove_duplicate=False)}
    model.to(device)
    for n, p in model.named_parameters(remove_duplicate=False):
        for t, v in tags.get(names.get(n, -1), {}).items():
            if not hasattr(p, t):
                setattr(p, t, v)
    retie_shared_weights(shared)
    return model


def get_model_sequential(model_fn: Callable[[], nn.Module], device, sequential_move_factor: int = 11,
                         meta_device_init: bool = False, param_init_fn=None) -> nn.Module:
    """Build / materialise the model ``sequential_move_factor`` local ranks at a time to bound peak
    host memory (reference :320-336).  With 180 GB HBM and direct-to-device init this is rarely the
    bottleneck on B200, so the stagger only applies to CPU-initialised models."""
    import torch.distributed as dist

    from ..parallel.utils import get_local_world_size

    local_world = get_local_world_size()
    rank = dist.get_rank() if dist.is_initialized() else 0
    local_rank = rank % max(local_world, 1)
    model = None
    rounds = max(1, (local_world + sequential_move_factor - 1) // max(sequential_move_factor, 1))
    needs_stagger = dist.is_initialized() and dist.get_world_size() > 1 and not meta_device_init
    for r in range(rounds if needs_stagger else 1):
        if (not needs_stagger) or local_rank // sequential_move_factor == r:
            model = model_fn()
            if meta_device_init:
                reinit_model(model, device, param_init_fn)
            else:
                move_model_to_device(model, device)
        if needs_stagger:
            dist.barrier()
    return model


def check_delay_tracing(nxd_config: dict) -> bool:
    # Delayed PP tracing is hard-disabled in the reference as well (model_utils.py:339-348).
    return False


def get_delay_tracing(arg) -> bool:
    return False


# ---- environment probes and attribute-preserving contexts (reference ``utils/model_utils.py:91-241``) -------------------
def is_hf_transformers_available() -> bool:
    try:
        import transformers  # noqa: F401
        return True
    except Exception:  # noqa: BLE001
        return False


def is_hf_accelerate_available() -> bool:
    try:
        import accelerate  # noqa: F401
        return True
    except Exception:  # noqa: BLE001
        return False


def is_hf_pretrained_model(model) -> bool:
    if not is_hf_transformers_available():
        return False
    from transformers import PreTrainedModel

    return isinstance(model, PreTrainedModel)


def is_nxdt_available() -> bool:
    """NeuronxDistributedTraining (Megatron-style modules) is a Neuron-only package."""
    return False


def is_nxdt_pretrained_model(model) -> bool:
    return False


@contextlib.contextmanager
def preserve_shared_weights(model: nn.Module, ignore_hf: bool = False):
    """Re-tie shared parameters (e.g. embedding / lm_head) after the body replaced them -- moving a model between
    devices or materialising it creates new Parameter objects and silently unties them."""
    hf = is_hf_pretrained_model(model) and not ignore_hf
    groups = None if hf else analyze_shared_parameters(model)
    try:
        yield
    finally:
        if hf:
            model.tie_weights()
        elif groups:
            retie_shared_weights(groups)


_PARALLEL_ATTRS = ("tensor_model_parallel", "partition_dim", "partition_stride", "num_partitions", "expert_model_parallel",
                   "sequence_parallel_enabled", "shared")


@contextlib.contextmanager
def preserve_parallel_attributes(model: nn.Module):
    """Parameters recreated inside the body (materialisation, ``.to()`` of meta tensors) lose the python attributes the
    framework hangs on them; record them by name and put them back."""
    saved = {n: {a: getattr(p, a) for a in _PARALLEL_ATTRS if hasattr(p, a)} for n, p in model.named_parameters()}
    try:
        yield
    finally:
        for n, p in model.named_parameters():
            for a, v in saved.get(n, {}).items():
                setattr(p, a, v)


def has_fake_tensors(model: nn.Module, ignored_params=()) -> bool:
    """Meta-device parameters are this framework's deferred-initialisation "fake tensors" (the reference uses
    torchdistx fakes)."""
    ignored = {id(p) for p in ignored_params}
    return any(p.device.type == "meta" and id(p) not in ignored for p in model.parameters())


def maybe_materalize_model(model: nn.Module, device=None) -> None:
    """Materialise a model built on the meta device (name kept from the reference, typo included)."""
    if not has_fake_tensors(model):
        return
    device = device or (torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu"))
    with preserve_parallel_attributes(model), preserve_shared_weights(model, ignore_hf=True):
        reinit_model(model, device)
