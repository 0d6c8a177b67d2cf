"""High-level trainer API: config dict, parallel model / optimizer construction.

Reference: ``src/neuronx_distributed/trainer/trainer.py:33-303``.
"""
from __future__ import annotations

from pprint import pformat
from typing import Any, Callable, Dict, Optional

import torch
import torch.distributed as dist

from ..parallel import parallel_state as ps
from ..parallel.pad import pad_model
from ..utils.activation_checkpoint import apply_activation_checkpointing
from ..utils.logger import get_logger
from ..utils.model_utils import get_model_sequential, init_on_device
from .model import NxDModel
from .optimizer import NxDOptimizer

logger = get_logger()


def _fill(cfg: Optional[dict], defaults: Dict[str, Any], name: str) -> dict:
    if cfg is None:
        return dict(defaults)
    assert isinstance(cfg, dict), f"{name} must be a dict."
    for k, v in defaults.items():
        if k not in cfg:
            if ps.is_global_rank_zero():
                logger.warning("%s is not set in %s, automatically set it to %s.", k, name, v)
            cfg[k] = v
    return cfg


def neuronx_distributed_config(tensor_parallel_size: int = 1, pipeline_parallel_size: int = 1,
                               expert_parallel_size: int = 1, pipeline_config: Optional[dict] = None,
                               optimizer_config: Optional[dict] = None, activation_checkpoint_config=None,
                               pad_model: bool = False, sequence_parallel: bool = False,
                               model_init_config: Optional[dict] = None, lora_config=None,
                               mixed_precision_config: Optional[dict] = None) -> Dict[str, Any]:
    """Build the plain-dict NxD config (keys and defaults as reference ``trainer.py:46-127``) and, when the
    default process group is up, create the model-parallel groups (``:129-134``)."""
    optimizer_config = _fill(optimizer_config, {"zero_one_enabled": False, "grad_clipping": True}, "optimizer_config")
    if optimizer_config["grad_clipping"] and "max_grad_norm" not in optimizer_config:
        optimizer_config["max_grad_norm"] = 1.0
    z1 = optimizer_config["zero_one_enabled"]
    mixed_precision_config = _fill(mixed_precision_config, {
        "use_master_weights": z1, "use_fp32_grad_acc": z1, "use_master_weights_in_ckpt": False,
    }, "mixed_precision_config")
    if model_init_config is not None and model_init_config.get("meta_device_init") \
            and "param_init_fn" not in model_init_config:
        raise ValueError("param_init_fn must be provided when meta_device_init is True")
    model_init_config = _fill(model_init_config, {
        "sequential_move_factor": 11, "meta_device_init": False, "param_init_fn": None,
    }, "model_init_config")
    config = {
        "tensor_parallel_size": tensor_parallel_size,
        "pipeline_parallel_size": pipeline_parallel_size,
        "expert_parallel_size": expert_parallel_size,
        "pipeline_config": pipeline_config,
        "optimizer_config": optimizer_config,
        "activation_checkpoint_config": activation_checkpoint_config,
        "pad_model": pad_model,
        "sequence_parallel": sequence_parallel,
        "model_init_config": model_init_config,
        "lora_config": lora_config,
        "mixed_precision_config": mixed_precision_config,
    }
    if dist.is_initialized() and not ps.model_parallel_is_initialized():
        ps.initialize_model_parallel(tensor_model_parallel_size=tensor_parallel_size,
                                     pipeline_model_parallel_size=pipeline_parallel_size,
                                     expert_model_parallel_size=expert_parallel_size)
    if dist.is_initialized() and ps.is_global_rank_zero():
        logger.info("NxD config: \n%s", pformat(config))
    return config


def _default_device() -> torch.device:
    if torch.cuda.is_available():
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def initialize_parallel_model(nxd_config: Dict[str, Any], model_fn: Callable[..., torch.nn.Module], *model_args,
                              device: Optional[torch.device] = None, **model_kwargs) -> NxDModel:
    """Phases as reference ``trainer.py:141-229``: build (optionally on ``meta``) -> pipeline wrap -> staggered
    materialise/move -> LoRA -> head padding -> ``NxDModel`` + activation checkpointing."""
    device = device or _default_device()
    init_cfg = nxd_config["model_init_config"]
    meta = bool(init_cfg.get("meta_device_init"))

    def build():
        if meta:
            with init_on_device(torch.device("meta")):
                return model_fn(*model_args, **model_kwargs)
        return model_fn(*model_args, **model_kwargs)

    def wrap(model):
        if nxd_config["pipeline_parallel_size"] > 1:
            from ..pipeline import NxDPPModel

            pc = dict(nxd_config["pipeline_config"] or {})
            model = NxDPPModel(model, **pc)
        return model

    model = get_model_sequential(lambda: wrap(build()), device, init_cfg.get("sequential_move_factor", 11), meta_device_init=meta,
                                 param_init_fn=init_cfg.get("param_init_fn"))
    if nxd_config.get("lora_config") is not None:
        from ..modules.lora import LoraModel

        model = LoraModel(model, nxd_config["lora_config"])
    if nxd_config.get("pad_model"):
        cfg = getattr(model, "config", None)
        n_heads = getattr(cfg, "num_attention_heads", None)
        if n_heads is not None:
            model = pad_model(model, nxd_config["tensor_parallel_size"], n_heads)
    nxd_model = NxDModel(model, nxd_config)
    ac = nxd_config.get("activation_checkpoint_config")
    if ac is not None:
        if ac == "full":
            from ..pipeline.model import NxDPPModel as _PP

            if isinstance(model, _PP):
                layer_cls = model.transformer_layer_cls
                apply_activation_checkpointing(nxd_model, check_fn=lambda m: isinstance(m, layer_cls))
            else:
                cfg = getattr(model, "config", None)
                if cfg is not None and hasattr(cfg, "activation_checkpointing"):
                    cfg.activation_checkpointing = True  # our model families checkpoint per decoder layer
                else:
                    raise ValueError('activation_checkpoint_config="full" needs a pipeline model or a model '
                                     "config with `activation_checkpointing`; pass the layer class instead")
        else:
            classes = ac if isinstance(ac, (tuple, list)) else (ac,)
            apply_activation_checkpointing(nxd_model, check_fn=lambda m: isinstance(m, tuple(classes)))
    return nxd_model


def initialize_optimizer_from_class(nxd_config, optimizer_class, parameters, model: Optional[NxDModel] = None,
                                    **defaults):
    oc = nxd_config["optimizer_config"]
    mp = nxd_config["mixed_precision_config"]
    if oc["zero_one_enabled"]:
        from ..optimizer import NeuronEPZero1Optimizer, NeuronZero1Optimizer

        zero_cls = NeuronEPZero1Optimizer if nxd_config["expert_parallel_size"] > 1 else NeuronZero1Optimizer
        return zero_cls(
            parameters, optimizer_class,
            optimizer_dtype=torch.float32 if mp["use_master_weights"] else None,
            pin_layout=False,
            sharding_groups=ps.get_data_parallel_group(as_list=True) if ps.model_parallel_is_initialized() else None,
            grad_norm_groups=ps.get_tensor_model_parallel_group(as_list=True)
            if ps.model_parallel_is_initialized() else None,
            grad_clipping=oc["grad_clipping"], max_norm=oc.get("max_grad_norm"),
            coalesce_cc=True, use_grad_acc_hook=mp["use_fp32_grad_acc"], higher_cc_precision=mp["use_fp32_grad_acc"],
            save_master_weights=mp["use_master_weights_in_ckpt"], lazy_init=False, **defaults)
    return optimizer_class(parameters, **defaults)


def initialize_parallel_optimizer(nxd_config: Dict[str, Any], optimizer_class, parameters, **defaults) -> NxDOptimizer:
    """Reference ``trainer.py:232-303``: ZeRO-1 wrapper when ``zero_one_enabled`` else the plain optimizer, inside
    ``NxDOptimizer`` (SP-grad all-reduce, DP bucket all-reduce, clipping)."""
    parameters = list(parameters)
    optimizer = initialize_optimizer_from_class(nxd_config, optimizer_class, parameters, **defaults)
    return NxDOptimizer(optimizer, nxd_config)


def filter_to_local_parameter_group(optimizer, model) -> None:
    """Keep only the parameters that live on THIS pipeline stage in the optimizer's param groups (reference
    ``:293-303``: with delayed materialisation the optimizer is built from meta parameters and re-pointed at the local,
    materialised ones)."""
    pmap = getattr(model, "meta_device_parameter_map", None)
    local = {id(p) for p in model.parameters()}
    for group in optimizer.param_groups:
        kept = []
        for p in group["params"]:
            q = pmap.get(p, p) if pmap is not None else p
            if q.device.type != "meta" and (pmap is not None or id(q) in local):
                kept.append(q)
        group["params"] = kept
