"""``LoraModel``: inject adapters into a model, freeze the rest, adapter-only checkpoints (reference
``src/neuronx_distributed/modules/lora/model.py:36-694``)."""
from __future__ import annotations

import os
import re
from typing import Tuple, Any, Dict, List, Optional

import torch
import torch.nn as nn

from ...parallel.layers import ColumnParallelLinear, RowParallelLinear
from ...utils.logger import get_logger
from ..qkv_linear import GQAQKVColumnParallelLinear
from .config import LoraConfig
from .layer import LoraConv2d, LoraEmbedding, LoraLayer, LoraLinear
from .tp_layer import LoraGQAQKVParallelLinear, LoraParallelLinear

logger = get_logger()

# default target modules per architecture (PEFT's table as used by the reference, ``model.py:36-68``) + our Llama
MODELS_TO_LORA_TARGET_MODULES_MAPPING = {
    "t5": ["q", "v"], "mt5": ["q", "v"], "bart": ["q_proj", "v_proj"], "gpt2": ["c_attn"],
    "bloom": ["query_key_value"], "opt": ["q_proj", "v_proj"], "gptj": ["q_proj", "v_proj"],
    "gpt_neox": ["query_key_value"], "gpt_neo": ["q_proj", "v_proj"], "bert": ["query", "value"],
    "roberta": ["query", "value"], "llama": ["qkv_proj", "q_proj", "v_proj"], "mistral": ["q_proj", "v_proj"],
    "mixtral": ["qkv_proj", "q_proj", "v_proj"], "falcon": ["query_key_value"], "codegen": ["qkv_proj"],
    "phi": ["q_proj", "v_proj", "fc1", "fc2"], "gemma": ["q_proj", "v_proj"],
}

LORA_PREFIX = "lora_"


def _wrap(module: nn.Module, cfg: LoraConfig) -> Optional[nn.Module]:
    if isinstance(module, GQAQKVColumnParallelLinear):
        return LoraGQAQKVParallelLinear(module, cfg)
    if isinstance(module, (ColumnParallelLinear, RowParallelLinear)):
        return LoraParallelLinear(module, cfg)
    if isinstance(module, nn.Linear):
        return LoraLinear(module, cfg)
    if isinstance(module, nn.Embedding):
        return LoraEmbedding(module, cfg)
    if isinstance(module, nn.Conv2d):
        return LoraConv2d(module, cfg)
    return None


class LoraModel(nn.Module):
    def __init__(self, module: nn.Module, config: Optional[LoraConfig] = None):
        super().__init__()
        self.module = module
        self.lora_config = config
        self.is_lora_enabled = False
        self.is_lora_merged = False
        self.modules_to_save: List[str] = list(config.modules_to_save or []) if config else []
        if config is not None and config.enable_lora:
            if config.load_lora_from_ckpt:
                self.load_lora(config)
            else:
                self.inject_adapter()

    # -- injection -----------------------------------------------------------------------------------------
    def _targets(self) -> Any:
        t = self.lora_config.target_modules
        if t is None:
            mt = getattr(getattr(self.get_base_model(), "config", None), "model_type", None)
            if mt not in MODELS_TO_LORA_TARGET_MODULES_MAPPING:
                raise ValueError("Please specify `target_modules` in `LoraConfig`")
            t = MODELS_TO_LORA_TARGET_MODULES_MAPPING[mt]
            self.lora_config.target_modules = t
        return t

    def _is_target(self, key: str) -> bool:
        t = self._targets()
        if isinstance(t, str):
            return re.fullmatch(t, key) is not None
        return any(key == s or key.endswith("." + s) for s in t)

    def get_base_model(self) -> nn.Module:
        m = self.module
        while hasattr(m, "module") and isinstance(getattr(m, "module"), nn.Module) and not hasattr(m, "config"):
            m = m.module
        return getattr(m, "original_torch_module", m)

    def inject_adapter(self) -> None:
        if self.is_lora_enabled:
            return
        root = self.module
        matched = 0
        for key, mod in list(root.named_modules()):
            if not key or isinstance(mod, LoraLayer) or not self._is_target(key):
                continue
            new = _wrap(mod, self.lora_config)
            if new is None:
                continue
            parent_name, _, child = key.rpartition(".")
            parent = root.get_submodule(parent_name) if parent_name else root
            setattr(parent, child, new)
            matched += 1
            if self.lora_config.lora_verbose:
                logger.info("LoRA: wrapped %s (%s)", key, type(mod).__name__)
        if matched == 0:
            raise ValueError(f"Target modules {self.lora_config.target_modules} not found in the base model. "
                             "Please check the target modules and try again.")
        self._mark_only_adapters_as_trainable()
        self.is_lora_enabled = True

    def _mark_only_adapters_as_trainable(self) -> None:
        bias_mode = self.lora_config.bias
        for n, p in self.module.named_parameters():
            keep = LORA_PREFIX in n or any(n.startswith(m + ".") or ("." + m + ".") in n for m in self.modules_to_save)
            p.requires_grad_(keep)
        if bias_mode == "all":
            for n, p in self.module.named_parameters():
                if "bias" in n:
                    p.requires_grad_(True)
        elif bias_mode == "lora_only":
            for m in self.module.modules():
                if isinstance(m, LoraLayer):
                    b = getattr(m.get_base_layer(), "bias", None)
                    if b is not None:
                        b.requires_grad_(True)

    # -- merge ---------------------------------------------------------------------------------------------
    def merge_lora(self) -> None:
        for m in self.module.modules():
            if isinstance(m, LoraLayer):
                m.merge()
        self.is_lora_merged = True

    def unmerge_lora(self) -> None:
        for m in self.module.modules():
            if isinstance(m, LoraLayer):
                m.unmerge()
        self.is_lora_merged = False

    # -- state dicts -----------------------------------------------------------------------------------------
    def _adapter_keys(self, sd: Dict[str, Any]) -> Dict[str, Any]:
        out = {}
        bias_mode = self.lora_config.bias if self.lora_config else "none"
        for k, v in sd.items():
            if LORA_PREFIX in k or any(k.startswith(m + ".") or ("." + m + ".") in k for m in self.modules_to_save):
                out[k] = v
            elif bias_mode == "all" and "bias" in k:
                out[k] = v
        return out

    def state_dict(self, *a, **kw):
        inner = self.module.local_state_dict(*a, **kw) if hasattr(self.module, "local_state_dict") \
            else self.module.state_dict(*a, **kw)
        if not self.is_lora_enabled:
            return inner
        if self.lora_config.save_lora_base:
            sd = dict(inner)
        else:
            sd = self._adapter_keys(inner)
        if self.lora_config.save_lora_config_adapter:
            sd["lora_config"] = self.lora_config.selected_fields_to_save()
        return sd

    def load_state_dict(self, state_dict, strict: bool = True):
        state_dict = dict(state_dict)
        saved_cfg = state_dict.pop("lora_config", None)
        if saved_cfg is not None and self.lora_config is not None and not self.is_lora_enabled:
            for k, v in saved_cfg.items():
                setattr(self.lora_config, k, v)
            self.inject_adapter()
        adapter_only = all(LORA_PREFIX in k or any(m in k for m in self.modules_to_save) or "bias" in k
                           for k in state_dict)
        res = self.module.load_state_dict(state_dict, strict=strict and not adapter_only)
        if adapter_only:
            missing = [k for k in res.missing_keys if LORA_PREFIX in k]
            if strict and (missing or res.unexpected_keys):
                raise RuntimeError(f"LoRA load: missing {missing}, unexpected {res.unexpected_keys}")
        return res

    # -- single-file save / load (reference ``:467-492,559-603``) ----------------------------------------------
    def save_lora(self, save_dir: Optional[str] = None, adapter_tag: Optional[str] = None, tag: Optional[str] = None) -> str:
        """``adapter_tag`` is the reference's argument name, ``tag`` the older alias; default file ``lora_adapter.pt``."""
        tag = adapter_tag or tag or "lora_adapter"
        save_dir = save_dir or self.lora_config.lora_save_dir
        assert save_dir, "lora_save_dir is not set"
        os.makedirs(save_dir, exist_ok=True)
        merged_here = False
        if self.lora_config.merge_lora and not self.is_lora_merged:
            self.merge_lora()
            merged_here = True
        sd = {k: (v.detach().cpu() if isinstance(v, torch.Tensor) else v) for k, v in self.state_dict().items()}
        path = os.path.join(save_dir, f"{tag}.pt")
        torch.save(sd, path)
        if merged_here:
            self.unmerge_lora()
        return path

    def load_lora(self, save_dir=None, adapter_tag: Optional[str] = None, ckpt_path: Optional[str] = None,
                  adapter_only: bool = True, config: Optional[LoraConfig] = None) -> None:
        """Single-file load (reference ``:559-591``): ``ckpt_path`` names the file directly, else
        ``<save_dir>/<adapter_tag>.pt`` with both defaulting to the LoRA config.  A ``LoraConfig`` as first argument is
        still accepted.  Whatever the file holds is loaded: the adapter (and ``modules_to_save``) always, base weights
        when it was saved with ``save_lora_base`` -- ``adapter_only`` is accepted for signature parity."""
        if isinstance(save_dir, LoraConfig):
            config, save_dir = save_dir, None
        config = config or self.lora_config
        path = ckpt_path or os.path.join(save_dir or config.lora_save_dir,
                                         f"{adapter_tag or config.lora_load_tag or 'lora_adapter'}.pt")
        sd = torch.load(path, map_location="cpu", weights_only=False)
        self.load_state_dict(sd, strict=False)

    # -- delegation ------------------------------------------------------------------------------------------
    def forward(self, *a, **kw):
        return self.module(*a, **kw)

    def __getattr__(self, name):
        try:
            return super().__getattr__(name)
        except AttributeError:
            return getattr(self.module, name)

    @property
    def dtype(self):
        return getattr(self.module, "dtype", None) or next(self.module.parameters()).dtype

    @property
    def config(self):
        return getattr(self.module, "config", None)

    def print_model_info(self) -> None:
        """Model, adapter configuration and trainable-parameter summary (reference ``:690-694``)."""
        if getattr(self.lora_config, "lora_verbose", False):
            logger.info("LoRA model: %s", self.module)
            logger.info("LoRA configuration: %s", self.lora_config)
        self.print_trainable_parameters()

    def get_nb_trainable_parameters(self) -> Tuple[int, int]:
        """``(trainable, all)`` parameter counts (reference ``:735-747``)."""
        t = sum(p.numel() for p in self.parameters() if p.requires_grad)
        return t, sum(p.numel() for p in self.parameters())

    def module_state_dict(self) -> Dict[str, Any]:
        """State dict of the wrapped model with the LoRA wrappers still in the key names (reference ``:641-645``)."""
        return self.module.local_state_dict() if hasattr(self.module, "local_state_dict") else self.module.state_dict()

    def update_state_dict_keys(self, state_dict: Dict[str, Any]) -> Dict[str, Any]:
        """Map the keys of a BASE-model checkpoint onto the wrapped model: ``...q_proj.weight`` becomes
        ``...q_proj.base_layer.weight`` for every module that now carries an adapter (reference ``:617-630``)."""
        wrapped = {name for name, m in self.module.named_modules() if hasattr(m, "base_layer")}
        out = {}
        for k, v in state_dict.items():
            if isinstance(k, str) and "." in k and LORA_PREFIX not in k and ".base_layer." not in k:
                mod, leaf = k.rsplit(".", 1)
                if mod in wrapped:
                    k = f"{mod}.base_layer.{leaf}"
            out[k] = v
        return out

    def save_config(self, save_dir: Optional[str] = None, filename: str = "adapter_config.json") -> str:
        """The adapter configuration as JSON next to the adapter weights (reference ``:446-465``)."""
        import json

        save_dir = save_dir or self.lora_config.lora_save_dir
        os.makedirs(save_dir, exist_ok=True)
        path = os.path.join(save_dir, filename)
        with open(path, "w") as f:
            json.dump(self.lora_config.selected_fields_to_save(), f, indent=2, sort_keys=True, default=str)
        return path

    def load_checkpoint(self, lora_config: Optional[LoraConfig] = None) -> Dict[str, Any]:
        """Read the adapter checkpoint named by ``lora_config`` (``lora_save_dir`` / ``lora_load_tag``) and adopt the
        configuration stored inside it, without touching the weights yet (reference ``:494-557``)."""
        config = lora_config or self.lora_config
        path = os.path.join(config.lora_save_dir, f"{config.lora_load_tag or 'lora_adapter'}.pt")
        sd = torch.load(path, map_location="cpu", weights_only=False)
        saved = sd.get("lora_config")
        if saved is not None and self.lora_config is not None:
            for k, v in saved.items():
                setattr(self.lora_config, k, v)
        self._loaded_adapter = sd
        return sd

    def load_lora_adapter(self) -> None:
        """Apply the adapter weights fetched by ``load_checkpoint`` (reference ``:605-615``)."""
        sd = getattr(self, "_loaded_adapter", None)
        if sd is None:
            sd = self.load_checkpoint()
        self.load_state_dict(sd, strict=False)
        self._loaded_adapter = None

    def generate(self, *args, **kwargs):
        """Delegate to the wrapped model's ``generate`` (reference ``:132-139``)."""
        return self.module.generate(*args, **kwargs)

    def print_trainable_parameters(self) -> str:
        t, n = self.get_nb_trainable_parameters()
        msg = f"trainable params: {t:,d} || all params: {n:,d} || trainable%: {100 * t / max(n, 1):.4f}"
        logger.info(msg)
        return msg
