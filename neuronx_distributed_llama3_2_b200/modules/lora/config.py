"""``LoraConfig`` (reference ``src/neuronx_distributed/modules/lora/config.py:5-145``)."""
from __future__ import annotations

from dataclasses import asdict, dataclass
from typing import List, Literal, Optional


@dataclass
class LoraConfig:
    enable_lora: bool = False
    lora_rank: int = 16
    target_modules: Optional[List[str]] = None          # module-name suffixes or a regex; None -> per-arch default
    lora_alpha: int = 8
    lora_dropout: float = 0.0
    bias: Literal["none", "all", "lora_only"] = "none"
    use_rslora: bool = False                             # scale = alpha / sqrt(r) instead of alpha / r
    init_lora_weights: Literal["default", "gaussian"] = "default"
    modules_to_save: Optional[List[str]] = None          # extra (non-LoRA) modules kept trainable and checkpointed
    lora_verbose: bool = False
    load_lora_from_ckpt: bool = False
    lora_load_tag: Optional[str] = None
    lora_save_dir: Optional[str] = None
    merge_lora: bool = False                             # fold A/B into the base weight when saving
    save_lora_base: bool = False
    save_lora_config_adapter: bool = True                # embed this config next to the adapter weights

    @staticmethod
    def get_selected_fields() -> List[str]:
        return ["lora_rank", "target_modules", "lora_alpha", "lora_dropout", "bias", "use_rslora", "init_lora_weights",
                "modules_to_save"]

    def selected_fields_to_save(self) -> dict:
        d = asdict(self)
        return {k: d[k] for k in self.get_selected_fields()}

    def __post_init__(self):
        if isinstance(self.target_modules, str):
            pass  # regex
        elif self.target_modules is not None:
            self.target_modules = list(self.target_modules)
        if self.lora_rank <= 0:
            raise ValueError("lora_rank must be positive")
