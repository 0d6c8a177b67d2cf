"""LoRA (reference ``src/neuronx_distributed/modules/lora/__init__.py:7-17``)."""
from .config import LoraConfig  # noqa: F401
from .layer import LoraConv2d, LoraEmbedding, LoraLayer, LoraLinear  # noqa: F401
from .model import LoraModel  # noqa: F401
from .tp_layer import LoraGQAQKVParallelLinear, LoraParallelLinear  # noqa: F401


def get_lora_model(model, lora_config: LoraConfig = None):
    """Wrap ``model`` with LoRA adapters when ``lora_config.enable_lora`` (no-op otherwise)."""
    if lora_config is None or not lora_config.enable_lora:
        return model
    return LoraModel(model, lora_config)
