"""B200-native model-parallel training / inference framework with the NeuronxDistributed API surface
(reference ``src/neuronx_distributed/__init__.py:1-13``).  Heavy sub-packages load lazily."""
from __future__ import annotations

import importlib

__version__ = "0.9.0+b200.1"

_LAZY = {
    # trainer API
    "neuronx_distributed_config": ".trainer", "initialize_parallel_model": ".trainer",
    "initialize_parallel_optimizer": ".trainer", "save_checkpoint": ".trainer", "load_checkpoint": ".trainer",
    "has_checkpoint": ".trainer", "finalize_checkpoint": ".trainer", "CheckpointIOState": ".trainer",
    "NxDModel": ".trainer", "NxDOptimizer": ".trainer",
    # sub-packages
    "parallel_layers": ".parallel", "parallel": ".parallel", "pipeline": ".pipeline", "optimizer": ".optimizer",
    "trainer": ".trainer", "modules": ".modules", "ops": ".ops", "models": ".models", "utils": ".utils",
    "quantization": ".quantization", "inference": ".inference", "trace": ".inference", "kernels": ".kernels",
    "lightning": ".lightning", "scripts": ".scripts",
}


def __getattr__(name):
    target = _LAZY.get(name)
    if target is None:
        raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
    mod = importlib.import_module(target, __name__)
    val = mod if target.lstrip(".") == name or name in ("parallel_layers", "trace") else getattr(mod, name)
    globals()[name] = val
    return val
