"""Inference: KV-cache model, ModelBuilder with shape-routed CUDA-graph buckets, sampler, runner (reference
``src/neuronx_distributed/trace/__init__.py:1-7`` + ``examples/inference``)."""
from .checkpoint_sharding import _mock_parallel_state, get_sharded_checkpoint, shard_children  # noqa: F401
from .config import InferenceConfig  # noqa: F401
from .kv_cache import KVCache  # noqa: F401
from .model import DbrxInferenceModel, KVCacheInferenceMixin, LlamaInferenceModel, MixtralInferenceModel  # noqa: F401
from .gqa import GQA, determine_sharding_strategy, get_shardable_head_counts  # noqa: F401
from .autobucketing import generate_buckets  # noqa: F401
from .checkpoint import load_state_dict, save_state_dict_safetensors  # noqa: F401
from .model_builder import BaseModelInstance, ModelBuilder, NxDModel  # noqa: F401
from .sampling import Sampler  # noqa: F401
from .trace import parallel_model_load, parallel_model_save, parallel_model_trace  # noqa: F401
from .benchmark import Benchmark, LatencyCollector, generate_report  # noqa: F401
from .runner import DbrxRunner, InferenceRunner, LlamaRunner, MixtralRunner  # noqa: F401
from .model_wrapper import ModelWrapper  # noqa: F401
