"""Trainer API (reference ``src/neuronx_distributed/trainer/__init__.py`` and ``src/neuronx_distributed/__init__.py:2-13``)."""
from .checkpoint import (CheckpointIOState, finalize_checkpoint, has_checkpoint, load_checkpoint,  # noqa: F401
                         save_checkpoint)
from .checkpoint_storage import (BaseCheckpointStorage, FilesysCheckpointStorage, S3CheckpointStorage,  # noqa: F401
                                 create_checkpoint_storage)
from .model import NxDModel  # noqa: F401
from .optimizer import NxDOptimizer  # noqa: F401
from .post_partition_hooks import PostPartitionHooks, hooks  # noqa: F401
from .trainer import (initialize_optimizer_from_class, initialize_parallel_model,  # noqa: F401
                      initialize_parallel_optimizer, neuronx_distributed_config)
