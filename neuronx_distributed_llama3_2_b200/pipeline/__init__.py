"""Pipeline parallelism (reference ``src/neuronx_distributed/pipeline/__init__.py``)."""
from .model import NxDPPModel  # noqa: F401
from .scheduler import (InferenceSchedule, Train1F1BSchedule, TrainInterleavedSchedule,  # noqa: F401
                        TrainSchedule)
