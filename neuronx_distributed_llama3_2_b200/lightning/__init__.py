"""PyTorch-Lightning integration (reference ``src/neuronx_distributed/lightning/*.py``).

Lightning is not part of this image.  The classes below implement the reference's roles (strategy, module, checkpoint
IO, precision plugin, accelerator, launcher, progress bar, TensorBoard logger) against Lightning's duck-typed protocol:
when ``lightning`` / ``pytorch_lightning`` is importable they subclass its base classes, otherwise they subclass small
local stand-ins and can be driven by ``NxDLightningLoop`` (a minimal fit loop used by the tests and examples).
"""
from .accelerator import CudaNxDAccelerator, NeuronXLAAccelerator  # noqa: F401
from .checkpoint_io import NeuronCheckpointIO  # noqa: F401
from .launcher import _NxDLauncher  # noqa: F401
from .logger import NeuronTensorBoardLogger  # noqa: F401
from .loop import NxDLightningLoop  # noqa: F401
from .module import NeuronLTModule  # noqa: F401
from .precision_plugin import NeuronXLAPrecisionPlugin  # noqa: F401
from .progress_bar import NeuronTQDMProgressBar  # noqa: F401
from .strategy import NeuronXLAStrategy  # noqa: F401
