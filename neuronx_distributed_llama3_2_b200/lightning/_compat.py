"""Optional Lightning base classes with local stand-ins."""
from __future__ import annotations

try:  # pragma: no cover - not installed in this image
    import lightning.pytorch as pl  # type: ignore
    HAVE_LIGHTNING = True
except Exception:  # noqa: BLE001
    try:  # pragma: no cover
        import pytorch_lightning as pl  # type: ignore
        HAVE_LIGHTNING = True
    except Exception:  # noqa: BLE001
        pl = None
        HAVE_LIGHTNING = False

import torch


class _Base:
    def __init__(self, *a, **kw):
        pass


class _ModuleBase(torch.nn.Module):
    """Stand-in for ``LightningModule``: manual-optimisation hooks only."""

    automatic_optimization = False

    def __init__(self):
        super().__init__()
        self.trainer = None
        self._logged = {}

    def log(self, name, value, **kw):
        self._logged[name] = value

    def optimizers(self):
        return self._optimizers

    def lr_schedulers(self):
        return self._schedulers


if HAVE_LIGHTNING:  # pragma: no cover
    from lightning.pytorch.accelerators import Accelerator as AcceleratorBase  # type: ignore
    from lightning.pytorch.callbacks import TQDMProgressBar as ProgressBarBase  # type: ignore
    from lightning.pytorch.plugins import CheckpointIO as CheckpointIOBase  # type: ignore
    from lightning.pytorch.plugins.precision import Precision as PrecisionBase  # type: ignore
    from lightning.pytorch.strategies import DDPStrategy as StrategyBase  # type: ignore
    ModuleBase = pl.LightningModule
else:
    AcceleratorBase = ProgressBarBase = CheckpointIOBase = PrecisionBase = StrategyBase = _Base
    ModuleBase = _ModuleBase
