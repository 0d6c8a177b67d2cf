"""Progress bar that only draws on global rank 0 (reference ``lightning/progress_bar.py:9-19``)."""
from __future__ import annotations

from ..parallel import parallel_state as ps
from ._compat import ProgressBarBase


class NeuronTQDMProgressBar(ProgressBarBase):
    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self._bar = None

    @property
    def is_enabled(self) -> bool:
        return ps.is_global_rank_zero()

    def on_train_start(self, total: int = None, *a, **kw):
        if self.is_enabled:
            from tqdm import tqdm

            self._bar = tqdm(total=total, desc="train")

    def on_train_batch_end(self, *a, **kw):
        if self._bar is not None:
            self._bar.update(1)

    def on_train_end(self, *a, **kw):
        if self._bar is not None:
            self._bar.close()
