"""Progress bar that only draws on global rank 0 (reference ``lightning/progress_bar.py:9-19``)."""
from __future__ import annotations

from ..parallel import parallel_state as ps
from ._compat import ProgressBarBase


class NeuronTQDMProgressBar(ProgressBarBase):
    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self._bar = None
        self._trainer = None
        self._off = False

    def setup(self, trainer, pl_module, stage: str = "fit") -> None:
        """Draw on ONE rank only: dp 0 / tp 0 of the first pipeline stage when the module logs from rank 0
        (``log_rank0``), else of the last stage, which is the one that holds the loss (reference ``progress_bar.py:10-19``)."""
        self._trainer = trainer
        if not ps.model_parallel_is_initialized():
            return
        want_pp = 0 if getattr(pl_module, "log_rank0", False) else ps.get_pipeline_model_parallel_size() - 1
        mine = (ps.get_data_parallel_rank() == 0 and ps.get_tensor_model_parallel_rank() == 0
                and ps.get_pipeline_model_parallel_rank() == want_pp)
        if not mine:
            self.disable()

    def disable(self) -> None:
        self._off = True

    def enable(self) -> None:
        self._off = False

    @property
    def is_enabled(self) -> bool:
        if self._trainer is not None:
            return not self._off
        return ps.is_global_rank_zero() and not self._off

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
