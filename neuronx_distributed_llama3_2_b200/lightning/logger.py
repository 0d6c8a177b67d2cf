"""TensorBoard logger that writes only from dp rank 0 / tp rank 0 / last pp rank (reference
``lightning/logger.py:24-154``: that rank owns the loss under pipeline parallelism)."""
from __future__ import annotations

import os
from typing import Dict, Optional

from ..parallel import parallel_state as ps


class NeuronTensorBoardLogger:
    def __init__(self, save_dir: str, name: str = "lightning_logs", version: Optional[str] = None, log_rank0: bool = False):
        self.save_dir, self.name, self.version, self.log_rank0 = save_dir, name, version or "version_0", log_rank0
        self._writer = None

    def should_print(self) -> bool:
        if not ps.model_parallel_is_initialized():
            return True
        if self.log_rank0:
            return ps.is_global_rank_zero()
        return (ps.get_data_parallel_rank() == 0 and ps.get_tensor_model_parallel_rank() == 0
                and ps.get_pipeline_model_parallel_rank() == ps.get_pipeline_model_parallel_size() - 1)

    @property
    def log_dir(self) -> str:
        return os.path.join(self.save_dir, self.name, self.version)

    @property
    def experiment(self):
        if self._writer is None and self.should_print():
            from torch.utils.tensorboard import SummaryWriter

            os.makedirs(self.log_dir, exist_ok=True)
            self._writer = SummaryWriter(self.log_dir)
        return self._writer

    def log_metrics(self, metrics: Dict[str, float], step: Optional[int] = None) -> None:
        w = self.experiment
        if w is None:
            return
        for k, v in metrics.items():
            w.add_scalar(k, float(v), step)

    def finalize(self, status: str = "success") -> None:
        if self._writer is not None:
            self._writer.flush()
            self._writer.close()
            self._writer = None
