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
        self._print_step = -1
        self.hparams: Dict[str, object] = {}

    @property
    def print_step(self) -> int:
        """Last step whose metrics were printed to the console (reference ``logger.py:31-36``)."""
        return self._print_step

    @print_step.setter
    def print_step(self, value: int = -1) -> None:
        self._print_step = value

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

    def log_hyperparams(self, params: Dict[str, object]) -> None:
        self.hparams.update(dict(params))

    def save(self) -> None:
        """Flush the event file and write ``hparams.yaml`` once, on the logging rank only (reference ``:61-75``)."""
        if not self.should_print():
            return
        if self._writer is not None:
            self._writer.flush()
        path = os.path.join(self.log_dir, "hparams.yaml")
        if os.path.isdir(self.log_dir) and not os.path.isfile(path):
            import yaml

            with open(path, "w") as f:
                yaml.safe_dump({k: (v if isinstance(v, (int, float, str, bool, type(None))) else str(v))
                                for k, v in self.hparams.items()}, f)

    def log_graph(self, model, input_array=None) -> None:
        """Model graph to TensorBoard, on the logging rank only (reference ``:100-125``).  Tracing a tensor-parallel
        model with ``add_graph`` would issue collectives on one rank only, so this is limited to un-initialised
        model parallelism; otherwise it is a no-op."""
        if not self.should_print() or input_array is None:
            return
        if ps.model_parallel_is_initialized() and ps.get_tensor_model_parallel_size() * ps.get_pipeline_model_parallel_size() > 1:
            return
        w = self.experiment
        if w is not None:
            w.add_graph(model, input_array)

    def finalize(self, status: str = "success") -> None:
        if self._writer is not None:
            self._writer.flush()
            self._writer.close()
            self._writer = None
