"""Strategy (reference ``lightning/strategy.py:31-214``): sets up the process group and the model-parallel groups from
the NxD config, hands DP-aware sampler kwargs to the data module, keeps batches on the host under pipeline parallelism
(every stage picks what it needs) and checkpoints through ``NeuronCheckpointIO``."""
from __future__ import annotations

from typing import Any, Dict, Optional

import torch
import torch.distributed as dist

from ..parallel import parallel_state as ps
from ._compat import StrategyBase
from .accelerator import CudaNxDAccelerator
from .checkpoint_io import NeuronCheckpointIO
from .launcher import _NxDLauncher


class NeuronXLAStrategy(StrategyBase):
    def __init__(self, nxd_config: Optional[Dict[str, Any]] = None, tensor_parallel_size: int = 1,
                 pipeline_parallel_size: int = 1, save_load_xser: bool = True, accelerator=None, checkpoint_io=None,
                 precision_plugin=None, expert_parallel_size: int = 1, debug: bool = False,
                 sync_module_states: bool = False, **kw):
        try:
            super().__init__(**kw)
        except TypeError:
            super().__init__()
        self.nxd_config = nxd_config
        self.tensor_parallel_size = nxd_config["tensor_parallel_size"] if nxd_config else tensor_parallel_size
        self.pipeline_parallel_size = nxd_config["pipeline_parallel_size"] if nxd_config else pipeline_parallel_size
        self.expert_parallel_size = nxd_config.get("expert_parallel_size", 1) if nxd_config else expert_parallel_size
        self.debug, self._sync_module_states = debug, sync_module_states
        self._accelerator = accelerator or CudaNxDAccelerator()
        self._checkpoint_io = checkpoint_io or NeuronCheckpointIO(save_xser=save_load_xser, load_xser=save_load_xser)
        self._precision = precision_plugin
        self._launcher = _NxDLauncher(self)

    @property
    def root_device(self) -> torch.device:
        return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")

    @property
    def checkpoint_io(self):
        return self._checkpoint_io

    def setup_distributed(self) -> None:
        self._launcher.launch(lambda: None)
        if dist.is_initialized() and not ps.model_parallel_is_initialized():
            ps.initialize_model_parallel(self.tensor_parallel_size, self.pipeline_parallel_size, self.expert_parallel_size)

    @property
    def distributed_sampler_kwargs(self) -> Dict[str, int]:
        return {"num_replicas": ps.get_data_parallel_size(), "rank": ps.get_data_parallel_rank()}

    def batch_to_device(self, batch, device=None, dataloader_idx: int = 0):
        if self.pipeline_parallel_size > 1:
            return batch   # NxDPPModel moves the micro-batches itself
        dev = device or self.root_device
        if isinstance(batch, dict):
            return {k: (v.to(dev, non_blocking=True) if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}
        if isinstance(batch, (list, tuple)):
            return type(batch)(v.to(dev, non_blocking=True) if isinstance(v, torch.Tensor) else v for v in batch)
        return batch.to(dev, non_blocking=True) if isinstance(batch, torch.Tensor) else batch

    def broadcast(self, obj, src: int = 0):
        """Every rank builds identical python objects from identical configuration (reference ``strategy.py:79-80``)."""
        return obj

    def model_to_device(self) -> None:
        """``initialize_parallel_model`` already placed the shards on this rank's GPU (reference ``:124-126``)."""

    def process_dataloader(self, dataloader):
        """The reference wraps loaders in an XLA device loader unless PP keeps batches on the host (``:178-196``); here
        ``batch_to_device`` issues the non-blocking H2D copy (and leaves batches on the host under PP, where the
        pipeline engine moves each micro-batch itself), so loaders pass through unchanged."""
        return getattr(dataloader, "_loader", dataloader) if self.pipeline_parallel_size > 1 else dataloader

    def reduce(self, tensor, group=None, reduce_op="mean"):
        if not dist.is_initialized():
            return tensor
        t = tensor.clone() if isinstance(tensor, torch.Tensor) else torch.tensor(float(tensor), device=self.root_device)
        dist.all_reduce(t)
        return t / dist.get_world_size() if reduce_op in ("mean", "avg") else t

    def save_checkpoint(self, checkpoint: Dict[str, Any], filepath: str, storage_options=None) -> None:
        self.checkpoint_io.save_checkpoint(checkpoint, filepath, storage_options)

    def load_checkpoint(self, checkpoint_path: str) -> Dict[str, Any]:
        return self.checkpoint_io.load_checkpoint(checkpoint_path)

    def teardown(self) -> None:
        if ps.model_parallel_is_initialized():
            ps.destroy_model_parallel()
