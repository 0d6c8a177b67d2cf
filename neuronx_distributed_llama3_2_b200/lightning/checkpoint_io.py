"""Checkpoint IO that routes Lightning's save/load to the sharded TP/PP checkpoint functions (reference
``lightning/checkpoint_io.py:13-64`` -> ``parallel_layers.checkpointing.save/load``)."""
from __future__ import annotations

from typing import Any, Dict, Optional

from ..parallel import checkpointing
from ._compat import CheckpointIOBase


class NeuronCheckpointIO(CheckpointIOBase):
    def __init__(self, save_xser: Optional[bool] = None, load_xser: Optional[bool] = None, save_load_xser: bool = False,
                 weights_only: bool = False):
        """``save_load_xser`` / ``weights_only`` are the reference's arguments (``checkpoint_io.py:14-17``); the separate
        ``save_xser`` / ``load_xser`` switches override the former when given."""
        super().__init__()
        self.save_load_xser, self.weights_only = save_load_xser, weights_only
        self.save_xser = save_load_xser if save_xser is None else save_xser
        self.load_xser = save_load_xser if load_xser is None else load_xser

    def save_checkpoint(self, checkpoint: Dict[str, Any], path: str, storage_options: Optional[Any] = None) -> None:
        checkpointing.save(checkpoint, path, save_xser=self.save_xser)

    def load_checkpoint(self, path: str, map_location=None) -> Dict[str, Any]:
        return checkpointing.load(path, load_xser=self.load_xser)

    def remove_checkpoint(self, path: str) -> None:
        import shutil

        shutil.rmtree(path, ignore_errors=True)
