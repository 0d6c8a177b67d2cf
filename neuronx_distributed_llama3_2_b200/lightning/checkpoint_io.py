"""Checkpoint IO that routes Lightning's save/load to the sharded TP/PP checkpoint functions (reference
``lightning/checkpoint_io.py:13-64`` -> ``parallel_layers.checkpointing.save/load``)."""
from __future__ import annotations

from typing import Any, Dict, Optional

from ..parallel import checkpointing
from ._compat import CheckpointIOBase


class NeuronCheckpointIO(CheckpointIOBase):
    def __init__(self, save_xser: bool = False, load_xser: bool = False):
        super().__init__()
        self.save_xser, self.load_xser = save_xser, load_xser

    def save_checkpoint(self, checkpoint: Dict[str, Any], path: str, storage_options: Optional[Any] = None) -> None:
        checkpointing.save(checkpoint, path, save_xser=self.save_xser)

    def load_checkpoint(self, path: str, map_location=None) -> Dict[str, Any]:
        return checkpointing.load(path, load_xser=self.load_xser)

    def remove_checkpoint(self, path: str) -> None:
        import shutil

        shutil.rmtree(path, ignore_errors=True)
