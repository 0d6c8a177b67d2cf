"""Launcher (reference ``lightning/launcher.py:12-91`` chooses xmp.spawn or torchrun): processes are created by
``torchrun``; this object only initialises ``torch.distributed`` from the environment when it is not up yet."""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


class _NxDLauncher:
    def __init__(self, strategy=None):
        self.strategy = strategy

    @property
    def is_interactive_compatible(self) -> bool:
        return False

    def launch(self, function, *args, trainer=None, **kwargs):
        if not dist.is_initialized() and int(os.environ.get("WORLD_SIZE", "1")) > 1:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
            if backend == "nccl":
                torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
            dist.init_process_group(backend)
        return function(*args, **kwargs)
