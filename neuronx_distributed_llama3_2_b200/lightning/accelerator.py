"""Accelerator (reference ``lightning/accelerator.py:15-100`` binds XLA devices; here: one CUDA device per process)."""
from __future__ import annotations

import os

import torch

from ._compat import AcceleratorBase


class CudaNxDAccelerator(AcceleratorBase):
    def setup_device(self, device: torch.device) -> None:
        if device.type == "cuda":
            torch.cuda.set_device(device)

    def get_device_stats(self, device):
        if torch.cuda.is_available():
            return {"allocated_bytes": torch.cuda.memory_allocated(device), "reserved_bytes": torch.cuda.memory_reserved(device)}
        return {}

    def teardown(self) -> None:
        pass

    @staticmethod
    def parse_devices(devices):
        return devices

    @staticmethod
    def get_parallel_devices(devices):
        n = devices if isinstance(devices, int) else len(devices)
        return [torch.device("cuda", i) for i in range(n)] if torch.cuda.is_available() else [torch.device("cpu")] * n

    @staticmethod
    def auto_device_count() -> int:
        return torch.cuda.device_count() if torch.cuda.is_available() else int(os.environ.get("WORLD_SIZE", 1))

    @staticmethod
    def is_available() -> bool:
        return True

    @classmethod
    def register_accelerators(cls, accelerator_registry) -> None:
        """``Trainer(accelerator="nxd_cuda")`` (and the reference's ``"neuron"``) resolve to this class (reference
        ``accelerator.py:68-70``).  The registry is Lightning's ``_AcceleratorRegistry`` or any mapping with ``register``."""
        for name in ("nxd_cuda", "neuron"):
            if hasattr(accelerator_registry, "register"):
                accelerator_registry.register(name, cls, description=cls.__name__)
            else:
                accelerator_registry[name] = cls


NeuronXLAAccelerator = CudaNxDAccelerator  # reference-compatible name
