"""``NxDModel``: uniform wrapper over a plain module or a pipeline-partitioned ``NxDPPModel``.

Reference: ``src/neuronx_distributed/trainer/model.py:8-116``.
"""
from __future__ import annotations

import torch


class NxDModel(torch.nn.Module):
    def __init__(self, module: torch.nn.Module, nxd_config: dict):
        super().__init__()
        self.module = module
        self.nxd_config = nxd_config
        self.pp_enabled = nxd_config["pipeline_parallel_size"] > 1

    def local_modules(self):
        """Modules that live on this rank: the local pipeline-stage modules under PP, else the whole model
        (reference ``trainer/model.py:24-27``)."""
        if hasattr(self.module, "local_stage_modules"):
            return self.module.local_stage_modules
        return [self.module]

    def __repr__(self):
        return "NxDModel(" + self.module.__repr__() + ")"

    def original_module(self):
        return self.module.original_torch_module if self.pp_enabled else self.module

    # -- execution -------------------------------------------------------------------------------
    def run_train(self, *args, **kwargs):
        """PP: run the 1F1B engine (fwd+bwd of all micro-batches).  Otherwise forward + ``loss.backward()``."""
        if self.pp_enabled:
            return self.module.run_train(*args, **kwargs)
        out = self.module(*args, **kwargs)
        loss = out if isinstance(out, torch.Tensor) else (out[0] if isinstance(out, (tuple, list)) else out.loss)
        loss.backward()
        return loss

    def run_eval(self, *args, **kwargs):
        if self.pp_enabled:
            return self.module.run_eval(*args, **kwargs)
        with torch.no_grad():
            return self.module(*args, **kwargs)

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    # -- parameter views route to the PP-local versions ------------------------------------------
    def parameters(self, *a, **kw):
        return self.module.local_parameters(*a, **kw) if self.pp_enabled else self.module.parameters(*a, **kw)

    def named_parameters(self, *a, **kw):
        return self.module.local_named_parameters(*a, **kw) if self.pp_enabled else \
            self.module.named_parameters(*a, **kw)

    def buffers(self, *a, **kw):
        return self.module.local_buffers(*a, **kw) if self.pp_enabled else self.module.buffers(*a, **kw)

    def named_buffers(self, *a, **kw):
        return self.module.local_named_buffers(*a, **kw) if self.pp_enabled else self.module.named_buffers(*a, **kw)

    def children(self):
        return self.module.local_children() if self.pp_enabled else self.module.children()

    def named_children(self):
        return self.module.local_named_children() if self.pp_enabled else self.module.named_children()

    def modules(self):
        return self.module.local_modules() if self.pp_enabled else self.module.modules()

    def named_modules(self, *a, **kw):
        return self.module.local_named_modules(*a, **kw) if self.pp_enabled else self.module.named_modules(*a, **kw)

    def state_dict(self, *a, **kw):
        return self.module.local_state_dict(*a, **kw) if self.pp_enabled else self.module.state_dict(*a, **kw)

    def load_state_dict(self, state_dict, strict: bool = True):
        if self.pp_enabled:
            return self.module.load_local_state_dict(state_dict, strict=strict)
        return self.module.load_state_dict(state_dict, strict=strict)

    def __getattr__(self, name):
        try:
            return super().__getattr__(name)
        except AttributeError:
            return getattr(self.module, name)

    # -- HF conveniences ---------------------------------------------------------------------------
    @property
    def dtype(self):
        m = self.original_module()
        if hasattr(m, "dtype"):
            return m.dtype
        return next(m.parameters()).dtype

    @property
    def config(self):
        return getattr(self.original_module(), "config", None)

    @property
    def name_or_path(self):
        return getattr(self.original_module(), "name_or_path", "")
