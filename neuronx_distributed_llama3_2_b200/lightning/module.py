"""``NeuronLTModule`` (reference ``lightning/module.py:24-320``): builds model and optimizer through the NxD trainer API
and runs manual optimisation with gradient accumulation; ``log`` only records on the rank that owns the loss."""
from __future__ import annotations

from typing import Any, Callable, Dict, Optional, Tuple

import torch

from .. import trainer as nxd
from ..parallel import parallel_state as ps
from ._compat import ModuleBase


class NeuronLTModule(ModuleBase):
    def __init__(self, nxd_config: Dict[str, Any], opt_cls: Callable, scheduler_cls: Optional[Callable] = None,
                 model_fn: Optional[Callable] = None, model_args: Tuple = (), model_kwargs: Optional[Dict] = None,
                 opt_args: Tuple = (), opt_kwargs: Optional[Dict] = None, scheduler_args: Tuple = (),
                 scheduler_kwargs: Optional[Dict] = None, grad_accum_steps: int = 1, log_rank0: bool = False,
                 manual_opt: bool = True, device=None, train_batch_size: int = 16, logging_interval: int = 1):
        super().__init__()
        self.train_batch_size, self.logging_interval = train_batch_size, logging_interval
        self.nxd_config, self.opt_cls, self.scheduler_cls = nxd_config, opt_cls, scheduler_cls
        self.model_fn, self.model_args, self.model_kwargs = model_fn, model_args, dict(model_kwargs or {})
        self.opt_args, self.opt_kwargs = opt_args, dict(opt_kwargs or {})
        self.scheduler_args, self.scheduler_kwargs = scheduler_args, dict(scheduler_kwargs or {})
        self.grad_accum_steps, self.log_rank0 = grad_accum_steps, log_rank0
        self.automatic_optimization = not manual_opt
        self._device = device
        self.model = None
        self.global_step_ = 0
        self._micro = 0
        self.loss: Optional[torch.Tensor] = None

    def setup(self, stage: Optional[str] = None) -> None:
        self.model = nxd.initialize_parallel_model(self.nxd_config, self.model_fn, *self.model_args,
                                                   **dict(self.model_kwargs, **({"device": self._device} if self._device else {})))

    def configure_optimizers(self):
        opt = nxd.initialize_parallel_optimizer(self.nxd_config, self.opt_cls, self.model.parameters(), *self.opt_args,
                                                **self.opt_kwargs)
        sched = self.scheduler_cls(opt, *self.scheduler_args, **self.scheduler_kwargs) if self.scheduler_cls else None
        self._optimizers, self._schedulers = opt, sched
        return ([opt], [sched]) if sched is not None else opt

    def forward(self, *a, **kw):
        return self.model(*a, **kw)

    # -- Lightning hooks the reference overrides (``module.py:88-140``) ----------------------------------------------------
    def configure_gradient_clipping(self, *args, **kwargs) -> None:
        """Clipping happens inside ``NxDOptimizer`` / the ZeRO-1 step (it needs the norm over TP, PP and the shards):
        Lightning's own clipping is switched off."""

    def clip_gradients(self, *args, **kwargs) -> None:
        """See ``configure_gradient_clipping``."""

    def on_train_batch_end(self, *args, **kwargs) -> None:
        """User hook."""

    def named_parameters(self, *args, **kwargs):
        if self.model is None:
            return iter(())
        return self.model.named_parameters(*args, **kwargs)

    def get_param_groups_by_weight_decay(self, weight_decay: float = 0.01):
        """Two optimizer groups: decayed weights, and biases / norm weights without decay.  Under pipeline parallelism
        only the parameters of the local stages are listed (ZeRO-1 takes its device from the first one)."""
        m = self.model
        inner = getattr(m, "module", m)
        if hasattr(inner, "local_named_parameters") and getattr(inner, "partitioned", True) and getattr(m, "pp_enabled", False):
            named = list(inner.local_named_parameters())
        else:
            named = list(m.named_parameters())
        no_decay = ("bias", "LayerNorm", "layernorm", "norm.weight")
        return [{"params": [p for n, p in named if not any(k in n for k in no_decay)], "weight_decay": weight_decay},
                {"params": [p for n, p in named if any(k in n for k in no_decay)], "weight_decay": 0.0}]

    def training_step(self, batch, batch_idx: int = 0):
        """Manual optimisation: backward every micro-step (inside ``run_train``), optimizer step every
        ``grad_accum_steps`` micro-steps."""
        opt, sched = self._optimizers, self._schedulers
        kwargs = batch if isinstance(batch, dict) else {"input_ids": batch, "labels": batch}
        if self._micro == 0:
            opt.zero_grad()
        loss = self.model.run_train(**kwargs)
        self._micro += 1
        if loss is not None:
            self.loss = loss.detach()
        if self._micro == self.grad_accum_steps:
            self._micro = 0
            opt.step()
            if sched is not None:
                sched.step()
            self.global_step_ += 1
            if self.should_print() and self.loss is not None:
                self.log("loss", float(self.loss))
                if opt.grad_norm is not None:
                    self.log("grad_norm", float(opt.grad_norm))
        return self.loss

    def should_print(self) -> bool:
        if not ps.model_parallel_is_initialized():
            return True
        if self.log_rank0:
            return ps.is_global_rank_zero()
        return (ps.get_data_parallel_rank() == 0 and ps.get_tensor_model_parallel_rank() == 0
                and ps.get_pipeline_model_parallel_rank() == ps.get_pipeline_model_parallel_size() - 1)

    def state_dict(self, *a, **kw):
        return self.model.state_dict(*a, **kw) if self.model is not None else {}
