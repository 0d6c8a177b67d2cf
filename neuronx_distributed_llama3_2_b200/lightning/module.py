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
                 manual_opt: bool = True, device=None):
        super().__init__()
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
