"""``NxDOptimizer``: wraps the inner optimizer with the model-parallel gradient plumbing.

Reference: ``src/neuronx_distributed/trainer/optimizer.py:10-154``.
"""
from __future__ import annotations

from typing import Any, Dict

import torch
import torch.distributed as dist

from ..parallel import grads
from ..parallel import parallel_state as ps


class NxDOptimizer(torch.optim.Optimizer):
    def __init__(self, optimizer: torch.optim.Optimizer, nxd_config: Dict[str, Any]):
        self.optimizer = optimizer
        self.nxd_config = nxd_config
        self._grad_norm = None
        self._optimizer_step_pre_hooks = getattr(optimizer, "_optimizer_step_pre_hooks", {})
        self._optimizer_step_post_hooks = getattr(optimizer, "_optimizer_step_post_hooks", {})

    @property
    def state(self):
        return self.optimizer.state

    # param_groups / defaults are the inner optimizer's objects: LR schedulers that assign a new list keep working
    @property
    def param_groups(self):
        return self.optimizer.param_groups

    @param_groups.setter
    def param_groups(self, param_groups):
        self.optimizer.param_groups = param_groups

    @property
    def defaults(self):
        return self.optimizer.defaults

    @defaults.setter
    def defaults(self, defaults):
        self.optimizer.defaults = defaults

    def add_param_group(self, param_group):
        self.optimizer.add_param_group(param_group)

    @property
    def params(self):
        return [p for g in self.optimizer.param_groups for p in g["params"]]

    @property
    def grad_norm(self):
        """Pre-clip global gradient norm of the last step (reference ``optimizer.py:26-28``)."""
        return self._grad_norm

    @property
    def zero1(self) -> bool:
        return bool(self.nxd_config["optimizer_config"]["zero_one_enabled"])

    def zero_grad(self, set_to_none: bool = True):
        if self.zero1:
            return self.optimizer.zero_grad()
        return self.optimizer.zero_grad(set_to_none=set_to_none)

    def step(self, closure=None):
        oc = self.nxd_config["optimizer_config"]
        if self.nxd_config["sequence_parallel"]:
            if hasattr(self.optimizer, "allreduce_sequence_parallel_gradients"):
                self.optimizer.allreduce_sequence_parallel_gradients()
            else:
                grads.allreduce_sequence_parallel_gradients(self)
        if not self.zero1:
            if dist.is_initialized() and ps.model_parallel_is_initialized() and ps.get_data_parallel_size() > 1:
                params = [p for p in self.params if p.grad is not None]
                non_ep = [p.grad for p in params if not getattr(p, "expert_model_parallel", False)]
                ep = [p.grad for p in params if getattr(p, "expert_model_parallel", False)]
                if ps.get_expert_model_parallel_size() > 1:
                    grads.bucket_allreduce_gradients(non_ep + ep)                       # expert-DP
                    grads.bucket_allreduce_gradients(non_ep, reduce_over_ep_group=True)  # then EP for non-expert
                else:
                    grads.bucket_allreduce_gradients(non_ep + ep)
            if oc["grad_clipping"]:
                self._grad_norm = grads.clip_grad_norm(self.params, oc["max_grad_norm"])
        ret = self.optimizer.step(closure=closure)
        if self.zero1:
            self._grad_norm = self.optimizer.grad_norm
        return ret

    # -- state dict: mark expert-parallel states so checkpoint IO can name files (reference :57-98) --
    def state_dict(self):
        sd = self.optimizer.state_dict()
        if not self.zero1 and isinstance(sd, dict) and "state" in sd:
            ep_ids = set()
            idx = 0
            for g in self.optimizer.param_groups:
                for p in g["params"]:
                    if getattr(p, "expert_model_parallel", False):
                        ep_ids.add(idx)
                    idx += 1
            for i, st in sd["state"].items():
                if i in ep_ids and isinstance(st, dict):
                    for v in st.values():
                        if isinstance(v, torch.Tensor):
                            setattr(v, "expert_model_parallel", True)
        return sd

    def load_state_dict(self, state_dict):
        return self.optimizer.load_state_dict(state_dict)

    # deprecated per-rank sharded save / load of the ZeRO-1 state (reference ``:141-154``; use nxd.save_checkpoint)
    def save_state_dict(self, output_dir: str, num_workers_per_step: int = 8) -> None:
        assert self.zero1, "save_state_dict needs optimizer_config['zero_one_enabled']"
        self.optimizer.save_sharded_state_dict(output_dir, num_workers_per_step)

    def load_state_dict_from(self, output_dir: str, num_workers_per_step: int = 8) -> None:
        assert self.zero1, "load_state_dict_from needs optimizer_config['zero_one_enabled']"
        self.optimizer.load_sharded_state_dict(output_dir, num_workers_per_step)

    def __getattr__(self, name):
        return getattr(self.__dict__["optimizer"], name)
