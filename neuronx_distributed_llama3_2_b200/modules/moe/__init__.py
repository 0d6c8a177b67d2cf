"""Mixture-of-Experts modules (reference ``src/neuronx_distributed/modules/moe/``)."""
from .expert_mlps import ExpertMLPs  # noqa: F401
from .experts import Experts  # noqa: F401
from .loss_function import load_balancing_loss_func  # noqa: F401
from .model import MoE  # noqa: F401
from .moe_parallel_layers import (ExpertFusedColumnParallelLinear, ExpertFusedRowParallelLinear,  # noqa: F401
                                  LinearRouter)
from .routing import RouterBase, RouterSinkhorn, RouterTopK  # noqa: F401

from ...utils.model_utils import ACT2FN  # noqa: E402,F401  (reference modules/moe/__init__.py:4)
