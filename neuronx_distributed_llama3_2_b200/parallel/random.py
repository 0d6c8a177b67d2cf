"""Named RNG streams for tensor parallelism.

Reference: ``src/neuronx_distributed/parallel_layers/random.py:20-127``
(``XLARNGStatesTracker``).  Two streams: the *default* stream is seeded
identically on all TP ranks (dropout on replicated activations, data order), the
*model-parallel* stream is seeded ``seed + 2718 + tp_rank`` (weight init and
dropout inside TP regions must differ per rank).

B200 design: each named stream owns a ``torch.Generator`` state for CPU and, when
CUDA is up, the CUDA RNG state of the current device; ``fork()`` swaps it in.
"""
from __future__ import annotations

import contextlib
from typing import Dict

import torch

from . import parallel_state as ps

_MODEL_PARALLEL_RNG_TRACKER_NAME = "model-parallel-rng"


def _get_state():
    st = {"cpu": torch.get_rng_state()}
    if torch.cuda.is_available() and torch.cuda.is_initialized():
        st["cuda"] = torch.cuda.get_rng_state()
    return st


def _set_state(st) -> None:
    torch.set_rng_state(st["cpu"])
    if "cuda" in st and torch.cuda.is_available():
        torch.cuda.set_rng_state(st["cuda"])


class RNGStatesTracker:
    """Tracks named RNG states; ``fork(name)`` runs a block under that state and
    saves the advanced state back."""

    def __init__(self):
        self.reset()

    def reset(self):
        self.states_: Dict[str, dict] = {}
        self.seeds_ = set()

    def get_states(self):
        return dict(self.states_)

    def set_states(self, states):
        self.states_ = states

    def add(self, name: str, seed: int) -> None:
        if seed in self.seeds_:
            raise Exception(f"seed {seed} already exists")
        if name in self.states_:
            raise Exception(f"rng state {name} already exists")
        self.seeds_.add(seed)
        orig = _get_state()
        torch.manual_seed(seed)
        self.states_[name] = _get_state()
        _set_state(orig)

    @contextlib.contextmanager
    def fork(self, name: str = _MODEL_PARALLEL_RNG_TRACKER_NAME):
        if name not in self.states_:
            # Not seeded through model_parallel_manual_seed: behave as a no-op fork.
            yield
            return
        orig = _get_state()
        _set_state(self.states_[name])
        try:
            yield
        finally:
            self.states_[name] = _get_state()
            _set_state(orig)


_TRACKER = RNGStatesTracker()


def get_rng_tracker() -> RNGStatesTracker:
    return _TRACKER


# reference-compatible names
XLARNGStatesTracker = RNGStatesTracker
get_xla_rng_tracker = get_rng_tracker


def model_parallel_manual_seed(seed: int) -> None:
    """default stream = ``seed`` (same inside a TP group); TP stream =
    ``seed + 2718 + tp_rank`` (reference random.py:100-127)."""
    offset = seed + 2718
    tp_seed = offset + ps.get_tensor_model_parallel_rank()
    _TRACKER.reset()
    torch.manual_seed(seed)
    _TRACKER.add(_MODEL_PARALLEL_RNG_TRACKER_NAME, tp_seed)


model_parallel_xla_manual_seed = model_parallel_manual_seed
