#!/usr/bin/env python
"""HF / Megatron full checkpoint <-> (tp, pp[, ep])-sharded training checkpoint -- counterpart of the reference's
``examples/training/llama/convert_checkpoints.py`` and ``examples/training/mixtral/convert_checkpoints.py``.

    # Llama: HF -> TP8 x PP1 shards with 4x KV replication (so 8 KV heads... 32 KV shards), fused QKV
    python examples/training/convert_checkpoints.py --model_family llama --input_dir ~/Llama-3.2-3B \
        --output_dir ~/ckpt/pretrained --config ~/Llama-3.2-3B/config.json --tp_size 8 --convert_from_full_state
    # Mixtral: HF -> TP4 x EP2 shards
    python examples/training/convert_checkpoints.py --model_family mixtral --input_dir ~/Mixtral-8x7B \
        --output_dir ~/ckpt/mixtral --config ~/Mixtral-8x7B/config.json --tp_size 4 --ep_size 2 --convert_from_full_state

The result directory is what ``--pretrained_weight`` of the training drivers takes.
"""
from __future__ import annotations

import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from neuronx_distributed_llama3_2_b200.scripts.checkpoint_converter import CheckpointConverterBase  # noqa: E402
from neuronx_distributed_llama3_2_b200.scripts.checkpoint_converter_moe import CheckpointConverterMixtral  # noqa: E402


class CheckpointConverterLlama(CheckpointConverterBase):
    """The Llama family needs nothing beyond the base class (as in the reference)."""


if __name__ == "__main__":
    argv = sys.argv[1:]
    family = "llama"
    if "--model_family" in argv:
        i = argv.index("--model_family")
        family = argv[i + 1]
        del argv[i:i + 2]
    {"llama": CheckpointConverterLlama, "mixtral": CheckpointConverterMixtral}[family].run(argv)
