#!/usr/bin/env python
"""Mixtral (sparse MoE) inference -- counterpart of the reference's ``examples/inference/run_mixtral.py``.  Context
encoding runs all experts without capacity limits, token generation loads only the selected experts.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/inference/run_mixtral.py --model mixtral-8x7b \
        --tp_degree 8 --batch_size 1 --max_context_length 1024 --seq_len 2048 --benchmark
"""
from __future__ import annotations

import _common as C

from neuronx_distributed_llama3_2_b200.inference import MixtralRunner
from neuronx_distributed_llama3_2_b200.models import (MixtralForCausalLM, convert_mixtral_hf_state_dict, mixtral_8x7b,
                                                      tiny_mixtral)

PRESETS = {"mixtral-8x7b": mixtral_8x7b, "tiny": tiny_mixtral}


def main():
    args = C.common_parser(PRESETS, "mixtral-8x7b").parse_args()
    C.init_distributed(args.tp_degree)
    cfg = C.model_config(args, PRESETS)
    cfg.capacity_factor = None                      # inference is dropless
    loader = C.hf_loader(args.checkpoint, lambda sd: convert_mixtral_hf_state_dict(sd, cfg))
    runner = MixtralRunner(cfg, C.inference_config(args), checkpoint_loader=loader)
    C.drive(runner, args, cfg, reference_model_fn=lambda: MixtralForCausalLM(cfg, device=runner.device))


if __name__ == "__main__":
    main()
