#!/usr/bin/env python
"""DBRX (16 fine-grained experts, top-4, clamped QKV, LayerNorm) inference -- counterpart of the reference's
``examples/inference/run_dbrx.py``.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/inference/run_dbrx.py --model dbrx-instruct \
        --tp_degree 8 --batch_size 1 --max_context_length 1024 --seq_len 2048 --benchmark
"""
from __future__ import annotations

import _common as C

from neuronx_distributed_llama3_2_b200.inference import DbrxRunner
from neuronx_distributed_llama3_2_b200.models import DbrxForCausalLM, convert_dbrx_hf_state_dict, dbrx_instruct, tiny_dbrx

PRESETS = {"dbrx-instruct": dbrx_instruct, "tiny": tiny_dbrx}


def main():
    args = C.common_parser(PRESETS, "dbrx-instruct").parse_args()
    C.init_distributed(args.tp_degree)
    cfg = C.model_config(args, PRESETS)
    cfg.capacity_factor = None
    loader = C.hf_loader(args.checkpoint, lambda sd: convert_dbrx_hf_state_dict(sd, cfg))
    runner = DbrxRunner(cfg, C.inference_config(args), checkpoint_loader=loader)
    C.drive(runner, args, cfg, reference_model_fn=lambda: DbrxForCausalLM(cfg, device=runner.device))


if __name__ == "__main__":
    main()
