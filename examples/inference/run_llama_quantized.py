#!/usr/bin/env python
"""Llama inference with int8 weight-only decoder linears -- counterpart of the reference's
``examples/inference/run_llama_quantized.py``: the float checkpoint is quantized on load (per tensor or per channel,
symmetric), ``o_proj`` / ``gate_up_proj`` / ``down_proj`` run as ``Quantized{Column,Row}Parallel``.

    python examples/inference/run_llama_quantized.py --model llama3.2-1b --quantization_type per_channel_symmetric \
        --save_quantized /tmp/llama_q.pt
"""
from __future__ import annotations

import torch

import _common as C
from run_llama import llama_hf_converter

from neuronx_distributed_llama3_2_b200.inference import LlamaRunner
from neuronx_distributed_llama3_2_b200.models import PRESETS, LlamaForCausalLM
from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps


def main():
    ap = C.common_parser(PRESETS, "llama3.2-1b")
    ap.add_argument("--quantization_type", default="per_tensor_symmetric",
                    choices=["per_tensor_symmetric", "per_channel_symmetric"])
    ap.add_argument("--save_quantized", default=None, help="also write the quantized full state dict here (torch.save)")
    args = ap.parse_args()
    C.init_distributed(args.tp_degree)
    cfg = C.model_config(args, PRESETS)
    loader = C.hf_loader(args.checkpoint, llama_hf_converter(cfg))
    if loader is None:                              # random weights still go through the quantizer
        def loader():
            with torch.device("cpu"):
                return {k: v for k, v in LlamaForCausalLM(cfg, device="cpu").state_dict().items()}
    ic = C.inference_config(args, quantized=True, quantization_type=args.quantization_type)
    runner = LlamaRunner(cfg, ic, checkpoint_loader=loader)
    if args.save_quantized and ps.is_global_rank_zero():
        torch.save(runner.generate_quantized_state_dict(loader()), args.save_quantized)
    C.drive(runner, args, cfg)


if __name__ == "__main__":
    main()
