#!/usr/bin/env python
"""Llama-3.2 inference: trace (bucketed context encoding + token generation CUDA graphs), generate, benchmark -- the
counterpart of ``examples/inference/run_llama.py`` / ``llama3_2_inference.ipynb`` of the reference.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/inference/run_llama.py --model llama3.2-1b \
        --tp_degree 8 --batch_size 4 --max_context_length 2048 --seq_len 4096 --max_new_tokens 256 --benchmark
"""
from __future__ import annotations

import _common as C

from neuronx_distributed_llama3_2_b200.inference import LlamaRunner
from neuronx_distributed_llama3_2_b200.models import PRESETS, LlamaForCausalLM
from neuronx_distributed_llama3_2_b200.scripts.checkpoint_converter import CheckpointConverterBase


def llama_hf_converter(cfg):
    conv = CheckpointConverterBase(cfg.num_hidden_layers, cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim,
                                   tie_word_embeddings=cfg.tie_word_embeddings)
    return conv.hf_to_nxd_full


def main():
    args = C.common_parser(PRESETS, "llama3.2-1b").parse_args()
    C.init_distributed(args.tp_degree)
    cfg = C.model_config(args, PRESETS)
    runner = LlamaRunner(cfg, C.inference_config(args), checkpoint_loader=C.hf_loader(args.checkpoint, llama_hf_converter(cfg)))
    C.drive(runner, args, cfg, reference_model_fn=lambda: LlamaForCausalLM(cfg, device=runner.device))


if __name__ == "__main__":
    main()
