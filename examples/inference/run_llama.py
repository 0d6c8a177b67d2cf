#!/usr/bin/env python
"""Llama-3.2 inference: trace (bucketed context encoding + token generation CUDA graphs), generate, benchmark -- the
counterpart of ``examples/inference/llama3_2_inference.ipynb`` / ``runner.py`` of the reference.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/inference/run_llama.py --model llama3.2-1b \
        --tp_degree 8 --batch_size 4 --max_context_length 2048 --seq_len 4096 --max_new_tokens 256 --benchmark
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from neuronx_distributed_llama3_2_b200.inference import InferenceConfig, InferenceRunner  # noqa: E402
from neuronx_distributed_llama3_2_b200.models import PRESETS  # noqa: E402
from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3.2-1b", choices=sorted(PRESETS))
    ap.add_argument("--tp_degree", type=int, default=1)
    ap.add_argument("--batch_size", type=int, default=1)
    ap.add_argument("--max_context_length", type=int, default=2048)
    ap.add_argument("--seq_len", type=int, default=4096)
    ap.add_argument("--max_new_tokens", type=int, default=128)
    ap.add_argument("--top_k", type=int, default=1)
    ap.add_argument("--checkpoint", default=None, help="HF safetensors dir (random weights when omitted)")
    ap.add_argument("--traced_model_path", default=None)
    ap.add_argument("--benchmark", action="store_true")
    ap.add_argument("--layers", type=int, default=0)
    args = ap.parse_args()
    cuda = torch.cuda.is_available()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        if cuda:
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
        dist.init_process_group("nccl" if cuda else "gloo")
        ps.initialize_model_parallel(tensor_model_parallel_size=args.tp_degree)
    cfg = PRESETS[args.model](sequence_parallel_enabled=False)
    if args.layers:
        cfg.num_hidden_layers = args.layers
    if not cuda:
        cfg.torch_dtype = torch.float32
    ic = InferenceConfig(tp_degree=args.tp_degree, batch_size=args.batch_size, max_context_length=args.max_context_length,
                         seq_len=args.seq_len, max_new_tokens=args.max_new_tokens, top_k=args.top_k,
                         do_sample=args.top_k > 1)
    loader = None
    if args.checkpoint:
        from safetensors.torch import load_file

        from neuronx_distributed_llama3_2_b200.scripts.checkpoint_converter import CheckpointConverterBase

        def loader():
            sd = {}
            for f in sorted(os.listdir(args.checkpoint)):
                if f.endswith(".safetensors"):
                    sd.update(load_file(os.path.join(args.checkpoint, f)))
            conv = CheckpointConverterBase(cfg.num_hidden_layers, cfg.num_attention_heads, cfg.num_key_value_heads,
                                           cfg.head_dim, tie_word_embeddings=cfg.tie_word_embeddings)
            return conv.hf_to_nxd_full(sd)

    runner = InferenceRunner(cfg, ic, checkpoint_loader=loader)
    runner.trace(args.traced_model_path)
    prompts = torch.randint(1, cfg.vocab_size, (args.batch_size, min(64, args.max_context_length)), device=runner.device)
    out = runner.generate(prompts, max_new_tokens=min(16, args.max_new_tokens))
    if ps.is_global_rank_zero():
        print("generated token ids:", out[0, -16:].tolist())
    if args.benchmark:
        rep = runner.benchmark_sampling()
        if ps.is_global_rank_zero():
            print(json.dumps(rep, indent=1))
    runner.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
