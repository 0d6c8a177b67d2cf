#!/usr/bin/env python
"""Draft / target speculative decoding -- counterpart of the reference's ``examples/inference/run_llama_speculative.py``:
a small draft model proposes ``--speculation_length`` tokens with its token-generation graph, the target verifies them
in one call of its ``speculation_model`` graph; the output is exactly the target's greedy continuation.

    python examples/inference/run_llama_speculative.py --model llama3.2-3b --draft_model llama3.2-1b --speculation_length 4
"""
from __future__ import annotations

import json
import time

import torch
import torch.distributed as dist

import _common as C
from run_llama import llama_hf_converter

from neuronx_distributed_llama3_2_b200.inference import LlamaRunner
from neuronx_distributed_llama3_2_b200.models import PRESETS
from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps


def main():
    ap = C.common_parser(PRESETS, "llama3.2-3b")
    ap.add_argument("--draft_model", default="llama3.2-1b", choices=sorted(PRESETS))
    ap.add_argument("--draft_checkpoint", default=None)
    ap.add_argument("--draft_layers", type=int, default=0)
    ap.add_argument("--speculation_length", type=int, default=4)
    args = ap.parse_args()
    assert args.batch_size == 1, "speculative decoding runs at batch 1 (as in the reference)"
    C.init_distributed(args.tp_degree)
    cfg = C.model_config(args, PRESETS)
    dcfg = PRESETS[args.draft_model](sequence_parallel_enabled=False, torch_dtype=cfg.torch_dtype)
    if args.draft_layers:
        dcfg.num_hidden_layers = args.draft_layers
    assert dcfg.vocab_size == cfg.vocab_size, "draft and target must share the vocabulary"
    target = LlamaRunner(cfg, C.inference_config(args, speculation_length=args.speculation_length),
                         checkpoint_loader=C.hf_loader(args.checkpoint, llama_hf_converter(cfg)))
    draft = LlamaRunner(dcfg, C.inference_config(args), checkpoint_loader=C.hf_loader(args.draft_checkpoint,
                                                                                       llama_hf_converter(dcfg)))
    target.trace(args.traced_model_path)
    draft.trace()
    prompt = torch.randint(1, cfg.vocab_size, (1, min(64, args.max_context_length)), device=target.device)
    n = min(64, args.max_new_tokens)
    ref = target.generate(prompt, max_new_tokens=n)
    if target.device.type == "cuda":
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = target.generate_speculative(draft, prompt, max_new_tokens=n)
    if target.device.type == "cuda":
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if ps.is_global_rank_zero():
        acc = target.last_acceptance
        print(json.dumps({"matches_target_greedy": bool(torch.equal(out, ref)), "new_tokens": n, "target_calls": len(acc),
                          "mean_accepted_per_call": sum(acc) / max(len(acc), 1), "seconds": round(dt, 4)}))
    target.close()
    draft.close()
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
