#!/usr/bin/env python
"""Serve a stream of requests of different lengths with continuous batching over a paged KV pool that is smaller than
``lines x pages-per-line``: the native scheduler admits requests while pages are free, finished requests return their
pages, starved ones are preempted and re-run (``inference/continuous_batching.py``, ``csrc/runtime/nxd_runtime.cpp``).

    python examples/inference/run_continuous_batching.py --model llama3.2-1b --max_running 16 --num_kv_pages 256 \
        --requests 64 --page_size 128
"""
from __future__ import annotations

import json
import random
import time

import torch

import _common as C

from neuronx_distributed_llama3_2_b200.inference import LlamaInferenceModel
from neuronx_distributed_llama3_2_b200.inference.continuous_batching import ContinuousBatcher
from neuronx_distributed_llama3_2_b200.models import PRESETS
from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps


def main():
    ap = C.common_parser(PRESETS, "llama3.2-1b")
    ap.add_argument("--max_running", type=int, default=8)
    ap.add_argument("--num_kv_pages", type=int, default=0, help="page pool size (default: half of lines x pages per line)")
    ap.add_argument("--page_size", type=int, default=128)
    ap.add_argument("--requests", type=int, default=32)
    ap.add_argument("--seed", type=int, default=0)
    args = ap.parse_args()
    C.init_distributed(args.tp_degree)
    cfg = C.model_config(args, PRESETS)
    per_line = (args.seq_len + args.page_size - 1) // args.page_size
    pages = args.num_kv_pages or max(per_line, args.max_running * per_line // 2)
    ic = C.inference_config(args, paged_kv=True, page_size=args.page_size, num_kv_pages=pages, is_continuous_batching=True)
    ic.batch_size = ic.ctx_batch_size = ic.tkg_batch_size = ic.max_batch_size = args.max_running
    device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    model = LlamaInferenceModel(cfg, ic, device=device)
    model.init_kv_cache(device=device)
    engine = ContinuousBatcher(model, max_running=args.max_running)
    rng = random.Random(args.seed)
    g = torch.Generator().manual_seed(args.seed)
    total_new = 0
    for _ in range(args.requests):
        T = rng.randint(8, max(8, min(args.max_context_length, args.seq_len // 2)))
        n = rng.randint(4, max(4, min(args.max_new_tokens, args.seq_len - T - 1)))
        total_new += n
        engine.submit(torch.randint(1, cfg.vocab_size, (T,), generator=g), n)
    t0 = time.perf_counter()
    out = engine.run()
    if device.type == "cuda":
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if ps.is_global_rank_zero():
        print(json.dumps({"requests": args.requests, "generated_tokens": sum(len(v) for v in out.values()),
                          "expected_tokens": total_new, "seconds": round(dt, 3), "kv_pages": pages,
                          "static_pages_would_be": args.max_running * per_line, **engine.stats}))


if __name__ == "__main__":
    main()
