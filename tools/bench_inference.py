#!/usr/bin/env python
"""Llama-3.2-1B/3B inference benchmark (BASELINE.json metric part 2: decode tokens/s + p50 TTFT; config #4).
    python tools/bench_inference.py --model llama3.2-1b --batch 1 4 16 64 --prompt 2048 --new 128
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/bench_inference.py ...
--impl baseline = the same layers with cuBLAS GEMMs, SDPA prefill attention and NCCL all-reduce (NXD_* switches).
Prompts are synthetic token ids, weights random-init (no network).  Latencies are CUDA-event timed per graph launch."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3.2-1b")
    ap.add_argument("--impl", default="ours", choices=["ours", "baseline"])
    ap.add_argument("--batch", type=int, nargs="+", default=[1, 4, 16, 64])
    ap.add_argument("--prompt", type=int, default=2048)
    ap.add_argument("--seq-len", type=int, default=4096)
    ap.add_argument("--new", type=int, default=128)
    ap.add_argument("--runs", type=int, default=5)
    ap.add_argument("--no-graphs", action="store_true")
    ap.add_argument("--profile", default="", help="diagnosis: write a kernel summary of one generate() to this file")
    args = ap.parse_args()
    if args.impl == "baseline":
        os.environ["NXD_GEMM_BACKEND"] = "cublas"
        os.environ["NXD_ATTN_BACKEND"] = "sdpa"
        os.environ["NXD_FUSED_COMM"] = "0"
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dev = torch.device("cuda", torch.cuda.current_device())
    from neuronx_distributed_llama3_2_b200.inference import InferenceConfig, InferenceRunner
    from neuronx_distributed_llama3_2_b200.models import PRESETS
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        ps.initialize_model_parallel(tensor_model_parallel_size=world)
    for B in args.batch:
        cfg = PRESETS[args.model](sequence_parallel_enabled=False)
        ic = InferenceConfig(tp_degree=world, batch_size=B, max_context_length=args.prompt, seq_len=args.seq_len,
                             buckets=[args.prompt], max_new_tokens=args.new, use_cuda_graphs=not args.no_graphs)
        runner = InferenceRunner(cfg, ic, device=dev)
        runner.trace()
        if args.profile:
            from torch.profiler import ProfilerActivity, profile

            ids = torch.randint(1, cfg.vocab_size, (B, args.prompt), device=dev)
            runner.generate(ids, max_new_tokens=args.new)
            torch.cuda.synchronize()
            with profile(activities=[ProfilerActivity.CUDA]) as prof:
                runner.generate(ids, max_new_tokens=args.new)
                torch.cuda.synchronize()
            if rank == 0:
                by = {}
                for e in prof.events():
                    if e.device_type == torch.autograd.DeviceType.CUDA:
                        d = by.setdefault(e.name[:100], [0.0, 0])
                        d[0] += e.time_range.end - e.time_range.start
                        d[1] += 1
                with open(args.profile, "a") as f:
                    f.write(f"# {args.model} tp{world} batch {B} prompt {args.prompt} new {args.new} impl {args.impl} (us, CUPTI)\n")
                    for n, (t, c) in sorted(by.items(), key=lambda kv: -kv[1][0])[:40]:
                        f.write(f"{t:10.0f} us {c:6d} x {t / c:8.2f}  {n}\n")
            if world > 1:
                dist.barrier()
            os._exit(0)          # CUPTI + communicator teardown can hang; the summary is on disk
        rep = runner.benchmark_sampling(prompt_len=args.prompt, new_tokens=args.new, n_runs=args.runs)
        tkg = rep["token_generation_model"]
        out = {"metric": f"{args.model} decode tokens/s + p50 TTFT", "impl": args.impl, "n_gpus": world, "batch": B,
               "prompt_len": args.prompt, "new_tokens": args.new, "ttft_ms_p50": rep["ttft_ms_p50"],
               "decode_tokens_per_s": rep["decode_tokens_per_s"], "decode_ms_per_token_p50": tkg["latency_ms_p50"],
               "decode_ms_per_token_p99": tkg["latency_ms_p99"], "e2e_ms_p50": rep["e2e_model"]["latency_ms_p50"],
               "cuda_graphs": not args.no_graphs, "dtype": "bf16", "data": "synthetic prompts, random-init weights"}
        if world > 1:
            t = torch.tensor([out["ttft_ms_p50"], out["decode_ms_per_token_p50"]], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            out["ttft_ms_p50"], out["decode_ms_per_token_p50"] = float(t[0]), float(t[1])
            out["decode_tokens_per_s"] = B * 1e3 / float(t[1])
        if rank == 0:
            print(json.dumps(out), flush=True)
        runner.close()
        del runner
        torch.cuda.empty_cache()
    if world > 1:
        ps.destroy_model_parallel()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
