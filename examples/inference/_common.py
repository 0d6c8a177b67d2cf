"""Shared driver of the inference examples (the reference keeps this in ``examples/inference/runner.py`` +
``run_*.py``): parse the common flags, bring up the process group, build the runner, generate, optionally check
accuracy / benchmark / profile."""
from __future__ import annotations

import argparse
import json
import os
import sys
from typing import Callable, Dict, Optional

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from neuronx_distributed_llama3_2_b200.inference import InferenceConfig  # noqa: E402
from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps  # noqa: E402


def common_parser(presets: Dict[str, Callable], default_model: str) -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default=default_model, choices=sorted(presets))
    ap.add_argument("--tp_degree", type=int, default=1)
    ap.add_argument("--batch_size", type=int, default=1)
    ap.add_argument("--max_context_length", type=int, default=2048)
    ap.add_argument("--seq_len", type=int, default=4096)
    ap.add_argument("--max_new_tokens", type=int, default=128)
    ap.add_argument("--top_k", type=int, default=1)
    ap.add_argument("--top_p", type=float, default=1.0)
    ap.add_argument("--temperature", type=float, default=1.0)
    ap.add_argument("--checkpoint", default=None, help="HF checkpoint directory (random weights when omitted)")
    ap.add_argument("--traced_model_path", default=None, help="write configs + per-rank sharded weights here")
    ap.add_argument("--benchmark", action="store_true")
    ap.add_argument("--check_accuracy", action="store_true", help="compare prefill logits with the training model")
    ap.add_argument("--profile", default=None, help="write a torch.profiler Chrome trace of one generate() here")
    ap.add_argument("--layers", type=int, default=0, help="truncate the model to this many layers (smoke runs)")
    ap.add_argument("--no_cuda_graphs", action="store_true")
    ap.add_argument("--paged_kv", action="store_true")
    return ap


def init_distributed(tp_degree: int) -> int:
    cuda = torch.cuda.is_available()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        if cuda:
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
        dist.init_process_group("nccl" if cuda else "gloo")
        ps.initialize_model_parallel(tensor_model_parallel_size=tp_degree)
    elif tp_degree != 1:
        raise SystemExit(f"--tp_degree {tp_degree} needs torchrun with {tp_degree} processes")
    return world


def inference_config(args, **extra) -> InferenceConfig:
    kw = dict(tp_degree=args.tp_degree, batch_size=args.batch_size, max_context_length=args.max_context_length,
              seq_len=args.seq_len, max_new_tokens=args.max_new_tokens, top_k=args.top_k, top_p=args.top_p,
              temperature=args.temperature, do_sample=args.top_k > 1, use_cuda_graphs=not args.no_cuda_graphs,
              paged_kv=args.paged_kv)
    kw.update(extra)
    return InferenceConfig(**kw)


def model_config(args, presets):
    cfg = presets[args.model](sequence_parallel_enabled=False)
    if args.layers:
        cfg.num_hidden_layers = args.layers
    if not torch.cuda.is_available():
        cfg.torch_dtype = torch.float32
    return cfg


def hf_loader(path: Optional[str], convert: Callable[[Dict[str, torch.Tensor]], Dict[str, torch.Tensor]]):
    """``checkpoint_loader`` for an HF directory (safetensors / bin, sharded or single file), or ``None``."""
    if not path:
        return None

    def load():
        from neuronx_distributed_llama3_2_b200.inference.checkpoint import load_state_dict

        return convert(load_state_dict(path))

    return load


def drive(runner, args, cfg, reference_model_fn: Optional[Callable] = None) -> None:
    """trace -> generate -> optional accuracy check / benchmark / profile -> close."""
    runner.trace(args.traced_model_path)
    T = min(64, args.max_context_length)
    prompts = torch.randint(1, cfg.vocab_size, (args.batch_size, T), device=runner.device)
    out = runner.generate(prompts, max_new_tokens=min(16, args.max_new_tokens))
    if ps.is_global_rank_zero():
        print("generated token ids:", out[0, T:].tolist())
    if args.check_accuracy and reference_model_fn is not None:
        ref = reference_model_fn()
        ref.load_state_dict(runner.model.state_dict(), strict=False)
        ok = runner.check_accuracy_logits(ref.eval(), prompts[:, :16])
        if ps.is_global_rank_zero():
            print("logit accuracy check:", "passed" if ok else "FAILED")
    if args.profile:
        runner.torch_profile(args.profile)
    if args.benchmark:
        rep = runner.benchmark_sampling()
        if ps.is_global_rank_zero():
            print(json.dumps(rep, indent=1))
    runner.close()
    if dist.is_initialized():
        dist.destroy_process_group()
