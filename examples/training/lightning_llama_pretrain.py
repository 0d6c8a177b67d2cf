#!/usr/bin/env python
"""Llama pre-training through the Lightning-protocol integration (``NeuronLTModule`` + ``NeuronXLAStrategy``), the
counterpart of ``examples/training/llama/lightning/run_llama_nxd_ptl.py``.  PyTorch Lightning itself is not installed
in this image; ``NxDLightningLoop`` drives the same module / strategy / checkpoint-IO / logger objects that a
``lightning.Trainer`` would.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/training/lightning_llama_pretrain.py \
        --model llama3.2-1b --tensor_parallel_size 8 --seq_len 4096 --max_steps 50
"""
from __future__ import annotations

import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import neuronx_distributed_llama3_2_b200 as nxd  # noqa: E402
from neuronx_distributed_llama3_2_b200.lightning import (NeuronLTModule, NeuronTensorBoardLogger, NeuronTQDMProgressBar,  # noqa: E402
                                                         NeuronXLAStrategy, NxDLightningLoop)
from neuronx_distributed_llama3_2_b200.models import PRESETS, LlamaForCausalLM  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.training_utils import CosineAnnealing, SyntheticTokens  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3.2-1b", choices=sorted(PRESETS) + ["mixtral-8x7b", "tiny-mixtral"],
                    help="a Llama preset, or a Mixtral one (the reference's mixtral_pretrain example is its Lightning driver)")
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--tensor_parallel_size", type=int, default=1)
    ap.add_argument("--expert_parallel_size", type=int, default=1)
    ap.add_argument("--seq_len", type=int, default=4096)
    ap.add_argument("--batch_size", type=int, default=1)
    ap.add_argument("--grad_accum_steps", type=int, default=1)
    ap.add_argument("--max_steps", type=int, default=10)
    ap.add_argument("--lr", type=float, default=3e-4)
    ap.add_argument("--output_dir", default="./lightning_out")
    ap.add_argument("--checkpoint_every", type=int, default=0)
    args = ap.parse_args()
    cuda = torch.cuda.is_available()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        if cuda:
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
        dist.init_process_group("nccl" if cuda else "gloo")
    device = torch.device("cuda", torch.cuda.current_device()) if cuda else torch.device("cpu")
    sp = args.tensor_parallel_size > 1
    nxd_config = nxd.neuronx_distributed_config(
        tensor_parallel_size=args.tensor_parallel_size, expert_parallel_size=args.expert_parallel_size,
        sequence_parallel=sp, optimizer_config={"zero_one_enabled": True, "grad_clipping": True, "max_grad_norm": 1.0})
    if "mixtral" in args.model:
        from neuronx_distributed_llama3_2_b200.models import MixtralForCausalLM, mixtral_8x7b, tiny_mixtral

        cfg = (tiny_mixtral if args.model.startswith("tiny") else mixtral_8x7b)(sequence_parallel_enabled=sp)
        model_cls = MixtralForCausalLM
    else:
        cfg, model_cls = PRESETS[args.model](sequence_parallel_enabled=sp), LlamaForCausalLM
    if args.layers:
        cfg.num_hidden_layers = args.layers
    if not cuda:
        cfg.torch_dtype = torch.float32
    module = NeuronLTModule(nxd_config, torch.optim.AdamW, CosineAnnealing,
                            model_fn=lambda **kw: model_cls(cfg, device=device),
                            opt_kwargs={"lr": args.lr, "betas": (0.9, 0.95), "weight_decay": 0.1},
                            scheduler_kwargs={"max_steps": args.max_steps, "warmup_steps": max(1, args.max_steps // 10)},
                            grad_accum_steps=args.grad_accum_steps, device=device)
    strategy = NeuronXLAStrategy(nxd_config=nxd_config, save_load_xser=True)
    loop = NxDLightningLoop(strategy, logger=NeuronTensorBoardLogger(args.output_dir), progress_bar=NeuronTQDMProgressBar(),
                            max_steps=args.max_steps, checkpoint_dir=os.path.join(args.output_dir, "checkpoints"),
                            checkpoint_every=args.checkpoint_every)
    loop.fit(module, SyntheticTokens(cfg.vocab_size, args.batch_size, args.seq_len, device))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
