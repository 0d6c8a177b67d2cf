#!/usr/bin/env python
"""Llama-3.x pre-training with tensor parallelism + sequence parallelism + ZeRO-1 (the B200 counterpart of the reference's
``examples/training/llama/tp_zero1_llama_hf_pretrain/tp_zero1_llama_hf_pretrain.py``).

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/training/tp_zero1_llama_pretrain.py \
        --model llama3.2-1b --tensor_parallel_size 8 --seq_len 8192 --batch_size 1 --grad_accum_usteps 4 --max_steps 100

Synthetic token ids stand in for a tokenised dataset (no network in this environment); plug a ``DataLoader`` with a
``DistributedSampler(num_replicas=dp_size, rank=dp_rank)`` in where ``SyntheticTokens`` is created.
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import neuronx_distributed_llama3_2_b200 as nxd  # noqa: E402
from neuronx_distributed_llama3_2_b200.models import PRESETS, LlamaForCausalLM, config_from_hf_json  # noqa: E402
from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps  # noqa: E402
from neuronx_distributed_llama3_2_b200.parallel.random import model_parallel_manual_seed  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.logger import get_logger  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.data import DevicePrefetcher, create_llama_pretraining_dataset  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.training_utils import (CosineAnnealing, SyntheticTokens, Throughput,  # noqa: E402
                                                                  TrainingLogger, TrainingMetrics, get_dtype,
                                                                  get_param_groups_by_weight_decay)

logger = get_logger()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3.2-1b", choices=sorted(PRESETS))
    ap.add_argument("--model_path", default=None, help="directory with an HF config.json (overrides --model)")
    ap.add_argument("--data_dir", default=None, help="tokenised dataset: HF datasets directory or flat .npy/.bin token "
                    "file (examples/training/get_dataset.py); synthetic tokens when omitted")
    ap.add_argument("--pretrained_weight", default=None, help="sharded checkpoint directory (checkpoint converter output) "
                    "to start from")
    ap.add_argument("--tensorboard", type=int, default=0)
    ap.add_argument("--tensor_parallel_size", type=int, default=1)
    ap.add_argument("--seq_len", type=int, default=4096)
    ap.add_argument("--batch_size", type=int, default=1, help="micro batch per DP rank")
    ap.add_argument("--grad_accum_usteps", type=int, default=1)
    ap.add_argument("--max_steps", type=int, default=20)
    ap.add_argument("--lr", type=float, default=3e-4)
    ap.add_argument("--warmup_steps", type=int, default=5)
    ap.add_argument("--weight_decay", type=float, default=0.1)
    ap.add_argument("--use_sequence_parallel", type=int, default=1)
    ap.add_argument("--selective_checkpoint", type=int, default=0)
    ap.add_argument("--full_activation_checkpoint", type=int, default=0)
    ap.add_argument("--use_fp32_grad_acc", type=int, default=0)
    ap.add_argument("--output_dir", default="./output")
    ap.add_argument("--checkpoint_dir", default=None)
    ap.add_argument("--checkpoint_freq", type=int, default=0)
    ap.add_argument("--num_kept_checkpoint", type=int, default=2)
    ap.add_argument("--loading_step", default="-1", help="step tag to resume from, or latest_if_exists")
    ap.add_argument("--use_xser", type=int, default=1)
    ap.add_argument("--async_checkpoint", type=int, default=1)
    ap.add_argument("--logging_interval", type=int, default=1)
    ap.add_argument("--seed", type=int, default=1234)
    args = ap.parse_args()

    world = int(os.environ.get("WORLD_SIZE", "1"))
    cuda = torch.cuda.is_available()
    if world > 1:
        if cuda:
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
        dist.init_process_group("nccl" if cuda else "gloo")
    device = torch.device("cuda", torch.cuda.current_device()) if cuda else torch.device("cpu")

    sp = bool(args.use_sequence_parallel) and args.tensor_parallel_size > 1
    nxd_config = nxd.neuronx_distributed_config(
        tensor_parallel_size=args.tensor_parallel_size, sequence_parallel=sp,
        optimizer_config={"zero_one_enabled": True, "grad_clipping": True, "max_grad_norm": 1.0},
        mixed_precision_config={"use_master_weights": True, "use_fp32_grad_acc": bool(args.use_fp32_grad_acc),
                                "use_master_weights_in_ckpt": False},
        activation_checkpoint_config="full" if args.full_activation_checkpoint else None)
    model_parallel_manual_seed(args.seed)
    kw = dict(sequence_parallel_enabled=sp, selective_checkpoint_enabled=bool(args.selective_checkpoint))
    cfg = config_from_hf_json(args.model_path, **kw) if args.model_path else PRESETS[args.model](**kw)
    if not cuda:
        cfg.torch_dtype = torch.float32
    model = nxd.initialize_parallel_model(nxd_config, lambda: LlamaForCausalLM(cfg, device=device), device=device)
    model.train()
    if args.pretrained_weight:
        w = os.path.abspath(args.pretrained_weight)      # <dir>/<tag> as written by the checkpoint converter
        nxd.load_checkpoint(os.path.dirname(w), tag=os.path.basename(w), model=model, strict=False)
    optimizer = nxd.initialize_parallel_optimizer(nxd_config, torch.optim.AdamW,
                                                  get_param_groups_by_weight_decay(model, args.weight_decay), lr=args.lr,
                                                  betas=(0.9, 0.95))
    scheduler = CosineAnnealing(optimizer, max_steps=args.max_steps, warmup_steps=args.warmup_steps, min_lr=args.lr * 0.1)

    step0 = 0
    if args.checkpoint_dir and args.loading_step != "-1":
        tag = None if args.loading_step == "latest_if_exists" else f"step_{args.loading_step}"
        if tag is not None or nxd.has_checkpoint(args.checkpoint_dir):
            user = nxd.load_checkpoint(args.checkpoint_dir, tag=tag, model=model, optimizer=optimizer, scheduler=scheduler)
            step0 = int((user or {}).get("total_steps", 0))
            logger.info("resumed from step %d", step0)

    dp = ps.get_data_parallel_size() if ps.model_parallel_is_initialized() else 1
    dp_rank = ps.get_data_parallel_rank() if ps.model_parallel_is_initialized() else 0
    if args.data_dir:
        loader, _ = create_llama_pretraining_dataset(args.data_dir, args.batch_size, dp, dp_rank, args.seed,
                                                     seq_len=args.seq_len)
        data = iter(DevicePrefetcher(loader, device))
    else:
        data = SyntheticTokens(cfg.vocab_size, args.batch_size, args.seq_len, device, seed=args.seed + dp_rank)
    tb = TrainingLogger(args, world, get_dtype(model)) if (args.tensorboard and ps.is_global_rank_zero()) else None
    throughput = Throughput(args.batch_size, dp, args.grad_accum_usteps, logging_interval=args.logging_interval)
    metrics = TrainingMetrics(os.path.join(args.output_dir, "metrics.json"))
    os.makedirs(args.output_dir, exist_ok=True)
    tps_hist = []
    t_start = time.time()
    for step in range(step0, args.max_steps):
        optimizer.zero_grad()
        running = 0.0
        for _ in range(args.grad_accum_usteps):
            batch = next(data)
            ids, labels = (batch["input_ids"], batch["labels"]) if isinstance(batch, dict) else (batch, batch)
            loss = model(ids, labels=labels) / args.grad_accum_usteps
            loss.backward()
            running += float(loss.detach())
        optimizer.step()
        scheduler.step()
        if (step + 1) % args.logging_interval == 0:
            seqs = throughput.get_throughput()
            tps_hist.append(seqs)
            if ps.is_global_rank_zero():
                logger.info("step %d loss %.4f grad-norm %.3f lr %.3e throughput %.2f seq/s (%.0f tokens/s)", step + 1,
                            running, float(optimizer.grad_norm), scheduler.get_last_lr()[0], seqs, seqs * args.seq_len)
                if tb is not None:
                    tb.log(0, step + 1, running, scheduler.get_last_lr()[0], seqs, float(optimizer.grad_norm))
        if args.checkpoint_dir and args.checkpoint_freq and (step + 1) % args.checkpoint_freq == 0:
            nxd.save_checkpoint(args.checkpoint_dir, f"step_{step + 1}", model=model, optimizer=optimizer,
                                scheduler=scheduler, user_content={"total_steps": step + 1},
                                use_xser=bool(args.use_xser), num_kept_ckpts=args.num_kept_checkpoint,
                                async_save=bool(args.async_checkpoint))
    nxd.finalize_checkpoint()
    if tb is not None:
        tb.close()
    if ps.is_global_rank_zero():
        skip = max(1, 10 // args.logging_interval)
        steady = tps_hist[skip:] or tps_hist
        metrics.store_parameters(vars(args))
        metrics.store_metrics([
            metrics.metric("Final loss", running, ""),
            metrics.metric("Time to train", (time.time() - t_start) / 60.0, "minutes"),
            metrics.metric("Average throughput", sum(steady) / len(steady), "seq/s"),
            metrics.metric("Peak throughput", max(tps_hist), "seq/s"),
        ])
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
