"""What the family training drivers share: common flags, process-group bring-up, data source, and the step loop with
gradient accumulation, LR schedule, logging, TensorBoard and checkpoint / resume (the reference repeats this in every
example script; see ``examples/training/llama/tp_zero1_llama_hf_pretrain/tp_zero1_llama_hf_pretrain.py:150-520``)."""
from __future__ import annotations

import argparse
import os
import sys
import time
from typing import Callable, Iterator, Optional

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import neuronx_distributed_llama3_2_b200 as nxd  # noqa: E402
from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.data import DevicePrefetcher, create_llama_pretraining_dataset  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.logger import get_logger  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.training_utils import (SyntheticTokens, Throughput, TrainingLogger,  # noqa: E402
                                                                  TrainingMetrics, get_dtype,
                                                                  get_learning_rate_scheduler,
                                                                  get_param_groups_by_weight_decay)

logger = get_logger()


def add_common_args(ap: argparse.ArgumentParser) -> argparse.ArgumentParser:
    ap.add_argument("--model_path", default=None, help="directory with an HF-style config.json (else the family preset)")
    ap.add_argument("--data_dir", default=None, help="tokenised dataset (get_dataset.py); synthetic tokens when omitted")
    ap.add_argument("--pretrained_weight", default=None, help="<dir>/<tag> written by convert_checkpoints.py")
    ap.add_argument("--tensor_parallel_size", type=int, default=1)
    ap.add_argument("--pipeline_parallel_size", type=int, default=1)
    ap.add_argument("--num_microbatches", type=int, default=4, help="pipeline parallelism: micro-batches per step "
                    "(--batch_size is then the batch of one step on a data-parallel rank)")
    ap.add_argument("--use_sequence_parallel", type=int, default=1)
    ap.add_argument("--layers", type=int, default=0, help="override the layer count (smoke runs)")
    ap.add_argument("--seq_len", type=int, default=2048)
    ap.add_argument("--batch_size", type=int, default=1, help="micro batch per data-parallel rank")
    ap.add_argument("--grad_accum_usteps", type=int, default=1)
    ap.add_argument("--max_steps", type=int, default=10)
    ap.add_argument("--lr", type=float, default=1e-4)
    ap.add_argument("--min_lr", type=float, default=0.0)
    ap.add_argument("--warmup_steps", type=int, default=0)
    ap.add_argument("--constant_steps", type=int, default=0)
    ap.add_argument("--weight_decay", type=float, default=0.01)
    ap.add_argument("--use_zero_1", type=int, default=1)
    ap.add_argument("--full_activation_checkpoint", type=int, default=0)
    ap.add_argument("--checkpoint_dir", default=None)
    ap.add_argument("--checkpoint_freq", type=int, default=0)
    ap.add_argument("--num_kept_checkpoint", type=int, default=2)
    ap.add_argument("--loading_step", default="-1", help="step tag to resume from, or latest_if_exists")
    ap.add_argument("--use_xser", type=int, default=1)
    ap.add_argument("--output_dir", default="./output")
    ap.add_argument("--tensorboard", type=int, default=0)
    ap.add_argument("--logging_interval", type=int, default=1)
    ap.add_argument("--seed", type=int, default=1234)
    return ap


def pipeline_kwargs(args, layer_cls, input_names) -> dict:
    """``pipeline_parallel_size`` / ``pipeline_config`` for ``neuronx_distributed_config`` (empty without PP)."""
    if args.pipeline_parallel_size <= 1:
        return {}
    return {"pipeline_parallel_size": args.pipeline_parallel_size,
            "pipeline_config": {"transformer_layer_cls": layer_cls, "num_microbatches": args.num_microbatches,
                                "auto_partition": True, "input_names": list(input_names),
                                "broadcast_and_average_loss": True, "use_zero1_optimizer": bool(args.use_zero_1)}}


def init_process_group() -> torch.device:
    cuda = torch.cuda.is_available()
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        if cuda:
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
        dist.init_process_group("nccl" if cuda else "gloo")
    return torch.device("cuda", torch.cuda.current_device()) if cuda else torch.device("cpu")


def dp_coords():
    if ps.model_parallel_is_initialized():
        return ps.get_data_parallel_size(), ps.get_data_parallel_rank()
    return 1, 0


def token_batches(args, vocab_size: int, device) -> Iterator:
    """Dict batches ``{"input_ids", "labels"}`` on ``device``, from ``--data_dir`` or synthetic."""
    dp, dp_rank = dp_coords()
    if args.data_dir:
        loader, _ = create_llama_pretraining_dataset(args.data_dir, args.batch_size, dp, dp_rank, args.seed,
                                                     seq_len=args.seq_len)
        return iter(DevicePrefetcher(loader, device))
    synth = SyntheticTokens(vocab_size, args.batch_size, args.seq_len, device, seed=args.seed + dp_rank)
    return ({"input_ids": ids, "labels": ids} for ids in synth)


def fit(args, nxd_config, model, batches: Iterator, step_fn: Optional[Callable] = None, teardown: bool = True) -> float:
    """Optimizer + schedule + (resume) + step loop.  ``step_fn(model, batch) -> loss`` runs forward AND backward of one
    micro batch, scaling the loss by ``1 / grad_accum_usteps`` before ``backward`` (default: ``model(**batch)``; pipeline
    models go through ``run_train``)."""
    if args.pretrained_weight:
        w = os.path.abspath(args.pretrained_weight)
        nxd.load_checkpoint(os.path.dirname(w), tag=os.path.basename(w), model=model, strict=False)
    optimizer = nxd.initialize_parallel_optimizer(nxd_config, torch.optim.AdamW,
                                                  get_param_groups_by_weight_decay(model, args.weight_decay), lr=args.lr,
                                                  betas=(0.9, 0.95))
    scheduler = get_learning_rate_scheduler(optimizer, args)
    step0 = 0
    if args.checkpoint_dir and args.loading_step != "-1":
        tag = None if args.loading_step == "latest_if_exists" else f"step_{args.loading_step}"
        if tag is not None or nxd.has_checkpoint(args.checkpoint_dir):
            user = nxd.load_checkpoint(args.checkpoint_dir, tag=tag, model=model, optimizer=optimizer, scheduler=scheduler)
            step0 = int((user or {}).get("total_steps", 0))
            logger.info("resumed from step %d", step0)
    dp, _ = dp_coords()
    world = dist.get_world_size() if dist.is_initialized() else 1
    meter = Throughput(args.batch_size, dp, args.grad_accum_usteps, logging_interval=args.logging_interval)
    os.makedirs(args.output_dir, exist_ok=True)
    tb = TrainingLogger(args, world, get_dtype(model)) if (args.tensorboard and ps.is_global_rank_zero()) else None
    accum = args.grad_accum_usteps
    pipelined = nxd_config["pipeline_parallel_size"] > 1
    assert not (pipelined and accum > 1), "with pipeline parallelism the micro-batches ARE the accumulation steps"

    def default_step(m, b):
        if pipelined:
            return m.run_train(**b)
        loss = m(**b)
        (loss / accum).backward()
        return loss

    step_fn = step_fn or default_step
    model.train()
    hist, running, t0 = [], float("nan"), time.time()
    for step in range(step0, args.max_steps):
        optimizer.zero_grad()
        running = 0.0
        for _ in range(args.grad_accum_usteps):
            loss = step_fn(model, next(batches))
            running += float(loss.detach()) / args.grad_accum_usteps
        optimizer.step()
        scheduler.step()
        if (step + 1) % args.logging_interval == 0:
            seqs = meter.get_throughput()
            hist.append(seqs)
            if ps.is_global_rank_zero():
                logger.info("step %d loss %.4f grad-norm %.3f lr %.2e throughput %.2f seq/s", step + 1, running,
                            float(optimizer.grad_norm), scheduler.get_last_lr()[0], seqs)
            if tb is not None:
                tb.log(0, step + 1, running, scheduler.get_last_lr()[0], seqs, float(optimizer.grad_norm))
        if args.checkpoint_dir and args.checkpoint_freq and (step + 1) % args.checkpoint_freq == 0:
            nxd.save_checkpoint(args.checkpoint_dir, f"step_{step + 1}", model=model, optimizer=optimizer,
                                scheduler=scheduler, user_content={"total_steps": step + 1},
                                use_xser=bool(args.use_xser), num_kept_ckpts=args.num_kept_checkpoint)
    nxd.finalize_checkpoint()
    if tb is not None:
        tb.close()
    if ps.is_global_rank_zero() and hist:
        m = TrainingMetrics(os.path.join(args.output_dir, "metrics.json"))
        steady = hist[max(1, 10 // args.logging_interval):] or hist
        m.store_parameters({k: (v if isinstance(v, (int, float, str, bool, type(None))) else str(v))
                            for k, v in vars(args).items()})
        m.store_metrics([m.metric("Final loss", running, ""), m.metric("Time to train", (time.time() - t0) / 60.0, "minutes"),
                         m.metric("Average throughput", sum(steady) / len(steady), "seq/s"),
                         m.metric("Peak throughput", max(hist), "seq/s")])
    if teardown and dist.is_initialized():      # teardown=False: the caller still needs the groups (e.g. evaluation)
        dist.destroy_process_group()
    return running
