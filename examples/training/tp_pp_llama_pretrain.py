#!/usr/bin/env python
"""Llama-3.x pre-training with TP x PP (NxDPPModel 1F1B) + activation checkpointing (BASELINE.json config #3; the
counterpart of the reference's ``examples/training/llama/tp_pp_llama_hf_pretrain/run_llama_nxd.py``).

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/training/tp_pp_llama_pretrain.py \
        --model llama3.2-3b --tensor_parallel_size 4 --pipeline_parallel_size 2 --num_microbatches 8 --seq_len 8192
"""
from __future__ import annotations

import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import neuronx_distributed_llama3_2_b200 as nxd  # noqa: E402
from neuronx_distributed_llama3_2_b200.models import PRESETS, LlamaForCausalLM, config_from_hf_json  # noqa: E402
from neuronx_distributed_llama3_2_b200.models.llama import LlamaDecoderLayer  # noqa: E402
from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps  # noqa: E402
from neuronx_distributed_llama3_2_b200.parallel.random import model_parallel_manual_seed  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.logger import get_logger  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.model_utils import default_param_init_fn  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.data import create_llama_pretraining_dataset  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.training_utils import (SyntheticTokens, Throughput, TrainingLogger,  # noqa: E402
                                                                  get_learning_rate_scheduler,
                                                                  get_param_groups_by_weight_decay)

logger = get_logger()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3.2-3b", choices=sorted(PRESETS))
    ap.add_argument("--training_config", default=None, help="directory with an HF config.json (overrides --model)")
    ap.add_argument("--training_dir", default=None, help="tokenised dataset (see get_dataset.py); synthetic when omitted")
    ap.add_argument("--tensor_parallel_size", type=int, default=1)
    ap.add_argument("--pipeline_parallel_size", type=int, default=2)
    ap.add_argument("--virtual_pipeline_size", type=int, default=1)
    ap.add_argument("--num_microbatches", type=int, default=4)
    ap.add_argument("--micro_batch_size", type=int, default=1)
    ap.add_argument("--seq_len", type=int, default=4096)
    ap.add_argument("--max_steps", type=int, default=10)
    ap.add_argument("--lr", type=float, default=3e-4)
    ap.add_argument("--min_lr", type=float, default=1e-6)
    ap.add_argument("--beta1", type=float, default=0.9)
    ap.add_argument("--beta2", type=float, default=0.95)
    ap.add_argument("--weight_decay", type=float, default=0.1)
    ap.add_argument("--warmup_steps", type=int, default=2)
    ap.add_argument("--constant_steps", type=int, default=0)
    ap.add_argument("--use_zero1_optimizer", type=int, default=1)
    ap.add_argument("--use_sequence_parallel", type=int, default=1)
    ap.add_argument("--activation_checkpointing", type=int, default=1)
    ap.add_argument("--use_selective_checkpoint", type=int, default=0)
    ap.add_argument("--use_meta_device_init", type=int, default=0, help="build on the meta device, materialise only the "
                    "local pipeline stage")
    ap.add_argument("--kv_replicator", type=int, default=1, help="KV-head replication when tp > num_key_value_heads")
    ap.add_argument("--pretrained_weight", default=None, help="<dir>/<tag> written by convert_checkpoints.py")
    ap.add_argument("--checkpoint_dir", default=None)
    ap.add_argument("--checkpoint_freq", type=int, default=0)
    ap.add_argument("--num_kept_checkpoint", type=int, default=-1)
    ap.add_argument("--loading_step", default="-1", help="step tag to resume from, or latest_if_exists")
    ap.add_argument("--save_load_xser", type=int, default=1)
    ap.add_argument("--tb_dir", default=None)
    ap.add_argument("--layers", type=int, default=0, help="override the layer count (smoke runs)")
    args = ap.parse_args()

    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    dist.init_process_group("nccl" if cuda else "gloo")
    device = torch.device("cuda", torch.cuda.current_device()) if cuda else torch.device("cpu")
    sp = bool(args.use_sequence_parallel) and args.tensor_parallel_size > 1
    nxd_config = nxd.neuronx_distributed_config(
        tensor_parallel_size=args.tensor_parallel_size, pipeline_parallel_size=args.pipeline_parallel_size,
        sequence_parallel=sp,
        pipeline_config={"transformer_layer_cls": LlamaDecoderLayer, "num_microbatches": args.num_microbatches,
                         "auto_partition": True, "input_names": ["input_ids", "labels"],
                         "virtual_pipeline_size": args.virtual_pipeline_size, "broadcast_and_average_loss": True,
                         "use_zero1_optimizer": bool(args.use_zero1_optimizer)},
        optimizer_config={"zero_one_enabled": bool(args.use_zero1_optimizer), "grad_clipping": True, "max_grad_norm": 1.0},
        activation_checkpoint_config="full" if args.activation_checkpointing else None,
        model_init_config={"meta_device_init": True, "param_init_fn": default_param_init_fn}
        if args.use_meta_device_init else None)
    model_parallel_manual_seed(1234)
    kw = dict(sequence_parallel_enabled=sp, selective_checkpoint_enabled=bool(args.use_selective_checkpoint),
              kv_shared_group_size=args.kv_replicator)
    cfg = config_from_hf_json(args.training_config, **kw) if args.training_config else PRESETS[args.model](**kw)
    if args.layers:
        cfg.num_hidden_layers = args.layers
    if not cuda:
        cfg.torch_dtype = torch.float32
    model = nxd.initialize_parallel_model(nxd_config, lambda: LlamaForCausalLM(cfg, device=device), device=device)
    if args.pretrained_weight:
        w = os.path.abspath(args.pretrained_weight)
        nxd.load_checkpoint(os.path.dirname(w), tag=os.path.basename(w), model=model, strict=False)
    optimizer = nxd.initialize_parallel_optimizer(nxd_config, torch.optim.AdamW,
                                                  get_param_groups_by_weight_decay(model, args.weight_decay), lr=args.lr,
                                                  betas=(args.beta1, args.beta2))
    scheduler = get_learning_rate_scheduler(optimizer, args)
    step0 = 0
    if args.checkpoint_dir and args.loading_step != "-1":
        tag = None if args.loading_step == "latest_if_exists" else f"step_{args.loading_step}"
        if tag is not None or nxd.has_checkpoint(args.checkpoint_dir):
            user = nxd.load_checkpoint(args.checkpoint_dir, tag=tag, model=model, optimizer=optimizer, scheduler=scheduler)
            step0 = int((user or {}).get("total_steps", 0))
    batch = args.micro_batch_size * args.num_microbatches
    dp, dp_rank = ps.get_data_parallel_size(), ps.get_data_parallel_rank()
    if args.training_dir:
        loader, _ = create_llama_pretraining_dataset(args.training_dir, batch, dp, dp_rank, 1234, seq_len=args.seq_len)

        def batches():
            epoch = 0
            while True:
                loader.sampler.set_epoch(epoch)
                for b in loader:
                    yield b["input_ids"], b["labels"]
                epoch += 1
        data = batches()
    else:
        synth = SyntheticTokens(cfg.vocab_size, batch, args.seq_len, "cpu", seed=dp_rank)
        data = ((ids, ids) for ids in synth)
    # the loss lives on the last pipeline stage: its dp 0 / tp 0 rank writes TensorBoard
    last_stage = ps.get_pipeline_model_parallel_rank() == ps.get_pipeline_model_parallel_size() - 1
    tb = None
    if args.tb_dir and last_stage and dp_rank == 0 and ps.get_tensor_model_parallel_rank() == 0:
        from types import SimpleNamespace

        tb = TrainingLogger(SimpleNamespace(output_dir=args.tb_dir, lr=args.lr, batch_size=batch, grad_accum_usteps=1,
                                            warmup_steps=args.warmup_steps, max_steps=args.max_steps), dist.get_world_size(),
                            str(cfg.torch_dtype))
    meter = Throughput(batch, dp, 1)
    for step in range(step0, args.max_steps):
        ids, labels = next(data)
        optimizer.zero_grad()
        loss = model.run_train(input_ids=ids, labels=labels)
        optimizer.step()
        scheduler.step()
        seqs = meter.get_throughput()
        if ps.is_global_rank_zero():
            logger.info("step %d loss %.4f grad-norm %.3f lr %.2e throughput %.2f seq/s", step + 1, float(loss),
                        float(optimizer.grad_norm), scheduler.get_last_lr()[0], seqs)
        if tb is not None:
            tb.log(0, step + 1, float(loss), scheduler.get_last_lr()[0], seqs, float(optimizer.grad_norm))
        if args.checkpoint_dir and args.checkpoint_freq > 0 and (step + 1) % args.checkpoint_freq == 0:
            nxd.save_checkpoint(args.checkpoint_dir, f"step_{step + 1}", model=model, optimizer=optimizer,
                                scheduler=scheduler, user_content={"total_steps": step + 1},
                                use_xser=bool(args.save_load_xser), num_kept_ckpts=args.num_kept_checkpoint)
    nxd.finalize_checkpoint()
    if tb is not None:
        tb.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
