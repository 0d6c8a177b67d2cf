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
from neuronx_distributed_llama3_2_b200.models import PRESETS, LlamaForCausalLM  # noqa: E402
from neuronx_distributed_llama3_2_b200.models.llama import LlamaDecoderLayer  # noqa: E402
from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps  # noqa: E402
from neuronx_distributed_llama3_2_b200.parallel.random import model_parallel_manual_seed  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.logger import get_logger  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.training_utils import SyntheticTokens, Throughput  # noqa: E402

logger = get_logger()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3.2-3b", choices=sorted(PRESETS))
    ap.add_argument("--tensor_parallel_size", type=int, default=1)
    ap.add_argument("--pipeline_parallel_size", type=int, default=2)
    ap.add_argument("--virtual_pipeline_size", type=int, default=1)
    ap.add_argument("--num_microbatches", type=int, default=4)
    ap.add_argument("--micro_batch_size", type=int, default=1)
    ap.add_argument("--seq_len", type=int, default=4096)
    ap.add_argument("--max_steps", type=int, default=10)
    ap.add_argument("--lr", type=float, default=3e-4)
    ap.add_argument("--use_zero1_optimizer", type=int, default=1)
    ap.add_argument("--use_sequence_parallel", type=int, default=1)
    ap.add_argument("--activation_checkpointing", type=int, default=1)
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
        activation_checkpoint_config="full" if args.activation_checkpointing else None)
    model_parallel_manual_seed(1234)
    cfg = PRESETS[args.model](sequence_parallel_enabled=sp)
    if args.layers:
        cfg.num_hidden_layers = args.layers
    if not cuda:
        cfg.torch_dtype = torch.float32
    model = nxd.initialize_parallel_model(nxd_config, lambda: LlamaForCausalLM(cfg, device=device), device=device)
    optimizer = nxd.initialize_parallel_optimizer(nxd_config, torch.optim.AdamW, model.parameters(), lr=args.lr,
                                                  betas=(0.9, 0.95), weight_decay=0.1)
    batch = args.micro_batch_size * args.num_microbatches
    data = SyntheticTokens(cfg.vocab_size, batch, args.seq_len, "cpu", seed=ps.get_data_parallel_rank())
    meter = Throughput(batch, ps.get_data_parallel_size(), 1)
    for step in range(args.max_steps):
        ids = next(data)
        optimizer.zero_grad()
        loss = model.run_train(input_ids=ids, labels=ids)
        optimizer.step()
        seqs = meter.get_throughput()
        if ps.is_global_rank_zero():
            logger.info("step %d loss %.4f grad-norm %.3f throughput %.2f seq/s", step + 1, float(loss),
                        float(optimizer.grad_norm), seqs)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
