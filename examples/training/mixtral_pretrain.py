#!/usr/bin/env python
"""Mixtral-style MoE pre-training with TP + SP + expert parallelism + EP-aware ZeRO-1 (BASELINE.json config #5; the
counterpart of ``examples/training/mixtral/``).

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/training/mixtral_pretrain.py \
        --tensor_parallel_size 2 --expert_parallel_size 4 --layers 4 --seq_len 4096
"""
from __future__ import annotations

import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import neuronx_distributed_llama3_2_b200 as nxd  # noqa: E402
from neuronx_distributed_llama3_2_b200.models import MixtralForCausalLM, mixtral_8x7b, tiny_mixtral  # noqa: E402
from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.logger import get_logger  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.training_utils import SyntheticTokens, Throughput  # noqa: E402

logger = get_logger()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tiny", type=int, default=0)
    ap.add_argument("--tensor_parallel_size", type=int, default=1)
    ap.add_argument("--expert_parallel_size", type=int, default=1)
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--seq_len", type=int, default=4096)
    ap.add_argument("--batch_size", type=int, default=1)
    ap.add_argument("--capacity_factor", type=float, default=2.0)
    ap.add_argument("--max_steps", type=int, default=10)
    ap.add_argument("--lr", type=float, default=1e-4)
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
    make = tiny_mixtral if args.tiny else mixtral_8x7b
    cfg = make(num_hidden_layers=args.layers, sequence_parallel_enabled=sp, capacity_factor=args.capacity_factor)
    if not cuda:
        cfg.torch_dtype = torch.float32
    model = nxd.initialize_parallel_model(nxd_config, lambda: MixtralForCausalLM(cfg, device=device), device=device)
    optimizer = nxd.initialize_parallel_optimizer(nxd_config, torch.optim.AdamW, model.parameters(), lr=args.lr)
    dp_rank = ps.get_data_parallel_rank() if ps.model_parallel_is_initialized() else 0
    dp = ps.get_data_parallel_size() if ps.model_parallel_is_initialized() else 1
    data = SyntheticTokens(cfg.vocab_size, args.batch_size, args.seq_len, device, seed=dp_rank)
    meter = Throughput(args.batch_size, dp, 1)
    for step in range(args.max_steps):
        ids = next(data)
        optimizer.zero_grad()
        loss = model.run_train(ids, labels=ids)
        optimizer.step()
        seqs = meter.get_throughput()
        if ps.is_global_rank_zero():
            logger.info("step %d loss %.4f grad-norm %.3f %.2f seq/s", step + 1, float(loss), float(optimizer.grad_norm), seqs)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
