#!/usr/bin/env python
"""Tensor-parallel x data-parallel pre-training of the other model families of the reference examples: GPT-NeoX-20B
(``examples/training/tp_dp_gpt_neox_hf_pretrain``), CodeGen2.5 (``examples/training/codegen25``) and BERT-large
(``examples/training/tp_dp_bert_hf_pretrain``), with ZeRO-1 over the data-parallel group.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/training/tp_dp_family_pretrain.py \
        --family neox --tensor_parallel_size 8 --seq_len 2048 --layers 4
    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/training/tp_dp_family_pretrain.py \
        --family bert --tensor_parallel_size 2 --seq_len 512 --batch_size 16
"""
from __future__ import annotations

import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import neuronx_distributed_llama3_2_b200 as nxd  # noqa: E402
from neuronx_distributed_llama3_2_b200.models import (BertForPreTraining, GPTNeoXForCausalLM, bert_large, codegen25_7b,  # noqa: E402
                                                     gpt_neox_20b, tiny_bert, tiny_gpt_neox)
from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.logger import get_logger  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.training_utils import Throughput  # noqa: E402

logger = get_logger()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--family", default="neox", choices=["neox", "codegen", "bert"])
    ap.add_argument("--tiny", type=int, default=0)
    ap.add_argument("--tensor_parallel_size", type=int, default=1)
    ap.add_argument("--use_sequence_parallel", type=int, default=1)
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--seq_len", type=int, default=2048)
    ap.add_argument("--batch_size", type=int, default=1)
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
    sp = bool(args.use_sequence_parallel) and args.tensor_parallel_size > 1
    nxd_config = nxd.neuronx_distributed_config(
        tensor_parallel_size=args.tensor_parallel_size, sequence_parallel=sp,
        optimizer_config={"zero_one_enabled": True, "grad_clipping": True, "max_grad_norm": 1.0})
    kw = dict(sequence_parallel_enabled=sp)
    if args.layers:
        kw["num_hidden_layers"] = args.layers
    if not cuda:
        kw["torch_dtype"] = torch.float32
    if args.family == "bert":
        cfg = (tiny_bert if args.tiny else bert_large)(**kw)
        build = lambda: BertForPreTraining(cfg, device=device)  # noqa: E731
    else:
        make = tiny_gpt_neox if args.tiny else (gpt_neox_20b if args.family == "neox" else codegen25_7b)
        if args.tiny and args.family == "codegen":
            kw.update(rotary_interleaved=True, shared_layernorm=True, qkv_bias=False, rotary_pct=0.5, hidden_act="gelu_new")
        cfg = make(**kw)
        build = lambda: GPTNeoXForCausalLM(cfg, device=device)  # noqa: E731
    model = nxd.initialize_parallel_model(nxd_config, build, device=device)
    optimizer = nxd.initialize_parallel_optimizer(nxd_config, torch.optim.AdamW, model.parameters(), lr=args.lr,
                                                  weight_decay=0.01)
    dp = ps.get_data_parallel_size() if ps.model_parallel_is_initialized() else 1
    dp_rank = ps.get_data_parallel_rank() if ps.model_parallel_is_initialized() else 0
    gen = torch.Generator().manual_seed(1234 + dp_rank)
    meter = Throughput(args.batch_size, dp, 1)
    for step in range(args.max_steps):
        ids = torch.randint(0, cfg.vocab_size, (args.batch_size, args.seq_len), generator=gen).to(device)
        optimizer.zero_grad()
        if args.family == "bert":
            labels = torch.full_like(ids, -100)
            pick = torch.rand(ids.shape, generator=gen).to(device) < 0.15
            labels[pick] = ids[pick]
            nsp = torch.randint(0, 2, (args.batch_size,), generator=gen).to(device)
            loss = model.run_train(ids, labels=labels, next_sentence_label=nsp)
        else:
            loss = model.run_train(ids, labels=ids)
        optimizer.step()
        seqs = meter.get_throughput()
        if ps.is_global_rank_zero():
            logger.info("step %d loss %.4f grad-norm %.3f %.2f seq/s", step + 1, float(loss), float(optimizer.grad_norm), seqs)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
