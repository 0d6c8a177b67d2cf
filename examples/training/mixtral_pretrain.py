#!/usr/bin/env python
"""Mixtral-style MoE pre-training with TP + SP + expert parallelism + EP-aware ZeRO-1 (BASELINE.json config #5; the
counterpart of ``examples/training/mixtral/``).

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/training/mixtral_pretrain.py \
        --tensor_parallel_size 2 --expert_parallel_size 4 --layers 4 --seq_len 4096

Flags shared with the other family drivers (data, schedule, checkpoints, TensorBoard): ``_train_common.add_common_args``.
"""
from __future__ import annotations

import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import _train_common as C  # noqa: E402  (puts the repo root on sys.path)

import neuronx_distributed_llama3_2_b200 as nxd  # noqa: E402
from neuronx_distributed_llama3_2_b200.models import (MixtralConfig, MixtralForCausalLM, config_from_json,  # noqa: E402
                                                      mixtral_8x7b, tiny_mixtral)
from neuronx_distributed_llama3_2_b200.models.mixtral import MixtralDecoderLayer  # noqa: E402


def main():
    ap = C.add_common_args(argparse.ArgumentParser())
    ap.add_argument("--tiny", type=int, default=0)
    ap.add_argument("--expert_parallel_size", type=int, default=1)
    ap.add_argument("--capacity_factor", type=float, default=2.0)
    ap.add_argument("--router_aux_loss_coef", type=float, default=0.02)
    args = ap.parse_args()
    device = C.init_process_group()
    sp = bool(args.use_sequence_parallel) and args.tensor_parallel_size > 1
    nxd_config = nxd.neuronx_distributed_config(
        tensor_parallel_size=args.tensor_parallel_size, expert_parallel_size=args.expert_parallel_size,
        sequence_parallel=sp,
        optimizer_config={"zero_one_enabled": bool(args.use_zero_1), "grad_clipping": True, "max_grad_norm": 1.0},
        activation_checkpoint_config="full" if args.full_activation_checkpoint else None,
        **C.pipeline_kwargs(args, MixtralDecoderLayer, ["input_ids", "labels"]))
    kw = dict(sequence_parallel_enabled=sp, capacity_factor=args.capacity_factor,
              router_aux_loss_coef=args.router_aux_loss_coef)
    if args.model_path:
        cfg = config_from_json(MixtralConfig, args.model_path, rope_scaling=None, **kw)
    else:
        cfg = (tiny_mixtral if args.tiny else mixtral_8x7b)(**kw)
    if args.layers:
        cfg.num_hidden_layers = args.layers
    if device.type != "cuda":
        cfg.torch_dtype = torch.float32
    model = nxd.initialize_parallel_model(nxd_config, lambda: MixtralForCausalLM(cfg, device=device), device=device)
    C.fit(args, nxd_config, model, C.token_batches(args, cfg.vocab_size, device))


if __name__ == "__main__":
    main()
