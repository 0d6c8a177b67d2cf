#!/usr/bin/env python
"""Tensor-parallel x data-parallel pre-training of the other model families of the reference examples: GPT-NeoX 20B / 6.9B
(``examples/training/tp_dp_gpt_neox_hf_pretrain``), CodeGen2.5 (``examples/training/codegen25``) and BERT-large
(``examples/training/tp_dp_bert_hf_pretrain``), with ZeRO-1 over the data-parallel group.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/training/tp_dp_family_pretrain.py \
        --family neox --tensor_parallel_size 8 --seq_len 2048 --layers 4
    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/training/tp_dp_family_pretrain.py \
        --family bert --tensor_parallel_size 2 --seq_len 512 --batch_size 16 --data_dir shards/   # .npz / .hdf5 shards

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
from neuronx_distributed_llama3_2_b200.models import (BertConfig, BertForPreTraining, GPTNeoXConfig,  # noqa: E402
                                                      GPTNeoXForCausalLM, bert_large, codegen25_7b, config_from_json,
                                                      gpt_neox_20b, tiny_bert, tiny_gpt_neox)
from neuronx_distributed_llama3_2_b200.models.bert import BertLayer  # noqa: E402
from neuronx_distributed_llama3_2_b200.models.gpt_neox import GPTNeoXLayer  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.data import BertPretrainingDataset  # noqa: E402


def bert_batches(args, cfg, device):
    """``(input_ids, token_type_ids, attention_mask, labels, next_sentence_label)`` dict batches: from pre-tokenised
    shards (one after the other, as the reference's HDF5 loop) or synthetic with 15 % masked positions."""
    dp, dp_rank = C.dp_coords()
    keys = ("input_ids", "token_type_ids", "attention_mask", "labels", "next_sentence_label")
    if args.data_dir:
        from torch.utils.data import DataLoader, DistributedSampler

        files = sorted(os.path.join(args.data_dir, f) for f in os.listdir(args.data_dir)
                       if f.endswith((".npz", ".hdf5", ".h5")))
        assert files, f"no .npz / .hdf5 shards under {args.data_dir}"
        epoch = 0
        while True:
            for f in files:
                ds = BertPretrainingDataset(f, max_pred_length=args.max_pred_length)
                sampler = DistributedSampler(ds, num_replicas=dp, rank=dp_rank, shuffle=True, seed=args.seed + epoch,
                                             drop_last=True)
                for row in DataLoader(ds, batch_size=args.batch_size, sampler=sampler, drop_last=True):
                    yield {k: v.to(device) for k, v in zip(keys, row)}
            epoch += 1
    gen = torch.Generator().manual_seed(args.seed + dp_rank)
    while True:
        ids = torch.randint(0, cfg.vocab_size, (args.batch_size, args.seq_len), generator=gen)
        labels = torch.full_like(ids, -100)
        pick = torch.rand(ids.shape, generator=gen) < 0.15
        labels[pick] = ids[pick]
        nsp = torch.randint(0, 2, (args.batch_size,), generator=gen)
        yield {"input_ids": ids.to(device), "labels": labels.to(device), "next_sentence_label": nsp.to(device)}


def main():
    ap = C.add_common_args(argparse.ArgumentParser())
    ap.add_argument("--family", default="neox", choices=["neox", "codegen", "bert"])
    ap.add_argument("--tiny", type=int, default=0)
    ap.add_argument("--max_pred_length", type=int, default=80, help="BERT: masked positions per sequence in the shards")
    args = ap.parse_args()
    device = C.init_process_group()
    sp = bool(args.use_sequence_parallel) and args.tensor_parallel_size > 1
    if args.family == "bert":
        # pipeline tracing fixes the inputs that are None (synthetic batches carry no segment ids / padding mask)
        layer_cls = BertLayer
        input_names = (["input_ids", "token_type_ids", "attention_mask", "labels", "next_sentence_label"] if args.data_dir
                       else ["input_ids", "labels", "next_sentence_label"])
    else:
        layer_cls, input_names = GPTNeoXLayer, ["input_ids", "labels"]
    nxd_config = nxd.neuronx_distributed_config(
        tensor_parallel_size=args.tensor_parallel_size, sequence_parallel=sp,
        optimizer_config={"zero_one_enabled": bool(args.use_zero_1), "grad_clipping": True, "max_grad_norm": 1.0},
        activation_checkpoint_config="full" if args.full_activation_checkpoint else None,
        **C.pipeline_kwargs(args, layer_cls, input_names))
    kw = dict(sequence_parallel_enabled=sp)
    if device.type != "cuda":
        kw["torch_dtype"] = torch.float32
    if args.family == "bert":
        cfg = (config_from_json(BertConfig, args.model_path, **kw) if args.model_path
               else (tiny_bert if args.tiny else bert_large)(**kw))
        build = lambda: BertForPreTraining(cfg, device=device)  # noqa: E731
    else:
        if args.model_path:
            cfg = config_from_json(GPTNeoXConfig, args.model_path, **kw)
        else:
            make = tiny_gpt_neox if args.tiny else (gpt_neox_20b if args.family == "neox" else codegen25_7b)
            if args.tiny and args.family == "codegen":
                kw.update(rotary_interleaved=True, shared_layernorm=True, qkv_bias=False, rotary_pct=0.5,
                          hidden_act="gelu_new")
            cfg = make(**kw)
        build = lambda: GPTNeoXForCausalLM(cfg, device=device)  # noqa: E731
    if args.layers:
        cfg.num_hidden_layers = args.layers
    model = nxd.initialize_parallel_model(nxd_config, build, device=device)
    batches = bert_batches(args, cfg, device) if args.family == "bert" else C.token_batches(args, cfg.vocab_size, device)
    C.fit(args, nxd_config, model, batches)


if __name__ == "__main__":
    main()
