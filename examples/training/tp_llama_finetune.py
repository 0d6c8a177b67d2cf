#!/usr/bin/env python
"""Instruction fine-tuning of a Llama model with tensor parallelism, ZeRO-1 and (optionally) LoRA adapters, followed by a
ROUGE evaluation of greedy generations -- counterpart of the reference's
``examples/training/llama/lightning/tp_llama_hf_finetune_ptl.py`` (LoRA rank 16 / alpha 32 / dropout 0.05 on the attention
projections, ``--do_eval`` with optional golden ROUGE scores).

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 examples/training/tp_llama_finetune.py \
        --model_path ~/Llama-3.2-3B --pretrained_weight ~/ckpt/pretrained --data_dir ~/data/dolly \
        --tokenizer ~/Llama-3.2-3B --tensor_parallel_size 8 --enable_lora 1 --do_eval 1

``--data_dir`` holds ``train.jsonl`` (and optionally ``eval.jsonl``) with ``instruction`` / ``input`` / ``output`` fields.
``--tokenizer bytes`` (UTF-8 bytes) needs no tokenizer files and is what the smoke test uses.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from collections import Counter
from typing import Dict, List, Sequence

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import _train_common as C  # noqa: E402  (puts the repo root on sys.path)

import neuronx_distributed_llama3_2_b200 as nxd  # noqa: E402
from neuronx_distributed_llama3_2_b200.models import PRESETS, LlamaForCausalLM, config_from_hf_json  # noqa: E402
from neuronx_distributed_llama3_2_b200.modules.lora import LoraConfig  # noqa: E402
from neuronx_distributed_llama3_2_b200.parallel import mappings  # noqa: E402
from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps  # noqa: E402
from neuronx_distributed_llama3_2_b200.utils.data import DevicePrefetcher, create_instruction_based_dataset  # noqa: E402


# -- ROUGE on token sequences (F-measure; the reference uses torchmetrics' ROUGEScore on decoded text) ------------------------
def _ngrams(x: Sequence, n: int) -> Counter:
    return Counter(tuple(x[i:i + n]) for i in range(len(x) - n + 1))


def _f(match: int, n_pred: int, n_ref: int) -> float:
    if match == 0 or n_pred == 0 or n_ref == 0:
        return 0.0
    p, r = match / n_pred, match / n_ref
    return 2 * p * r / (p + r)


def rouge_n(pred: Sequence, ref: Sequence, n: int) -> float:
    a, b = _ngrams(pred, n), _ngrams(ref, n)
    return _f(sum((a & b).values()), sum(a.values()), sum(b.values()))


def rouge_l(pred: Sequence, ref: Sequence) -> float:
    m, n = len(pred), len(ref)
    row = [0] * (n + 1)
    for i in range(1, m + 1):
        prev = 0
        for j in range(1, n + 1):
            cur = row[j]
            row[j] = prev + 1 if pred[i - 1] == ref[j - 1] else max(row[j], row[j - 1])
            prev = cur
    return _f(row[n], m, n)


def rouge_scores(preds: List[Sequence], refs: List[Sequence]) -> Dict[str, float]:
    k = max(len(preds), 1)
    return {"rouge1_fmeasure": sum(rouge_n(p, r, 1) for p, r in zip(preds, refs)) / k,
            "rouge2_fmeasure": sum(rouge_n(p, r, 2) for p, r in zip(preds, refs)) / k,
            "rougeL_fmeasure": sum(rouge_l(p, r) for p, r in zip(preds, refs)) / k}


@torch.no_grad()
def greedy_generate(model, prompt: torch.Tensor, max_new_tokens: int, eos: int = -1) -> List[int]:
    """Greedy continuation by re-running the training graph on the growing sequence (evaluation sets are small; the
    KV-cache path is ``inference.InferenceRunner``).  Vocab-parallel logits are gathered before the argmax."""
    seq = prompt.reshape(1, -1)
    tp = ps.get_tensor_model_parallel_size() if ps.model_parallel_is_initialized() else 1
    out: List[int] = []
    for _ in range(max_new_tokens):
        pad = (-seq.shape[1]) % tp                     # sequence parallelism needs S divisible by tp
        ids = torch.nn.functional.pad(seq, (0, pad)) if pad else seq
        logits = model(ids)                            # [S, 1, V / tp]
        last = logits[seq.shape[1] - 1, 0]
        if tp > 1:
            last = mappings.gather_from_tensor_model_parallel_region(last.reshape(1, -1)).reshape(-1)
        tok = int(last.argmax())
        if tok == eos:
            break
        out.append(tok)
        seq = torch.cat([seq, torch.tensor([[tok]], device=seq.device)], dim=1)
    return out


def evaluate(model, loader, max_new_tokens: int, golden_path=None, tol: float = 0.02) -> Dict[str, float]:
    model.eval()
    preds, refs = [], []
    for batch in loader:
        for ids, labels, mask in zip(batch["input_ids"], batch["labels"], batch["attention_mask"]):
            n_prompt = int((labels == -100).long().cumprod(0).sum())           # leading masked positions = the prompt
            ref = labels[n_prompt:int(mask.sum())].tolist()
            preds.append(greedy_generate(model, ids[:n_prompt], min(max_new_tokens, max(len(ref), 1))))
            refs.append(ref)
    scores = rouge_scores(preds, refs)
    if ps.is_global_rank_zero():
        print(json.dumps({"eval_samples": len(preds), **scores}))
        if golden_path:
            golden = json.load(open(golden_path))
            for k, v in golden.items():
                assert abs(v - scores[k]) <= tol, f"{k}: {scores[k]:.4f} vs golden {v:.4f}"
            print("Evaluation rouge scores matched goldens!")
    model.train()
    return scores


def load_tokenizer(spec: str):
    if spec == "bytes":
        return lambda s: list(s.encode("utf-8"))
    from transformers import AutoTokenizer

    return AutoTokenizer.from_pretrained(spec)


def main():
    ap = C.add_common_args(argparse.ArgumentParser())
    ap.add_argument("--model", default="llama3.2-1b", choices=sorted(PRESETS))
    ap.add_argument("--tokenizer", default="bytes")
    ap.add_argument("--enable_lora", type=int, default=1)
    ap.add_argument("--lora_rank", type=int, default=16)
    ap.add_argument("--lora_alpha", type=int, default=32)
    ap.add_argument("--lora_dropout", type=float, default=0.05)
    ap.add_argument("--lora_target_modules", default="qkv_proj,o_proj")
    ap.add_argument("--do_eval", type=int, default=0)
    ap.add_argument("--eval_max_new_tokens", type=int, default=32)
    ap.add_argument("--golden_rouge_score_path", default=None)
    args = ap.parse_args()
    assert args.data_dir, "--data_dir with train.jsonl is required"
    device = C.init_process_group()
    sp = bool(args.use_sequence_parallel) and args.tensor_parallel_size > 1
    lora = LoraConfig(enable_lora=True, lora_rank=args.lora_rank, lora_alpha=args.lora_alpha,
                      lora_dropout=args.lora_dropout, target_modules=args.lora_target_modules.split(","),
                      lora_verbose=True, save_lora_base=False, merge_lora=False) if args.enable_lora else None
    nxd_config = nxd.neuronx_distributed_config(
        tensor_parallel_size=args.tensor_parallel_size, sequence_parallel=sp,
        optimizer_config={"zero_one_enabled": bool(args.use_zero_1), "grad_clipping": True, "max_grad_norm": 1.0},
        lora_config=lora)
    kw = dict(sequence_parallel_enabled=sp)
    cfg = config_from_hf_json(args.model_path, **kw) if args.model_path else PRESETS[args.model](**kw)
    if args.layers:
        cfg.num_hidden_layers = args.layers
    if device.type != "cuda":
        cfg.torch_dtype = torch.float32
    model = nxd.initialize_parallel_model(nxd_config, lambda: LlamaForCausalLM(cfg, device=device), device=device)
    dp, dp_rank = C.dp_coords()
    train_dl, eval_dl = create_instruction_based_dataset(args.data_dir, args.batch_size, dp, dp_rank, args.seed,
                                                         tokenizer=load_tokenizer(args.tokenizer), seq_len=args.seq_len)
    batches = ({"input_ids": b["input_ids"], "labels": b["labels"]} for b in DevicePrefetcher(train_dl, device))
    eval_batches = list(DevicePrefetcher(eval_dl, device, loop=False)) if (args.do_eval and eval_dl is not None) else None

    C.fit(args, nxd_config, model, batches, teardown=False)
    if eval_batches:
        evaluate(model, eval_batches, args.eval_max_new_tokens, args.golden_rouge_score_path)
    import torch.distributed as dist

    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
