#!/usr/bin/env python
"""Tokenise a text corpus into fixed-length training windows -- counterpart of the reference's
``examples/training/llama/get_dataset.py`` (which downloads wikicorpus; there is no network here, so the corpus is a local
``.txt`` / ``.jsonl`` file or directory).

    python examples/training/get_dataset.py --input corpus.txt --tokenizer /path/to/hf_tokenizer --block_size 8192 \
        --output ~/data/corpus_tokenized            # HF datasets directory (default)
    python examples/training/get_dataset.py --input corpus.txt --tokenizer bytes --format npy --output tokens.npy

``--tokenizer`` is a local Hugging Face tokenizer directory, a SentencePiece ``.model`` file, or ``bytes`` (UTF-8 bytes,
vocabulary 256 -- for smoke tests).  Documents are separated by the EOS token, concatenated and cut into ``block_size``
windows; ``labels`` equal ``input_ids`` (the model shifts them).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from typing import Callable, Iterator, List, Optional, Tuple

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from neuronx_distributed_llama3_2_b200.utils.data import pack_dataset  # noqa: E402


def load_tokenizer(spec: str) -> Tuple[Callable[[str], List[int]], Optional[int], int]:
    """``(encode, eos_id, vocab_size)``."""
    if spec == "bytes":
        return (lambda s: list(s.encode("utf-8"))), 0, 256
    if spec.endswith(".model"):
        import sentencepiece as spm

        sp = spm.SentencePieceProcessor(model_file=spec)
        return (lambda s: sp.encode(s)), (sp.eos_id() if sp.eos_id() >= 0 else None), sp.get_piece_size()
    from transformers import AutoTokenizer

    tok = AutoTokenizer.from_pretrained(spec)
    return (lambda s: tok.encode(s, add_special_tokens=False)), tok.eos_token_id, len(tok)


def documents(path: str, text_key: str) -> Iterator[str]:
    files = [path] if os.path.isfile(path) else sorted(os.path.join(path, f) for f in os.listdir(path))
    for f in files:
        if f.endswith((".jsonl", ".json")):
            with open(f) as fh:
                for line in fh:
                    if line.strip():
                        yield json.loads(line)[text_key]
        elif f.endswith(".txt"):
            with open(f) as fh:
                doc: List[str] = []
                for line in fh:                       # blank line = document boundary
                    if line.strip():
                        doc.append(line)
                    elif doc:
                        yield "".join(doc)
                        doc = []
                if doc:
                    yield "".join(doc)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--input", required=True)
    ap.add_argument("--output", required=True)
    ap.add_argument("--tokenizer", required=True)
    ap.add_argument("--block_size", type=int, default=8192)
    ap.add_argument("--text_key", default="text")
    ap.add_argument("--format", choices=["datasets", "npy", "bin"], default="datasets")
    a = ap.parse_args(argv)
    encode, eos, vocab = load_tokenizer(a.tokenizer)

    def token_lists():
        for d in documents(a.input, a.text_key):
            ids = encode(d)
            yield ids + ([eos] if eos is not None else [])

    windows = pack_dataset(token_lists(), a.block_size)
    if not windows:
        raise SystemExit(f"the corpus has fewer than {a.block_size} tokens")
    if a.format == "datasets":
        import datasets

        ds = datasets.Dataset.from_dict({"input_ids": windows, "labels": windows})
        ds.save_to_disk(a.output)
    else:
        arr = np.asarray(windows, dtype=np.uint16 if vocab <= 65536 else np.int32).reshape(-1)
        if a.format == "npy":
            np.save(a.output, arr)
        else:
            arr.tofile(a.output)
            json.dump({"dtype": str(arr.dtype), "block_size": a.block_size}, open(a.output + ".json", "w"))
    print(f"{len(windows)} windows of {a.block_size} tokens -> {a.output}")


if __name__ == "__main__":
    main()
