"""Datasets and the host->device input pipeline of the training drivers (reference
``examples/training/llama/training_utils.py:29-208`` -- ``pack_dataset``, ``create_llama_pretraining_dataset``,
``create_instruction_based_dataset`` -- ``get_dataset.py`` and torch-xla's ``MpDeviceLoader``).

Storage formats, in the order they are probed:
  * a Hugging Face ``datasets`` directory (``load_from_disk``) with ``input_ids`` (and optionally ``labels``) columns --
    what the reference's ``get_dataset.py`` writes;
  * a flat token file ``*.npy`` / ``*.bin`` (uint16 / int32 ids, one long stream) cut into ``seq_len`` windows --
    what ``examples/training/get_dataset.py`` of this repo writes when ``datasets`` is not wanted.
Batches are dicts ``{"input_ids", "labels"[, "attention_mask"]}`` of int64 tensors.
"""
from __future__ import annotations

import json
import os
from typing import Any, Callable, Dict, Iterable, Iterator, List, Optional, Sequence

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset, DistributedSampler

IGNORE_INDEX = -100


def pack_dataset(token_lists: Iterable[Sequence[int]], chunk_length: int = 2048) -> List[List[int]]:
    """Concatenate tokenised documents and cut the stream into ``chunk_length`` windows; the tail shorter than a window
    is dropped (the reference carries it into the next ``map`` batch, ``training_utils.py:29-62``; one pass over the whole
    corpus makes that a no-op here)."""
    out: List[List[int]] = []
    buf: List[int] = []
    for toks in token_lists:
        buf.extend(toks)
        while len(buf) >= chunk_length:
            out.append(buf[:chunk_length])
            buf = buf[chunk_length:]
    return out


class TokenFileDataset(Dataset):
    """Windows of ``seq_len`` tokens over one flat token file (memory-mapped)."""

    def __init__(self, path: str, seq_len: int, dtype: Optional[str] = None):
        if path.endswith(".npy"):
            self.tokens = np.load(path, mmap_mode="r")
        else:
            meta = path + ".json"
            if dtype is None and os.path.exists(meta):
                dtype = json.load(open(meta)).get("dtype")
            self.tokens = np.memmap(path, dtype=np.dtype(dtype or "uint16"), mode="r")
        self.seq_len = seq_len
        self.n = len(self.tokens) // seq_len
        if self.n == 0:
            raise ValueError(f"{path}: {len(self.tokens)} tokens are fewer than one window of {seq_len}")

    def __len__(self) -> int:
        return self.n

    def __getitem__(self, i: int) -> Dict[str, torch.Tensor]:
        ids = torch.from_numpy(np.asarray(self.tokens[i * self.seq_len:(i + 1) * self.seq_len]).astype(np.int64))
        return {"input_ids": ids, "labels": ids.clone()}


class _HFDatasetView(Dataset):
    def __init__(self, ds):
        self.ds = ds

    def __len__(self) -> int:
        return len(self.ds)

    def __getitem__(self, i: int) -> Dict[str, torch.Tensor]:
        row = self.ds[int(i)]
        out = {k: torch.as_tensor(v, dtype=torch.long) for k, v in row.items()
               if k in ("input_ids", "labels", "attention_mask", "token_type_ids")}
        out.setdefault("labels", out["input_ids"].clone())
        return out


def _collate(rows: List[Dict[str, torch.Tensor]]) -> Dict[str, torch.Tensor]:
    return {k: torch.stack([r[k] for r in rows]) for k in rows[0]}


def open_token_dataset(data_dir: str, seq_len: Optional[int] = None) -> Dataset:
    if os.path.isdir(data_dir):
        if any(os.path.exists(os.path.join(data_dir, f)) for f in ("dataset_info.json", "state.json", "dataset_dict.json")):
            import datasets

            ds = datasets.load_from_disk(data_dir)
            if hasattr(ds, "keys") and "train" in ds:
                ds = ds["train"]
            return _HFDatasetView(ds)
        files = sorted(f for f in os.listdir(data_dir) if f.endswith((".npy", ".bin")))
        if not files:
            raise FileNotFoundError(f"no dataset found under {data_dir}")
        data_dir = os.path.join(data_dir, files[0])
    if seq_len is None:
        raise ValueError("a flat token file needs seq_len")
    return TokenFileDataset(data_dir, seq_len)


def _loader(ds: Dataset, mini_batch_size: int, dp_size: int, dp_rank: int, seed: int, shuffle: bool) -> DataLoader:
    sampler = DistributedSampler(ds, num_replicas=dp_size, rank=dp_rank, shuffle=shuffle, seed=seed, drop_last=True)
    return DataLoader(ds, batch_size=mini_batch_size, sampler=sampler, collate_fn=_collate, num_workers=0, drop_last=True,
                      pin_memory=torch.cuda.is_available())


def create_llama_pretraining_dataset(data_dir: str, mini_batch_size: int, dp_size: int, dp_rank: int, seed: int,
                                     seq_len: Optional[int] = None):
    """``(train_dataloader, None)`` as the reference: un-shuffled ``DistributedSampler`` over the DP group, drop-last,
    pinned batches (``training_utils.py:99-127``)."""
    return _loader(open_token_dataset(data_dir, seq_len), mini_batch_size, dp_size, dp_rank, seed, shuffle=False), None


class InstructionDataset(Dataset):
    """Prompt / response pairs for fine-tuning: the loss is taken on the response only (prompt and padding labels are
    ``-100``), sequences are right-padded / truncated to ``seq_len`` (reference ``training_utils.py:130-208``)."""

    PROMPT = "Below is an instruction that describes a task. Write a response that appropriately completes the request.\n\n" \
             "### Instruction:\n{instruction}\n\n### Response:\n"
    PROMPT_WITH_INPUT = "Below is an instruction that describes a task, paired with an input that provides further context. " \
                        "Write a response that appropriately completes the request.\n\n### Instruction:\n{instruction}\n\n" \
                        "### Input:\n{input}\n\n### Response:\n"

    def __init__(self, records: Sequence[Dict[str, str]], tokenize: Callable[[str], List[int]], seq_len: int,
                 pad_token_id: int = 0, eos_token_id: Optional[int] = None):
        self.rows = []
        for r in records:
            tmpl = self.PROMPT_WITH_INPUT if r.get("input") else self.PROMPT
            prompt = tokenize(tmpl.format(instruction=r["instruction"], input=r.get("input", "")))
            resp = tokenize(r.get("output", r.get("response", "")))
            if eos_token_id is not None:
                resp = resp + [eos_token_id]
            ids = (prompt + resp)[:seq_len]
            labels = ([IGNORE_INDEX] * len(prompt) + resp)[:seq_len]
            pad = seq_len - len(ids)
            self.rows.append({"input_ids": torch.tensor(ids + [pad_token_id] * pad),
                              "labels": torch.tensor(labels + [IGNORE_INDEX] * pad),
                              "attention_mask": torch.tensor([1] * len(ids) + [0] * pad)})

    def __len__(self) -> int:
        return len(self.rows)

    def __getitem__(self, i: int) -> Dict[str, torch.Tensor]:
        return self.rows[i]


def create_instruction_based_dataset(data_dir: str, mini_batch_size: int, dp_size: int, dp_rank: int, seed: int,
                                     tokenizer=None, task=None, seq_len: int = 2048):
    """``(train_dataloader, eval_dataloader)`` from ``train.jsonl`` / ``eval.jsonl`` (or one ``*.jsonl``/``*.json``,
    split 95/5) with ``instruction`` / ``input`` / ``output`` fields.  ``tokenizer`` is an HF tokenizer or any callable
    ``str -> list[int]``."""
    def read(path):
        with open(path) as f:
            txt = f.read().strip()
        return json.loads(txt) if txt.startswith("[") else [json.loads(line) for line in txt.splitlines() if line.strip()]

    if os.path.isdir(data_dir):
        train = read(os.path.join(data_dir, "train.jsonl"))
        ev_path = os.path.join(data_dir, "eval.jsonl")
        ev = read(ev_path) if os.path.exists(ev_path) else []
    else:
        rows = read(data_dir)
        cut = max(1, int(len(rows) * 0.95))
        train, ev = rows[:cut], rows[cut:]
    if hasattr(tokenizer, "encode"):
        tok = lambda s: tokenizer.encode(s, add_special_tokens=False)  # noqa: E731
        pad = tokenizer.pad_token_id if getattr(tokenizer, "pad_token_id", None) is not None else 0
        eos = getattr(tokenizer, "eos_token_id", None)
    else:
        tok, pad, eos = tokenizer, 0, None
    make = lambda rows: InstructionDataset(rows, tok, seq_len, pad, eos)  # noqa: E731
    train_dl = _loader(make(train), mini_batch_size, dp_size, dp_rank, seed, shuffle=True)
    eval_dl = _loader(make(ev), mini_batch_size, dp_size, dp_rank, seed, shuffle=False) if len(ev) >= mini_batch_size * dp_size else None
    return train_dl, eval_dl


class DevicePrefetcher:
    """Wraps any iterable of (dicts / tuples of) host tensors: batch ``i + 1`` is copied to the device on a side stream
    while step ``i`` computes -- the counterpart of torch-xla's ``MpDeviceLoader`` (reference
    ``tp_zero1_llama_hf_pretrain.py`` wraps its DataLoader in one).  Restarts the underlying loader when it is exhausted
    (``set_epoch`` is forwarded to a ``DistributedSampler``) unless ``loop=False``."""

    def __init__(self, loader: Iterable, device, loop: bool = True):
        self.loader, self.device, self.loop = loader, torch.device(device), loop
        self.cuda = self.device.type == "cuda"
        self.stream = torch.cuda.Stream(device=self.device) if self.cuda else None
        self.epoch = 0
        self._it: Optional[Iterator] = None
        self._next: Any = None

    def _to_device(self, x):
        if isinstance(x, torch.Tensor):
            if self.cuda and not x.is_pinned():
                x = x.pin_memory()
            return x.to(self.device, non_blocking=True)
        if isinstance(x, dict):
            return {k: self._to_device(v) for k, v in x.items()}
        if isinstance(x, (list, tuple)):
            return type(x)(self._to_device(v) for v in x)
        return x

    def _fetch(self) -> None:
        try:
            host = next(self._it)
        except StopIteration:
            if not self.loop:
                self._next = None
                return
            self.epoch += 1
            sampler = getattr(self.loader, "sampler", None)
            if hasattr(sampler, "set_epoch"):
                sampler.set_epoch(self.epoch)
            self._it = iter(self.loader)
            host = next(self._it)
        if self.cuda:
            with torch.cuda.stream(self.stream):
                self._next = self._to_device(host)
        else:
            self._next = self._to_device(host)

    def __iter__(self):
        self._it = iter(self.loader)
        self._fetch()
        return self

    def __next__(self):
        if self._it is None:
            iter(self)
        if self._next is None:
            raise StopIteration
        if self.cuda:
            torch.cuda.current_stream(self.device).wait_stream(self.stream)
        out = self._next
        self._fetch()
        return out


class BertPretrainingDataset(Dataset):
    """One shard of pre-tokenised BERT pre-training samples with the columns of the reference's HDF5 shards
    (``tp_dp_bert_large_hf_pretrain_hdf5.py`` ``pretraining_dataset``): ``input_ids, input_mask, segment_ids,
    masked_lm_positions, masked_lm_ids, next_sentence_labels``.  Reads ``.npz`` (always) or ``.hdf5`` (when ``h5py`` is
    installed) and returns ``(input_ids, segment_ids, input_mask, masked_lm_labels, next_sentence_label)`` where
    ``masked_lm_labels`` is ``-100`` except at the masked positions."""

    KEYS = ("input_ids", "input_mask", "segment_ids", "masked_lm_positions", "masked_lm_ids", "next_sentence_labels")

    def __init__(self, input_file: str, max_pred_length: int = 80):
        self.max_pred_length = max_pred_length
        if input_file.endswith((".hdf5", ".h5")):
            try:
                import h5py
            except ImportError as e:  # pragma: no cover - h5py is not in this image
                raise ImportError("reading .hdf5 shards needs h5py; convert them to .npz with the same keys") from e
            with h5py.File(input_file, "r") as f:
                self.inputs = [np.asarray(f[k][:]) for k in self.KEYS]
        else:
            z = np.load(input_file)
            self.inputs = [z[k] for k in self.KEYS]

    def __len__(self) -> int:
        return len(self.inputs[0])

    def __getitem__(self, index: int):
        ids, mask, seg, pos, mids, nsl = [torch.from_numpy(np.asarray(a[index]).astype(np.int64)) for a in self.inputs]
        labels = torch.full_like(ids, IGNORE_INDEX)
        n = self.max_pred_length
        pad = (pos == 0).nonzero()
        if len(pad) != 0:
            n = int(pad[0])
        labels[pos[:n]] = mids[:n]
        return ids, seg, mask, labels, nsl


__all__ = ["BertPretrainingDataset", "DevicePrefetcher", "IGNORE_INDEX", "InstructionDataset", "TokenFileDataset",
           "create_instruction_based_dataset", "create_llama_pretraining_dataset", "open_token_dataset", "pack_dataset"]
