"""HuggingFace-style checkpoint directories (reference ``examples/inference/modules/checkpoint.py:20-149``):
``model.safetensors`` | ``model.safetensors.index.json`` + shards | ``pytorch_model.bin`` | ``pytorch_model.bin.index.json``
+ shards.  ``.safetensors`` files are read by the C++ runtime (mmap + copy threads) when it is available."""
from __future__ import annotations

import json
import os
import re
from typing import Callable, Dict, List

import torch

SAFETENSORS = "model.safetensors"
SAFETENSORS_INDEX = "model.safetensors.index.json"
PYTORCH_BIN = "pytorch_model.bin"
PYTORCH_BIN_INDEX = "pytorch_model.bin.index.json"


def _read_safetensors(path: str) -> Dict[str, torch.Tensor]:
    try:
        from ..runtime import available, load_safetensors

        if available():
            return load_safetensors(path)
    except (OSError, KeyError, RuntimeError):
        pass
    from safetensors.torch import load_file

    return load_file(path)


def _torch_load(path: str) -> Dict[str, torch.Tensor]:
    return torch.load(path, map_location="cpu", weights_only=True)


def _load_from_files(filenames: List[str], state_dict_dir: str, load_func: Callable) -> Dict[str, torch.Tensor]:
    state: Dict[str, torch.Tensor] = {}
    for name in filenames:
        part = load_func(os.path.join(state_dict_dir, name))
        dup = set(part) & set(state)
        if dup:
            raise ValueError(f"{name} repeats keys of an earlier shard: {sorted(dup)[:3]}")
        state.update(part)
    return state


def _shards_of(index_path: str) -> List[str]:
    with open(index_path) as f:
        index = json.load(f)
    return sorted(set(index["weight_map"].values()))


def load_safetensors(state_dict_dir: str) -> Dict[str, torch.Tensor]:
    return _read_safetensors(os.path.join(state_dict_dir, SAFETENSORS))


def load_safetensors_sharded(state_dict_dir: str) -> Dict[str, torch.Tensor]:
    return _load_from_files(_shards_of(os.path.join(state_dict_dir, SAFETENSORS_INDEX)), state_dict_dir, _read_safetensors)


def load_pytorch_model_bin(state_dict_dir: str) -> Dict[str, torch.Tensor]:
    return _torch_load(os.path.join(state_dict_dir, PYTORCH_BIN))


def load_pytorch_model_bin_sharded(state_dict_dir: str) -> Dict[str, torch.Tensor]:
    return _load_from_files(_shards_of(os.path.join(state_dict_dir, PYTORCH_BIN_INDEX)), state_dict_dir, _torch_load)


def load_state_dict(state_dict_dir: str) -> Dict[str, torch.Tensor]:
    """Whatever format the directory holds, preferring safetensors (reference ``:20-52``)."""
    for marker, fn in ((SAFETENSORS, load_safetensors), (SAFETENSORS_INDEX, load_safetensors_sharded),
                       (PYTORCH_BIN, load_pytorch_model_bin), (PYTORCH_BIN_INDEX, load_pytorch_model_bin_sharded)):
        if os.path.exists(os.path.join(state_dict_dir, marker)):
            return fn(state_dict_dir)
    raise FileNotFoundError(f"no model.safetensors / pytorch_model.bin (or their index files) in {state_dict_dir}")


def _parse_size(size) -> int:
    if isinstance(size, int):
        return size
    m = re.fullmatch(r"\s*([\d.]+)\s*([KMGT]?)I?B\s*", str(size).upper())
    if not m:
        raise ValueError(f"cannot parse shard size {size!r}")
    return int(float(m.group(1)) * {"": 1, "K": 1e3, "M": 1e6, "G": 1e9, "T": 1e12}[m.group(2)])


def save_state_dict_safetensors(state_dict: Dict[str, torch.Tensor], state_dict_dir: str, max_shard_size="10GB") -> None:
    """One ``model.safetensors`` or numbered shards + ``model.safetensors.index.json`` (reference ``:133-149``).  Tied
    tensors (same storage) are written once under their first name."""
    from safetensors.torch import save_file

    os.makedirs(state_dict_dir, exist_ok=True)
    limit = _parse_size(max_shard_size)
    seen, shards, cur, cur_bytes = {}, [], {}, 0
    for k, v in state_dict.items():
        key = (v.data_ptr(), tuple(v.shape), v.dtype) if v.numel() else None
        if key is not None and key in seen:
            continue
        if key is not None:
            seen[key] = k
        nbytes = v.numel() * v.element_size()
        if cur and cur_bytes + nbytes > limit:
            shards.append(cur)
            cur, cur_bytes = {}, 0
        cur[k] = v.detach().contiguous().cpu()
        cur_bytes += nbytes
    if cur or not shards:
        shards.append(cur)
    if len(shards) == 1:
        save_file(shards[0], os.path.join(state_dict_dir, SAFETENSORS), metadata={"format": "pt"})
        return
    weight_map, total = {}, 0
    for i, shard in enumerate(shards, start=1):
        name = f"model-{i:05d}-of-{len(shards):05d}.safetensors"
        save_file(shard, os.path.join(state_dict_dir, name), metadata={"format": "pt"})
        for k, v in shard.items():
            weight_map[k] = name
            total += v.numel() * v.element_size()
    with open(os.path.join(state_dict_dir, SAFETENSORS_INDEX), "w") as f:
        json.dump({"metadata": {"total_size": total}, "weight_map": weight_map}, f, indent=2, sort_keys=True)
