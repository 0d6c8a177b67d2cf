"""Native host runtime (csrc/runtime/nxd_runtime.cpp): safetensors parallel reader and bucket router."""
from __future__ import annotations

import os
import tempfile

import pytest
import torch

from neuronx_distributed_llama3_2_b200.runtime import NativeRouter, SafetensorsFile, available, load_safetensors

pytestmark = pytest.mark.skipif(not available(), reason="no C++ compiler on this machine")


def _make(path):
    from safetensors.torch import save_file

    torch.manual_seed(0)
    tensors = {"model.embed_tokens.weight": torch.randn(512, 64).bfloat16(),
               "model.layers.0.mlp.down_proj.weight": torch.randn(64, 96),
               "a.scalar": torch.tensor(3.5), "b.int": torch.arange(40, dtype=torch.int64).reshape(5, 8),
               "c.half": torch.randn(3, 5, 7).half(), "d.big": torch.randn(3000, 1500)}   # > one 8 MB copy piece
    save_file(tensors, path, metadata={"format": "pt", "note": 'has "quotes" and {braces}'})
    return tensors


def test_safetensors_reader_matches_python_package():
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "tp0_sharded_checkpoint.safetensors")
        ref = _make(path)
        with SafetensorsFile(path) as f:
            assert sorted(f.keys()) == sorted(ref)
            for threads in (1, 5):
                got = f.read(threads=threads)
                for k, v in ref.items():
                    assert got[k].dtype == v.dtype and got[k].shape == v.shape
                    assert torch.equal(got[k], v), k
            part = f.read(["b.int", "c.half"])
            assert set(part) == {"b.int", "c.half"}
            rows = f.read_rows("model.embed_tokens.weight", 128, 256)       # dim-0 shard of tp rank 1 of 4
            assert torch.equal(rows, ref["model.embed_tokens.weight"][128:256])
            with pytest.raises(RuntimeError):
                f.read_rows("b.int", 2, 9)
        full = load_safetensors(path)
        assert torch.equal(full["d.big"], ref["d.big"])
    with pytest.raises(OSError):
        SafetensorsFile("/nonexistent/file.safetensors")


def test_reader_rejects_corrupt_header():
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "bad.safetensors")
        with open(path, "wb") as f:
            f.write((1 << 40).to_bytes(8, "little") + b"{}")
        with pytest.raises(OSError):
            SafetensorsFile(path)
        with open(path, "wb") as f:
            hdr = b'{"x":{"dtype":"F32","shape":[4],"data_offsets":[0,4096]}}'
            f.write(len(hdr).to_bytes(8, "little") + hdr + bytes(16))
        with pytest.raises(OSError):
            SafetensorsFile(path)


def test_router_exact_and_padded():
    r = NativeRouter()
    for i, t in enumerate((128, 512, 2048)):
        r.add("context_encoding_model", 4, t, i)
    r.add("token_generation_model", 4, 1, 3)
    assert r.route_exact(4, 512) == 1 and r.route_exact(4, 1) == 3 and r.route_exact(2, 512) == -1
    assert r.route_padded("context_encoding_model", 4, 100) == (0, 4, 128)
    assert r.route_padded("context_encoding_model", 3, 129) == (1, 4, 512)
    assert r.route_padded("context_encoding_model", 4, 4096)[0] == -1
    assert r.route_padded("token_generation_model", 1, 1) == (3, 4, 1)
    with pytest.raises(ValueError):
        r.add("token_generation_model", 4, 1, 9)


def test_nxd_model_loads_per_rank_safetensors_through_native_reader():
    from neuronx_distributed_llama3_2_b200.inference import (BaseModelInstance, InferenceConfig, LlamaInferenceModel,
                                                             ModelBuilder)
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama

    torch.manual_seed(0)
    train = LlamaForCausalLM(tiny_llama(sequence_parallel_enabled=False))
    full_sd = {k: v.clone() for k, v in train.state_dict().items()}

    def make():
        m = LlamaInferenceModel(tiny_llama(sequence_parallel_enabled=False),
                                InferenceConfig(batch_size=2, max_context_length=32, seq_len=64))
        m.init_kv_cache(device=torch.device("cpu"))
        return m

    builder = ModelBuilder(tp_degree=1, checkpoint_loader=lambda: full_sd, use_cuda_graphs=False)
    inst = BaseModelInstance(make)
    ex = lambda t: (torch.zeros(2, t, dtype=torch.long), torch.zeros(2, t, dtype=torch.long))  # noqa: E731
    builder.add("context_encoding_model", inst, [ex(16), ex(32)], priority_model_idx=0)
    builder.add("token_generation_model", inst, [ex(1)])
    model = builder.trace(initialize_model_weights=False)          # random weights ...
    assert model.route_padded("context_encoding_model", 2, 20) == (2, 32)
    assert model.route_padded("context_encoding_model", 2, 33) is None
    with tempfile.TemporaryDirectory() as d:
        builder.shard_checkpoint(d)
        model.load_sharded_checkpoint(d, rank=0)                    # ... replaced by the sharded checkpoint
    ids = torch.randint(1, 512, (2, 16))
    pos = torch.arange(16).unsqueeze(0).expand(2, 16)
    with torch.no_grad():
        ref = train(ids).transpose(0, 1)[:, -1].argmax(-1)
    assert torch.equal(model(ids, pos), ref)
