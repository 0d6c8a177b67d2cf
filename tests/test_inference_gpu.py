"""Inference path on B200: KV-cache generation through the CUDA kernels (prefill flash attention, in-place KV write,
split-K decode attention, skinny GEMM, fused GEMM->all-reduce) and CUDA-graph buckets against full recomputation."""
from __future__ import annotations

import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _cfg(**kw):
    from neuronx_distributed_llama3_2_b200.models import tiny_llama

    base = dict(vocab_size=1024, hidden_size=512, intermediate_size=1024, num_hidden_layers=2, num_attention_heads=8,
                num_key_value_heads=2, head_dim=64, torch_dtype=torch.bfloat16, sequence_parallel_enabled=False)
    base.update(kw)
    return tiny_llama(**base)


def _oracle_tokens(train, prompts, lens, n_new):
    """Greedy decoding by recomputing the whole sequence with the training model (fp32 copy)."""
    outs = []
    for b in range(prompts.shape[0]):
        seq = prompts[b, :lens[b]].clone()
        for _ in range(n_new):
            with torch.no_grad():
                logits = train(seq.unsqueeze(0)).transpose(0, 1)[0, -1]
            seq = torch.cat([seq, logits.argmax().reshape(1)])
        outs.append(seq)
    return outs


@pytest.mark.parametrize("paged", [False, True])
def test_generation_matches_recompute(paged):
    from neuronx_distributed_llama3_2_b200.inference import InferenceConfig, LlamaInferenceModel
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM

    torch.manual_seed(0)
    dev = torch.device("cuda")
    train = LlamaForCausalLM(_cfg(), device=dev)
    inf = LlamaInferenceModel(_cfg(), InferenceConfig(batch_size=2, max_context_length=128, seq_len=256, paged_kv=paged,
                                                      page_size=64), device=dev)
    inf.load_state_dict(train.state_dict())
    prompts = torch.randint(1, 1024, (2, 64), device=dev)
    mask = torch.ones_like(prompts)
    mask[1, 40:] = 0
    ids = prompts.clone()
    pos = torch.arange(64, device=dev).unsqueeze(0).expand(2, 64)
    inf.init_kv_cache(device=dev)
    lens = mask.sum(1)
    tok, logits = inf.forward_infer(ids, pos, last_token_index=lens - 1, return_logits=True)
    with torch.no_grad():
        ref = train(ids).transpose(0, 1).float()
    ref_last = torch.stack([ref[b, lens[b] - 1] for b in range(2)])
    err = (logits.float() - ref_last).abs().max().item()
    assert err < 0.08, err
    # decode steps: logits of the next position equal the recomputed ones within bf16 noise
    cur = lens.clone()
    seqs = [torch.cat([prompts[b, :lens[b]], tok[b].reshape(1)]) for b in range(2)]
    for _ in range(4):
        tok, logits = inf.forward_infer(tok.reshape(2, 1), cur.reshape(2, 1), return_logits=True)
        for b in range(2):
            with torch.no_grad():
                r = train(seqs[b].unsqueeze(0)).transpose(0, 1)[0, -1].float()
            assert (logits[b].float() - r).abs().max().item() < 0.08
            seqs[b] = torch.cat([seqs[b], tok[b].reshape(1)])
        cur = cur + 1


@pytest.mark.parametrize("B,paged,head_dim", [(12, False, 64), (40, True, 64), (72, False, 128)])
def test_batched_decode_layer_matches_library_arm(B, paged, head_dim):
    """9 .. 128 sequences take the small-M tensor-core layer (QKV GEMM + RoPE + cache write, gate/up + SwiGLU, residual
    epilogues): its logits and its KV cache must agree with the library arm (cuBLAS + SDPA, unfused) step by step."""
    from neuronx_distributed_llama3_2_b200.inference import InferenceConfig, LlamaInferenceModel
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM
    from neuronx_distributed_llama3_2_b200.ops import backend

    torch.manual_seed(3)
    dev = torch.device("cuda")
    cfg = _cfg(head_dim=head_dim, hidden_size=8 * head_dim)
    train = LlamaForCausalLM(cfg, device=dev)
    prompts = torch.randint(1, 1024, (B, 32), device=dev)
    pos = torch.arange(32, device=dev).unsqueeze(0).expand(B, 32)
    runs = []
    for arm in (backend.NATIVE_ARM, backend.LIBRARY_ARM):
        with backend.use_backends(**arm):
            inf = LlamaInferenceModel(cfg, InferenceConfig(batch_size=B, max_context_length=64, seq_len=128,
                                                           paged_kv=paged, page_size=32), device=dev)
            inf.load_state_dict(train.state_dict())
            inf.init_kv_cache(device=dev)
            tok, _ = inf.forward_infer(prompts, pos, return_logits=True)
            if arm is backend.NATIVE_ARM:
                layer = next(iter(inf._decoder_layers()))
                assert inf._fused_decode_ok(layer, torch.empty(B, 1, cfg.hidden_size, device=dev,
                                                               dtype=torch.bfloat16), batched=True)
            cur = torch.full((B,), 32, device=dev)
            steps = []
            feed = prompts[:, -1]
            for _ in range(3):          # the same fed tokens in both arms: compare logits, not trajectories
                _, logits = inf.forward_infer(feed.reshape(B, 1), cur.reshape(B, 1), return_logits=True)
                steps.append(logits.float())
                feed = (feed * 7 + 1) % 1024
                cur = cur + 1
            runs.append(steps)
    for a, b in zip(*runs):
        assert (a - b).abs().max().item() < 0.12, (a - b).abs().max().item()
        assert ((a - b).norm() / b.norm()).item() < 2e-2


@pytest.mark.parametrize("batch", [2, 12, 40])     # 2: fused GEMV layer, 12 / 40: batched small-M tensor-core layer
def test_runner_cuda_graphs_match_eager(batch):
    from neuronx_distributed_llama3_2_b200.inference import InferenceConfig, InferenceRunner
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM

    torch.manual_seed(1)
    dev = torch.device("cuda")
    train = LlamaForCausalLM(_cfg(), device=dev)
    sd = {k: v.detach().cpu().clone() for k, v in train.state_dict().items()}
    outs, caches = [], []
    for graphs in (False, True):
        ic = InferenceConfig(batch_size=batch, max_context_length=128, seq_len=256, buckets=[64, 128], max_new_tokens=8,
                             use_cuda_graphs=graphs)
        runner = InferenceRunner(_cfg(), ic, checkpoint_loader=lambda: sd, device=dev)
        runner.trace()
        g = torch.Generator(device="cpu").manual_seed(5)
        prompts = torch.randint(1, 1024, (batch, 50), generator=g).to(dev)
        outs.append(runner.generate(prompts, max_new_tokens=8))
        outs.append(runner.generate(prompts, max_new_tokens=8))     # replay: same result again
        torch.cuda.synchronize()
        caches.append([tuple(t[:, :, :57].float().clone() for t in runner.model.kv_cache.layer(i)) for i in range(2)])
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[2], outs[3])
    assert torch.equal(outs[0], outs[2]), (outs[0][:, -8:], outs[2][:, -8:])
    # the tokens of a random model barely react to a misplaced cache row: compare the caches themselves (a graph that
    # replays the positions of its capture writes every new K / V row to the wrong place)
    for (ke, ve), (kg, vg) in zip(*caches):
        assert torch.equal(ke, kg) and torch.equal(ve, vg)
    rep = runner.benchmark_sampling(prompt_len=64, new_tokens=8, n_runs=2)
    assert rep["ttft_ms_p50"] > 0 and rep["decode_tokens_per_s"] > 0


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _tp_worker(rank, world, port, modes, batch=2):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    from neuronx_distributed_llama3_2_b200.inference import InferenceConfig, InferenceRunner
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    # full (tp = 1) weights, identical on every rank, built BEFORE the process group exists (sizes are all 1 then);
    # the runner shards them
    torch.manual_seed(3)
    full = LlamaForCausalLM(_cfg(), device=torch.device("cpu"))
    sd = {k: v.detach().clone() for k, v in full.state_dict().items()}
    g = torch.Generator(device="cpu").manual_seed(9)
    prompts = torch.randint(1, 1024, (batch, 60), generator=g).to(dev)
    with torch.no_grad():   # single-process forward of the unsharded model = the oracle for the prefill logits
        ref = full.to(dev)(prompts).transpose(0, 1)[:, -1].float()
    del full
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        ps.initialize_model_parallel(tensor_model_parallel_size=world)
        res = {}
        table = {"nccl_eager": ("0", False), "fused_eager": ("1", False), "fused_graphs": ("1", True)}
        for name in modes:
            fused, graphs = table[name]
            from neuronx_distributed_llama3_2_b200.ops import backend as _be

            _be.set_backends(fused_comm=(fused == "1"))
            ic = InferenceConfig(tp_degree=world, batch_size=batch, max_context_length=128, seq_len=256, buckets=[64, 128],
                                 max_new_tokens=8, use_cuda_graphs=graphs)
            runner = InferenceRunner(_cfg(), ic, checkpoint_loader=lambda: sd, device=dev)
            runner.trace()
            res[name] = runner.generate(prompts, max_new_tokens=8)
            again = runner.generate(prompts, max_new_tokens=8)
            assert torch.equal(res[name], again), name
            # prefill logits against the unsharded model
            pos = torch.arange(60, device=dev).unsqueeze(0).expand(batch, 60)
            _, logits = runner.model.forward_infer(prompts, pos, return_logits=True)
            assert (logits.float() - ref).abs().max().item() < 0.1, name
            gathered = [torch.empty_like(res[name]) for _ in range(world)]
            dist.all_gather(gathered, res[name])
            assert all(torch.equal(gathered[0], t) for t in gathered), f"{name}: ranks generated different tokens"
            runner.close()      # graphs that captured NCCL collectives must be gone before the group is destroyed
        if "fused_eager" in res and "fused_graphs" in res:
            assert torch.equal(res["fused_eager"], res["fused_graphs"])
        if "fused_eager" in res and "nccl_eager" in res:   # greedy tokens may differ only through bf16 ties
            assert (res["fused_eager"] == res["nccl_eager"]).float().mean().item() > 0.9
        torch.cuda.synchronize()
        dist.barrier()
    finally:
        ps.destroy_model_parallel()
        dist.destroy_process_group()


@pytest.mark.parametrize("batch", [2, 12], ids=["b2", "b12"])     # 12 rows: the batched small-M layer + one-shot all-reduce
@pytest.mark.parametrize("modes", [("nccl_eager", "fused_eager"), ("fused_eager", "fused_graphs")],
                         ids=["eager", "graphs"])
def test_tp2_inference_fused_allreduce(modes, batch):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from conftest import spawn_gpu_workers

    spawn_gpu_workers(_tp_worker, (2, _free_port(), modes, batch), 2, deadline_s=90.0)
