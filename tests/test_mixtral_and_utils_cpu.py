"""Mixtral-style MoE model (dense + TP=2 x EP=2 over gloo) and the training utilities."""
from __future__ import annotations

import json
import os
import tempfile

import torch

from conftest import run_distributed


def test_tiny_mixtral_trains_single_process():
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.models import MixtralForCausalLM, tiny_mixtral

    torch.manual_seed(0)
    cfg = nxd.neuronx_distributed_config(optimizer_config={"zero_one_enabled": True, "grad_clipping": True,
                                                           "max_grad_norm": 1.0})
    model = nxd.initialize_parallel_model(cfg, lambda: MixtralForCausalLM(tiny_mixtral(sequence_parallel_enabled=False)),
                                          device=torch.device("cpu"))
    opt = nxd.initialize_parallel_optimizer(cfg, torch.optim.AdamW, model.parameters(), lr=5e-3)
    ids = torch.randint(0, 512, (2, 32))
    losses = []
    for _ in range(8):
        opt.zero_grad()
        losses.append(float(model.run_train(ids, labels=ids)))
        opt.step()
    assert losses[-1] < losses[0] - 0.3, losses


def _moe_ep_worker(rank, world, tmp):
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.models import MixtralForCausalLM, tiny_mixtral
    from neuronx_distributed_llama3_2_b200.optimizer import NeuronEPZero1Optimizer
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    cfg = nxd.neuronx_distributed_config(tensor_parallel_size=2, expert_parallel_size=2, sequence_parallel=True,
                                         optimizer_config={"zero_one_enabled": True, "grad_clipping": True,
                                                           "max_grad_norm": 1.0})
    assert (ps.get_tensor_model_parallel_size(), ps.get_expert_model_parallel_size(), ps.get_data_parallel_size()) == (2, 2, 2)
    torch.manual_seed(0)
    model = nxd.initialize_parallel_model(cfg, lambda: MixtralForCausalLM(tiny_mixtral(capacity_factor=1.5)),
                                          device=torch.device("cpu"))
    opt = nxd.initialize_parallel_optimizer(cfg, torch.optim.AdamW, model.parameters(), lr=5e-3)
    assert isinstance(opt.optimizer, NeuronEPZero1Optimizer)
    g = torch.Generator().manual_seed(10 + ps.get_data_parallel_rank())
    ids = torch.randint(0, 512, (2, 32), generator=g)
    losses = []
    for _ in range(6):
        opt.zero_grad()
        losses.append(float(model.run_train(ids, labels=ids)))
        opt.step()
    assert losses[-1] < losses[0] - 0.2, losses
    assert float(opt.grad_norm) > 0

    # checkpoint with expert parallelism: one model file per (ep, tp) coordinate, EP-aware ZeRO-1 state; round trip
    import os

    from neuronx_distributed_llama3_2_b200.scripts.checkpoint_converter_moe import CheckpointConverterMixtral

    ep_rank, tp_rank = ps.get_expert_model_parallel_rank(), ps.get_tensor_model_parallel_rank()
    nxd.save_checkpoint(tmp, "s", model=model, optimizer=opt, use_xser=False)
    files = sorted(os.listdir(os.path.join(tmp, "s", "model")))
    # the reference layout: ONE complete model file per (EP rank, TP rank), dense tensors included (checkpoint.py:655)
    assert files == sorted([f"dp_rank_00_ep_rank_{e:02d}_tp_rank_{t:02d}_pp_rank_00.pt" for e in range(2)
                            for t in range(2)]), files
    before = {k: v.clone() for k, v in model.state_dict().items()}
    import copy

    opt_before = copy.deepcopy(opt.state_dict())
    opt.zero_grad()
    model.run_train(ids, labels=ids)
    opt.step()                                   # weights, master copies and moments all move away from the checkpoint
    nxd.load_checkpoint(tmp, "s", model=model, optimizer=opt)
    for k, v in model.state_dict().items():
        assert torch.equal(v, before[k]), k
    def tensors(obj, prefix=""):
        if isinstance(obj, torch.Tensor):
            yield prefix, obj
        elif isinstance(obj, dict):
            for k, v in obj.items():
                yield from tensors(v, f"{prefix}/{k}")
        elif isinstance(obj, (list, tuple)):
            for i, v in enumerate(obj):
                yield from tensors(v, f"{prefix}/{i}")

    a, b = dict(tensors(opt_before)), dict(tensors(opt.state_dict()))
    assert a and set(a) == set(b) and all(torch.equal(a[k], b[k]) for k in a)
    # the offline converter merges those (tp, ep) shards into HF names and cuts them again: this rank's shard comes back
    mcfg = tiny_mixtral(capacity_factor=1.5)
    conv = CheckpointConverterMixtral(mcfg, tp_size=2, ep_size=2)
    hf = conv.convert_to_full_state(os.path.join(tmp, "s"))
    assert hf["model.layers.0.block_sparse_moe.experts.3.w1.weight"].shape == (mcfg.intermediate_size, mcfg.hidden_size)
    mine = conv.shard(conv.to_model_names(dict(hf)), tp_rank, ep_rank)
    for k, v in before.items():
        torch.testing.assert_close(mine[k], v, msg=lambda m: f"{k}: {m}")


def test_mixtral_tp2_ep2_sp_zero1_trains(tmp_path):
    run_distributed(_moe_ep_worker, 4, str(tmp_path))


def test_training_utils():
    from neuronx_distributed_llama3_2_b200.utils.training_utils import (CosineAnnealing, SyntheticTokens, Throughput,
                                                                      TrainingMetrics)

    t = Throughput(batch_size=2, world_size=4, grad_accum_usteps=3, moving_avg_window_size=4)
    assert t.seqs_per_iteration == 24
    for _ in range(6):
        v = t.get_throughput()
    assert v > 0 and t.moving_avg_window.qsize() == 4
    with tempfile.TemporaryDirectory() as d:
        f = os.path.join(d, "m.json")
        tm = TrainingMetrics(f)
        tm.store_parameters({"lr": 1e-4})
        tm.store_metrics([tm.metric("Final loss", 2.5, "")])
        tm.store_metrics([tm.metric("Final loss", 2.0, ""), tm.metric("Average throughput", 10.0, "seq/s")])
        data = json.load(open(f))
        assert data["parameters"]["lr"] == 1e-4 and len(data["metrics"]) == 2
        assert {m["MetricName"]: m["MeasuredValue"] for m in data["metrics"]}["Final loss"] == 2.0
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([p], lr=1.0)
    sch = CosineAnnealing(opt, max_steps=20, warmup_steps=4, min_lr=0.1)
    lrs = []
    for _ in range(20):
        lrs.append(opt.param_groups[0]["lr"])
        opt.step()
        sch.step()
    assert lrs[0] < lrs[3] <= 1.0 and abs(max(lrs) - 1.0) < 1e-6 and lrs[-1] < 0.2
    it = SyntheticTokens(100, 2, 8, "cpu", n_batches=2)
    a, b, c = next(it), next(it), next(it)
    assert a.shape == (2, 8) and torch.equal(a, c) and not torch.equal(a, b)


def test_utils_namespace_matches_reference_modules():
    """``utils/{timeline,sampling,speculative_decoding,medusa_utils,tensor_utils}`` exist with the reference names."""
    import torch

    from neuronx_distributed_llama3_2_b200.modules.moe import ACT2FN
    from neuronx_distributed_llama3_2_b200.utils import medusa_utils, sampling, speculative_decoding, tensor_utils, timeline

    assert "silu" in ACT2FN and callable(ACT2FN["silu"])
    assert timeline.Event("fwd", 1, 10, 25).as_chrome_event()["dur"] == 15 and hasattr(timeline, "Timeline")
    assert hasattr(sampling, "Sampler") and hasattr(speculative_decoding, "NeuronSpeculation")
    x = torch.arange(12).reshape(3, 4)
    assert torch.equal(tensor_utils.cumsum(x, dim=1), torch.cumsum(x, 1))
    assert medusa_utils.pad_path([1, 2, 3], 5) == [1, 2, 3, -2, -2] and medusa_utils.pad_path([1, 2], 1) == [1, 2]
    bufs = medusa_utils.generate_medusa_buffers([[0], [1], [0, 0], [0, 1], [1, 0]])
    ri = bufs["retrieve_indices"]
    assert bufs["tree_indices"].shape[0] == 6 and ri.shape[1] == 3
    # accept two speculated tokens of path 0: ids grow by 3, logits are cut to the position after the accepted prefix
    V, heads = 7, 2
    ids = torch.tensor([[5, 6, 7]])
    cands = torch.tensor([[1, 2, 3], [1, 4, 0], [1, 5, 6]])
    logits = torch.randn(3, 3, V)
    mlog = torch.randn(heads, 3, 3, V)
    ids2, l2, m2, n_new, sel = medusa_utils.update_inference_inputs(ids, cands, 0, 2, ri, None, logits, mlog, 0)
    assert ids2.tolist() == [[5, 6, 7, 1, 2, 3]] and n_new == 3
    assert l2.shape == (1, 1, V) and torch.equal(l2[0, 0], logits[0, 2]) and m2.shape == (heads, 1, 1, V)
    assert torch.equal(sel, ri[0, :3] + 3)
    # the reference's token-id convention: base top-k ids [1, 1, k], head top-k ids [heads, 1, 1, topk]
    base_ids = torch.tensor([[[9, 8]]])
    head_ids = torch.tensor([[[[11, 12]]], [[[21, 22]]]])
    cart, tree = medusa_utils.generate_candidates(head_ids, base_ids, bufs["tree_indices"], ri)
    assert tree.shape == (1, 6) and tree[0, 0] == 9 and cart[0].tolist() == [9, 11, 21]
    verifier = torch.zeros(cart.shape[0], 3, 2, dtype=torch.long)
    verifier[:, 0, 0], verifier[:, 1, 0] = 11, 22                  # agrees with head 0's best guess, then with 22
    best, acc = medusa_utils.evaluate_posterior(verifier, cart)
    assert int(acc) == 2 and cart[int(best)].tolist() == [9, 11, 22]
    best0, acc0 = medusa_utils.evaluate_posterior(torch.full_like(verifier, 3), cart)
    assert int(acc0) == 0 and int(best0) == 0
