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


def _moe_ep_worker(rank, world):
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


def test_mixtral_tp2_ep2_sp_zero1_trains():
    run_distributed(_moe_ep_worker, 4)


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
