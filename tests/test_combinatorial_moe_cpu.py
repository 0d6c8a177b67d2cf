"""The Mixtral counterpart of ``test_combinatorial_cpu``: TP x SP x EP x ZeRO-1.  Dropless routing and no auxiliary loss,
so the loss of a global batch does not depend on how its tokens are spread over the expert-parallel ranks; the loss
trajectory and the pre-clip global gradient norm (expert shards summed over EP, dense parameters counted once) must match
the single-process run."""
from __future__ import annotations

import pytest
import torch

from conftest import run_distributed

STEPS, LR, BATCH, SEQ = 3, 1e-2, 4, 32


def _batch():
    return torch.randint(0, 512, (BATCH, SEQ), generator=torch.Generator().manual_seed(78))


def _config(sp):
    from neuronx_distributed_llama3_2_b200.models import tiny_mixtral

    return tiny_mixtral(sequence_parallel_enabled=bool(sp), capacity_factor=None, router_aux_loss_coef=0.0)


def _golden():
    from neuronx_distributed_llama3_2_b200.scripts.checkpoint_converter_moe import mock_parallel_state
    from neuronx_distributed_llama3_2_b200.models import MixtralForCausalLM

    with mock_parallel_state(1, 0, 1, 0):
        torch.manual_seed(0)
        model = MixtralForCausalLM(_config(0))
        opt = torch.optim.AdamW(model.parameters(), lr=LR, betas=(0.9, 0.95), weight_decay=0.01)
        start = {k: v.detach().clone() for k, v in model.state_dict().items()}
        ids, out, norms = _batch(), [], []
        for _ in range(STEPS):
            opt.zero_grad()
            loss = torch.stack([model(c, labels=c) for c in ids.chunk(BATCH, 0)]).mean()
            loss.backward()
            norms.append(float(torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)))
            opt.step()
            out.append(float(loss.detach()))
        return out, norms, start


def _worker(rank, world, tp, sp, ep, zero1, steps):
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.models import MixtralForCausalLM
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps
    from neuronx_distributed_llama3_2_b200.scripts.checkpoint_converter_moe import CheckpointConverterMixtral

    golden, golden_norms, start = _golden()
    tag = dict(tp=tp, sp=sp, ep=ep, zero1=zero1)
    dp = world // tp
    per_rank = BATCH // dp
    cfg = nxd.neuronx_distributed_config(
        tensor_parallel_size=tp, expert_parallel_size=ep, sequence_parallel=bool(sp),
        optimizer_config={"zero_one_enabled": bool(zero1), "grad_clipping": True, "max_grad_norm": 1.0})
    assert (ps.get_data_parallel_size(), ps.get_expert_model_parallel_size()) == (dp, ep)

    def build():
        torch.manual_seed(0)
        return MixtralForCausalLM(_config(sp))

    model = nxd.initialize_parallel_model(cfg, build, device=torch.device("cpu"))
    # the stacked expert weights are initialised per shard, so the starting point is the single-process model's state cut
    # for this (tp, ep) coordinate by the offline converter -- which this also exercises against the live layout
    conv = CheckpointConverterMixtral(_config(sp), tp_size=tp, ep_size=ep)
    shard = conv.shard(start, ps.get_tensor_model_parallel_rank(), ps.get_expert_model_parallel_rank())
    res = model.load_state_dict(shard, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    opt = nxd.initialize_parallel_optimizer(cfg, torch.optim.AdamW, model.parameters(), lr=LR, betas=(0.9, 0.95),
                                            weight_decay=0.01)
    mine = _batch().chunk(dp, 0)[ps.get_data_parallel_rank()]
    losses, norm_errors = [], []
    for _ in range(steps):
        opt.zero_grad()
        loss = torch.stack([model(c, labels=c) for c in mine.chunk(per_rank, 0)]).mean()
        loss.backward()
        opt.step()
        losses.append(float(loss.detach()))
        gn, want_gn = float(opt.grad_norm), golden_norms[len(losses) - 1]
        norm_errors.append(abs(gn - want_gn) / want_gn)
    t = torch.tensor(losses, dtype=torch.float64)
    torch.distributed.all_reduce(t)
    t /= world
    for step, (got, want) in enumerate(zip(t.tolist(), golden)):
        assert abs(got - want) < 2e-4, (tag, step, t.tolist(), golden, norm_errors)
    if steps > 1:
        assert max(norm_errors) < 2e-3, (tag, norm_errors)


# (world, tp, sp, ep, zero1)
CONFIGS = [(2, 2, 1, 1, 0), (2, 1, 0, 2, 0), (2, 1, 0, 2, 1), (4, 2, 1, 2, 1)]


@pytest.mark.parametrize("world,tp,sp,ep,zero1", [CONFIGS[0], CONFIGS[3]])
def test_moe_first_step_loss_is_independent_of_the_parallel_layout(world, tp, sp, ep, zero1):
    """Forward parity from a converter-cut checkpoint: the loss before any update equals the single-process loss."""
    run_distributed(_worker, world, tp, sp, ep, zero1, 1)


@pytest.mark.parametrize("world,tp,sp,ep,zero1", CONFIGS)
def test_moe_loss_trajectory_is_independent_of_the_parallel_layout(world, tp, sp, ep, zero1):
    run_distributed(_worker, world, tp, sp, ep, zero1, STEPS)


def _ep_distinct_worker(rank, world):
    """ep = world = 2, expert-DP size 1: the expert shards of the two ranks are DIFFERENT tensors.  With lr = 0 a step must
    leave each rank's expert weights exactly where they were (no reduce-scatter / all-gather across EP ranks)."""
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.models import MixtralForCausalLM
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    cfg = nxd.neuronx_distributed_config(
        tensor_parallel_size=1, expert_parallel_size=2,
        optimizer_config={"zero_one_enabled": True, "grad_clipping": True, "max_grad_norm": 1.0})
    assert ps.get_expert_data_parallel_size() == 1

    def build():
        torch.manual_seed(0)
        return MixtralForCausalLM(_config(0))

    model = nxd.initialize_parallel_model(cfg, build, device=torch.device("cpu"))
    experts = {n: p for n, p in model.named_parameters() if getattr(p, "expert_model_parallel", False)}
    assert experts
    with torch.no_grad():
        for p in experts.values():
            p.fill_(float(rank + 1))
    opt = nxd.initialize_parallel_optimizer(cfg, torch.optim.AdamW, model.parameters(), lr=0.0, weight_decay=0.0)
    ids = _batch().chunk(2, 0)[rank]
    opt.zero_grad()
    model(ids, labels=ids).backward()
    opt.step()
    for n, p in experts.items():
        assert torch.all(p == float(rank + 1)), (rank, n, p.flatten()[:4])


def test_expert_shards_stay_local_to_their_ep_rank_under_zero1():
    run_distributed(_ep_distinct_worker, 2)


def _ep_resume_worker(rank, world, tmp, zero1):
    """Save after step 1, run step 2; rebuild everything, load, run step 2 again: parameters must agree exactly.  Under
    EP the expert Adam moments are different tensors on every EP rank, so each EP rank needs its own optimizer file."""
    import os

    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.models import MixtralForCausalLM

    cfg = nxd.neuronx_distributed_config(
        tensor_parallel_size=1, expert_parallel_size=2,
        optimizer_config={"zero_one_enabled": bool(zero1), "grad_clipping": True, "max_grad_norm": 1.0})

    def build():
        torch.manual_seed(0)
        model = nxd.initialize_parallel_model(cfg, lambda: MixtralForCausalLM(_config(0)), device=torch.device("cpu"))
        opt = nxd.initialize_parallel_optimizer(cfg, torch.optim.AdamW, model.parameters(), lr=LR, betas=(0.9, 0.95),
                                                weight_decay=0.01)
        return model, opt

    ids = _batch().chunk(2, 0)[rank]

    def step(model, opt):
        opt.zero_grad()
        model(ids, labels=ids).backward()
        opt.step()

    model, opt = build()
    step(model, opt)
    nxd.save_checkpoint(tmp, "s1", model=model, optimizer=opt)
    step(model, opt)
    want = {k: v.detach().clone() for k, v in model.state_dict().items()}
    if rank == 0:
        names = sorted(os.listdir(os.path.join(tmp, "s1", "model")))
        assert names == ["dp_rank_00_ep_rank_00_tp_rank_00_pp_rank_00.pt", "dp_rank_00_ep_rank_01_tp_rank_00_pp_rank_00.pt"], names
        onames = sorted(os.listdir(os.path.join(tmp, "s1", "optim")))
        if zero1:
            assert onames == ["dp_rank_00_tp_rank_00_pp_rank_00.pt", "dp_rank_01_tp_rank_00_pp_rank_00.pt"], onames
        else:
            assert onames == ["dp_rank_00_ep_rank_00_tp_rank_00_pp_rank_00.pt",
                              "dp_rank_00_ep_rank_01_tp_rank_00_pp_rank_00.pt"], onames
    model2, opt2 = build()
    with torch.no_grad():
        for p in model2.parameters():
            p.add_(1.0)                       # make sure the load does the work
    nxd.load_checkpoint(tmp, "s1", model=model2, optimizer=opt2)
    step(model2, opt2)
    for k, v in model2.state_dict().items():
        assert torch.equal(v, want[k]), (rank, k, float((v - want[k]).abs().max()))


@pytest.mark.parametrize("zero1", [0, 1])
def test_ep_checkpoint_resume_keeps_every_ep_ranks_optimizer_state(tmp_path, zero1):
    run_distributed(_ep_resume_worker, 2, str(tmp_path), zero1)
