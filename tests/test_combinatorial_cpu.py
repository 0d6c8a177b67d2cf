"""One model, one global batch, every parallel configuration: the loss trajectory of three optimizer steps AND the
pre-clip global gradient norm of every step must be the same as in the single-process run (the reference's
``test/integration/combinatorial_tests`` sweeps TP x SP x selective checkpointing x PP x ZeRO-1 and compares loss curves
with golden metrics; here the golden curve is computed in place)."""
from __future__ import annotations

import pytest
import torch

from conftest import run_distributed

STEPS, LR, BATCH, SEQ = 3, 1e-2, 4, 32


def _batch():
    return torch.randint(0, 512, (BATCH, SEQ), generator=torch.Generator().manual_seed(77))


def _golden():
    from neuronx_distributed_llama3_2_b200.inference import _mock_parallel_state
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama

    with _mock_parallel_state(1, 0):
        torch.manual_seed(0)
        model = LlamaForCausalLM(tiny_llama(num_hidden_layers=4, sequence_parallel_enabled=False))
        opt = torch.optim.AdamW(model.parameters(), lr=LR, betas=(0.9, 0.95), weight_decay=0.01)
        ids, out, norms = _batch(), [], []
        for _ in range(STEPS):
            opt.zero_grad()
            loss = torch.stack([model(c, labels=c) for c in ids.chunk(BATCH, 0)]).mean()
            loss.backward()
            norms.append(float(torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)))
            opt.step()
            out.append(float(loss))
        return out, norms


def _worker(rank, world, tp, sp, sc, pp, zero1):
    from neuronx_distributed_llama3_2_b200 import trainer as nxd
    from neuronx_distributed_llama3_2_b200.models import LlamaForCausalLM, tiny_llama
    from neuronx_distributed_llama3_2_b200.models.llama import LlamaDecoderLayer
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps

    golden, golden_norms = _golden()
    tag = dict(tp=tp, sp=sp, sc=sc, pp=pp, zero1=zero1)
    dp = world // (tp * pp)
    per_rank = BATCH // dp
    pipe = {"pipeline_parallel_size": pp,
            "pipeline_config": {"transformer_layer_cls": LlamaDecoderLayer, "num_microbatches": per_rank,
                                "auto_partition": True, "input_names": ["input_ids", "labels"],
                                "broadcast_and_average_loss": True, "use_zero1_optimizer": bool(zero1)}} if pp > 1 else {}
    cfg = nxd.neuronx_distributed_config(
        tensor_parallel_size=tp, sequence_parallel=bool(sp),
        optimizer_config={"zero_one_enabled": bool(zero1), "grad_clipping": True, "max_grad_norm": 1.0}, **pipe)
    assert ps.get_data_parallel_size() == dp

    def build():
        torch.manual_seed(0)          # identical full-size master weights on every rank before they are sliced
        return LlamaForCausalLM(tiny_llama(num_hidden_layers=4, sequence_parallel_enabled=bool(sp),
                                           selective_checkpoint_enabled=bool(sc)))

    model = nxd.initialize_parallel_model(cfg, build, device=torch.device("cpu"))
    opt = nxd.initialize_parallel_optimizer(cfg, torch.optim.AdamW, model.parameters(), lr=LR, betas=(0.9, 0.95),
                                            weight_decay=0.01)
    mine = _batch().chunk(dp, 0)[ps.get_data_parallel_rank()]
    losses = []
    for _ in range(STEPS):
        opt.zero_grad()
        if pp > 1:
            loss = model.run_train(input_ids=mine, labels=mine)
        else:   # several forward passes, ONE backward: every weight receives several gradient contributions in one pass
            loss = torch.stack([model(c, labels=c) for c in mine.chunk(per_rank, 0)]).mean()
            loss.backward()
        opt.step()
        losses.append(float(loss.detach()))
        # AdamW barely reacts to a wrong clipping factor, so the pre-clip global gradient norm is checked on its own
        # (TP-duplicated parameters once, tied weights that live on two pipeline stages once, ZeRO-1 shards summed)
        gn, want_gn = float(opt.grad_norm), golden_norms[len(losses) - 1]
        assert abs(gn - want_gn) < 2e-3 * want_gn, (tag, len(losses), gn, want_gn)
    # the global loss is the mean over the data-parallel replicas (PP already broadcasts / averages inside its replica)
    t = torch.tensor(losses, dtype=torch.float64)
    torch.distributed.all_reduce(t)
    t /= world
    for step, (got, want) in enumerate(zip(t.tolist(), golden)):
        assert abs(got - want) < 2e-4, (tag, step, t.tolist(), golden)


# (world, tp, sp, selective checkpoint, pp, zero1) -- the reference's sweep in miniature, plus data parallelism
CONFIGS = [(2, 2, 1, 0, 1, 0), (2, 2, 1, 1, 1, 1), (2, 1, 0, 0, 2, 0), (2, 1, 0, 0, 1, 1),
           (4, 2, 0, 0, 2, 0), (4, 2, 1, 0, 2, 1), (4, 1, 0, 0, 2, 1)]


@pytest.mark.parametrize("world,tp,sp,sc,pp,zero1", CONFIGS)
def test_loss_trajectory_is_independent_of_the_parallel_layout(world, tp, sp, sc, pp, zero1):
    run_distributed(_worker, world, tp, sp, sc, pp, zero1)
