"""Topology maths, mappings and parallel layers on CPU/gloo (reference ``test/integration/parallel_layers/
test_parallel_state.py``, ``test_layers.py``, ``test_loss_functions.py``, ``test_grads.py``)."""
from __future__ import annotations

import pytest
import torch
import torch.nn.functional as F

from conftest import run_distributed


def test_rank_meshes_match_reference_layout():
    from neuronx_distributed_llama3_2_b200.parallel.parallel_state import compute_rank_meshes

    m = compute_rank_meshes(world_size=16, tp=2, pp=2, ep=2)
    # rank tensor [PP, DP, TP] row-major, TP contiguous
    assert m["tp"][:3] == [[0, 1], [2, 3], [4, 5]]
    assert m["dp"][0] == [0, 2, 4, 6] and m["dp"][1] == [1, 3, 5, 7]
    assert m["pp"][0] == [0, 8] and m["pp"][7] == [7, 15]
    # expert view [PP, DP_exp, EP, TP]: EP groups are strided by tp inside a DP group, EXP-DP by tp*ep
    assert m["ep"][0] == [0, 2] and m["ep"][1] == [1, 3]
    assert m["exp_dp"][0] == [0, 4] and m["exp_dp"][1] == [1, 5]
    for tp, pp, ep in [(1, 1, 1), (8, 1, 1), (4, 2, 1), (2, 2, 4), (1, 4, 2)]:
        mm = compute_rank_meshes(32, tp, pp, ep)
        for key, size in (("tp", tp), ("pp", pp), ("dp", 32 // (tp * pp)), ("ep", ep), ("exp_dp", 32 // (tp * pp * ep))):
            groups = mm[key]
            assert all(len(g) == size for g in groups)
            assert sorted(r for g in groups for r in g) == list(range(32))      # a partition of the world


def _layers_worker(rank, world):
    import torch.distributed as dist

    from neuronx_distributed_llama3_2_b200.parallel import (ColumnParallelLinear, InputChannelParallelConv2d,
                                                            OutputChannelParallelConv2d, ParallelEmbedding,
                                                            RowParallelLinear, mappings, parallel_cross_entropy)
    from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps
    from neuronx_distributed_llama3_2_b200.parallel.grads import clip_grad_norm, get_grad_norm
    from neuronx_distributed_llama3_2_b200.parallel.layers import create_local_weight

    ps.initialize_model_parallel(tensor_model_parallel_size=2)
    assert ps.get_tensor_model_parallel_rank() == rank and ps.get_data_parallel_size() == 1
    assert ps.rmsg("x").startswith(f"[rank_{rank}_pp0_tp{rank}_dp0]")
    torch.manual_seed(0)

    # ---- mappings: forward/backward collectives pair up
    x = torch.arange(8.0).reshape(2, 4).requires_grad_()
    y = mappings.scatter_to_tensor_model_parallel_region(x)
    assert torch.equal(y, x[:, rank * 2:(rank + 1) * 2])
    y.sum().backward()
    assert torch.equal(x.grad, torch.ones_like(x))                               # scatter bwd = all-gather
    z = torch.full((4, 3), float(rank + 1), requires_grad=True)
    g = mappings.gather_from_sequence_parallel_region(z)                          # [8, 3]; bwd = reduce-scatter
    assert g.shape == (8, 3) and torch.equal(g[:4], torch.ones(4, 3)) and torch.equal(g[4:], 2 * torch.ones(4, 3))
    (g * (rank + 1)).sum().backward()
    assert torch.equal(z.grad, torch.full((4, 3), 3.0))                           # sum over ranks of (rank + 1)
    r = mappings.reduce_scatter_to_sequence_parallel_region(torch.ones(8, 2) * (rank + 1))
    assert torch.equal(r, torch.full((4, 2), 3.0))

    # ---- Column -> Row MLP against the un-sharded computation (no SP and SP)
    full_up = torch.randn(16, 8)
    full_down = torch.randn(8, 16)
    inp = torch.randn(6, 2, 8)
    ref_in = inp.clone().requires_grad_()
    ref = F.linear(torch.tanh(F.linear(ref_in, full_up)), full_down)
    ref.pow(2).sum().backward()
    for sp in (False, True):
        up = ColumnParallelLinear(8, 16, bias=False, gather_output=False, sequence_parallel_enabled=sp)
        down = RowParallelLinear(16, 8, bias=False, input_is_parallel=True, sequence_parallel_enabled=sp)
        with torch.no_grad():
            up.weight.copy_(create_local_weight(full_up, 0, 8, 1))
            down.weight.copy_(create_local_weight(full_down, 1, 8, 1))
        xin = (inp[rank * 3:(rank + 1) * 3] if sp else inp).clone().requires_grad_()
        out = down(torch.tanh(up(xin)))
        want = ref[rank * 3:(rank + 1) * 3] if sp else ref
        torch.testing.assert_close(out, want.detach(), atol=1e-5, rtol=1e-5)
        # every rank holds (a slice of) the same logical output: sum the local loss pieces
        loss = out.pow(2).sum() if sp else out.pow(2).sum() / 1.0
        loss.backward()
        want_gin = ref_in.grad[rank * 3:(rank + 1) * 3] if sp else ref_in.grad
        torch.testing.assert_close(xin.grad, want_gin, atol=1e-4, rtol=1e-4)
        wg_up = torch.zeros(16, 8)
        wg_up[rank * 8:(rank + 1) * 8] = up.weight.grad
        dist.all_reduce(wg_up)
        ref_up_grad = torch.autograd.grad(F.linear(torch.tanh(F.linear(inp, full_up.requires_grad_())), full_down)
                                          .pow(2).sum(), full_up)[0]
        torch.testing.assert_close(wg_up, ref_up_grad, atol=1e-4, rtol=1e-4)

    # ---- embeddings: vocab-sharded and dim-sharded
    table = torch.randn(10, 6)
    ids = torch.tensor([[0, 3, 9], [5, 4, 7]])
    e = ParallelEmbedding(10, 6)
    with torch.no_grad():
        e.weight.copy_(table[rank * 5:(rank + 1) * 5])
    torch.testing.assert_close(e(ids), table[ids])
    e2 = ParallelEmbedding(10, 6, shard_across_embedding=True)
    with torch.no_grad():
        e2.weight.copy_(table[:, rank * 3:(rank + 1) * 3])
    torch.testing.assert_close(e2(ids), table[ids])

    # ---- vocab-parallel cross entropy incl. label smoothing (vs F.cross_entropy)
    logits = torch.randn(7, 12)
    tgt = torch.randint(0, 12, (7,))
    for ls in (0.0, 0.1):
        lg = logits[:, rank * 6:(rank + 1) * 6].clone().requires_grad_()
        loss = parallel_cross_entropy(lg, tgt, label_smoothing=ls)
        ref_l = logits.clone().requires_grad_()
        want = F.cross_entropy(ref_l, tgt, reduction="none", label_smoothing=ls)
        torch.testing.assert_close(loss, want, atol=1e-5, rtol=1e-5)
        loss.sum().backward()
        want.sum().backward()
        torch.testing.assert_close(lg.grad, ref_l.grad[:, rank * 6:(rank + 1) * 6], atol=1e-5, rtol=1e-5)

    # ---- grad norm: TP-sharded grads summed over ranks, replicated ones counted once; clipping scales all
    a = torch.nn.Parameter(torch.zeros(3))
    b = torch.nn.Parameter(torch.zeros(2))
    a.tensor_model_parallel = True
    a.grad = torch.full((3,), float(rank + 1))        # shard
    b.grad = torch.full((2,), 2.0)                    # replicated
    norm = get_grad_norm([a, b])
    assert abs(float(norm) - (3 * 1 + 3 * 4 + 2 * 4) ** 0.5) < 1e-5
    clip_grad_norm([a, b], max_norm=1.0)
    assert abs(float(b.grad[0]) - 2.0 / (float(norm) + 1e-6)) < 1e-5

    # ---- parallel conv2d
    conv_ref = torch.nn.Conv2d(4, 6, 3, padding=1)
    img = torch.randn(2, 4, 5, 5)
    oc = OutputChannelParallelConv2d(4, 6, 3, padding=1, gather_output=True)
    ic = InputChannelParallelConv2d(4, 6, 3, padding=1)
    with torch.no_grad():
        oc.weight.copy_(conv_ref.weight[rank * 3:(rank + 1) * 3])
        oc.bias.copy_(conv_ref.bias[rank * 3:(rank + 1) * 3])
        ic.weight.copy_(conv_ref.weight[:, rank * 2:(rank + 1) * 2])
        ic.bias.copy_(conv_ref.bias)
    torch.testing.assert_close(oc(img), conv_ref(img), atol=1e-5, rtol=1e-5)
    torch.testing.assert_close(ic(img), conv_ref(img), atol=1e-5, rtol=1e-5)


def test_mappings_layers_loss_grads_tp2():
    run_distributed(_layers_worker, 2)


def test_pad_model_and_rng_tracker():
    from neuronx_distributed_llama3_2_b200.parallel.pad import get_number_of_extra_heads
    from neuronx_distributed_llama3_2_b200.parallel.random import get_rng_tracker, model_parallel_manual_seed

    assert get_number_of_extra_heads(12, 8) == 4 and get_number_of_extra_heads(16, 8) == 0
    model_parallel_manual_seed(123)
    tr = get_rng_tracker()
    with tr.fork():
        a = torch.rand(3)
    b = torch.rand(3)
    model_parallel_manual_seed(123)
    with tr.fork():
        a2 = torch.rand(3)
    assert torch.equal(a, a2) and not torch.equal(a, b)
    with pytest.raises(Exception):
        tr.add("model-parallel-rng", 1)


def test_legacy_checkpointing_layout(tmp_path):
    from neuronx_distributed_llama3_2_b200.parallel import checkpointing

    sd = {"model": {"w": torch.arange(4.0)}, "step": 3}
    for xs in (False, True):
        d = str(tmp_path / f"x{int(xs)}")
        checkpointing.save(sd, d, save_xser=xs)
        base = tmp_path / f"x{int(xs)}" / "tp_rank_00_pp_rank_00"
        assert (base / "checkpoint.pt").exists() if not xs else (base.exists() and (tmp_path / f"x{int(xs)}" / "tp_rank_00_pp_rank_00.tensors").is_dir())
        back = checkpointing.load(d, load_xser=xs)
        assert back["step"] == 3 and torch.equal(back["model"]["w"], sd["model"]["w"])
